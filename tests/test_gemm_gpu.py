"""tcgen05 GEMM numerics vs fp32 PyTorch (all operand layouts, tails, epilogues, autograd)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _mk(rows, cols, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(rows, cols, device="cuda", generator=g) * scale).to(torch.bfloat16)


def _ref(a, b, a_mn, b_mn):
    A = (a.t() if a_mn else a).float()
    B = (b if b_mn else b.t()).float()
    return A @ B


SHAPES = [(128, 256, 64), (128, 128, 128), (256, 512, 256), (304, 520, 200), (64, 72, 8), (1000, 1000, 1000), (2048, 4096, 1024), (520, 136, 4104)]


@pytest.mark.parametrize("a_mn", [False, True])
@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("shape", SHAPES)
def test_gemm_layouts(shape, a_mn, b_mn):
    from determined_b200.ops import gemm as G

    M, N, K = shape
    a = _mk(K, M, 1) if a_mn else _mk(M, K, 1)
    b = _mk(K, N, 2) if b_mn else _mk(N, K, 2)
    want = _ref(a, b, a_mn, b_mn)
    got = G.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32)
    torch.cuda.synchronize()
    torch.testing.assert_close(got, want, rtol=2e-3, atol=2e-3 * (K ** 0.5))
    got16 = G.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn)
    torch.testing.assert_close(got16.float(), want, rtol=1.6e-2, atol=2e-2 * (K ** 0.5))


@pytest.mark.parametrize("b_mn", [False, True])
def test_gemm_odd_row_count(b_mn):
    from determined_b200.ops import gemm as G

    M, N, K = 301, 520, 200
    a = _mk(M, K, 1)
    b = _mk(K, N, 2) if b_mn else _mk(N, K, 2)
    got = G.gemm(a, b, b_mn_major=b_mn, out_dtype=torch.float32)
    torch.testing.assert_close(got, _ref(a, b, False, b_mn), rtol=2e-3, atol=3e-2)


def test_gemm_strided_views_and_accumulate():
    from determined_b200.ops import gemm as G

    big_a = _mk(300, 512, 3)
    a = big_a[:, 64:64 + 200]               # row stride 512, 16B-aligned offset
    b = _mk(264, 200, 4)
    out = torch.randn(300, 264, device="cuda")
    want = out + a.float() @ b.float().t()
    G.gemm(a, b, out=out, accumulate=True)
    torch.testing.assert_close(out, want, rtol=2e-3, atol=3e-2)
    o16 = torch.randn(300, 264, device="cuda").to(torch.bfloat16)
    want16 = o16.float() + a.float() @ b.float().t()
    G.gemm(a, b, out=o16, accumulate=True)
    torch.testing.assert_close(o16.float(), want16, rtol=1.6e-2, atol=0.3)


@pytest.mark.parametrize("bias_dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("act", ["none", "gelu", "gelu_tanh", "relu"])
def test_gemm_bias_activation_epilogues(act, bias_dtype):
    from determined_b200.ops import gemm as G

    M, N, K = 384, 520, 256
    a, b = _mk(M, K, 5, 0.5), _mk(N, K, 6, 0.1)
    bias = torch.randn(N, device="cuda").to(bias_dtype)
    pre_want = a.float() @ b.float().t() + bias.float()
    f = {"none": lambda z: z, "gelu": F.gelu, "gelu_tanh": lambda z: F.gelu(z, approximate="tanh"), "relu": F.relu}[act]
    aux = torch.empty(M, N, device="cuda", dtype=torch.bfloat16) if act != "none" else None
    got = G.gemm(a, b, bias=bias, epilogue=act, aux=aux, out_dtype=torch.float32)
    torch.testing.assert_close(got, f(pre_want), rtol=2e-3, atol=2e-2)
    if aux is not None:
        torch.testing.assert_close(aux.float(), pre_want, rtol=1e-2, atol=2e-2)
        # derivative epilogue: (dy @ W) * act'(pre)
        dy = _mk(M, N, 7)
        w2 = _mk(N, 264, 8, 0.1)            # [N(k), 264(n)] read MN-major
        got_d = G.gemm(dy, w2, b_mn_major=True, epilogue="d" + act, aux=_mk(M, 264, 9), out_dtype=torch.float32)
        pre2 = _mk(M, 264, 9).float().requires_grad_(True)
        (f(pre2)).sum().backward()
        torch.testing.assert_close(got_d, (dy.float() @ w2.float()) * pre2.grad, rtol=3e-3, atol=3e-2)


def test_linear_and_mlp_autograd_match_torch():
    from determined_b200.ops import gemm as G

    torch.manual_seed(0)
    x = torch.randn(8, 96, 256, device="cuda", requires_grad=True)
    mlp = G.FusedMLP(256, 1024).cuda()
    lin = G.FusedLinear(256, 520).cuda()
    y = lin(mlp(x))
    y.float().square().mean().backward()
    got = [x.grad.clone()] + [p.grad.clone() for p in list(mlp.parameters()) + list(lin.parameters())]
    x.grad = None
    for p in list(mlp.parameters()) + list(lin.parameters()):
        p.grad = None
    h = F.linear(F.gelu(F.linear(x, mlp.fc1.weight, mlp.fc1.bias)), mlp.fc2.weight, mlp.fc2.bias)
    yr = F.linear(h, lin.weight, lin.bias)
    torch.testing.assert_close(y.float(), yr, rtol=3e-2, atol=3e-2)
    yr.square().mean().backward()
    want = [x.grad] + [p.grad for p in list(mlp.parameters()) + list(lin.parameters())]
    for g, w in zip(got, want):
        assert g.dtype == w.dtype and g.shape == w.shape
        rel = (g.float() - w).norm() / (w.norm() + 1e-12)
        assert rel < 2e-2, rel


def test_gemm_rejects_misaligned():
    from determined_b200.ops import gemm as G

    a = _mk(128, 72, 1)[:, 1:65]
    with pytest.raises(ValueError):
        G.gemm(a, _mk(128, 64, 2))
    with pytest.raises(TypeError):
        G.gemm(torch.randn(128, 64, device="cuda"), _mk(128, 64, 2))


def test_zero3_layers_single_rank():
    """The gather- and reduce-scatter-fused GEMMs with a world of one (no peers to wait for): partitioned linear and MLP
    reproduce the plain fused layers, forward and backward."""
    from determined_b200.ops import comm, gemm_comm
    from determined_b200.ops import gemm as G

    tp = gemm_comm.TensorParallelGemm(comm.PeerGroup.from_torch_distributed(torch.device("cuda", 0)))
    torch.manual_seed(0)
    lin = G.FusedLinear(512, 1024, activation="gelu").cuda()
    mlp = G.FusedMLP(512, 2048, activation="gelu_tanh").cuda()
    box = torch.nn.ModuleDict({"lin": lin, "mlp": mlp})
    import copy

    box3 = copy.deepcopy(box)
    assert gemm_comm.partition_linears_(box3, tp, min_params=1) == 3
    x = _mk(384, 512, 3).requires_grad_(True)
    x3 = x.detach().clone().requires_grad_(True)
    y = box["mlp"](x) + box["lin"](x)[:, :512]
    y3 = box3["mlp"](x3) + box3["lin"](x3)[:, :512]
    dy = _mk(384, 512, 4)
    y.backward(dy)
    y3.backward(dy)
    torch.cuda.synchronize()
    torch.testing.assert_close(y3.float(), y.float(), rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(x3.grad.float(), x.grad.float(), rtol=2e-2, atol=5e-2)
    for a, b in [(box["lin"].weight, box3["lin"].weight_shard), (box["mlp"].fc1.weight, box3["mlp"].fc1.weight_shard),
                 (box["mlp"].fc2.weight, box3["mlp"].fc2.weight_shard), (box["mlp"].fc1.bias, box3["mlp"].fc1.bias)]:
        rel = float((a.grad - b.grad).norm() / a.grad.norm())
        assert rel < 2e-2, rel


def test_zero3_engine_single_rank():
    """ZeRO stage 3 through the engine on one GPU: the partitioning pass, the local shard update and the checkpoint
    consolidation, against the stage-1 engine on the same data."""
    from determined_b200.models import gpt
    from determined_b200.parallel import zero

    def run(stage):
        torch.manual_seed(0)
        model = gpt.build("tiny", native=True, n_embd=256, n_head=4, n_layer=2, vocab_size=1024, block_size=128, rotary_pct=0.25,
                          parallel_residual=True, tie_embeddings=False)
        cfg = {"train_micro_batch_size_per_gpu": 4, "gradient_clipping": 1.0, "bf16": {"enabled": True},
               "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.1}},
               "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 1000}}
        eng, _, _, _ = zero.initialize(model=model, config=cfg)
        g = torch.Generator().manual_seed(5)
        losses = []
        for _ in range(6):
            idx = torch.randint(0, 1024, (4, 128), generator=g).cuda()
            _, loss = eng(idx, idx)
            eng.backward(loss)
            eng.step()
            losses.append(float(loss))
        return eng, losses

    e1, l1 = run(1)
    e3, l3 = run(3)
    assert e3.partitioned == 9, e3.partitioned          # 4 per block + the LM head (1024 rows)
    assert l3[-1] < l3[0]
    for a, b in zip(l1, l3):
        assert abs(a - b) < 3e-2 * abs(a), (l1, l3)
    sd1, sd3 = e1.consolidated_state_dict(), e3.consolidated_state_dict()
    assert set(sd1) == set(sd3)
    assert float((sd1["blocks.1.qkv.weight"] - sd3["blocks.1.qkv.weight"]).abs().max()) < 2e-2
