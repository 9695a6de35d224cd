"""Numerics of every native kernel against a plain PyTorch fp32 reference of the same op."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from determined_b200.ops import _lib

    return _lib


@pytest.mark.parametrize("kind", ["sgd", "sgd_mom", "sgd_nesterov", "adam", "adamw", "adadelta"])
@pytest.mark.parametrize("n", [1, 7, 4096 + 3, 1 << 20])
def test_fused_optimizers_match_torch(kind, n):
    from determined_b200.ops import optim
    from determined_b200.parallel.flat import FlatLayout

    torch.manual_seed(0)
    dev = torch.device("cuda")
    shapes = [(n,)] if n < 1000 else [(n // 2,), (n - n // 2 - 5,), (5,)]
    mk = lambda: [torch.nn.Parameter(torch.randn(s, device=dev)) for s in shapes]  # noqa: E731
    ref_p, fus_p = mk(), None
    fus_p = [torch.nn.Parameter(p.detach().clone()) for p in ref_p]
    def build(ps):
        if kind == "sgd":
            return torch.optim.SGD(ps, lr=0.1, weight_decay=0.01)
        if kind == "sgd_mom":
            return torch.optim.SGD(ps, lr=0.1, momentum=0.9, dampening=0.1)
        if kind == "sgd_nesterov":
            return torch.optim.SGD(ps, lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-3)
        if kind == "adam":
            return torch.optim.Adam(ps, lr=1e-2, weight_decay=0.01)
        if kind == "adamw":
            return torch.optim.AdamW(ps, lr=1e-2, weight_decay=0.1, betas=(0.9, 0.95))
        return torch.optim.Adadelta(ps, lr=1.0, rho=0.9, weight_decay=1e-3)
    ref, fus = build(ref_p), build(fus_p)
    lay = FlatLayout(fus.param_groups, dev)
    assert optim.fuse_optimizer_(fus, lay)
    for step in range(4):
        grads = [torch.randn_like(p) for p in ref_p]
        for p, g in zip(ref_p, grads):
            p.grad = g.clone()
        for p, g in zip(fus_p, grads):
            p.grad.copy_(g)
        ref.step()
        fus.step(zero_grad=True)
        for a, b in zip(ref_p, fus_p):
            torch.testing.assert_close(b.detach(), a.detach(), rtol=2e-5, atol=2e-6)
        assert float(lay.flat_grad.abs().max()) == 0.0
    # state dict round trip through the stock optimizer
    sd = fus.state_dict()
    ref2 = build([torch.nn.Parameter(p.detach().clone()) for p in ref_p])
    sd.pop("detb200_fused")
    ref2.load_state_dict(sd)


def test_adam_grad_scale_and_skip():
    from determined_b200.ops import optim
    from determined_b200.parallel.flat import FlatLayout

    dev = torch.device("cuda")
    p = torch.nn.Parameter(torch.randn(1000, device=dev))
    q = torch.nn.Parameter(p.detach().clone())
    ref, fus = torch.optim.AdamW([p], lr=1e-2), torch.optim.AdamW([q], lr=1e-2)
    lay = FlatLayout(fus.param_groups, dev)
    optim.fuse_optimizer_(fus, lay)
    g = torch.randn_like(p)
    p.grad = g * 0.25
    q.grad.copy_(g)
    scale = torch.tensor([0.5], device=dev)
    ref.step()
    fus.step(grad_scale=scale, static_scale=0.5)
    torch.testing.assert_close(q.detach(), p.detach(), rtol=2e-5, atol=2e-6)
    before = q.detach().clone()
    q.grad.copy_(g)
    skip = torch.ones(1, dtype=torch.int32, device=dev)
    fus.step(skip=skip, zero_grad=True)
    torch.testing.assert_close(q.detach(), before, rtol=0, atol=0)
    assert float(q.grad.abs().max()) == 0.0


@pytest.mark.parametrize("n", [5, 1024, (1 << 22) + 1])
def test_sumsq_and_clip_coef(n):
    from determined_b200.ops import optim

    dev = torch.device("cuda")
    x = torch.randn(n + 3, device=dev)[:n] if False else torch.randn(n, device=dev)
    sc = optim.NormScratch(dev)
    for _ in range(2):  # the scratch ticket must re-arm itself
        out = optim.sumsq(x, sc)
        torch.testing.assert_close(out[0].double(), x.double().pow(2).sum(), rtol=1e-5, atol=1e-6)
    optim.sumsq(x, sc, accumulate=True)
    torch.testing.assert_close(sc.sumsq[0].double(), 2 * x.double().pow(2).sum(), rtol=1e-5, atol=1e-6)
    optim.sumsq(x, sc)
    coef = optim.clip_coef(sc.sumsq, 1.0, sc, extra_scale=2.0)
    norm = x.double().norm()
    torch.testing.assert_close(coef[0].double(), 2.0 * torch.clamp(1.0 / (norm + 1e-6), max=1.0), rtol=1e-5, atol=1e-7)
    torch.testing.assert_close(sc.norm[0].double(), norm, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,cols", [(1, 2), (37, 1024), (513, 2560), (64, 8192), (300, 770), (1001, 768), (130, 1536), (9, 256), (77, 4096), (5, 8)])
@pytest.mark.parametrize("rms", [False, True])
def test_norm_fwd_bwd(dtype, rows, cols, rms):
    from determined_b200.ops import norm

    torch.manual_seed(1)
    dev = torch.device("cuda")
    x = torch.randn(rows, cols, device=dev, dtype=dtype, requires_grad=True)
    g = (torch.randn(cols, device=dev) * 0.5 + 1).to(dtype).requires_grad_()
    b = torch.randn(cols, device=dev).to(dtype).requires_grad_()
    dy = torch.randn(rows, cols, device=dev, dtype=dtype)
    xr, gr, br = (t.detach().float().requires_grad_() for t in (x, g, b))
    if rms:
        y = norm.rms_norm(x, g, 1e-6)
        yr = xr * torch.rsqrt(xr.pow(2).mean(-1, keepdim=True) + 1e-6) * gr
    else:
        y = norm.layer_norm(x, g, b, 1e-5)
        yr = torch.nn.functional.layer_norm(xr, (cols,), gr, br, 1e-5)
    y.backward(dy)
    yr.backward(dy.float())
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(y.float(), yr, **tol)
    torch.testing.assert_close(x.grad.float(), xr.grad, **tol)
    gtol = dict(rtol=3e-2, atol=3e-2 * (rows ** 0.5)) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(g.grad.float(), gr.grad, **gtol)
    if not rms:
        torch.testing.assert_close(b.grad.float(), br.grad, **gtol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,vocab", [(1, 10), (33, 50257), (128, 50304), (7, 131)])
def test_xent_fwd_bwd(dtype, rows, vocab):
    from determined_b200.ops import xent

    torch.manual_seed(2)
    dev = torch.device("cuda")
    logits = (torch.randn(rows, vocab, device=dev) * 3).to(dtype).requires_grad_()
    tgt = torch.randint(0, vocab, (rows,), device=dev)
    if rows > 2:
        tgt[1] = -100
    ref_logits = logits.detach().float().requires_grad_()
    loss = xent.cross_entropy(logits, tgt)
    ref = torch.nn.functional.cross_entropy(ref_logits, tgt, ignore_index=-100)
    (loss * 2).backward()
    (ref * 2).backward()
    torch.testing.assert_close(loss, ref, rtol=1e-4, atol=1e-4)
    tol = dict(rtol=2e-2, atol=1e-3) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(logits.grad.float(), ref_logits.grad, **tol)
    # in-place backward gives the same gradient
    l2 = logits.detach().clone().requires_grad_()
    xent.cross_entropy(l2 * 1.0, tgt, inplace_backward=True).mul(2).backward()
    torch.testing.assert_close(l2.grad.float(), ref_logits.grad, **tol)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_allreduce_fused_emulated_ranks(world):
    """Single-GPU emulation: all `world` rank buffers live on this device and the reduce phase is
    launched once per emulated rank (no inter-kernel spinning — see B200_PROFILING.md)."""
    from determined_b200.ops import _lib, comm

    lib = _lib.load()
    dev = torch.device("cuda")
    n = 4 * world * 1000
    groups = [comm.PeerGroup(0, 1, None, dev) for _ in range(world)]
    bufs = [comm.SymmetricBuffer(4 * n, dev) for _ in range(world)]
    tens = [b.tensor(torch.float32, n) for b in bufs]
    torch.manual_seed(3)
    for t in tens:
        t.copy_(torch.randn(n, device=dev))
    expect = torch.stack([t.clone() for t in tens]).double().sum(0) * (0.5 / world)
    ptrs = (ctypes.c_void_p * world)(*[b.ptr for b in bufs])
    ctrls = (ctypes.c_void_p * world)(*[g.ctrl.ptr for g in groups])
    found = torch.zeros(1, dtype=torch.int32, device=dev)
    for r in range(world):
        _lib.check(lib.detb200_allreduce_fused(ptrs, ctrls, r, world, 0, n, 0.5 / world, None, found.data_ptr(), 1, 2,
                                               0.0, _lib.stream_ptr()))
    torch.cuda.synchronize()
    for t in tens:
        torch.testing.assert_close(t.double(), expect, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("world,cpr", [(2, 3), (4, 2), (8, 1), (8, 4)])
def test_allreduce_full_protocol_emulated_in_one_cooperative_launch(world, cpr):
    """The whole collective (start barrier, reduce + scatter, statistics exchange, end barrier, epoch advance) for `world`
    ranks on ONE GPU: the CTAs of all ranks form one cooperative grid, so the flag spins are between co-resident CTAs of
    one kernel.  Several buckets back to back and two steps exercise the epoch counters and the running statistics."""
    from determined_b200.ops import _lib, comm

    lib = _lib.load()
    dev = torch.device("cuda")
    n = 4 * world * 3001
    groups = [comm.PeerGroup(0, 1, None, dev) for _ in range(world)]
    bufs = [comm.SymmetricBuffer(4 * n, dev) for _ in range(world)]
    tens = [b.tensor(torch.float32, n) for b in bufs]
    ptrs = (ctypes.c_void_p * world)(*[b.ptr for b in bufs])
    ctrls = (ctypes.c_void_p * world)(*[g.ctrl.ptr for g in groups])
    found = torch.zeros(world, dtype=torch.int32, device=dev)
    inv = torch.full((1,), 0.25, device=dev)
    scale = 1.0 / world
    cut = (n // 3) // (4 * world) * (4 * world)
    for step in range(2):
        torch.manual_seed(10 + step)
        for t in tens:
            t.copy_(torch.randn(n, device=dev))
        expect = torch.stack([t.clone() for t in tens]).double().sum(0) * (scale * 0.25)
        for bi, (off, cnt) in enumerate([(0, cut), (cut, n - cut)]):
            _lib.check(lib.detb200_allreduce_fused_emulated(ptrs, ctrls, world, off, cnt, scale, inv.data_ptr(), found.data_ptr(),
                                                            int(bi == 0), cpr, 20.0, _lib.stream_ptr()), "allreduce_emulated")
        torch.cuda.synchronize()
        for t in tens:
            torch.testing.assert_close(t.double(), expect, rtol=1e-5, atol=1e-6)
        # every rank's table holds every rank's slice statistics: their sum is the squared norm of the result
        for g in groups:
            tot = float(g.stats_tensor()[0, :world].sum())
            torch.testing.assert_close(torch.tensor(tot), expect.pow(2).sum().float().cpu(), rtol=1e-3, atol=1e-3)
        assert int(found.sum()) == 0
    # a non-finite value in ONE rank's buffer raises found_inf on EVERY rank
    for t in tens:
        t.zero_()
    tens[world - 1][7] = float("inf")
    _lib.check(lib.detb200_allreduce_fused_emulated(ptrs, ctrls, world, 0, n, scale, None, found.data_ptr(), 1, cpr, 20.0,
                                                    _lib.stream_ptr()), "allreduce_emulated")
    torch.cuda.synchronize()
    assert found.tolist() == [1] * world


def test_resnet_step_fused_vs_torch():
    """One optimizer's worth of ResNet-18 params: fused engine (world=1) must track torch SGD."""
    from determined_b200 import core, pytorch
    from determined_b200.models.resnet_cifar import resnet18_cifar

    torch.manual_seed(0)
    torch.backends.cudnn.allow_tf32 = False     # exact fp32 on both sides
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda")
    ref_model = resnet18_cifar().to(dev)
    x = torch.randn(16, 3, 32, 32, device=dev)
    y = torch.randint(0, 10, (16,), device=dev)
    with pytorch.init(hparams={"global_batch_size": 16}) as ctx:
        import copy

        model = ctx.wrap_model(copy.deepcopy(ref_model))
        opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4))
        ref_opt = torch.optim.SGD(ref_model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
        assert type(opt).__name__ == "FusedSGD"
        ctx._current_batch_idx = 0
        for i in range(3):
            ctx._current_batch_idx = i
            loss = torch.nn.functional.cross_entropy(model(x), y)
            ctx.backward(loss)
            ctx.step_optimizer(opt)
            rl = torch.nn.functional.cross_entropy(ref_model(x), y)
            ref_opt.zero_grad()
            rl.backward()
            ref_opt.step()
            torch.testing.assert_close(loss, rl, rtol=1e-3, atol=1e-3)
        for a, b in zip(model.parameters(), ref_model.parameters()):
            torch.testing.assert_close(a, b, rtol=5e-3, atol=1e-3)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(4, 64, 8, 8), (32, 128, 16, 16), (7, 512, 4, 4), (256, 64, 32, 32), (3, 8, 5, 7)])
def test_batchnorm_train_fwd_bwd(dtype, shape):
    from determined_b200.ops import batchnorm

    torch.manual_seed(4)
    dev = torch.device("cuda")
    N, C, H, W = shape
    x = (torch.randn(shape, device=dev) * 2 + 0.7).to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_()
    ref = torch.nn.BatchNorm2d(C).to(dev)
    with torch.no_grad():
        ref.weight.uniform_(0.5, 1.5)
        ref.bias.uniform_(-0.5, 0.5)
    import copy

    mine = copy.deepcopy(ref)
    assert batchnorm.convert_batchnorm_(mine) == 1
    dy = torch.randn(shape, device=dev).to(dtype).contiguous(memory_format=torch.channels_last)
    xr = x.detach().float().requires_grad_()
    for _ in range(2):   # twice: running stats accumulate and the workspace ticket must re-arm
        y = mine(x)
        yr = ref(xr)
    y.backward(dy)
    yr.backward(dy.float())
    tol = dict(rtol=2e-2, atol=3e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(y.float(), yr, **tol)
    torch.testing.assert_close(x.grad.float(), xr.grad, **tol)
    gtol = dict(rtol=3e-2, atol=0.05 * (N * H * W) ** 0.5) if dtype == torch.bfloat16 else dict(rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(mine.weight.grad, ref.weight.grad, **gtol)
    torch.testing.assert_close(mine.bias.grad, ref.bias.grad, **gtol)
    torch.testing.assert_close(mine.running_mean, ref.running_mean, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(mine.running_var, ref.running_var, rtol=1e-3, atol=1e-3)
    assert int(mine.num_batches_tracked) == int(ref.num_batches_tracked) == 2
    # eval mode goes through the stock path and must agree exactly
    mine.eval(); ref.eval()
    torch.testing.assert_close(mine(x.detach().float()), ref(x.detach().float()), rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_batchnorm_fused_relu_and_residual(dtype):
    from determined_b200.ops import batchnorm

    torch.manual_seed(3)
    dev = torch.device("cuda")
    N, C, H, W = 16, 64, 8, 8
    x = torch.randn(N, C, H, W, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last).requires_grad_()
    res = torch.randn(N, C, H, W, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last).requires_grad_()
    dy = torch.randn(N, C, H, W, device=dev, dtype=dtype).contiguous(memory_format=torch.channels_last)
    for use_res in (False, True):
        bn = batchnorm.FusedBatchNorm2d(C).to(dev)
        ref = torch.nn.BatchNorm2d(C).to(dev)
        with torch.no_grad():
            bn.weight.uniform_(0.5, 1.5)
            bn.bias.uniform_(-0.5, 0.5)
            ref.load_state_dict(bn.state_dict())
        y = bn.forward_fused(x, res if use_res else None, True)
        y.backward(dy)
        gx, gr = x.grad.clone(), (res.grad.clone() if use_res else None)
        gw, gb = bn.weight.grad.clone(), bn.bias.grad.clone()
        x.grad = None
        res.grad = None
        xr, rr = x.detach().float().requires_grad_(), res.detach().float().requires_grad_()
        z = ref(xr)
        yr = torch.relu(z + rr) if use_res else torch.relu(z)
        yr.backward(dy.float())
        tol = dict(rtol=3e-2, atol=3e-2) if dtype == torch.bfloat16 else dict(rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(y.float(), yr, **tol)
        # entries within rounding of the ReLU kink may be gated differently in bf16: compare in norm
        rel = lambda a, b: float((a.float() - b.float()).norm() / (b.float().norm() + 1e-9))  # noqa: E731
        lim = 3e-2 if dtype == torch.bfloat16 else 1e-4
        assert rel(gx, xr.grad) < lim and rel(gw, ref.weight.grad) < lim and rel(gb, ref.bias.grad) < lim
        if use_res:
            assert rel(gr, rr.grad) < lim
        torch.testing.assert_close(bn.running_var, ref.running_var, rtol=1e-2, atol=1e-3)


def test_residual_block_fusion_is_verified_and_matches():
    from determined_b200.models.resnet_cifar import BasicBlock, resnet18_cifar
    from determined_b200.ops import batchnorm

    torch.manual_seed(4)
    dev = torch.device("cuda")
    model = resnet18_cifar().to(dev).to(memory_format=torch.channels_last)
    ref = resnet18_cifar().to(dev).to(memory_format=torch.channels_last)
    ref.load_state_dict(model.state_dict())
    assert batchnorm.convert_batchnorm_(model) == 20
    assert batchnorm.fuse_residual_blocks_(model) == 8

    class Odd(BasicBlock):                 # same children, different math: must not be patched
        def forward(self, x):
            return super().forward(x) * 2.0

    odd = Odd(64, 64).to(dev)
    batchnorm.convert_batchnorm_(odd)
    assert batchnorm.fuse_residual_blocks_(odd) == 0
    x = torch.randn(32, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        out, out_ref = model(x), ref(x)
        torch.testing.assert_close(out, out_ref, rtol=1e-3, atol=1e-3)
        out.square().mean().backward()
        out_ref.square().mean().backward()
        for (n, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
            rel = float((a.grad - b.grad).norm() / (b.grad.norm() + 1e-9))
            assert rel < 1e-2, (n, rel)      # cuDNN picks its algorithms per run: 5e-3 was within run-to-run noise
    finally:
        torch.backends.cudnn.allow_tf32 = tf32


def test_fused_glue_kernels():
    from determined_b200.ops import fused

    torch.manual_seed(6)
    x = torch.randn(3001, 520, device="cuda").to(torch.bfloat16)
    torch.testing.assert_close(fused.colsum(x), x.float().sum(0), rtol=1e-3, atol=5e-2)
    view = torch.randn(777, 1024, device="cuda").to(torch.bfloat16)[:, 128:128 + 260]
    out = torch.ones(260, device="cuda")
    fused.colsum(view, out=out, accumulate=True)
    torch.testing.assert_close(out, 1 + view.float().sum(0), rtol=1e-3, atol=5e-2)
    a, b, c = (torch.randn(4, 100, 64, device="cuda").to(torch.bfloat16) for _ in range(3))
    torch.testing.assert_close(fused.add(a, b, c).float(), a.float() + b.float() + c.float(), rtol=1e-2, atol=3e-2)
    torch.testing.assert_close(fused.add(a, b), a + b)


def test_accum_multi_adds_many_tensors_in_one_launch():
    """fp32 destinations inside one flat buffer += bf16 / fp32 sources of assorted sizes (vector bodies, scalar tails,
    misaligned starts, more tensors than one launch holds) against a plain PyTorch loop."""
    from determined_b200.ops import _lib

    lib = _lib.load()
    dev = torch.device("cuda")
    torch.manual_seed(5)
    sizes = [1, 7, 8, 9, 64, 1000, 4099, 36864, 3 * 3 * 64 * 128 + 5] * 7        # 63 tensors > 48 per launch
    flat = torch.randn(sum(s + 3 for s in sizes) + 16, device=dev)
    want = flat.clone()
    srcs, dsts = [], []
    off = 1                                                                          # misaligned on purpose
    for i, n in enumerate(sizes):
        dt = torch.bfloat16 if i % 3 else torch.float32
        g = (torch.randn(n + 1, device=dev) * 3).to(dt)[1:] if i % 5 == 0 else (torch.randn(n, device=dev) * 3).to(dt)
        srcs.append(g)
        dsts.append(flat[off:off + n])
        want[off:off + n] += g.float()
        off += n + 3
    k = len(srcs)
    sp = (ctypes.c_void_p * k)(*[g.data_ptr() for g in srcs])
    dp = (ctypes.c_void_p * k)(*[d.data_ptr() for d in dsts])
    nn_ = (ctypes.c_int64 * k)(*[g.numel() for g in srcs])
    kinds = bytes(1 if g.dtype == torch.bfloat16 else 0 for g in srcs)
    _lib.check(lib.detb200_accum_multi(sp, dp, nn_, kinds, k, _lib.stream_ptr()), "accum_multi")
    torch.cuda.synchronize()
    torch.testing.assert_close(flat, want, rtol=1e-6, atol=1e-6)


def test_conv_pass_bf16_shadow_and_gradient_sink_match_autocast():
    """A small conv net under bf16 autocast: the engine's path (bf16 weight mirror, weight gradients through the
    multi-tensor accumulation) against stock modules + AccumulateGrad, over aggregation, an optimizer step, a weight clamp
    written through ``.data`` between steps (WGAN style) and an eval forward."""
    import copy

    from determined_b200 import pytorch
    from determined_b200.ops import conv as conv_ops

    torch.manual_seed(0)
    dev = torch.device("cuda")

    def make():
        return torch.nn.Sequential(torch.nn.Conv2d(3, 32, 3, padding=1, bias=False), torch.nn.BatchNorm2d(32), torch.nn.ReLU(),
                                   torch.nn.Conv2d(32, 64, 3, stride=2, padding=1, bias=False), torch.nn.BatchNorm2d(64), torch.nn.ReLU(),
                                   torch.nn.Conv2d(64, 64, 1, bias=True), torch.nn.AdaptiveAvgPool2d(1), torch.nn.Flatten(),
                                   torch.nn.Linear(64, 10)).to(dev).to(memory_format=torch.channels_last)

    base = make()
    ref = copy.deepcopy(base)
    ro = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
    with pytorch.init(hparams={"global_batch_size": 16}, aggregation_frequency=2) as ctx:
        model = ctx.wrap_model(copy.deepcopy(base))
        opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9))
        convs = [m for m in model.modules() if isinstance(m, conv_ops.FusedConv2d)]
        assert len(convs) == 3 and convs[0]._detb200_lowp_weight is not None and convs[0]._detb200_lowp_weight.dtype == torch.bfloat16
        eng = ctx._get_engine()
        rec = next(iter(eng.records.values()))
        assert rec.sink is not None and rec.shadow is not None
        g = torch.Generator().manual_seed(3)
        for i in range(6):
            ctx._current_batch_idx = i
            x = torch.randn(16, 3, 16, 16, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
            y = torch.randint(0, 10, (16,), generator=g).to(dev)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = torch.nn.functional.cross_entropy(model(x).float(), y)
                rl = torch.nn.functional.cross_entropy(ref(x).float(), y)
            ctx.backward(loss)
            (rl / 2).backward()
            if i % 2 == 1:
                # compare the accumulated gradients before they are consumed (ours are averaged by the step: x2 here)
                for (n1, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
                    torch.testing.assert_close(p.grad / 2, q.grad, rtol=3e-2, atol=3e-3, msg=lambda m, n1=n1: f"{n1}: {m}")
                ctx.step_optimizer(opt)
                ro.step()
                ro.zero_grad()
                with torch.no_grad():               # WGAN-style clamp through .data: invisible to version counters
                    for p, q in zip(model.parameters(), ref.parameters()):
                        p.data.clamp_(-0.5, 0.5)
                        q.data.clamp_(-0.5, 0.5)
            else:
                ctx.step_optimizer(opt)             # not a boundary: a no-op
            torch.testing.assert_close(loss, rl, rtol=2e-2, atol=2e-2)
        for p, q in zip(model.parameters(), ref.parameters()):
            torch.testing.assert_close(p, q, rtol=2e-2, atol=3e-3)
        # the mirror is what the convolutions read: after the clamp + refresh it equals the rounded parameters
        model.eval()
        ref.eval()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            xe = torch.randn(4, 3, 16, 16, device=dev).contiguous(memory_format=torch.channels_last)
            torch.testing.assert_close(model(xe).float(), ref(xe).float(), rtol=3e-2, atol=3e-2)
        torch.testing.assert_close(convs[0]._detb200_lowp_weight.float(), convs[0].weight.detach().to(torch.bfloat16).float())
        # outside autocast, and for a bare loss.backward(), the module is a plain Conv2d again
        model.train()
        out = model(xe)
        assert out.dtype == torch.float32
        out.sum().backward()
