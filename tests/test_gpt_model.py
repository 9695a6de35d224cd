"""GPT family: CPU path (torch math behind the same ops) and GPU path (native kernels) agree."""
import pytest
import torch

from determined_b200.models import gpt


@pytest.mark.parametrize("name,kw", [("tiny", {}), ("tiny", {"rotary_pct": 0.25, "parallel_residual": True, "tie_embeddings": False})])
def test_gpt_cpu_trains(name, kw):
    torch.manual_seed(0)
    m = gpt.build(name, **kw)
    opt = torch.optim.AdamW(m.parameters(), lr=3e-3)
    idx = torch.randint(0, m.cfg.vocab_size, (4, 64))
    losses = []
    for _ in range(8):
        _, loss = m(idx[:, :-1], idx[:, 1:])
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0] - 0.3, losses
    logits, _ = m(idx[:, :-1])
    assert logits.shape == (4, 63, m.cfg.vocab_size)
    if kw.get("tie_embeddings", True):
        assert m.lm_head.weight is m.wte.weight
    assert m.flops_per_token(64) > 6 * m.num_params() * 0.5


def test_presets_have_expected_sizes():
    with torch.device("meta"):
        assert abs(gpt.build("gpt2").num_params() / 124e6 - 1) < 0.05
        assert abs(gpt.build("neox-1.3b").num_params() / 1.4e9 - 1) < 0.1


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [{}, {"rotary_pct": 0.5, "parallel_residual": True, "tie_embeddings": False}])
def test_gpt_gpu_matches_cpu(kw):
    torch.manual_seed(0)
    cfg = dict(n_layer=2, n_head=2, n_embd=128, block_size=256, vocab_size=1024)
    cfg.update(kw)
    ref = gpt.GPT(gpt.GPTConfig(**cfg))
    dev = gpt.GPT(gpt.GPTConfig(**cfg)).cuda()
    dev.load_state_dict(ref.state_dict())
    idx = torch.randint(0, 1024, (2, 257))
    _, l_ref = ref(idx[:, :-1], idx[:, 1:])
    l_ref.backward()
    _, l_dev = dev(idx[:, :-1].cuda(), idx[:, 1:].cuda())
    l_dev.backward()
    assert abs(float(l_ref) - float(l_dev)) < 3e-2, (float(l_ref), float(l_dev))
    for (n, a), (_, b) in zip(ref.named_parameters(), dev.named_parameters()):
        rel = (a.grad - b.grad.cpu().float()).norm() / (a.grad.norm() + 1e-8)
        assert rel < 8e-2, (n, float(rel))


def test_native_and_torch_twins_agree_on_cpu():
    """The op-based model and the stock-torch twin are the same function (rotary convention, packed QKV
    layout, parallel residual) and exchange state dicts."""
    torch.manual_seed(3)
    cfg = gpt.GPTConfig(n_layer=2, n_head=4, n_embd=128, block_size=64, vocab_size=256, rotary_pct=0.5, parallel_residual=True, tie_embeddings=False)
    a, b = gpt.GPT(cfg, native=True), gpt.GPT(cfg, native=False)
    b.load_state_dict(a.state_dict())
    idx = torch.randint(0, 256, (3, 65))
    _, la = a(idx[:, :-1], idx[:, 1:])
    _, lb = b(idx[:, :-1], idx[:, 1:])
    torch.testing.assert_close(la, lb, rtol=1e-3, atol=1e-3)        # the op-based path rounds GEMM operands to bf16
    la.backward()
    lb.backward()
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        rel = float((p.grad - q.grad).norm() / (q.grad.norm() + 1e-12))
        assert rel < 2e-2, (n, rel)
