"""tcgen05 flash attention forward vs fp32 PyTorch (layouts, causal, GQA, tails), plus autograd wiring."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _ref(q, k, v, causal, scale):
    # q,k,v: [B,H,S,D] fp32
    if k.shape[1] != q.shape[1]:
        rep = q.shape[1] // k.shape[1]
        k, v = k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1)
    s = (q @ k.transpose(-1, -2)) * scale
    if causal:
        Sq, Sk = q.shape[2], k.shape[2]
        mask = torch.ones(Sq, Sk, dtype=torch.bool, device=q.device).tril(Sk - Sq)
        s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, -1)
    return torch.softmax(s, -1) @ v, lse


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("cfg", [(2, 4, 4, 256, 256, 128), (1, 8, 2, 384, 384, 64), (2, 3, 3, 200, 200, 128), (1, 2, 2, 1000, 1000, 64),
                                 (1, 4, 4, 128, 512, 128), (1, 2, 1, 77, 333, 64), (1, 16, 16, 2048, 2048, 128)])
def test_attention_forward_matches_reference(cfg, causal):
    from determined_b200.ops import attention as A

    B, H, Hkv, Sq, Skv, D = cfg
    torch.manual_seed(0)
    q = torch.randn(B, H, Sq, D, device="cuda").to(torch.bfloat16)
    k = torch.randn(B, Hkv, Skv, D, device="cuda").to(torch.bfloat16)
    v = torch.randn(B, Hkv, Skv, D, device="cuda").to(torch.bfloat16)
    scale = 1.0 / math.sqrt(D)
    want, want_lse = _ref(q.float(), k.float(), v.float(), causal, scale)
    got, lse = A.attention_forward(q, k, v, causal, scale, "bhsd")
    torch.cuda.synchronize()
    torch.testing.assert_close(got.float(), want, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(lse, want_lse, rtol=1e-3, atol=2e-3)


def test_attention_packed_qkv_bshd_and_autograd():
    from determined_b200.ops import attention as A

    B, S, H, D = 2, 320, 6, 64
    torch.manual_seed(1)
    qkv = torch.randn(B, S, 3, H, D, device="cuda").to(torch.bfloat16).requires_grad_(True)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]                # strided views, consumed in place
    o = A.flash_attention(q, k, v, causal=True, layout="bshd")
    assert o.shape == (B, S, H, D)
    do = torch.randn_like(o)
    o.backward(do)
    g = qkv.grad.clone()
    qkv.grad = None
    qf = qkv.detach().float().requires_grad_(True)
    want, _ = _ref(qf[:, :, 0].transpose(1, 2), qf[:, :, 1].transpose(1, 2), qf[:, :, 2].transpose(1, 2), True, 1.0 / math.sqrt(D))
    want.transpose(1, 2).backward(do.float())
    torch.testing.assert_close(o.float(), want.transpose(1, 2), rtol=2e-2, atol=2e-2)
    rel = (g.float() - qf.grad).norm() / qf.grad.norm()
    assert rel < 3e-2, rel


def test_attention_large_scores_stay_finite():
    from determined_b200.ops import attention as A

    torch.manual_seed(2)
    q = (torch.randn(1, 2, 256, 128, device="cuda") * 6).to(torch.bfloat16)
    k = (torch.randn(1, 2, 256, 128, device="cuda") * 6).to(torch.bfloat16)
    v = torch.randn(1, 2, 256, 128, device="cuda").to(torch.bfloat16)
    got, lse = A.attention_forward(q, k, v, False, 1.0, "bhsd")
    want, want_lse = _ref(q.float(), k.float(), v.float(), False, 1.0)
    assert torch.isfinite(got.float()).all() and torch.isfinite(lse).all()
    torch.testing.assert_close(got.float(), want, rtol=3e-2, atol=3e-2)
    torch.testing.assert_close(lse, want_lse, rtol=1e-3, atol=1e-2)


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("cfg", [(2, 256, 2, 128, 0), (1, 384, 3, 128, 32), (2, 200, 2, 128, 64), (1, 1024, 4, 128, 0),
                                 (2, 256, 3, 64, 0), (1, 333, 2, 64, 16), (1, 1024, 12, 64, 0)])
def test_attention_qkvpacked_native_backward(cfg, causal):
    """Packed QKV + in-place rotary + tcgen05 forward and backward vs fp32 math on the same inputs."""
    from determined_b200.models import gpt as G
    from determined_b200.ops import attention as A

    B, S, H, D, rot = cfg
    torch.manual_seed(5)
    qkv0 = (torch.randn(B, S, 3, H, D, device="cuda") * 0.7).to(torch.bfloat16)
    do = torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16)
    rope = None
    if rot:
        cos_f, sin_f, cos_h, sin_h = G._rotary_tables(S, rot, 10000.0, torch.device("cuda"))
        rope = (cos_h, sin_h)
    x = qkv0.clone().requires_grad_(True)
    o = A.attention_qkvpacked(x * 1.0, rope, rot, causal=causal)          # x * 1.0: the op consumes (rotates) its input buffer
    o.backward(do)
    got = x.grad.float()
    # reference in fp32
    xr = qkv0.float().requires_grad_(True)
    q, k, v = xr[:, :, 0], xr[:, :, 1], xr[:, :, 2]
    if rot:
        q, k = G.apply_rotary(q, cos_f, sin_f, rot), G.apply_rotary(k, cos_f, sin_f, rot)
    want, _ = _ref(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), causal, 1.0 / math.sqrt(D))
    want = want.transpose(1, 2)
    want.backward(do.float())
    torch.testing.assert_close(o.float(), want, rtol=2e-2, atol=2e-2)
    for i, name in enumerate("qkv"):
        g, w = got[:, :, i], xr.grad[:, :, i]
        rel = float((g - w).norm() / (w.norm() + 1e-9))
        assert rel < 2.5e-2, (name, rel)


@pytest.mark.parametrize("D", [64, 128])
def test_forward_survives_late_v_tiles(D):
    """Regression: the softmax may finish P(j) and P(j+1) before V(j) lands (S(j+1) only needs K(j+1)).  With a single
    "P ready" barrier that put the producer two phases ahead of the MMA thread's parity wait and the kernel hung; seen on
    8-GPU runs where peers' NVLink pushes slow the V loads.  The debug hook holds every V tile back to force that order."""
    from determined_b200.ops import _lib
    from determined_b200.ops import attention as A

    lib = _lib.load()
    torch.manual_seed(0)
    B, S, H = 2, 1024, 4
    q, k, v = (torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16) for _ in range(3))
    want = F.scaled_dot_product_attention(q.transpose(1, 2).float(), k.transpose(1, 2).float(), v.transpose(1, 2).float(),
                                          is_causal=True).transpose(1, 2)
    try:
        for ns in (3000, 20000):
            lib.detb200_attn_debug_delay_v(ns)
            o = A.flash_attention(q, k, v, causal=True, layout="bshd")
            torch.cuda.synchronize()
            torch.testing.assert_close(o.float(), want, rtol=2e-2, atol=2e-2)
    finally:
        lib.detb200_attn_debug_delay_v(0)


def test_large_and_growing_logits():
    """Row maxima that keep jumping (the lazy-rescale path) and logits far outside the unit range; non-finite inputs must
    not hang either."""
    from determined_b200.ops import attention as A

    torch.manual_seed(1)
    B, S, H, D = 1, 1024, 4, 128
    for scale, ramp in [(12.0, 0.0), (3.0, 20.0)]:
        q = torch.randn(B, S, H, D, device="cuda") * scale
        k = torch.randn(B, S, H, D, device="cuda") * scale
        if ramp:
            k = k + (torch.arange(S, device="cuda").float() / S * ramp)[None, :, None, None] * torch.sign(q.mean(1, keepdim=True))
        v = torch.randn(B, S, H, D, device="cuda")
        q, k, v = (t.to(torch.bfloat16) for t in (q, k, v))
        for causal in (True, False):
            o = A.flash_attention(q, k, v, causal=causal, layout="bshd")
            want = F.scaled_dot_product_attention(q.transpose(1, 2).float(), k.transpose(1, 2).float(), v.transpose(1, 2).float(),
                                                  is_causal=causal).transpose(1, 2)
            torch.testing.assert_close(o.float(), want, rtol=3e-2, atol=4e-2)
    q = torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16)
    q[0, 100, 0, 0] = float("nan")
    k = torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16)
    k[0, 7, 3, 5] = float("inf")
    A.flash_attention(q, k, torch.randn_like(q), causal=True, layout="bshd")
    torch.cuda.synchronize()
