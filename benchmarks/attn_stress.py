"""Attention forward/backward under large-magnitude, growing and non-finite logits (exercises the lazy-rescale path)."""
import sys

import torch

from determined_b200.ops import attention as A


def ref(q, k, v, causal):
    return torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2).float(), k.transpose(1, 2).float(), v.transpose(1, 2).float(),
                                                            is_causal=causal).transpose(1, 2)


def main():
    torch.manual_seed(0)
    B, S, H, D = 2, 2048, 8, 128
    worst = 0.0
    for trial, (scale, ramp) in enumerate([(1.0, 0.0), (4.0, 0.0), (12.0, 0.0), (1.0, 6.0), (3.0, 20.0), (30.0, 0.0), (0.02, 0.0)]):
        for causal in (True, False):
            q = torch.randn(B, S, H, D, device="cuda") * scale
            k = torch.randn(B, S, H, D, device="cuda") * scale
            if ramp:          # keys whose scores grow along the sequence: the running max keeps jumping
                k = k + (torch.arange(S, device="cuda").float() / S * ramp)[None, :, None, None] * torch.sign(q.mean(1, keepdim=True))
            v = torch.randn(B, S, H, D, device="cuda")
            q, k, v = (t.to(torch.bfloat16).requires_grad_(True) for t in (q, k, v))
            o = A.flash_attention(q, k, v, causal=causal, layout="bshd")
            torch.cuda.synchronize()
            want = ref(q.detach(), k.detach(), v.detach(), causal)
            err = float((o.float() - want).abs().max())
            worst = max(worst, err)
            print(f"scale {scale} ramp {ramp} causal {causal}: max err {err:.4f} finite {bool(torch.isfinite(o).all())}", flush=True)
            qkv = torch.stack([q.detach(), k.detach(), v.detach()], 2).requires_grad_(True)      # [B,S,3,H,D]
            o2 = A.attention_qkvpacked(qkv, None, 0, causal=causal)
            o2.backward(torch.randn_like(o2))
            torch.cuda.synchronize()
            print("   packed fwd+bwd ok, dqkv finite", bool(torch.isfinite(qkv.grad).all()), flush=True)
    # non-finite inputs must not hang (values are garbage by definition)
    for bad in (float("nan"), float("inf")):
        q = torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16)
        q[0, 100, 0, 0] = bad
        k = torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16)
        k[1, 7, 3, 5] = bad
        v = torch.randn(B, S, H, D, device="cuda").to(torch.bfloat16)
        o = A.flash_attention(q, k, v, causal=True, layout="bshd")
        torch.cuda.synchronize()
        print("non-finite input", bad, "completed", flush=True)
    print("ATTN_STRESS worst", worst)
    return 0


if __name__ == "__main__":
    sys.exit(main())
