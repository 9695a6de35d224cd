"""Small-shape sweep over the native kernels for compute-sanitizer runs:
    compute-sanitizer --tool memcheck python benchmarks/sanitize_target.py"""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from determined_b200.ops import attention as A, batchnorm, fused, gemm as G, norm, xent  # noqa: E402


def main() -> None:
    dev = torch.device("cuda")
    torch.manual_seed(0)
    # GEMM: 1-CTA (M <= 128), CTA pair, tails, every layout, epilogues, reduce-add
    for (M, N, K) in [(128, 136, 72), (520, 264, 200), (300, 520, 264)]:
        for a_mn in (False, True):
            for b_mn in (False, True):
                if a_mn and M % 8:
                    continue
                a = torch.randn((K, M) if a_mn else (M, K), device=dev).to(torch.bfloat16)
                b = torch.randn((K, N) if b_mn else (N, K), device=dev).to(torch.bfloat16)
                G.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn)
                out = torch.zeros(M, N, device=dev)
                G.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out=out, accumulate=True)
    x = torch.randn(384, 256, device=dev).to(torch.bfloat16)
    w = torch.randn(520, 256, device=dev).to(torch.bfloat16)
    aux = torch.empty(384, 520, device=dev, dtype=torch.bfloat16)
    G.gemm(x, w, bias=torch.randn(520, device=dev), epilogue="gelu", aux=aux)
    G.gemm(torch.randn(384, 520, device=dev).to(torch.bfloat16), torch.randn(520, 264, device=dev).to(torch.bfloat16), b_mn_major=True, epilogue="dgelu",
           aux=torch.randn(384, 264, device=dev).to(torch.bfloat16))
    # attention forward + backward (both head dims, tails, causal)
    for (B, S, H, D) in [(1, 200, 2, 128), (2, 333, 2, 64)]:
        qkv = torch.randn(B, S, 3, H, D, device=dev).to(torch.bfloat16).requires_grad_(True)
        o = A.attention_qkvpacked(qkv * 1.0, None, 0, causal=True)
        o.backward(torch.randn_like(o))
    q = torch.randn(1, 4, 77, 64, device=dev).to(torch.bfloat16)
    k = torch.randn(1, 2, 333, 64, device=dev).to(torch.bfloat16)
    A.attention_forward(q, k, k, False, 1 / math.sqrt(64), "bhsd")
    # norms, cross-entropy, glue
    for cols in (256, 1536, 4096, 770):
        t = torch.randn(37, cols, device=dev, dtype=torch.bfloat16, requires_grad=True)
        g = torch.ones(cols, device=dev, requires_grad=True)
        norm.layer_norm(t, g, torch.zeros(cols, device=dev, requires_grad=True), 1e-5).sum().backward()
        norm.rms_norm(t, g, 1e-6).sum().backward()
    lg = torch.randn(33, 50257, device=dev, dtype=torch.bfloat16, requires_grad=True)
    xent.cross_entropy(lg, torch.randint(0, 50257, (33,), device=dev)).backward()
    fused.colsum(torch.randn(301, 520, device=dev).to(torch.bfloat16))
    # BatchNorm with fused ReLU / residual
    bn = batchnorm.FusedBatchNorm2d(64).to(dev)
    xi = torch.randn(8, 64, 8, 8, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    res = torch.randn_like(xi).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    bn.forward_fused(xi, res, True).sum().backward()
    torch.cuda.synchronize()
    print("SANITIZE_TARGET_OK")


if __name__ == "__main__":
    main()
