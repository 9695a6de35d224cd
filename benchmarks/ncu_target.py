"""Tiny single-kernel drivers for `ncu` captures:  python benchmarks/ncu_target.py {norm|gemm|attn}"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from determined_b200.ops import _lib  # noqa: E402


def norm():
    lib = _lib.load()
    rows_, cols = 16384, 4096
    x = torch.randn(rows_, cols, device="cuda", dtype=torch.bfloat16)
    dy = torch.randn_like(x)
    y, dx = torch.empty_like(x), torch.empty_like(x)
    g = torch.ones(cols, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(cols, device="cuda", dtype=torch.bfloat16)
    dg, db = torch.empty_like(g), torch.empty_like(g)
    mean, rstd = torch.empty(rows_, device="cuda"), torch.empty(rows_, device="cuda")
    ws = torch.empty(int(lib.detb200_norm_bwd_workspace_elems(rows_, cols, 0)), device="cuda")
    s = _lib.stream_ptr()
    for _ in range(3):
        lib.detb200_norm_fwd(x.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), rows_, cols, 1e-5, 1, 0, s)
        lib.detb200_norm_bwd(dy.data_ptr(), x.data_ptr(), g.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(), dg.data_ptr(), db.data_ptr(),
                             ws.data_ptr(), rows_, cols, 1, 0, s)
    torch.cuda.synchronize()


def gemm():
    from determined_b200.ops import gemm as G

    a = torch.randn(8192, 8192, device="cuda").to(torch.bfloat16)
    b = torch.randn(8192, 8192, device="cuda").to(torch.bfloat16)
    out = torch.empty(8192, 8192, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        G.gemm(a, b, out=out)
    torch.cuda.synchronize()


def attn():
    from determined_b200.ops import attention as A

    q = torch.randn(4, 16, 2048, 128, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        A.flash_attention(q, q, q, causal=True)
    torch.cuda.synchronize()


if __name__ == "__main__":
    {"norm": norm, "gemm": gemm, "attn": attn}[sys.argv[1]]()
