"""tcgen05 flash-attention forward vs PyTorch fused SDPA (CUDA events, L2 flushed).
    python benchmarks/attn_bench.py [--json out.json]"""
import argparse
import json
import math
import pathlib
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from determined_b200.ops import attention as A  # noqa: E402


def timeit(fn, flush, iters=15, warmup=4):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json")
    args = ap.parse_args()
    p = pathlib.Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"
    peak = float(json.loads(p.read_text()).get("bf16_tflops", 1654.0)) if p.exists() else 1654.0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    for (B, H, S, D) in [(8, 16, 2048, 128), (4, 16, 4096, 128), (2, 16, 8192, 128), (16, 12, 1024, 64), (8, 32, 2048, 64)]:
        for causal in (False, True):
            q, k, v = (torch.randn(B, H, S, D, device="cuda").to(torch.bfloat16) for _ in range(3))
            scale = 1.0 / math.sqrt(D)
            t_ours = timeit(lambda: A.attention_forward(q, k, v, causal, scale, "bhsd"), flush)
            t_ref = timeit(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=causal), flush)
            fl = 4.0 * B * H * S * S * D * (0.5 if causal else 1.0)
            row = {"B": B, "H": H, "S": S, "D": D, "causal": causal, "ours_ms": t_ours, "sdpa_ms": t_ref, "ours_tflops": fl / t_ours / 1e9,
                   "sdpa_tflops": fl / t_ref / 1e9, "ours_frac_of_measured_peak": fl / t_ours / 1e9 / peak}
            if D == 128:
                # backward only (2.5x the forward FLOPs): packed QKV through our kernel vs SDPA's backward
                qkv = torch.randn(B, S, 3, H, D, device="cuda").to(torch.bfloat16)
                o, lse = A.attention_forward(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal, scale, "bshd")
                do = torch.randn_like(o)
                dqkv = torch.empty_like(qkv)
                t_b = timeit(lambda: A.attention_backward_packed(qkv, o, lse, do, dqkv, causal, scale), flush)
                qr, kr, vr = (t.clone().requires_grad_(True) for t in (q, k, v))
                out_r = F.scaled_dot_product_attention(qr, kr, vr, is_causal=causal)
                do_r = torch.randn_like(out_r)
                t_rb = timeit(lambda: torch.autograd.grad(out_r, (qr, kr, vr), do_r, retain_graph=True), flush)
                row.update({"ours_bwd_ms": t_b, "sdpa_bwd_ms": t_rb, "ours_bwd_tflops": 2.5 * fl / t_b / 1e9, "sdpa_bwd_tflops": 2.5 * fl / t_rb / 1e9})
            rows.append(row)
            print(json.dumps(rows[-1]))
    if args.json:
        pathlib.Path(args.json).write_text(json.dumps({"peak_tflops": peak, "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
