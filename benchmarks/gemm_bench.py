"""tcgen05 GEMM throughput vs cuBLAS (torch.matmul) — CUDA events, L2 flushed between timed runs.

    python benchmarks/gemm_bench.py [--json out.json]
"""
import argparse
import json
import pathlib
import sys

import torch

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from determined_b200.ops import gemm as G  # noqa: E402


def timeit(fn, flush, iters=20, warmup=5):
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
    peaks = {}
    p = pathlib.Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"
    if p.exists():
        peaks = json.loads(p.read_text())
    peak = float(peaks.get("bf16_tflops", peaks.get("cublas_bf16_tflops", 1654.0)))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    shapes = [(8192, 8192, 8192), (4096, 4096, 4096), (16384, 4096, 4096), (8192, 16384, 4096), (8192, 4096, 16384), (2048, 2048, 2048),
              (8192, 2304, 768), (8192, 768, 3072), (8192, 50304, 768)]
    for (M, N, K) in shapes:
        for name, a_mn, b_mn in (("TN(fwd)", False, False), ("NN(dgrad)", False, True), ("NT(wgrad)", True, True)):
            a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(torch.bfloat16)
            b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(torch.bfloat16)
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            A = a.t() if a_mn else a
            B = b if b_mn else b.t()
            t_ours = timeit(lambda: G.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, out=out), flush)
            t_ref = timeit(lambda: torch.matmul(A, B, out=out), flush)
            fl = 2.0 * M * N * K
            rows.append({"M": M, "N": N, "K": K, "layout": name, "ours_ms": t_ours, "cublas_ms": t_ref,
                         "ours_tflops": fl / t_ours / 1e9, "cublas_tflops": fl / t_ref / 1e9, "ours_frac_of_measured_peak": fl / t_ours / 1e9 / peak})
            print(json.dumps(rows[-1]))
    if args.json:
        pathlib.Path(args.json).write_text(json.dumps({"peak_tflops": peak, "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
