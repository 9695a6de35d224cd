"""Gradient all-reduce alone: the fused two-shot P2P kernel (sum * scale + sum-of-squares + non-finite check in one pass) vs
NCCL all-reduce followed by the same scaling and norm as separate kernels, over bucket sizes; device-timed, max over ranks.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/allreduce_bench.py --json out.json

Roofline: a two-shot all-reduce moves 2 * (W-1)/W * bytes per GPU over NVLink in each direction; the reported
``frac_of_nvlink`` is that traffic over 900 GB/s per direction.
"""
import argparse
import json
import os
import pathlib
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import comm  # noqa: E402

NVLINK_GBS = 900.0


def timed(fn, iters, warmup, dev):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", default="1,4,16,25,64,256")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--json")
    a = ap.parse_args()
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl")
    group = comm.PeerGroup.from_torch_distributed(dev)
    sizes = [int(float(x) * (1 << 20)) // 4 for x in a.sizes_mb.split(",")]
    n_max = max(sizes)
    align = 4 * world
    n_max = (n_max + align - 1) // align * align
    buf = comm.SymmetricBuffer(4 * n_max, dev)
    peers = group.map(buf)
    mine = buf.tensor(torch.float32, n_max)
    nccl_buf = torch.empty(n_max, device=dev)
    rows = []
    for n in sizes:
        n = n // align * align
        mine[:n].normal_()
        nccl_buf[:n].copy_(mine[:n])

        def ours():
            group.begin_step()
            comm.allreduce_fused(group, peers, 0, n, 1.0 / world, found_inf=group.found_inf, timeout_s=20)

        def nccl():
            t = nccl_buf[:n]
            dist.all_reduce(t)
            t.mul_(1.0 / world)
            torch.linalg.vector_norm(t)
            torch.isfinite(t).all()

        ms_o = timed(ours, a.iters, a.warmup, dev)
        ms_n = timed(nccl, a.iters, a.warmup, dev)
        wire = 2.0 * (world - 1) / world * n * 4
        rows.append({"mbytes": n * 4 / 2**20, "fused_ms": ms_o, "nccl_plus_scale_norm_check_ms": ms_n, "speedup": ms_n / ms_o,
                     "fused_algo_GBs": n * 4 / ms_o / 1e6, "frac_of_nvlink": wire / (ms_o * 1e-3) / (NVLINK_GBS * 1e9)})
    out = {"world": world, "rows": rows}
    if rank == 0:
        print("ALLREDUCE_BENCH", json.dumps(out), flush=True)
        if a.json:
            pathlib.Path(a.json).write_text(json.dumps(out, indent=1))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
