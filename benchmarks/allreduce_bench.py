"""Gradient all-reduce alone: the fused kernel (sum * scale + sum-of-squares + non-finite check in one pass) in its two
transports — NVSwitch in-switch reduction (multimem.ld_reduce / multimem.st on a multicast mapping) and two-shot
peer-to-peer loads/stores — over bucket sizes and CTA budgets, against NCCL all-reduce followed by the same scaling and
norm as separate kernels.  Device-timed (CUDA events), max over ranks, 5 warm-ups; every size is larger than nothing in
particular, so between timed iterations a 256 MB buffer is rewritten to flush the 126 MB L2.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/allreduce_bench.py --json out.json

Roofline (B200_PROFILING.md: measured peer copy 770 GB/s per direction per GPU): the bytes one GPU must send are
``(W-1)/W * N`` for the scatter half and the same again for the gather half with peer-to-peer traffic (2 (W-1)/W N), and about
``N`` with the in-switch reduction (each GPU's whole buffer is read by the switch once; the multicast write costs the
sender N/W).  ``frac_of_nvlink`` = (those bytes / 770 GB/s) / measured time.
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

NVLINK_GBS = 770.0


def timed(fn, iters, warmup, dev, flush):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.add_(1.0)                     # 256 MB rewritten: nothing of the bucket survives in L2
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    t = torch.tensor([tot / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", default="1,4,16,25,45,64,256")
    ap.add_argument("--ctas", default="4,8,16,32,64,148")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--json")
    a = ap.parse_args()
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl")
    group = comm.PeerGroup.from_torch_distributed(dev)
    sizes = [int(float(x) * (1 << 20)) // 4 for x in a.sizes_mb.split(",")]
    ctas = [int(x) for x in a.ctas.split(",")]
    n_max = max(sizes)
    align = 8 * world
    n_max = (n_max + align - 1) // align * align
    buf = group.alloc(4 * n_max)
    peers = group.map(buf)
    mc = group.multicast(buf)
    mine = buf.tensor(torch.float32, n_max)
    stage_buf = group.alloc(2 * n_max)
    peer_stage = group.map(stage_buf)
    mc_stage = group.multicast(stage_buf)
    stage = stage_buf.tensor(torch.bfloat16, n_max)
    nccl_buf = torch.empty(n_max, device=dev)
    flush = torch.zeros(64 << 20, device=dev)
    rows = []
    for n in sizes:
        n = n // align * align
        mine[:n].normal_()
        nccl_buf[:n].copy_(mine[:n])

        def nccl():
            t = nccl_buf[:n]
            dist.all_reduce(t)
            t.mul_(1.0 / world)
            torch.linalg.vector_norm(t)
            torch.isfinite(t).all()

        def nccl_only():
            dist.all_reduce(nccl_buf[:n])

        ms_n = timed(nccl, a.iters, a.warmup, dev, flush)
        ms_n0 = timed(nccl_only, a.iters, a.warmup, dev, flush)
        row = {"mbytes": n * 4 / 2**20, "nccl_plus_scale_norm_check_ms": ms_n, "nccl_allreduce_only_ms": ms_n0, "variants": []}
        for transport, mcp in (("nvls", mc), ("p2p", None)):
            if transport == "nvls" and not mc:
                continue
            for c in ctas:
                def ours():
                    group.begin_step()
                    comm.allreduce_fused(group, peers, 0, n, 1.0 / world, found_inf=group.found_inf, timeout_s=20, mc_ptr=mcp, max_ctas=c)

                ms = timed(ours, a.iters, a.warmup, dev, flush)
                wire = (n * 4) if transport == "nvls" else 2.0 * (world - 1) / world * n * 4
                row["variants"].append({"transport": transport, "ctas": c, "ms": ms, "algo_GBs": n * 4 / ms / 1e6,
                                        "frac_of_nvlink": wire / (ms * 1e-3) / (NVLINK_GBS * 1e9), "speedup_vs_nccl_fused_work": ms_n / ms,
                                        "speedup_vs_nccl_allreduce_only": ms_n0 / ms})

        def ours_bf16():
            group.begin_step()
            comm.narrow_bf16(mine[:n], stage[:n])
            comm.allreduce_bf16(group, peer_stage, 0, n, 1.0 / world, found_inf=group.found_inf, timeout_s=20, mc_ptr=mc_stage)
            comm.widen_bf16(stage[:n], mine[:n])

        row["bf16_wire_ms"] = timed(ours_bf16, a.iters, a.warmup, dev, flush)
        best = min(row["variants"], key=lambda v: v["ms"])
        row["best"] = best
        rows.append(row)
        if rank == 0:
            print(f"{row['mbytes']:8.1f} MB  nccl+work {ms_n:.3f} ms  nccl {ms_n0:.3f} ms  best {best['transport']}/{best['ctas']} "
                  f"{best['ms']:.3f} ms ({best['frac_of_nvlink']:.2f} of NVLink)  bf16 wire {row['bf16_wire_ms']:.3f} ms", flush=True)
    out = {"world": world, "nvls": bool(mc), "nvlink_GBs_per_direction_measured": NVLINK_GBS, "timing": "CUDA events, max over ranks, "
           f"{a.warmup} warm-up, {a.iters} timed, L2 flushed (256 MB write) between iterations", "rows": rows}
    if rank == 0:
        print("ALLREDUCE_BENCH", json.dumps(out), flush=True)
        if a.json:
            pathlib.Path(a.json).parent.mkdir(parents=True, exist_ok=True)
            pathlib.Path(a.json).write_text(json.dumps(out, indent=1))
    dist.barrier()
    group.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
