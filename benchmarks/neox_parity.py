"""Loss-curve parity of the N-GPU ZeRO stage-3 run against a 1-GPU stage-1 run on the SAME data and global batch:
step s trains on sequences s*N .. s*N+N-1 of a fixed synthetic pool; the N-GPU run gives one sequence to each rank
(micro batch 1, accumulation 1), the 1-GPU run accumulates the same N sequences (micro batch 1, accumulation N).

    python benchmarks/neox_parity.py --role single --world 8 --out gpurun_out/parity_single.json          # 1 GPU
    torchrun --nproc-per-node 8 benchmarks/neox_parity.py --role multi --compare gpurun_out/parity_single.json --out ...
"""
import argparse
import json
import os
import pathlib
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from determined_b200.models import gpt  # noqa: E402
from determined_b200.parallel import zero  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--role", choices=["single", "multi"], required=True)
    ap.add_argument("--world", type=int, default=8, help="single: number of ranks to stand in for")
    ap.add_argument("--model", default="neox-1.3b")
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--compare")
    ap.add_argument("--out")
    a = ap.parse_args()
    multi = a.role == "multi"
    if multi:
        rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
        torch.cuda.set_device(local)
        dist.init_process_group("nccl")
    else:
        rank, local, world = 0, 0, a.world
        torch.cuda.set_device(0)
    dev = torch.device("cuda", local)
    vocab = (50304 + 256 * world - 1) // (256 * world) * (256 * world)
    g = torch.Generator().manual_seed(2024)
    pool = torch.randint(0, 50304, (a.steps * world, a.seq + 1), generator=g)
    torch.manual_seed(0)
    model = gpt.build(a.model, block_size=a.seq, vocab_size=vocab)
    cfg = {"train_micro_batch_size_per_gpu": 1, "gradient_accumulation_steps": 1 if multi else world, "bf16": {"enabled": True},
           "gradient_clipping": 1.0, "optimizer": {"type": "AdamW", "params": {"lr": 1e-4, "weight_decay": 0.1, "betas": [0.9, 0.95]}},
           "zero_optimization": {"stage": 3 if multi else 1}}
    eng, _, _, _ = zero.initialize(model=model, config=cfg)
    losses = []
    for s in range(a.steps):
        if multi:
            idx = pool[s * world + rank: s * world + rank + 1].to(dev)
            _, loss = eng(idx[:, :-1], idx[:, 1:])
            eng.backward(loss)
            eng.step()
            t = loss.detach().float().clone()
            dist.all_reduce(t)
            losses.append(float(t) / world)
        else:
            tot = 0.0
            for r in range(world):
                idx = pool[s * world + r: s * world + r + 1].to(dev)
                _, loss = eng(idx[:, :-1], idx[:, 1:])
                eng.backward(loss)
                eng.step()
                tot += float(loss)
            losses.append(tot / world)
    out = {"role": a.role, "world": world, "model": a.model, "seq": a.seq, "steps": a.steps, "losses": losses,
           "zero_stage": 3 if multi else 1, "partitioned_weights": eng.partitioned}
    if multi and a.compare and rank == 0:
        ref = json.loads(pathlib.Path(a.compare).read_text())["losses"]
        rel = [abs(x - y) / max(abs(y), 1e-9) for x, y in zip(losses, ref)]
        out.update(reference_losses=ref, max_rel_diff=max(rel), rel_diff_last=rel[-1], within_3e_2=max(rel) <= 3e-2)
        print("NEOX_PARITY max_rel_diff", max(rel), "first/last", losses[0], losses[-1], "ref", ref[0], ref[-1], flush=True)
    if rank == 0 and a.out:
        pathlib.Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        pathlib.Path(a.out).write_text(json.dumps(out, indent=1))
    if multi:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
