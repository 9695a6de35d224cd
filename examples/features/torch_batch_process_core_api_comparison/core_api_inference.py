"""Variant B: the same job written directly against the Core API — the script shards the batches, decides when to
checkpoint, reads the checkpoint back on restart and keeps the workers in step itself."""
import json
import os
import pathlib
import sys

import torch
import torch.distributed as dist

import determined_b200 as det
import model as m
from determined_b200 import core

BATCH, EVERY = 50, 4

if __name__ == "__main__":
    info = det.get_cluster_info()
    distributed = None
    if "RANK" in os.environ:
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
        distributed = core.DistributedContext.from_torch_distributed()
    with core.init(distributed=distributed) as ctx:
        rank, size = ctx.distributed.rank, ctx.distributed.size
        net, data = m.get_model(), m.Images()
        batches = [range(i, min(i + BATCH, len(data))) for i in range(0, len(data), BATCH)]
        mine = batches[rank::size]
        start = 0
        if info is not None and info.latest_checkpoint:
            with ctx.checkpoint.restore_path(info.latest_checkpoint) as p:
                start = int((pathlib.Path(p) / "steps").read_text())
        out = pathlib.Path("comparison_out/core_api")
        out.mkdir(parents=True, exist_ok=True)
        steps = max(ctx.distributed.allgather(len(mine)))
        for i in range(start, steps):
            if i < len(mine):
                with torch.no_grad():
                    pred = net(torch.stack([data[j] for j in mine[i]])).argmax(1).tolist()
                (out / f"rank{rank}-batch{i:05d}.json").write_text(json.dumps(pred))
            if (i + 1) % EVERY == 0 or i + 1 == steps:
                ctx.distributed.allgather(i)                # nobody checkpoints "i + 1 done" before everybody is there
                if rank == 0:
                    with ctx.checkpoint.store_path({"steps_completed": i + 1}) as (p, _):
                        (pathlib.Path(p) / "steps").write_text(str(i + 1))
                if ctx.preempt.should_preempt():
                    sys.exit(0)
        ctx.distributed.allgather(0)
        if rank == 0:
            n = sum(len(json.loads(f.read_text())) for f in out.glob("*.json"))
            print(f"core api variant scored {n} records")
