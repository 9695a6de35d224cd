"""Distributed inference with the Core API: every worker scores its shard of the test set, per-shard results are written
as the job goes and progress is checkpointed, so a preempted job picks up at the first batch it had not finished
(reference examples/features/inference_mnist_pytorch/inference.py).

    python3 inference.py                                                     # one process, no master
    det experiment create distributed_inference.yaml .                      # two workers on a cluster
"""
import json
import logging
import os
import pathlib
import sys

import torch
import torch.distributed as dist
from torch import nn

import determined_b200 as det
from determined_b200 import core


def model_for_inference() -> nn.Module:
    """A trained model would come from a checkpoint: ``client.get_checkpoint(uuid).download()`` + ``load_state_dict``
    (hyperparameter ``checkpoint_uuid``); without one the script scores with fixed random weights."""
    torch.manual_seed(7)
    return nn.Sequential(nn.Flatten(), nn.Linear(784, 64), nn.ReLU(), nn.Linear(64, 10)).eval()


def main(ctx: core.Context, hp: dict, latest_checkpoint) -> None:
    rank, size = ctx.distributed.rank, ctx.distributed.size
    dev = torch.device(f"cuda:{ctx.distributed.local_rank}" if torch.cuda.is_available() else "cpu")
    model = model_for_inference().to(dev)
    if hp.get("checkpoint_uuid"):
        from determined_b200.experimental import client

        path = client.get_checkpoint(hp["checkpoint_uuid"]).download()
        model.load_state_dict(torch.load(pathlib.Path(path) / "model.pt", map_location=dev))
    g = torch.Generator().manual_seed(1)
    images = torch.randn(int(hp.get("n_records", 1000)), 1, 28, 28, generator=g)      # stand-in for the MNIST test images
    bs = int(hp.get("batch_size", 100))
    batches = [images[i:i + bs] for i in range(0, len(images), bs)]
    mine = list(range(rank, len(batches), size))                                       # batch b belongs to worker b % size
    done = 0
    if latest_checkpoint is not None:
        with ctx.checkpoint.restore_path(latest_checkpoint) as p:
            done = int(json.loads((pathlib.Path(p) / "progress.json").read_text())["batches_per_worker"])
    out_dir = pathlib.Path(hp.get("output_dir", "inference_out"))
    out_dir.mkdir(parents=True, exist_ok=True)
    steps = max(ctx.distributed.allgather(len(mine)))       # every worker takes part in every synchronisation point
    for i in range(done, steps):
        if i < len(mine):
            with torch.no_grad():
                pred = model(batches[mine[i]].to(dev)).argmax(1).cpu().tolist()
            (out_dir / f"batch-{mine[i]:05d}.json").write_text(json.dumps(pred))
        if (i + 1) % int(hp.get("checkpoint_every", 2)) == 0 or i + 1 == steps:
            counts = ctx.distributed.gather(i + 1)
            if rank == 0:
                ctx.train.report_validation_metrics(steps_completed=i + 1, metrics={"batches_scored": float(sum(min(c, len(range(r, len(batches), size))) for r, c in enumerate(counts)))})
                with ctx.checkpoint.store_path({"steps_completed": i + 1}) as (p, uuid):
                    (pathlib.Path(p) / "progress.json").write_text(json.dumps({"batches_per_worker": i + 1}))
            if ctx.preempt.should_preempt():
                return
    ctx.distributed.allgather(0)                             # everybody has written its last file
    if rank == 0:
        n = sum(len(json.loads(f.read_text())) for f in out_dir.glob("batch-*.json"))
        print(f"scored {n} records into {out_dir}/")


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO, format=det.LOG_FORMAT)
    info = det.get_cluster_info()
    distributed = None
    if "RANK" in os.environ:
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
        distributed = core.DistributedContext.from_torch_distributed()
    with core.init(distributed=distributed) as context:
        main(context, dict(info.trial.hparams) if info is not None else {}, info.latest_checkpoint if info is not None else None)
    sys.exit(0)
