"""Stage 4: the same loop on several workers (reference 4_distributed.py).  Every worker adds its share, the chief gathers
the results, reports metrics and writes the checkpoint; all workers agree on preemption.

    python3 -m determined_b200.launch.torch_distributed python3 4_distributed.py      (what 4_distributed.yaml runs)
"""
import logging
import pathlib
import sys
from typing import Optional

import torch.distributed as dist

import determined_b200 as det
from determined_b200 import core


def main(ctx: core.Context, latest_checkpoint: Optional[str], trial_id: int, increment_by: int) -> None:
    x, start = 0, 0
    if latest_checkpoint is not None:
        with ctx.checkpoint.restore_path(latest_checkpoint) as path:
            x, start, ckpt_trial = (int(v) for v in (pathlib.Path(path) / "state").read_text().split(","))
            start = start if ckpt_trial == trial_id else 0
    batch = start
    for op in ctx.searcher.operations():
        while batch < op.length:
            # each worker contributes; the sum over workers is what one batch adds
            x += sum(ctx.distributed.allgather(increment_by))
            batch += 1
            if batch % 10 == 0:
                if ctx.distributed.rank == 0:
                    ctx.train.report_training_metrics(steps_completed=batch, metrics={"x": x})
                    op.report_progress(batch)
                    with ctx.checkpoint.store_path({"steps_completed": batch}) as (path, uuid):
                        (pathlib.Path(path) / "state").write_text(f"{x},{batch},{trial_id}")
                if ctx.preempt.should_preempt():      # the chief asks the master and tells the others
                    return
        if ctx.distributed.rank == 0:
            ctx.train.report_validation_metrics(steps_completed=batch, metrics={"x": x})
            op.report_completed(x)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    info = det.get_cluster_info()
    latest = info.latest_checkpoint if info is not None else None
    tid = info.trial.trial_id if info is not None else -1
    inc = int(info.trial.hparams["increment_by"]) if info is not None else 1
    distributed = None
    if dist.is_available() and "RANK" in __import__("os").environ:
        dist.init_process_group("gloo")
        distributed = core.DistributedContext.from_torch_distributed()
    with core.init(distributed=distributed) as context:
        main(context, latest, tid, inc)
    sys.exit(0)
