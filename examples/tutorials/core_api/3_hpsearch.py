"""Stage 3: let the searcher decide how long to train and with which hyperparameters (reference 3_hpsearch.py)."""
import logging
import pathlib
import sys
from typing import Optional

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
            x += increment_by
            batch += 1
            if batch % 10 == 0:
                ctx.train.report_training_metrics(steps_completed=batch, metrics={"x": x})
                op.report_progress(batch)
                with ctx.checkpoint.store_path({"steps_completed": batch}) as (path, uuid):
                    (pathlib.Path(path) / "state").write_text(f"{x},{batch},{trial_id}")
                if ctx.preempt.should_preempt():
                    return
        ctx.train.report_validation_metrics(steps_completed=batch, metrics={"x": x})
        op.report_completed(x)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    info = det.get_cluster_info()
    latest = info.latest_checkpoint if info is not None else None
    tid = info.trial.trial_id if info is not None else -1
    inc = int(info.trial.hparams["increment_by"]) if info is not None else 1
    with core.init() as context:
        main(context, latest, tid, inc)
    sys.exit(0)
