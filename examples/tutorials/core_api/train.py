"""Core API walk-through: metrics, checkpoints, preemption and searcher operations in a plain loop
(reference ``examples/tutorials/core_api/{1_metrics,2_checkpoints,3_hpsearch,4_distributed}.py`` in one file)."""
import logging
import pathlib
import sys

import determined_b200 as det
from determined_b200 import core


def load_state(path: pathlib.Path) -> int:
    return int((path / "state").read_text())


def main(ctx: core.Context, latest_checkpoint: str, increment_by: int) -> None:
    x = 0
    start = 0
    if latest_checkpoint is not None:
        with ctx.checkpoint.restore_path(latest_checkpoint) as p:
            x = load_state(pathlib.Path(p))
            start = x // max(increment_by, 1)
    for op in ctx.searcher.operations():
        batch = start
        while batch < op.length:
            x += increment_by
            batch += 1
            if batch % 10 == 0:
                ctx.train.report_training_metrics(steps_completed=batch, metrics={"x": x})
                op.report_progress(batch)
                with ctx.checkpoint.store_path({"steps_completed": batch}) as (p, uuid):
                    (pathlib.Path(p) / "state").write_text(str(x))
                if ctx.preempt.should_preempt():
                    return
        ctx.train.report_validation_metrics(steps_completed=batch, metrics={"x": x})
        op.report_completed(x)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    info = det.get_cluster_info()
    latest = info.latest_checkpoint if info is not None else None
    inc = int(info.trial.hparams["increment_by"]) if info is not None else 1
    with core.init() as context:
        main(context, latest, inc)
    sys.exit(0)
