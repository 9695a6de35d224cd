"""Stage 2: save checkpoints, resume from the latest one, stop when preempted (reference 2_checkpoints.py).
Pause and re-activate the experiment to see it continue where it stopped."""
import logging
import pathlib
import sys
from typing import Optional

import determined_b200 as det
from determined_b200 import core


def save_state(x: int, steps: int, trial_id: int, path: pathlib.Path) -> None:
    (path / "state").write_text(f"{x},{steps},{trial_id}")


def load_state(trial_id: int, path: pathlib.Path):
    x, steps, ckpt_trial = (int(v) for v in (path / "state").read_text().split(","))
    # a checkpoint of another trial (a forked experiment) starts this trial's step count from zero
    return (x, steps) if ckpt_trial == trial_id else (x, 0)


def main(ctx: core.Context, latest_checkpoint: Optional[str], trial_id: int, increment_by: int) -> None:
    x, start = 0, 0
    if latest_checkpoint is not None:
        with ctx.checkpoint.restore_path(latest_checkpoint) as path:
            x, start = load_state(trial_id, pathlib.Path(path))
    steps = start
    for batch in range(start, 100):
        x += increment_by
        steps = batch + 1
        if steps % 10 == 0:
            ctx.train.report_training_metrics(steps_completed=steps, metrics={"x": x})
            with ctx.checkpoint.store_path({"steps_completed": steps}) as (path, uuid):
                save_state(x, steps, trial_id, pathlib.Path(path))
            if ctx.preempt.should_preempt():
                return
    ctx.train.report_validation_metrics(steps_completed=steps, metrics={"x": x})


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    info = det.get_cluster_info()
    latest = info.latest_checkpoint if info is not None else None
    tid = info.trial.trial_id if info is not None else -1
    with core.init() as context:
        main(context, latest, tid, increment_by=1)
    sys.exit(0)
