"""An unmanaged run that saves checkpoints and finds itself again: start it, stop it (Ctrl-C), start it again with the same
``external ids`` and it continues in the same trial from its last checkpoint (reference examples/features/unmanaged/2_checkpoints.py).

    DET_MASTER=http://127.0.0.1:8080 python 2_checkpoints.py [--steps N] [--run-id NAME] [--checkpoint-dir DIR]
"""
import argparse
import pathlib
import random

from determined_b200.experimental import core_v2


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--run-id", default="unmanaged-2-checkpoints")
    ap.add_argument("--checkpoint-dir", default="/tmp/unmanaged-checkpoints")
    a = ap.parse_args()
    core_v2.init(
        defaults=core_v2.DefaultConfig(name="unmanaged-2-checkpoints", hparams={"lr": 0.1}, checkpoint_storage=a.checkpoint_dir),
        unmanaged=core_v2.UnmanagedConfig(external_experiment_id=a.run_id, external_trial_id=a.run_id),
    )
    start, loss = 0, 1.0
    latest = core_v2.info.latest_checkpoint if core_v2.info is not None else None
    if latest is not None:
        with core_v2.checkpoint.restore_path(latest) as path:
            s, l = (pathlib.Path(path) / "state").read_text().split(",")
            start, loss = int(s), float(l)
        print(f"resumed from checkpoint {latest} at step {start}")
    for step in range(start, a.steps):
        loss *= 0.97 + 0.02 * random.random()
        if step % 10 == 9:
            core_v2.train.report_training_metrics(steps_completed=step + 1, metrics={"loss": loss})
            with core_v2.checkpoint.store_path({"steps_completed": step + 1}) as (path, uuid):
                (pathlib.Path(path) / "state").write_text(f"{step + 1},{loss}")
    print("experiment page:", core_v2.url_reverse_webui_exp_view())
    core_v2.close()


if __name__ == "__main__":
    main()
