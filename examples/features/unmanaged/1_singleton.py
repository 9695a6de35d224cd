"""An unmanaged ("detached") experiment: this script was not launched by the master, it just reports to one.

    DET_MASTER=http://127.0.0.1:8080 python 1_singleton.py

Reference counterpart: examples/features/unmanaged/1_singleton.py.
"""
import random

from determined_b200.experimental import core_v2


def main() -> None:
    core_v2.init(
        defaults=core_v2.DefaultConfig(name="unmanaged-1-singleton", hparams={"lr": 0.1}, labels=["unmanaged"]),
        # unmanaged=core_v2.UnmanagedConfig(workspace="...", project="..."),
    )
    loss = 1.0
    for step in range(100):
        loss *= 0.97 + 0.02 * random.random()
        if step % 10 == 9:
            core_v2.train.report_training_metrics(steps_completed=step + 1, metrics={"loss": loss})
        if step % 50 == 49:
            core_v2.train.report_validation_metrics(steps_completed=step + 1, metrics={"val_loss": loss * 1.1})
    print("experiment page:", core_v2.url_reverse_webui_exp_view())
    core_v2.close()


if __name__ == "__main__":
    main()
