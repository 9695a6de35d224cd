"""Stage 1: report training and validation metrics through the Core API (reference 1_metrics.py)."""
import logging
import sys

from determined_b200 import core


def main(ctx: core.Context, increment_by: int) -> None:
    x = 0
    for batch in range(100):
        x += increment_by
        steps = batch + 1
        if steps % 10 == 0:
            ctx.train.report_training_metrics(steps_completed=steps, metrics={"x": x})
    ctx.train.report_validation_metrics(steps_completed=steps, metrics={"x": x})


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    with core.init() as context:
        main(context, increment_by=1)
    sys.exit(0)
