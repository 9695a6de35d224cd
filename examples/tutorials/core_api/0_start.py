"""Stage 0: a script that knows nothing about the platform (reference examples/tutorials/core_api/0_start.py).
Run it as an experiment with 0_start.yaml: the master schedules it, captures its output, and that is all."""
import logging
import sys
import time


def main(increment_by: int) -> None:
    x = 0
    for batch in range(100):
        x += increment_by
        time.sleep(0.01)
        print("x is now", x)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    main(increment_by=1)
    sys.exit(0)
