"""``python -m determined.launch.deepspeed [--trial mod:Class] [--] script…``

Parity: reference ``harness/determined/launch/deepspeed.py:1-411``.  The reference builds a
hostfile, starts sshd on workers and shells out to the ``deepspeed`` launcher over pdsh; on one box
the same ranks are started locally with the DeepSpeed environment contract (``RANK, LOCAL_RANK,
WORLD_SIZE, LOCAL_SIZE, CROSS_RANK, CROSS_SIZE``) and ``USE_DEEPSPEED=1`` (``:307``)."""
import argparse
import sys
from typing import List

from determined_b200 import util
from determined_b200.launch import _common


def parse_args(argv: List[str]):
    p = argparse.ArgumentParser(prog="python3 -m determined.launch.deepspeed")
    p.add_argument("--trial", help="legacy trial class entrypoint module:Class")
    known, rest = p.parse_known_args(argv)
    overrides, script = _common.split_args(rest)
    if known.trial is not None:
        if script:
            p.error("script is mutually exclusive with --trial")
        script = util.legacy_trial_entrypoint_to_script(known.trial)
    elif not script:
        p.error("a script (or --trial) is required")
    return overrides, script


def main(argv: List[str]) -> int:
    overrides, script = parse_args(argv)
    return _common.launch(script, None, extra_env={"USE_DEEPSPEED": "1"})


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
