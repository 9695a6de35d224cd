"""``python -m determined.launch.torch_distributed [torchrun overrides] -- script args…``

Parity: reference ``harness/determined/launch/torch_distributed.py:15-163`` (``--trial mod:Class``
legacy form supported)."""
import argparse
import sys
from typing import List

from determined_b200 import util
from determined_b200.launch import _common


def parse_args(argv: List[str]):
    p = argparse.ArgumentParser(prog="python3 -m determined.launch.torch_distributed", usage="%(prog)s [torch overrides] [--] script [args]")
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
    return _common.launch(script, overrides)


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
