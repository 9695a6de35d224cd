"""``python -m determined.launch.horovod [--autohorovod] [--trial mod:Class] [--] script…``

Parity: reference ``harness/determined/launch/horovod.py:1-274`` + ``horovod.py:99-163``
(``horovodrun`` flags).  Horovod is not part of the B200 stack: gradient exchange belongs to the
harness' fused engine, so this launcher accepts the reference's arguments (fusion threshold,
cycle time, autotune — recorded in the environment for the engine's bucket sizing) and starts the
ranks with the torch contract."""
import argparse
import os
import sys
from typing import List

import determined_b200 as det
from determined_b200 import util
from determined_b200.launch import _common


def parse_args(argv: List[str]):
    p = argparse.ArgumentParser(prog="python3 -m determined.launch.horovod")
    p.add_argument("--autohorovod", action="store_true")
    p.add_argument("--trial", help="legacy trial class entrypoint module:Class")
    known, rest = p.parse_known_args(argv)
    overrides, script = _common.split_args(rest)
    if known.trial is not None:
        if script:
            p.error("script is mutually exclusive with --trial")
        script = util.legacy_trial_entrypoint_to_script(known.trial)
    elif not script:
        p.error("a script (or --trial) is required")
    return known, overrides, script


def main(argv: List[str]) -> int:
    known, overrides, script = parse_args(argv)
    env = {}
    info = det.get_cluster_info()
    if info is not None and info.task_type == "TRIAL":
        opt = det.ExperimentConfig(info.trial._config).get_optimizations_config()
        env["DETB200_BUCKET_CAP_MB"] = str(opt.get("tensor_fusion_threshold", 64))
    if known.autohorovod and _common.num_local_ranks() == 1:
        import subprocess

        p = subprocess.Popen(script)
        with util.forward_signals(p):
            return p.wait()
    return _common.launch(script, None, extra_env=env)


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
