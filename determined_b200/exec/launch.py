"""``python -m determined.exec.launch`` — first Python process of an allocation: turns the
experiment's ``entrypoint`` into the user's process tree.

Parity: reference ``harness/determined/exec/launch.py:1-96``: legacy ``module:Class`` entrypoints
become ``determined.launch.torch_distributed --trial …`` (the reference defaulted to Horovod);
string entrypoints run under ``sh -c``; checkpoint storage is validated first (``:85-94``);
SIGTERM is converted into a pending-preemption signal for Slurm-style terminations (``:18-26``).
"""
import logging
import os
import signal
import subprocess
import sys
from typing import List, Optional

import determined_b200 as det
from determined_b200 import util
from determined_b200.common import api, storage

logger = logging.getLogger("determined.exec.launch")


def mask_and_log_config(cfg: dict) -> None:
    logger.info("experiment config: %s", util.json_encode(util.mask_config_dict(cfg)))


def entrypoint_to_cmd(entrypoint) -> List[str]:
    if isinstance(entrypoint, str):
        if util.match_legacy_trial_class(entrypoint):
            return [sys.executable, "-m", "determined_b200.launch.torch_distributed", "--trial", entrypoint]
        return ["sh", "-c", _alias_modules(entrypoint)]
    if isinstance(entrypoint, list) and entrypoint:
        return [_alias_modules(x) if i < 4 else x for i, x in enumerate(entrypoint)]
    raise ValueError("experiment config has no entrypoint")


def _alias_modules(s: str) -> str:
    """Experiment configs written for the reference say ``-m determined.launch.…``; run ours."""
    for name in ("launch.torch_distributed", "launch.deepspeed", "launch.horovod", "launch.wrap_rank", "exec.harness"):
        s = s.replace(f"determined.{name}", f"determined_b200.{name}")
    return s


def launch(experiment_config) -> int:  # noqa: ANN001
    """Run the experiment's entrypoint as a child process and return its exit code (reference ``exec/launch.py:launch``)."""
    cfg = experiment_config if isinstance(experiment_config, dict) else dict(experiment_config)
    cmd = entrypoint_to_cmd(cfg.get("entrypoint"))
    logger.info("launching: %s", cmd)
    p = subprocess.Popen(cmd)
    return _wait_forwarding_signals(p)


def _wait_forwarding_signals(p, on_term=None) -> int:  # noqa: ANN001
    try:
        signal.signal(signal.SIGTERM, on_term or (lambda s, f: util.signal_process_tree(p, signal.SIGTERM)))
        signal.signal(signal.SIGINT, lambda s, f: util.signal_process_tree(p, signal.SIGINT))
    except ValueError:       # not the main thread: the caller owns signal handling
        pass
    return p.wait()


def main() -> int:
    logging.basicConfig(level=logging.INFO, format=det.LOG_FORMAT)
    info = det.get_cluster_info()
    if info is None or info.task_type != "TRIAL":
        print("determined.exec.launch must run inside a trial allocation", file=sys.stderr)
        return 2
    cfg = info.trial._config
    if cfg.get("debug"):
        logging.getLogger().setLevel(logging.DEBUG)
    mask_and_log_config(cfg)
    try:
        storage.validate_config(cfg["checkpoint_storage"], None)
    except Exception as e:
        logger.error("checkpoint storage validation failed: %s", e)
        return 1
    cmd = entrypoint_to_cmd(cfg.get("entrypoint"))
    logger.info("launching: %s", cmd)
    p = subprocess.Popen(cmd)

    def on_term(signum, frame) -> None:   # Slurm-style SIGTERM: ask for a graceful preemption first
        try:
            sess = api.Session(info.master_url, None, info.session_token, None, max_retries=1)
            sess.post(f"/api/v1/allocations/{info.allocation_id}/signals/pending_preemption")
        except Exception:
            pass
        util.signal_process_tree(p, signal.SIGTERM)

    return _wait_forwarding_signals(p, on_term)


if __name__ == "__main__":
    sys.exit(main())
