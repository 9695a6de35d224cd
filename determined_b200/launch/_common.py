import logging
import os
import pathlib
import shutil
import signal
import subprocess
import sys
from typing import Dict, List, Optional

import determined_b200 as det
from determined_b200 import constants, util

logger = logging.getLogger("determined.launch")
NATIVE_LAUNCHER = pathlib.Path(__file__).resolve().parent.parent / "_native" / "detb200-launch"


def num_local_ranks() -> int:
    info = det.get_cluster_info()
    if info is not None:
        return max(len(info.slot_ids), 1)
    if os.environ.get("DET_SLOT_IDS"):
        import json

        return max(len(json.loads(os.environ["DET_SLOT_IDS"])), 1)
    try:
        import torch

        return max(torch.cuda.device_count(), 1)
    except Exception:
        return 1


def c10d_port() -> int:
    base = int(os.environ.get("C10D_PORT", constants.C10D_PORT))
    # several trials share the box: offset by the first slot id so rendezvous ports never clash
    info = det.get_cluster_info()
    off = 0
    if info is not None and info.slot_ids:
        off = int(info.slot_ids[0])
    elif os.environ.get("DET_SLOT_IDS"):
        import json

        ids = json.loads(os.environ["DET_SLOT_IDS"])
        off = int(ids[0]) if ids else 0
    return base + off


def native_cmd(nproc: int, script: List[str], port: Optional[int] = None) -> List[str]:
    return [str(NATIVE_LAUNCHER), "--nproc", str(nproc), "--master-addr", "127.0.0.1", "--master-port",
            str(port or c10d_port()), "--"] + script


def torchrun_cmd(nproc: int, script: List[str], override_args: Optional[List[str]] = None, port: Optional[int] = None) -> List[str]:
    """The reference's command shape (``launch/torch_distributed.py:15-38``): torchrun with
    ``--max_restarts 0`` running every rank under pid_client + wrap_rank."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes", "1", "--nproc_per_node", str(nproc), "--node_rank", "0",
           "--max_restarts", "0", "--master_addr", "127.0.0.1", "--master_port", str(port or c10d_port())]
    cmd += list(override_args or [])
    cmd += ["--module", "determined_b200.launch.wrap_rank", "RANK", "--"]
    return cmd + script


def pid_server_wrap(nproc: int, cmd: List[str], addr: str) -> List[str]:
    return [sys.executable, "-m", "determined_b200.exec.pid_server", "--on-fail", "SIGTERM", "--on-exit", "WAIT", "--grace-period", "10",
            addr, str(nproc), "--"] + cmd


def launch(script: List[str], override_args: Optional[List[str]] = None, prefer_native: bool = True,
           extra_env: Optional[Dict[str, str]] = None) -> int:
    nproc = num_local_ranks()
    env = dict(os.environ)
    env["USE_TORCH_DISTRIBUTED"] = "1"
    env.setdefault("DET_CHIEF_IP", "127.0.0.1")
    env.update(extra_env or {})
    if nproc == 1 and not override_args:
        # nothing to rendezvous: run the script directly (still sets the single-rank contract)
        env.update({"RANK": "0", "LOCAL_RANK": "0", "WORLD_SIZE": "1", "LOCAL_WORLD_SIZE": "1"})
        env.pop("USE_TORCH_DISTRIBUTED", None)
        p = subprocess.Popen(script, env=env)
        with util.forward_signals(p):
            return p.wait()
    if prefer_native and NATIVE_LAUNCHER.exists() and not override_args:
        cmd = native_cmd(nproc, script)
    else:
        cmd = torchrun_cmd(nproc, script, override_args)
    logger.info("launching %d ranks: %s", nproc, " ".join(cmd))
    p = subprocess.Popen(cmd, env=env)
    with util.forward_signals(p):
        return p.wait()


def split_args(argv: List[str]) -> (List[str], List[str]):
    """``[launcher overrides] -- script...``; with no ``--`` everything is the script."""
    if "--" in argv:
        i = argv.index("--")
        return argv[:i], argv[i + 1:]
    return [], argv
