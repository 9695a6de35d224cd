"""``det deploy local``: the reference (``harness/determined/deploy/local/cluster_utils.py``) starts
master + Postgres + agent containers with docker; here the "cluster" is one master process that owns
the box's GPUs, so cluster-up/master-up start it (detached, pid file under ``~/.local/share``) and
agent-up/agent-down are accepted no-ops (the master *is* the agent)."""
import os
import pathlib
import signal
import subprocess
import sys
import time
from typing import Optional

from determined_b200.deploy import healthcheck

STATE_DIR = pathlib.Path(os.path.expanduser(os.environ.get("DET_LOCAL_STATE_DIR", "~/.local/share/determined_b200")))


def _pid_file(cluster_name: str) -> pathlib.Path:
    return STATE_DIR / f"{cluster_name}.pid"


def master_up(port: int = 8080, cluster_name: str = "determined", slots: Optional[int] = None, scheduler: Optional[str] = None,
              db_path: Optional[str] = None, detach: bool = True, wait: bool = True) -> Optional[int]:
    STATE_DIR.mkdir(parents=True, exist_ok=True)
    cmd = [sys.executable, "-m", "determined_b200.master", "--port", str(port), "--db", db_path or str(STATE_DIR / f"{cluster_name}.db")]
    if slots is not None:
        cmd += ["--slots", str(slots)]
    if scheduler:
        cmd += ["--scheduler", scheduler]
    if not detach:
        return subprocess.call(cmd)
    log = open(STATE_DIR / f"{cluster_name}.log", "ab")
    p = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, start_new_session=True)
    _pid_file(cluster_name).write_text(str(p.pid))
    if wait:
        healthcheck.wait_for_master(f"http://127.0.0.1:{port}")
    return p.pid


def master_down(cluster_name: str = "determined", delete_db: bool = False) -> bool:
    pf = _pid_file(cluster_name)
    if not pf.exists():
        return False
    pid = int(pf.read_text())
    try:
        os.kill(pid, signal.SIGTERM)
        for _ in range(50):
            os.kill(pid, 0)
            time.sleep(0.1)
        os.kill(pid, signal.SIGKILL)
    except ProcessLookupError:
        pass
    pf.unlink(missing_ok=True)
    if delete_db:
        (STATE_DIR / f"{cluster_name}.db").unlink(missing_ok=True)
    return True


def cluster_up(**kw) -> Optional[int]:
    return master_up(**kw)


def cluster_down(cluster_name: str = "determined", delete_db: bool = False) -> bool:
    return master_down(cluster_name, delete_db)


def agent_up(*_, **__) -> None:
    """The single-box master schedules its own GPUs; nothing to start."""


def agent_down(*_, **__) -> None:
    pass


def logs(cluster_name: str = "determined", follow: bool = False) -> None:
    path = STATE_DIR / f"{cluster_name}.log"
    if not path.exists():
        return
    with path.open() as f:
        sys.stdout.write(f.read())
        while follow:
            line = f.readline()
            if line:
                sys.stdout.write(line)
            else:
                time.sleep(0.3)
