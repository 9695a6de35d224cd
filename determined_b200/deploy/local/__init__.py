"""``det deploy local``: the reference (``harness/determined/deploy/local/cluster_utils.py``) starts
master + Postgres + agent containers with docker; here the "cluster" is one master process that owns
the box's GPUs, so cluster-up/master-up start it (detached, pid file under ``~/.local/share``).
``agent-up`` starts an extra agent daemon (``python -m determined_b200.agent``) that offers more
slots to a master — ``master-up --slots 0`` plus agents is the reference's master/agent split."""
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


def agent_up(master_url: str = "http://127.0.0.1:8080", agent_name: str = "determined-agent", slots: Optional[int] = None,
             slot_type: str = "auto", resource_pool: str = "default", detach: bool = True) -> Optional[int]:
    """Start an agent daemon (reference ``det deploy local agent-up MASTER_HOST``)."""
    STATE_DIR.mkdir(parents=True, exist_ok=True)
    cmd = [sys.executable, "-m", "determined_b200.agent", "--master-url", master_url, "--agent-id", agent_name, "--slot-type", slot_type,
           "--resource-pool", resource_pool, "--work-dir", str(STATE_DIR / f"agent-{agent_name}")]
    if slots is not None:
        cmd += ["--slots", str(slots)]
    if not detach:
        return subprocess.call(cmd)
    log = open(STATE_DIR / f"agent-{agent_name}.log", "ab")
    p = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, start_new_session=True)
    _pid_file(f"agent-{agent_name}").write_text(str(p.pid))
    return p.pid


def agent_down(agent_name: str = "determined-agent") -> bool:
    """SIGINT: the agent kills its tasks and exits (SIGTERM would leave them running for re-attachment)."""
    pf = _pid_file(f"agent-{agent_name}")
    if not pf.exists():
        return False
    try:
        os.kill(int(pf.read_text()), signal.SIGINT)
    except ProcessLookupError:
        pass
    pf.unlink(missing_ok=True)
    return True


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
