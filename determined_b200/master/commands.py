"""Generic tasks: commands, shells, notebooks, tensorboards ("NTSC").

Parity: reference ``master/internal/command/*`` (command actors, ``CommandService``) and the
``api_command.go`` / ``api_shell.go`` / ``api_notebook.go`` / ``api_tensorboard.go`` handlers:
a task is a config (``entrypoint``, ``resources.slots``, ``environment.environment_variables``,
``description``) scheduled through the same resource pool as trials (non-preemptible), launched
with the task env contract, its output shipped to task logs, killable.

Single-box differences: the process runs directly on the host (no container), so shells have no
sshd to attach to and notebooks need ``jupyterlab`` importable; both are still scheduled, logged
and killable like commands.  Tensorboards serve the experiments' tfevents through
``determined_b200.exec.tensorboard``.
"""
import json
import logging
import os
import pathlib
import secrets
import shlex
import signal
import subprocess
import sys
import tempfile
import threading
import time
import uuid
from typing import Any, Dict, List, Optional

from determined_b200.master import scheduler

logger = logging.getLogger("determined.master.commands")


def _iso(t: Optional[float]) -> Optional[str]:
    if t is None:
        return None
    return time.strftime("%Y-%m-%dT%H:%M:%S", time.gmtime(t)) + (".%06dZ" % int((t % 1) * 1e6))

TYPES = {"command": "COMMAND", "shell": "SHELL", "notebook": "NOTEBOOK", "tensorboard": "TENSORBOARD"}


def default_entrypoint(kind: str, body: Dict[str, Any]) -> List[str]:
    if kind == "NOTEBOOK":
        return [sys.executable, "-m", "jupyterlab", "--no-browser", "--ip=127.0.0.1", "--ServerApp.token="]
    if kind == "SHELL":
        # no sshd on a single box: keep a login shell alive so the slot reservation, logs and kill path behave
        return ["/bin/bash", "-lc", "echo shell task ready on $(hostname); exec sleep infinity"]
    if kind == "TENSORBOARD":
        ids = [str(i) for i in body.get("experimentIds") or []]
        tids = [str(i) for i in body.get("trialIds") or []]
        return [sys.executable, "-m", "determined_b200.exec.tensorboard", "--experiments", ",".join(ids), "--trials", ",".join(tids)]
    raise ValueError("a command needs an entrypoint")


def parse_duration(v: Any) -> Optional[float]:
    """``90`` / ``"90s"`` / ``"30m"`` / ``"1h30m"`` -> seconds; None / empty -> None (Go duration strings of the reference configs)."""
    if v in (None, "", 0):
        return None
    if isinstance(v, (int, float)):
        return float(v)
    import re

    parts = re.findall(r"(\d+(?:\.\d+)?)(ms|s|m|h|d)?", str(v).strip())
    if not parts or "".join(a + b for a, b in parts) != str(v).strip():
        raise ValueError(f"cannot parse duration {v!r}")
    unit = {"ms": 1e-3, "s": 1.0, "m": 60.0, "h": 3600.0, "d": 86400.0, "": 1.0}
    return sum(float(a) * unit[b] for a, b in parts)


class CommandTask:
    def __init__(self, mgr: "CommandManager", kind: str, config: Dict[str, Any], owner_id: int, owner_name: str, body: Dict[str, Any]) -> None:
        self.mgr, self.kind, self.config = mgr, kind, config
        self.id = str(uuid.uuid4())
        self.owner_id, self.owner_name = owner_id, owner_name
        self.state = "QUEUED"
        self.start_time = time.time()
        self.end_time: Optional[float] = None
        self.exit_code: Optional[int] = None
        self.slots = int((config.get("resources") or {}).get("slots", 0) or 0)
        self.description = config.get("description") or f"{kind.lower()} ({owner_name})"
        self.proc: Optional[subprocess.Popen] = None
        self.slot_ids: List[int] = []
        self.token = secrets.token_hex(16)
        self.killed = False
        # the context directory shipped with the task (``det cmd run --context DIR`` / ``--include``): v1 file entries
        self.files: List[Dict[str, Any]] = list(body.get("files") or [])
        self.context: Optional[bytes] = body.get("_context")
        self.paused = False
        ep = config.get("entrypoint")
        if not ep:
            ep = default_entrypoint(kind, body)
        self.argv: List[str] = ep if isinstance(ep, list) else ["/bin/bash", "-lc", str(ep)]
        self.run_dir: Optional[str] = None
        # tasks that serve HTTP get a port from the master's registry and are reachable through /proxy/<task id>/
        self.service_port: Optional[int] = None
        self.service_host = "127.0.0.1"
        # reference notebook / tensorboard configs: ``idle_timeout`` kills a service nobody has talked to for that long
        self.idle_timeout: Optional[float] = parse_duration(config.get("idle_timeout"))
        if self.idle_timeout is None:                # the master's notebook_timeout / tensorboard_timeout, when configured
            dflt = (getattr(mgr, "default_idle_timeout", None) or {}).get(kind)
            self.idle_timeout = float(dflt) if dflt else None
        self.last_activity = time.time()

    @property
    def pool_key(self) -> str:
        return f"cmd-{self.id}"

    def to_json(self) -> Dict[str, Any]:
        return {"id": self.id, "description": self.description, "state": "STATE_" + self.state, "startTime": _iso(self.start_time), "endTime": _iso(self.end_time),
                "username": self.owner_name, "userId": self.owner_id, "resourcePool": (self.config.get("resources") or {}).get("resource_pool") or "default", "exitStatus": self.exit_code,
                "jobId": self.id, "type": self.kind, "slots": self.slots, "slotIds": self.slot_ids, "config": self.config,
                "serviceAddress": f"/proxy/{self.id}/" if self.service_port else None}

    def start(self, slot_ids: List[int], agent_id: Optional[str] = None) -> None:
        m = self.mgr.master
        self.slot_ids = slot_ids
        self.agent_id = agent_id or m.agent_id
        task_env: Dict[str, str] = {}
        if self.kind in ("TENSORBOARD", "NOTEBOOK") or (self.config.get("proxy_ports") or self.config.get("service")):
            self.service_port = m.ports.acquire()
            ra = m.agents.get(self.agent_id)
            if ra is not None and ra.addresses:
                self.service_host = ra.addresses[0]
            task_env["DET_SERVICE_PORT"] = str(self.service_port)
            if self.kind == "TENSORBOARD" and "--port" not in self.argv:
                self.argv = self.argv + ["--port", str(self.service_port)]
        for kv in ((self.config.get("environment") or {}).get("environment_variables") or []):
            if isinstance(kv, str) and "=" in kv:
                k, v = kv.split("=", 1)
                task_env[k] = v
        task_env.update({**m.master_env(), "DET_CLUSTER_ID": m.cluster_id,
                         "DET_AGENT_ID": self.agent_id, "DET_SLOT_IDS": json.dumps(slot_ids), "DET_TASK_ID": self.id,
                         "DET_ALLOCATION_ID": self.id + ".1", "DET_SESSION_TOKEN": self.token, "DET_TASK_TYPE": self.kind,
                         "DET_USER": self.owner_name, "PYTHONUNBUFFERED": "1"})
        if m.agents.is_remote(agent_id):
            from determined_b200.master import agents as agents_mod

            m.db.exec("INSERT INTO sessions(token,user_id,expiry,task_id) VALUES(?,?,?,?)", (self.token, self.owner_id, time.time() + 14 * 86400, self.id))
            self.state, self.last_activity = "RUNNING", time.time()
            m.db.exec("UPDATE tasks SET state=? WHERE id=?", (self.state, self.id))
            argv = [agents_mod.PYTHON_PLACEHOLDER if a == sys.executable else a for a in self.argv]
            if self.files and not self.context:
                from determined_b200.common import context as det_context

                self.context = det_context.v1_context_to_tarball(self.files)
            m.agents.start(self.agent_id, self.id, self, argv, task_env, slot_ids, archive=self.context, workdir="model",
                           container=agents_mod.container_spec(self.config))
            return
        self.run_dir = tempfile.mkdtemp(prefix=f"detb200-{self.kind.lower()}-{self.id[:8]}-", dir=m.work_dir)
        if self.files:
            from determined_b200.common import context as det_context

            det_context.write_v1_context(self.files, pathlib.Path(self.run_dir))
        env = dict(os.environ)
        env.update(task_env)
        env["DET_RUN_DIR"] = self.run_dir
        if slot_ids:
            env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in slot_ids)
        elif m.num_gpus:
            env["CUDA_VISIBLE_DEVICES"] = ""
        root = str(pathlib.Path(__file__).resolve().parents[2])
        env["PYTHONPATH"] = os.pathsep.join([root] + ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
        m.db.exec("INSERT INTO sessions(token,user_id,expiry,task_id) VALUES(?,?,?,?)", (self.token, self.owner_id, time.time() + 14 * 86400, self.id))
        self.state, self.last_activity = "RUNNING", time.time()
        m.db.exec("UPDATE tasks SET state=? WHERE id=?", (self.state, self.id))
        logger.info("starting %s %s: %s", self.kind, self.id, " ".join(shlex.quote(a) for a in self.argv))
        self.proc = subprocess.Popen(self.argv, cwd=self.run_dir, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, start_new_session=True,
                                     text=True, bufsize=1, errors="replace")
        threading.Thread(target=self._pump, daemon=True, name=f"cmdlogs-{self.id[:8]}").start()

    def _pump(self) -> None:
        assert self.proc is not None and self.proc.stdout is not None
        m = self.mgr.master
        batch: List[Any] = []
        lock = threading.Lock()
        done = threading.Event()

        def flush() -> None:
            with lock:
                rows, batch[:] = list(batch), []
            if rows:
                m.db.add_logs(rows)

        def ticker() -> None:          # a line must not wait for the next one to be written: flush on a clock as well
            while not done.wait(0.3):
                flush()

        threading.Thread(target=ticker, daemon=True, name=f"cmdflush-{self.id[:8]}").start()
        for line in self.proc.stdout:
            with lock:
                batch.append((self.id, self.id + ".1", time.time(), "INFO", None, "stdout", "task", line if line.endswith("\n") else line + "\n"))
                full = len(batch) >= 100
            if full:
                flush()
        done.set()
        flush()
        code = self.proc.wait()
        self.finish(code)

    # sink interface of master/agents.py (task running under a remote agent)
    def on_logs(self, lines: List[str]) -> None:
        now = time.time()
        self.mgr.master.db.add_logs([(self.id, self.id + ".1", now, "INFO", None, "stdout", "task", ln if ln.endswith("\n") else ln + "\n") for ln in lines])

    def on_exit(self, code: Optional[int]) -> None:
        if self.state != "TERMINATED":
            self.finish(code)

    def finish(self, code: Optional[int]) -> None:
        m = self.mgr.master
        self.exit_code, self.end_time, self.state = code, time.time(), "TERMINATED"
        m.db.exec("UPDATE tasks SET state=?, end_time=? WHERE id=?", (self.state, self.end_time, self.id))
        m.db.exec("DELETE FROM sessions WHERE token=?", (self.token,))
        m.ports.release(self.service_port)
        with m.lock:
            m.release_request(self.pool_key)
        m.kick()

    def kill(self) -> None:
        self.killed = True
        m = self.mgr.master
        if m.agents.is_remote(getattr(self, "agent_id", None)) and self.state == "RUNNING":
            m.agents.kill(self.agent_id, self.id)
            return
        if self.proc is not None and self.proc.poll() is None:
            try:
                os.killpg(self.proc.pid, signal.SIGTERM)
            except ProcessLookupError:
                pass

            def hard(p: subprocess.Popen = self.proc) -> None:
                time.sleep(5)
                if p.poll() is None:
                    try:
                        os.killpg(p.pid, signal.SIGKILL)
                    except ProcessLookupError:
                        pass

            threading.Thread(target=hard, daemon=True).start()
        elif self.state == "QUEUED":
            self.finish(None)


class CommandManager:
    def __init__(self, master: Any) -> None:
        self.master = master
        self.tasks: Dict[str, CommandTask] = {}

    def launch(self, kind: str, config: Dict[str, Any], owner_id: int, owner_name: str, body: Dict[str, Any]) -> CommandTask:
        t = CommandTask(self, TYPES[kind], config, owner_id, owner_name, body)
        m = self.master
        res = config.get("resources") or {}
        pool = m.pool_for(res)                   # unknown pool: refused before anything is recorded
        with m.lock:
            self.tasks[t.id] = t
            m.db.exec("INSERT INTO tasks(id,type,start_time,job_id,config,state,owner_id) VALUES(?,?,?,?,?,?,?)",
                      (t.id, t.kind, t.start_time, t.id, json.dumps(config), t.state, owner_id))
            pool.allocate(scheduler.AllocateRequest(t.pool_key, t.id, t.slots, priority=res.get("priority"), weight=float(res.get("weight", 1) or 1),
                                                      preemptible=False, name=t.description, submit_time=time.time()))
        m.kick()
        return t

    def on_assigned(self, pool_key: str, slot_ids: List[int], agent_id: Optional[str] = None) -> bool:
        """Called by the scheduler loop for requests whose id starts with ``cmd-``."""
        t = self.tasks.get(pool_key[4:])
        if t is None or t.state != "QUEUED":
            return False
        if not self.master.launch_enabled:
            t.slot_ids, t.state = slot_ids, "RUNNING"
            return True
        try:
            t.start(slot_ids, agent_id)
        except Exception as e:
            logger.exception("failed to start %s %s", t.kind, t.id)
            self.master.db.add_logs([(t.id, t.id + ".1", time.time(), "ERROR", None, "stdout", "master", f"failed to start: {e}\n")])
            t.exit_code, t.end_time, t.state = 127, time.time(), "TERMINATED"
            self.master.db.exec("UPDATE tasks SET state=?, end_time=? WHERE id=?", (t.state, t.end_time, t.id))
            return False
        return True

    def get(self, task_id: str, kind: Optional[str] = None) -> CommandTask:
        t = self.tasks.get(task_id)
        if t is None or (kind is not None and t.kind != TYPES[kind]):
            raise KeyError(task_id)
        return t

    def list(self, kind: str) -> List[CommandTask]:
        return sorted((t for t in self.tasks.values() if t.kind == TYPES[kind]), key=lambda t: t.start_time)

    def reap_idle(self, now: Optional[float] = None) -> List[str]:
        """Kill running tasks whose ``idle_timeout`` has passed without proxy traffic; called from the scheduler loop."""
        now = now if now is not None else time.time()
        killed = []
        for t in list(self.tasks.values()):
            if t.state == "RUNNING" and t.idle_timeout and now - t.last_activity > t.idle_timeout:
                logger.info("%s %s idle for %.0fs (timeout %.0fs): killing", t.kind, t.id, now - t.last_activity, t.idle_timeout)
                self.master.db.add_logs([(t.id, t.id + ".1", now, "INFO", None, "stdout", "master", f"idle for more than {t.idle_timeout:.0f}s: terminating\n")])
                t.kill()
                killed.append(t.id)
        return killed

    def stop_all(self) -> None:
        for t in list(self.tasks.values()):
            if t.state != "TERMINATED":
                t.kill()
