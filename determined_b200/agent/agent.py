"""The agent daemon: offers this host's slots to a master and runs the tasks it is handed.

Parity: reference ``agent/internal/agent.go`` (connect, announce devices, reconnect with the list
of surviving containers), ``agent/internal/containers/manager.go`` + ``container/container.go``
(start / monitor / kill, log forwarding, state persisted for re-attachment after a restart) and
``agent/internal/detect`` (slot detection).  What is different: every task is a process group that
``ProcessRuntime`` supervises; with ``--container-runtime docker|podman|apptainer|singularity`` that
process is the runtime's foreground client (``agent/runtimes.py`` builds its command line), without
it the task runs directly on the host.  The link to the master is a REST long-poll instead of a
websocket (``master/agents.py``).

``python -m determined_b200.agent --master-url http://127.0.0.1:8080 --agent-id gpu-box-2``
"""
import base64
import io
import json
import logging
import os
import pathlib
import shutil
import signal
import socket
import subprocess
import sys
import tarfile
import tempfile
import threading
import time
from typing import Any, Dict, List, Optional

import requests

from determined_b200 import __version__, gpu
from determined_b200.agent import runtimes

logger = logging.getLogger("determined.agent")

PYTHON_PLACEHOLDER = "{python}"
KILL_GRACE_S = 10.0


def detect_slots(limit: Optional[int] = None, slot_type: str = "auto") -> List[Dict[str, Any]]:
    """One entry per schedulable device: the visible NVIDIA GPUs (``nvidia-smi --query-gpu=index,name,uuid``, reference
    ``detect/nvidia.go:19-117``), else CPU slots (``slot_type: cpu`` gives ``limit`` of them; reference ``detect/cpu.go``)."""
    gpus, _ = gpu.get_gpus() if slot_type in ("auto", "cuda", "gpu") else ([], "")
    if gpus:
        out = [{"type": "DEVICE_TYPE_CUDA", "brand": g.name, "uuid": g.uuid, "index": g.id} for g in gpus]
    else:
        out = [{"type": "DEVICE_TYPE_CPU", "brand": "cpu", "uuid": f"cpu-{i}", "index": i} for i in range(limit if limit is not None else 1)]
    return out[:limit] if limit is not None else out


class _Task:
    def __init__(self, cid: str, proc: Optional[subprocess.Popen], run_dir: Optional[str], pid: int) -> None:
        self.cid, self.proc, self.run_dir, self.pid = cid, proc, run_dir, pid
        self.kill_sent: Optional[float] = None
        self.container: Optional[str] = None     # the runtime's name for it, when a container runtime is in use


class ProcessRuntime:
    """start / kill / wait for process groups; survives agent restarts through a small state file."""

    def __init__(self, work_dir: str, emit: Any, driver: Optional[runtimes.Driver] = None, default_image: str = "",
                 container_args: Optional[List[str]] = None, slots: Optional[List[Dict[str, Any]]] = None) -> None:
        self.work_dir = work_dir
        self.emit = emit                         # emit(event dict): queued for the next poll
        self.driver = driver                     # None: tasks run directly on the host
        self.default_image = default_image
        self.container_args = list(container_args or [])
        self.slots = list(slots or [])
        self.tasks: Dict[str, _Task] = {}
        self.lock = threading.Lock()
        self.state_path = os.path.join(work_dir, "tasks.json")

    # -- persistence (re-attachment) -----------------------------------------------------------------
    def _save(self) -> None:
        with self.lock:
            data = {cid: {"pid": t.pid, "run_dir": t.run_dir, "container": t.container} for cid, t in self.tasks.items()}
        tmp = self.state_path + ".tmp"
        with open(tmp, "w") as f:
            json.dump(data, f)
        os.replace(tmp, self.state_path)

    def reattach(self) -> List[str]:
        """Tasks of a previous agent life that are still alive.  Their pipes are gone with the old agent, so their
        remaining output is lost; exits are detected by polling the process group leader."""
        try:
            with open(self.state_path) as f:
                old = json.load(f)
        except (OSError, ValueError):
            return []
        alive = []
        for cid, rec in old.items():
            pid = int(rec.get("pid", 0))
            if pid > 0 and _alive(pid):
                t = _Task(cid, None, rec.get("run_dir"), pid)
                t.container = rec.get("container")
                self.tasks[cid] = t
                threading.Thread(target=self._watch_pid, args=(t,), daemon=True, name=f"watch-{cid[:8]}").start()
                alive.append(cid)
        self._save()
        return alive

    def _watch_pid(self, t: _Task) -> None:
        while _alive(t.pid):
            time.sleep(0.5)
        self._done(t, None)

    # -- verbs ---------------------------------------------------------------------------------------
    def start(self, action: Dict[str, Any], agent_id: str) -> None:
        cid = str(action["container_id"])
        try:
            run_dir = tempfile.mkdtemp(prefix=f"task-{cid[:12]}-", dir=self.work_dir)
            model_dir = os.path.join(run_dir, "model")
            info_dir = os.path.join(run_dir, "info")
            os.makedirs(model_dir, exist_ok=True)
            os.makedirs(info_dir, exist_ok=True)
            if action.get("archive"):
                with tarfile.open(fileobj=io.BytesIO(base64.b64decode(action["archive"])), mode="r:*") as tf:
                    tf.extractall(model_dir, filter="data")
            task_env = {str(k): str(v) for k, v in (action.get("env") or {}).items()}
            if task_env.get("DET_MASTER_CERT_FILE") and not os.path.exists(task_env["DET_MASTER_CERT_FILE"]) and getattr(self, "master_cert_file", None):
                task_env["DET_MASTER_CERT_FILE"] = self.master_cert_file   # the master named ITS file; tasks here use this host's copy
            slot_ids = [int(s) for s in action.get("slot_ids") or []]
            task_env.update({"DET_AGENT_ID": agent_id, "DET_RUN_DIR": run_dir, "DET_INFO_DIR": info_dir, "DET_MODEL_DIR": model_dir,
                             "DET_SLOT_IDS": json.dumps(slot_ids)})
            task_env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in slot_ids)
            root = str(pathlib.Path(__file__).resolve().parents[2])
            cwd = os.path.join(run_dir, action.get("workdir") or "")
            name: Optional[str] = None
            if self.driver is None:
                env = dict(os.environ)
                env.pop("DET_LATEST_CHECKPOINT", None)
                env.update(task_env)
                env["PYTHONPATH"] = os.pathsep.join([model_dir, root] + ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
                argv = [sys.executable if a == PYTHON_PLACEHOLDER else str(a) for a in action["argv"]]
            else:
                argv, env, name = self._containerize(action, cid, task_env, cwd, run_dir, model_dir, root, slot_ids)
            proc = subprocess.Popen(argv, cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, start_new_session=True,
                                    text=True, bufsize=1, errors="replace")
        except Exception as e:  # noqa: BLE001
            logger.exception("could not start task %s", cid)
            self.emit({"type": "start_failed", "container_id": cid, "error": repr(e)})
            return
        t = _Task(cid, proc, run_dir, proc.pid)
        t.container = name
        with self.lock:
            self.tasks[cid] = t
        self._save()
        logger.info("task %s started: pid %d, slots %s", cid, proc.pid, slot_ids)
        threading.Thread(target=self._pump, args=(t,), daemon=True, name=f"task-{cid[:8]}").start()

    def _containerize(self, action: Dict[str, Any], cid: str, task_env: Dict[str, str], cwd: str, run_dir: str, model_dir: str,
                      root: str, slot_ids: List[int]) -> Any:
        """The host command line (and the client's environment) that runs this task inside a container."""
        drv = self.driver
        assert drv is not None
        c = action.get("container") or {}
        task_env["PYTHONPATH"] = os.pathsep.join([model_dir, root])
        spec = runtimes.build_spec(action, cid=cid, argv=[str(c.get("python") or "python3") if a == PYTHON_PLACEHOLDER else str(a) for a in action["argv"]],
                                   env=task_env, cwd=cwd, run_dir=run_dir, repo_root=root, slots=self.slots, default_image=self.default_image,
                                   extra_args=self.container_args)
        if not spec.image:
            raise RuntimeError(f"no image for the {drv.name} runtime: set environment.image or start the agent with --default-image")
        if isinstance(drv, runtimes._OCI) and spec.gpu_ids:
            spec.env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(len(spec.gpu_ids)))   # the runtime renumbers what it passes in
        auth = c.get("registry_auth") or None
        runtimes.run_prepare(drv.pull(spec.image, bool(c.get("force_pull_image")), auth), password=(auth or {}).get("password"))
        argv, extra = drv.run(spec)
        env = dict(os.environ)
        env.update(extra)
        if auth and isinstance(drv, runtimes.Apptainer):
            env[f"{drv.name.upper()}_DOCKER_USERNAME"], env[f"{drv.name.upper()}_DOCKER_PASSWORD"] = str(auth.get("username", "")), str(auth.get("password", ""))
        self.emit({"type": "log", "container_id": cid, "lines": [f"{drv.name}: {spec.image} as {spec.name}"]})
        return argv, env, spec.name

    def _pump(self, t: _Task) -> None:
        """Forward the task's output in batches: at 100 lines, or 0.3 s after the first unsent line (a ticker flushes, so a
        lone line is not held back until the next one arrives)."""
        assert t.proc is not None and t.proc.stdout is not None
        batch: List[str] = []
        lock = threading.Lock()
        over = threading.Event()

        def flush() -> None:
            with lock:
                lines, batch[:] = list(batch), []
            if lines:
                self.emit({"type": "log", "container_id": t.cid, "lines": lines})

        def tick() -> None:
            while not over.wait(0.3):
                flush()

        threading.Thread(target=tick, daemon=True, name=f"flush-{t.cid[:8]}").start()
        for line in t.proc.stdout:
            with lock:
                batch.append(line.rstrip("\n"))
                full = len(batch) >= 100
            if full:
                flush()
        over.set()
        flush()
        self._done(t, t.proc.wait())

    def _done(self, t: _Task, code: Optional[int]) -> None:
        with self.lock:
            self.tasks.pop(t.cid, None)
        self._save()
        if t.run_dir and not os.environ.get("DET_KEEP_RUN_DIRS"):
            shutil.rmtree(t.run_dir, ignore_errors=True)
        logger.info("task %s exited with %s", t.cid, code)
        self.emit({"type": "exit", "container_id": t.cid, "exit_code": code})

    def kill(self, cid: str, sig: str = "TERM") -> None:
        with self.lock:
            t = self.tasks.get(cid)
        if t is None:
            return
        self._signal(t, "KILL" if sig == "KILL" else "TERM")
        if sig != "KILL" and t.kill_sent is None:
            t.kill_sent = time.time()

            def hard() -> None:
                time.sleep(KILL_GRACE_S)
                if _alive(t.pid):
                    self._signal(t, "KILL")
            threading.Thread(target=hard, daemon=True).start()

    def _signal(self, t: _Task, sig: str) -> None:
        """A container gets the signal through its runtime (killing a ``docker run`` client with KILL would orphan the
        container); the client's process group gets it too, which is the whole story for host tasks and apptainer."""
        cmd = self.driver.signal(t.container, sig) if self.driver is not None and t.container else None
        if cmd:
            try:
                subprocess.run(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=30)
            except (OSError, subprocess.SubprocessError):
                logger.warning("%s failed", runtimes.describe(cmd))
            if sig != "KILL":
                return                           # the client relays the container's exit; KILL below is the backstop
        _signal_group(t.pid, signal.SIGKILL if sig == "KILL" else signal.SIGTERM)

    def kill_all(self) -> None:
        for cid in list(self.tasks):
            self.kill(cid, "KILL")

    def running(self) -> List[str]:
        with self.lock:
            return sorted(self.tasks)


def _alive(pid: int) -> bool:
    try:
        os.kill(pid, 0)
    except ProcessLookupError:
        return False
    except PermissionError:
        return True
    # a zombie child of ours still answers signal 0: reap it if it is ours
    try:
        done, _ = os.waitpid(pid, os.WNOHANG)
        return done == 0
    except ChildProcessError:
        return True


def _signal_group(pid: int, sig: int) -> None:
    try:
        os.killpg(pid, sig)
    except (ProcessLookupError, PermissionError):
        pass


class Agent:
    def __init__(self, master_url: str, agent_id: Optional[str] = None, slots: Optional[int] = None, slot_type: str = "auto",
                 resource_pool: str = "default", work_dir: Optional[str] = None, token: Optional[str] = None,
                 master_cert_file: Optional[str] = None, master_cert_name: Optional[str] = None, container_runtime: str = "process",
                 default_image: str = "", container_exe: Optional[str] = None, container_args: Optional[List[str]] = None) -> None:
        self.master_url = master_url.rstrip("/")
        # an https master: the certificate to trust (a PEM file, or "noverify"), and the name it carries if not the URL's
        self.master_cert_file = master_cert_file or os.environ.get("DET_MASTER_CERT_FILE")
        self.master_cert_name = master_cert_name or os.environ.get("DET_MASTER_CERT_NAME")
        self.id = agent_id or socket.gethostname()
        self.slots = detect_slots(slots, slot_type)
        self.resource_pool = resource_pool
        self.token = token or os.environ.get("DET_AGENT_TOKEN")
        self.work_dir = work_dir or tempfile.mkdtemp(prefix=f"detb200-agent-{self.id}-")
        os.makedirs(self.work_dir, exist_ok=True)
        self._events: List[Dict[str, Any]] = []
        self._ev_lock = threading.Lock()
        self._ev_flag = threading.Event()
        self._stop = threading.Event()
        self.runtime = ProcessRuntime(self.work_dir, self._emit, runtimes.get(container_runtime, container_exe), default_image, container_args, self.slots)
        self.http = requests.Session()
        self._verify_kw: Dict[str, Any] = {}
        if self.master_url.startswith("https://"):
            from determined_b200.common.api import certs

            cert = certs.default_load(self.master_url, self.master_cert_file, self.master_cert_name)
            if cert.bundle is not None:
                self._verify_kw = {"verify": cert.bundle}
            if cert.name:
                self.http.mount("https://", certs.host_name_adapter(cert.name))
        self.runtime.master_cert_file = self.master_cert_file
        self.poll_timeout = 10.0
        self.registered = False

    def _emit(self, ev: Dict[str, Any]) -> None:
        with self._ev_lock:
            self._events.append(ev)
        self._ev_flag.set()

    def _post(self, path: str, body: Dict[str, Any], timeout: float) -> Dict[str, Any]:
        if self.token:
            body = {**body, "token": self.token}
        r = self.http.post(self.master_url + path, json=body, timeout=timeout, **self._verify_kw)
        if r.status_code >= 400:
            raise RuntimeError(f"{path}: HTTP {r.status_code}: {r.text[:200]}")
        return r.json()

    def register(self) -> Dict[str, Any]:
        body = {"id": self.id, "slots": self.slots, "resourcePool": self.resource_pool,
                "version": __version__ + (f"+{self.runtime.driver.name}" if self.runtime.driver else ""),
                "addresses": [_my_address()], "containers": self.runtime.running()}
        resp = self._post("/api/v1/agents/register", body, 15)
        self.poll_timeout = float(resp.get("pollTimeoutSeconds", 10))
        self.registered = True
        logger.info("registered with %s as %s (%d slots, re-attached: %s)", self.master_url, self.id, len(self.slots), resp.get("reattached"))
        return resp

    def _act(self, resp: Dict[str, Any]) -> None:
        for act in resp.get("actions") or []:
            if act.get("type") == "start":
                self.runtime.start(act, self.id)
            elif act.get("type") == "kill":
                self.runtime.kill(str(act.get("container_id")), str(act.get("signal") or "TERM"))

    def step(self, ship: bool = True) -> None:
        """One poll: ship pending events (unless the shipper thread owns them), execute the actions that come back."""
        events: List[Dict[str, Any]] = []
        if ship:
            with self._ev_lock:
                events, self._events = self._events, []
        try:
            resp = self._post(f"/api/v1/agents/{self.id}/poll", {"events": events, "timeoutSeconds": 0 if events else self.poll_timeout},
                              self.poll_timeout + 15)
        except Exception:
            with self._ev_lock:                      # keep them for the next attempt
                self._events = events + self._events
            raise
        if resp.get("reregister"):
            with self._ev_lock:
                self._events = events + self._events
            self.register()
            return
        self._act(resp)

    def _ship_loop(self) -> None:
        """Events (task output, exits) go out as they happen, beside the long poll that waits for actions: a task's exit
        reaches the master in one round trip instead of waiting for the poll in flight to time out."""
        while not self._stop.is_set():
            self._ev_flag.wait(timeout=1.0)
            self._ev_flag.clear()
            if not self.registered:
                continue
            with self._ev_lock:
                events, self._events = self._events, []
            if not events:
                continue
            try:
                resp = self._post(f"/api/v1/agents/{self.id}/poll", {"events": events, "timeoutSeconds": 0}, 30)
                if resp.get("reregister"):
                    raise RuntimeError("the master does not know this agent")
                self._act(resp)
            except Exception as e:  # noqa: BLE001
                with self._ev_lock:
                    self._events = events + self._events
                logger.debug("events not delivered (%s); will retry", e)
                self._stop.wait(0.5)
                self._ev_flag.set()

    def run(self) -> None:
        self.runtime.reattach()
        threading.Thread(target=self._ship_loop, daemon=True, name="agent-events").start()
        backoff = 0.5
        while not self._stop.is_set():
            try:
                if not self.registered:
                    self.register()
                self.step(ship=False)
                backoff = 0.5
            except Exception as e:  # noqa: BLE001
                if self._stop.is_set():
                    break
                logger.warning("master unreachable (%s); retrying in %.1fs", e, backoff)
                self.registered = False
                self._stop.wait(backoff)
                backoff = min(backoff * 2, 10.0)

    def stop(self, kill_tasks: bool = True) -> None:
        self._stop.set()
        if kill_tasks:
            self.runtime.kill_all()


def _my_address() -> str:
    try:
        return socket.gethostbyname(socket.gethostname())
    except OSError:
        return "127.0.0.1"
