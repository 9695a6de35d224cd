"""Slurm backend: a resource pool whose "agent" submits every task as a batch job.

Parity: reference ``master/internal/rm/dispatcherrm`` + ``hpc-ard-launcher-go`` (the generated client of HPE's launcher
service, which turns task specs into Slurm / PBS jobs).  That service is not open; here the master talks to Slurm's own
command line — ``sbatch --parsable`` to submit, ``squeue`` to watch, ``scancel`` to stop — which is all the dispatcher needs
from it.  Master and compute nodes must share ``work_dir`` (normal on a cluster with Slurm).

The backend registers itself with the :class:`~determined_b200.master.agents.AgentRegistry` as an in-process agent, so the
scheduler, allocations and commands treat it like any other agent: they queue ``start`` / ``kill`` actions, it answers with
``log`` / ``exit`` events.  Its ``slots`` are a quota (how many GPUs the master may hold in Slurm's queue at once), not
devices it owns; placement inside the partition is Slurm's business.

Pool config (``master.yaml``)::

    resource_pools:
      - pool_name: hpc
        provider: {type: slurm, slots: 64, partition: gpu, gres: "gpu:{slots}", sbatch_args: ["--time=04:00:00"], poll_seconds: 2}
"""
import io
import logging
import os
import shlex
import subprocess
import sys
import tarfile
import tempfile
import threading
import time
from typing import Any, Dict, List, Optional

from determined_b200.master import agents as agents_mod

logger = logging.getLogger("determined.master.slurm")


class _Job:
    def __init__(self, cid: str, job_id: str, run_dir: str, out_file: str = "slurm.out") -> None:
        self.cid, self.job_id, self.run_dir = cid, job_id, run_dir
        self.out_path = os.path.join(run_dir, out_file)
        self.exit_path = os.path.join(run_dir, "exit_code")
        self.offset = 0
        self.cancelled = False
        self.gone_since: Optional[float] = None


class SlurmBackend:
    def __init__(self, master: Any, pool_name: str, cfg: Dict[str, Any]) -> None:
        self.master, self.pool_name, self.cfg = master, pool_name, dict(cfg)
        self.id = str(cfg.get("agent_id") or f"{self.NAME.lower()}-{pool_name}")
        self.slots = int(cfg.get("slots", 8))
        self.poll_s = float(cfg.get("poll_seconds", 2.0))
        self.work_dir = str(cfg.get("work_dir") or os.path.join(master.work_dir, f"{self.NAME.lower()}-{pool_name}"))
        os.makedirs(self.work_dir, exist_ok=True)
        self.jobs: Dict[str, _Job] = {}
        self._stop = threading.Event()
        kind = self.NAME.lower()
        info = {"id": self.id, "resourcePool": pool_name, "version": kind, "addresses": [kind],
                "slots": [{"type": "DEVICE_TYPE_CUDA", "brand": f"{kind}:{cfg.get('partition') or cfg.get('queue') or 'default'}", "uuid": f"{self.id}-{i}"}
                          for i in range(self.slots)]}
        master.agents.register(info)
        self.agent = master.agents.get(self.id)
        self._thread = threading.Thread(target=self._run, daemon=True, name=f"{kind}-{pool_name}")
        self._thread.start()

    # -- Slurm command line ----------------------------------------------------------------------------
    def _cmd(self, name: str) -> str:
        return str(self.cfg.get(name) or name)

    NAME = "Slurm"
    OUT_FILE = "slurm.out"
    DONE_STATES = ("COMPLETED", "FAILED", "CANCELLED", "TIMEOUT", "NODE_FAIL", "OUT_OF_MEMORY")

    def directives(self, action: Dict[str, Any], run_dir: str, n: int) -> List[str]:
        """The ``#SBATCH`` header of a job that needs ``n`` slots."""
        lines = [f"#SBATCH --job-name=det-{action['container_id'][:24]}", f"#SBATCH --output={os.path.join(run_dir, self.OUT_FILE)}",
                 "#SBATCH --open-mode=append", "#SBATCH --nodes=1"]
        if self.cfg.get("partition"):
            lines.append(f"#SBATCH --partition={self.cfg['partition']}")
        if n and self.cfg.get("gres", "gpu:{slots}"):
            lines.append("#SBATCH --gres=" + str(self.cfg.get("gres", "gpu:{slots}")).format(slots=n))
        mine = (action.get("hpc") or {}).get("slurm") or {}          # the experiment's own ``slurm:`` section
        if mine.get("gpu_type") and n:
            lines = [l for l in lines if not l.startswith("#SBATCH --gres=")] + [f"#SBATCH --gres=gpu:{mine['gpu_type']}:{n}"]
        for a in list(self.cfg.get("sbatch_args") or []) + list(mine.get("sbatch_args") or []):
            lines.append(f"#SBATCH {a}")
        return lines

    def submit(self, path: str) -> str:
        r = subprocess.run([self._cmd("sbatch"), "--parsable", path], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60)
        if r.returncode != 0:
            raise RuntimeError(f"sbatch failed ({r.returncode}): {r.stderr.strip() or r.stdout.strip()}")
        return r.stdout.strip().split(";")[0]

    def cancel(self, job_id: str) -> None:
        subprocess.run([self._cmd("scancel"), job_id], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=60)

    def script(self, action: Dict[str, Any], run_dir: str) -> str:
        n = len(action.get("slot_ids") or [])
        lines = ["#!/bin/bash"] + self.directives(action, run_dir, n)
        env = {str(k): str(v) for k, v in (action.get("env") or {}).items()}
        model_dir, info_dir = os.path.join(run_dir, "model"), os.path.join(run_dir, "info")
        env.update({"DET_AGENT_ID": self.id, "DET_RUN_DIR": run_dir, "DET_INFO_DIR": info_dir, "DET_MODEL_DIR": model_dir,
                    "DET_SLOT_IDS": "[" + ", ".join(str(i) for i in range(n)) + "]"})      # devices are numbered from 0 inside the job
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        lines += [f"export {k}={shlex.quote(v)}" for k, v in sorted(env.items())]
        lines.append(f"export PYTHONPATH={shlex.quote(model_dir)}:{shlex.quote(root)}${{PYTHONPATH:+:$PYTHONPATH}}")
        python = str(self.cfg.get("python") or sys.executable)
        argv = [python if a == agents_mod.PYTHON_PLACEHOLDER else str(a) for a in action["argv"]]
        lines += [f"cd {shlex.quote(os.path.join(run_dir, action.get('workdir') or ''))}", " ".join(shlex.quote(a) for a in argv),
                  f"echo $? > {shlex.quote(os.path.join(run_dir, 'exit_code'))}"]
        return "\n".join(lines) + "\n"

    def _start(self, action: Dict[str, Any]) -> None:
        cid = str(action["container_id"])
        try:
            run_dir = tempfile.mkdtemp(prefix=f"job-{cid[:12]}-", dir=self.work_dir)
            os.makedirs(os.path.join(run_dir, "model"), exist_ok=True)
            os.makedirs(os.path.join(run_dir, "info"), exist_ok=True)
            if action.get("archive"):
                import base64

                with tarfile.open(fileobj=io.BytesIO(base64.b64decode(action["archive"])), mode="r:*") as tf:
                    tf.extractall(os.path.join(run_dir, "model"), filter="data")
            path = os.path.join(run_dir, "job.sh")
            with open(path, "w") as f:
                f.write(self.script(action, run_dir))
            job_id = self.submit(path)
            self.jobs[cid] = _Job(cid, job_id, run_dir, self.OUT_FILE)
            self._emit({"type": "log", "container_id": cid, "lines": [f"submitted {self.NAME} job {job_id}"]})
        except Exception as e:  # noqa: BLE001
            logger.exception("could not submit %s", cid)
            self._emit({"type": "start_failed", "container_id": cid, "error": str(e)})

    def _kill(self, cid: str) -> None:
        j = self.jobs.get(cid)
        if j is not None:
            j.cancelled = True
            self.cancel(j.job_id)

    def _in_queue(self, job_ids: List[str]) -> Dict[str, str]:
        """job id -> Slurm state for the jobs Slurm still knows (PENDING, RUNNING, COMPLETING...)."""
        if not job_ids:
            return {}
        r = subprocess.run([self._cmd("squeue"), "-h", "-j", ",".join(job_ids), "-o", "%i %T"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=60)
        out = {}
        for line in r.stdout.splitlines():
            parts = line.split()
            if len(parts) >= 2:
                out[parts[0]] = parts[1]
        return out

    # -- event pump ------------------------------------------------------------------------------------
    def _emit(self, ev: Dict[str, Any]) -> None:
        self.master.agents._event(self.agent, ev)

    def _ship_output(self, j: _Job) -> None:
        try:
            with open(j.out_path, "rb") as f:
                f.seek(j.offset)
                chunk = f.read()
        except OSError:
            return
        keep = chunk.rfind(b"\n") + 1           # only whole lines; the rest comes with the next poll
        if keep:
            j.offset += keep
            self._emit({"type": "log", "container_id": j.cid, "lines": chunk[:keep].decode(errors="replace").splitlines()})

    def _poll_jobs(self) -> None:
        if not self.jobs:
            return
        states = self._in_queue([j.job_id for j in self.jobs.values()])
        for cid, j in list(self.jobs.items()):
            self._ship_output(j)
            if j.job_id in states and states[j.job_id] not in self.DONE_STATES:
                j.gone_since = None
                continue
            # out of the queue: give the shared filesystem a moment to show the exit file, then settle
            j.gone_since = j.gone_since or time.time()
            code: Optional[int] = None
            try:
                code = int(open(j.exit_path).read().strip())
            except (OSError, ValueError):
                if time.time() - j.gone_since < float(self.cfg.get("exit_file_grace_seconds", 5.0)):
                    continue
            self._ship_output(j)
            if code is None:
                self._emit({"type": "log", "container_id": cid, "lines": [f"{self.NAME} job {j.job_id} left the queue as {states.get(j.job_id, 'gone')} without an exit code"]})
                code = 143 if j.cancelled else 1
            del self.jobs[cid]
            self._emit({"type": "exit", "container_id": cid, "exit_code": code})

    def _run(self) -> None:
        ag = self.agent
        while not self._stop.is_set():
            ag.last_seen = time.time()
            with ag.cv:
                if not ag.actions:
                    ag.cv.wait(timeout=self.poll_s)
                actions = list(ag.actions)
                ag.actions.clear()
            for act in actions:
                if act.get("type") == "start":
                    self._start(act)
                elif act.get("type") == "kill":
                    self._kill(str(act.get("container_id")))
            try:
                self._poll_jobs()
            except Exception:  # noqa: BLE001
                logger.exception("slurm poll failed")

    def stop(self) -> None:
        self._stop.set()
        for cid in list(self.jobs):
            self._kill(cid)
