"""The single-box master: experiment + trial state machines, allocation service, checkpoint GC.

Parity (behavioural), reduced from the Go master to one process on one 8xB200 node
(BASELINE.json north star):
  * experiment actor — reference ``master/internal/experiment.go``: start/initial ops ``:245-319``,
    ``processOperations`` ``:763-887``, ``TrialCompleteOperation`` ``:321-355``, pause / activate /
    cancel / kill ``:597-658``, snapshot/restore ``:980-998`` and ``core.go:821-873``;
  * trial state machine — ``master/internal/trial.go``: ``maybeAllocateTask`` ``:365-477``,
    ``handleAllocationExit`` (``max_restarts``, reschedule from latest checkpoint) ``:534-673``;
  * allocation lifecycle — ``master/internal/task/allocation.go`` (start, preempt ``:953``, kill
    ``:968``, exit ``:1146``), rendezvous ``task/rendezvous.go``, preemption long-poll + ack
    ``task/preemptible``, master-mediated all-gather ``task/allgather``;
  * trial env-var contract — ``master/pkg/tasks/task_trial.go:113-125``;
  * checkpoint GC — ``master/internal/checkpoint_gc.go`` + ``harness/determined/exec/gc_checkpoints.py``.

No containers and no agent processes: an allocation is a local process group started by
:mod:`determined_b200.launch` with ``CUDA_VISIBLE_DEVICES`` set to the slots the scheduler picked.
"""
import base64
import io
import json
import logging
import os
import pathlib
import secrets
import shutil
import signal
import subprocess
import sys
import tarfile
import tempfile
import threading
import time
import uuid
from typing import Any, Callable, Dict, List, Optional, Tuple

from determined_b200 import constants, gpu, util
from determined_b200.master import db as dbmod
from determined_b200.master import agents as agents_mod
from determined_b200.master import expconf, scheduler
from determined_b200.master import searcher as S

logger = logging.getLogger("determined.master")

ACTIVE, PAUSED = "ACTIVE", "PAUSED"
STOPPING_COMPLETED, STOPPING_CANCELED, STOPPING_KILLED, STOPPING_ERROR = (
    "STOPPING_COMPLETED", "STOPPING_CANCELED", "STOPPING_KILLED", "STOPPING_ERROR")
COMPLETED, CANCELED, ERROR = "COMPLETED", "CANCELED", "ERROR"
TERMINAL = (COMPLETED, CANCELED, ERROR)
STOPPING_TO_TERMINAL = {STOPPING_COMPLETED: COMPLETED, STOPPING_CANCELED: CANCELED, STOPPING_KILLED: CANCELED,
                        STOPPING_ERROR: ERROR}
PREEMPT_KILL_DEADLINE_S = 3600.0       # reference preemptible.go:22
RENDEZVOUS_TIMEOUT_S = 600.0


class NotFound(KeyError):
    pass


class Conflict(ValueError):
    pass


class Allocation:
    """One run of a trial's process group on a set of slots."""

    def __init__(self, master: "Master", trial: "Trial", alloc_id: str, slots: int) -> None:
        self.master, self.trial, self.id, self.slots = master, trial, alloc_id, slots
        self.slot_ids: List[int] = []
        self.proc: Optional[subprocess.Popen] = None
        self.state = "PENDING"
        self.preempt_requested = False
        self.preempt_acked = False
        self.preempt_cv = threading.Condition()
        self.killed = False
        self.exit_code: Optional[int] = None
        self.token = secrets.token_hex(16)
        self.run_dir: Optional[str] = None
        self.allgathers: Dict[str, Dict[int, Any]] = {}
        self.allgather_cv = threading.Condition()
        self.preempt_deadline: Optional[float] = None
        self._threads: List[threading.Thread] = []

    # -- lifecycle ---------------------------------------------------------------------------------
    def start(self, slot_ids: List[int], agent_id: Optional[str] = None) -> None:
        m, t = self.master, self.trial
        self.slot_ids = slot_ids
        self.agent_id = agent_id or m.agent_id
        self.state = "RUNNING"
        exp = t.experiment
        remote = m.agents.is_remote(agent_id)
        # Where the new run starts is where the trial's latest checkpoint ends, not how far metrics were reported before a
        # crash (reference master/internal/trial.go:489-520 buildTaskSpecifier: StepsCompleted comes with LatestCheckpoint);
        # a trial restarted before its first checkpoint starts from zero.
        steps_done = 0
        if t.latest_checkpoint:
            row = m.db.one("SELECT steps_completed FROM checkpoints WHERE uuid=?", (t.latest_checkpoint,))
            steps_done = int(row["steps_completed"] or 0) if row is not None else 0
        if t.total_batches != steps_done:
            t.total_batches = steps_done
            t.persist()
        # what the master decides; the host running the task (this process, or an agent) layers it over its own environment
        task_env: Dict[str, str] = dict(exp.environment_variables())
        task_env.update({
            **m.master_env(),
            "DET_CLUSTER_ID": m.cluster_id, "DET_AGENT_ID": self.agent_id, "DET_SLOT_IDS": json.dumps(slot_ids),
            "DET_TASK_ID": t.task_id, "DET_ALLOCATION_ID": self.id, "DET_SESSION_TOKEN": self.token,
            "DET_TASK_TYPE": "TRIAL", "DET_TRIAL_ID": str(t.id), "DET_EXPERIMENT_ID": str(exp.id),
            "DET_TRIAL_SEED": str(t.seed), "DET_HPARAMS": util.json_encode(t.hparams),
            "DET_EXPERIMENT_CONFIG": util.json_encode(exp.config), "DET_STEPS_COMPLETED": str(steps_done),
            "DET_TRIAL_RUN_ID": str(t.run_id), "DET_RESOURCES_ID": f"{self.id}-box", "DET_CONTAINER_ID": self.id,
            "DET_USER": exp.owner_name, "PYTHONUNBUFFERED": "1",
        })
        ckpt = t.latest_checkpoint or t.warm_start_checkpoint
        if ckpt:
            task_env["DET_LATEST_CHECKPOINT"] = ckpt
        m.db.exec("INSERT INTO sessions(token,user_id,expiry,task_id) VALUES(?,?,?,?)",
                  (self.token, exp.owner_id, time.time() + 14 * 86400, t.task_id))
        m.db.exec("INSERT INTO allocations(id,task_id,trial_id,state,slots,slot_ids,start_time,resource_pool) VALUES(?,?,?,?,?,?,?,?)",
                  (self.id, t.task_id, t.id, "RUNNING", self.slots, json.dumps(slot_ids), time.time(), getattr(self, "pool_name", "default")))
        if remote:
            row = m.db.one("SELECT model_definition FROM experiments WHERE id=?", (exp.id,))
            m.agents.start(self.agent_id, self.id, self, [agents_mod.PYTHON_PLACEHOLDER, "-m", "determined_b200.exec.launch"], task_env, slot_ids,
                           archive=row["model_definition"] if row is not None else None, workdir="model",
                           hpc={k: exp.config[k] for k in ("slurm", "pbs") if isinstance(exp.config.get(k), dict)},
                           container=agents_mod.container_spec(exp.config))
            return
        self.run_dir = tempfile.mkdtemp(prefix=f"detb200-alloc-{self.id[:8]}-", dir=m.work_dir)
        model_dir = os.path.join(self.run_dir, "model")
        os.makedirs(model_dir, exist_ok=True)
        exp.extract_model_def(model_dir)
        info_dir = os.path.join(self.run_dir, "info")
        os.makedirs(info_dir, exist_ok=True)
        env = dict(os.environ)
        env.pop("DET_LATEST_CHECKPOINT", None)
        env.update(task_env)
        env.update({"DET_RUN_DIR": self.run_dir, "DET_INFO_DIR": info_dir, "DET_MODEL_DIR": model_dir})
        if slot_ids:
            env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in slot_ids)
        elif m.num_gpus:
            env["CUDA_VISIBLE_DEVICES"] = ""
        root = str(pathlib.Path(__file__).resolve().parents[2])
        env["PYTHONPATH"] = os.pathsep.join([model_dir, root] + ([env["PYTHONPATH"]] if env.get("PYTHONPATH") else []))
        cmd = [sys.executable, "-m", "determined_b200.exec.launch"]
        self.proc = subprocess.Popen(cmd, cwd=model_dir, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                                     start_new_session=True, text=True, bufsize=1, errors="replace")
        th = threading.Thread(target=self._pump_logs, daemon=True, name=f"logs-{self.id[:8]}")
        th.start()
        self._threads.append(th)
        th = threading.Thread(target=self._wait, daemon=True, name=f"wait-{self.id[:8]}")
        th.start()
        self._threads.append(th)

    def _log_rows(self, lines: List[str]) -> List[Tuple[Any, ...]]:
        rows: List[Tuple[Any, ...]] = []
        for line in lines:
            rank = None
            txt = line.rstrip("\n")
            if txt.startswith("[rank="):
                try:
                    end = txt.index("]")
                    rank = int(txt[6:end])
                    txt = txt[end + 2:] if len(txt) > end + 1 else ""
                except ValueError:
                    pass
            level = "INFO"
            for lv in ("DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"):
                if txt.startswith(lv + ":"):
                    level = lv
                    break
            rows.append((self.trial.task_id, self.id, time.time(), level, rank, "stdout", "task", txt + "\n"))
            self.master.log_policies(self.trial, txt)
            self.master.webhook_log_line(self.trial.task_id, txt)
        return rows

    def _pump_logs(self) -> None:
        assert self.proc is not None and self.proc.stdout is not None
        batch: List[Tuple[Any, ...]] = []
        last = time.time()
        for line in self.proc.stdout:
            batch.extend(self._log_rows([line]))
            if len(batch) >= 200 or time.time() - last > 0.5:
                self.master.db.add_logs(batch)
                batch, last = [], time.time()
        if batch:
            self.master.db.add_logs(batch)

    def _wait(self) -> None:
        assert self.proc is not None
        code = self.proc.wait()
        for th in self._threads[:1]:
            th.join(timeout=5)
        self._finish(code)

    # sink interface of master/agents.py: the same two things, reported by a remote agent
    def on_logs(self, lines: List[str]) -> None:
        rows = self._log_rows(lines)
        if rows:
            self.master.db.add_logs(rows)

    def on_exit(self, code: Optional[int]) -> None:
        if self.state != "TERMINATED":
            self._finish(code if code is not None else 137)

    def _finish(self, code: int) -> None:
        self.exit_code = code
        self.state = "TERMINATED"
        self.master.db.exec("UPDATE allocations SET state='TERMINATED', end_time=?, exit_code=?, exit_reason=? WHERE id=?",
                            (time.time(), code, "killed" if self.killed else ("preempted" if self.preempt_acked else "exited"), self.id))
        self.master.db.exec("DELETE FROM sessions WHERE token=?", (self.token,))
        with self.preempt_cv:
            self.preempt_cv.notify_all()
        with self.allgather_cv:
            self.allgather_cv.notify_all()
        if self.run_dir and not os.environ.get("DET_KEEP_RUN_DIRS"):
            shutil.rmtree(self.run_dir, ignore_errors=True)
        self.master.on_allocation_exit(self)

    def preempt(self) -> None:
        with self.preempt_cv:
            if not self.preempt_requested:
                self.preempt_requested = True
                self.preempt_deadline = time.time() + PREEMPT_KILL_DEADLINE_S
                self.preempt_cv.notify_all()

    def kill(self) -> None:
        self.killed = True
        if self.master.agents.is_remote(getattr(self, "agent_id", None)):
            if self.state == "RUNNING":
                self.master.agents.kill(self.agent_id, self.id)        # the agent escalates to SIGKILL itself
            return
        if self.proc is not None and self.proc.poll() is None:
            try:
                os.killpg(self.proc.pid, signal.SIGTERM)
            except ProcessLookupError:
                return

            def hard() -> None:
                time.sleep(10)
                if self.proc is not None and self.proc.poll() is None:
                    try:
                        os.killpg(self.proc.pid, signal.SIGKILL)
                    except ProcessLookupError:
                        pass
            threading.Thread(target=hard, daemon=True).start()

    def wait_preemption(self, timeout: float) -> bool:
        with self.preempt_cv:
            if not self.preempt_requested and self.state == "RUNNING":
                self.preempt_cv.wait(timeout=max(timeout, 0))
            return self.preempt_requested

    def all_gather(self, request_uuid: str, rank: int, num_peers: int, data: Any, timeout: float = RENDEZVOUS_TIMEOUT_S) -> List[Any]:
        with self.allgather_cv:
            slot = self.allgathers.setdefault(request_uuid, {})
            slot[rank] = data
            self.allgather_cv.notify_all()
            deadline = time.time() + timeout
            while len(slot) < num_peers and self.state == "RUNNING":
                left = deadline - time.time()
                if left <= 0:
                    raise TimeoutError("all_gather timed out")
                self.allgather_cv.wait(timeout=left)
            return [slot[r] for r in sorted(slot)]


class Trial:
    def __init__(self, exp: "Experiment", row: Dict[str, Any]) -> None:
        self.experiment = exp
        self.id: int = row["id"]
        self.request_id: str = row["request_id"]
        self.state: str = row["state"]
        self.hparams: Dict[str, Any] = row["hparams"]
        self.seed: int = row["seed"]
        self.restarts: int = row.get("restarts", 0)
        self.run_id: int = row.get("run_id", 0)
        self.latest_checkpoint: Optional[str] = row.get("latest_checkpoint")
        self.warm_start_checkpoint: Optional[str] = row.get("warm_start_checkpoint")
        self.total_batches: int = row.get("total_batches", 0)
        self.task_id: str = row.get("task_id") or f"trial-{self.id}"
        ss = row.get("searcher_state") or {}
        self.ops: List[int] = list(ss.get("ops", []))            # absolute target lengths, FIFO
        self.target: int = int(ss.get("target", 0))               # cumulative units requested so far
        self.completed_units: int = int(ss.get("completed_units", 0))
        self.close_requested: bool = bool(ss.get("close_requested", False))
        self.early_exit_reason: Optional[str] = ss.get("early_exit_reason")
        self.allocation: Optional[Allocation] = None
        self.best_validation: Optional[float] = row.get("best_validation")
        self.runner_state: str = row.get("runner_state") or ""
        self.latest_validation_batches: Optional[int] = ss.get("latest_validation_batches")
        self.cancel_retries = False

    def persist(self) -> None:
        ss = {"ops": self.ops, "target": self.target, "completed_units": self.completed_units,
              "close_requested": self.close_requested, "early_exit_reason": self.early_exit_reason,
              "latest_validation_batches": self.latest_validation_batches}
        self.experiment.master.db.exec(
            "UPDATE trials SET state=?, restarts=?, run_id=?, latest_checkpoint=?, total_batches=?, searcher_state=?, "
            "best_validation=?, runner_state=?, last_activity=?, end_time=COALESCE(end_time, ?) WHERE id=?",
            (self.state, self.restarts, self.run_id, self.latest_checkpoint, self.total_batches, dbmod.J(ss),
             self.best_validation, self.runner_state, time.time(),
             time.time() if self.state in TERMINAL else None, self.id))

    def current_op(self) -> Optional[int]:
        return self.ops[0] if self.ops else None

    def wants_resources(self) -> bool:
        return (self.state == ACTIVE and self.allocation is None and bool(self.ops)
                and self.experiment.state == ACTIVE)


class Experiment:
    def __init__(self, master: "Master", row: Dict[str, Any]) -> None:
        self.master = master
        self.id: int = row["id"]
        self.state: str = row["state"]
        self.config: Dict[str, Any] = row["config"]
        self.owner_id: int = row.get("owner_id") or 1
        self.owner_name: str = row.get("owner_name") or "determined"
        self.job_id = row.get("job_id") or f"exp-{self.id}"
        seed = int((self.config.get("reproducibility") or {}).get("experiment_seed") or 0)
        self.searcher = S.Searcher(seed, S.new_search_method(self.config["searcher"]), self.config.get("hyperparameters") or {})
        self.trials: Dict[str, Trial] = {}        # by request id
        self.trials_by_id: Dict[int, Trial] = {}
        self.lock = master.lock
        self.shutdown_requested = False
        self.best_metric: Optional[float] = None

    # -- helpers -----------------------------------------------------------------------------------
    def environment_variables(self) -> Dict[str, str]:
        ev = (self.config.get("environment") or {}).get("environment_variables") or []
        if isinstance(ev, dict):
            ev = ev.get("cuda") or ev.get("cpu") or []
        out = {}
        for item in ev:
            if isinstance(item, str) and "=" in item:
                k, v = item.split("=", 1)
                out[k] = v
        return out

    def extract_model_def(self, dst: str) -> None:
        row = self.master.db.one("SELECT model_definition FROM experiments WHERE id=?", (self.id,))
        blob = row["model_definition"] if row is not None else None
        if blob:
            with tarfile.open(fileobj=io.BytesIO(blob), mode="r:*") as tf:
                tf.extractall(dst, filter="data")

    def slots_per_trial(self) -> int:
        return int((self.config.get("resources") or {}).get("slots_per_trial", 1))

    def smaller_is_better(self) -> bool:
        return bool(self.config["searcher"].get("smaller_is_better", True))

    def persist(self) -> None:
        self.master.db.exec("UPDATE experiments SET state=?, progress=?, searcher_snapshot=?, end_time=COALESCE(end_time, ?) WHERE id=?",
                            (self.state, self.searcher.progress(), dbmod.J(self.snapshot()),
                             time.time() if self.state in TERMINAL else None, self.id))

    def snapshot(self) -> Dict[str, Any]:
        return {"searcher": self.searcher.snapshot(), "shutdown_requested": self.shutdown_requested,
                "best_metric": self.best_metric}

    # -- searcher plumbing ---------------------------------------------------------------------------
    def start(self) -> None:
        self.process_operations(self.searcher.initial_operations())

    def process_operations(self, ops: List[Any]) -> None:
        """Apply searcher operations; may recurse through ``trial_created`` etc. (reference
        ``experiment.go:763-887``).  Snapshots afterwards so a master restart resumes here."""
        queue = list(ops)
        while queue:
            op = queue.pop(0)
            if isinstance(op, S.Create):
                t = self._new_trial(op)
                queue.extend(self.searcher.trial_created(op.request_id))
                logger.info("experiment %d: created trial %d (%s)", self.id, t.id, op.request_id[:8])
            elif isinstance(op, S.ValidateAfter):
                t = self.trials.get(op.request_id)
                if t is None or t.state in TERMINAL:
                    continue
                t.target += op.length
                t.ops.append(t.target)
                t.persist()
            elif isinstance(op, S.Close):
                t = self.trials.get(op.request_id)
                if t is None:
                    continue
                t.close_requested = True
                t.persist()
                self._maybe_finish_trial(t, queue)
            elif isinstance(op, S.Shutdown):
                self.shutdown_requested = True
                if op.failure:
                    self._set_state(STOPPING_ERROR)
                elif op.cancel:
                    self._set_state(STOPPING_CANCELED)
                elif self.state in (ACTIVE, PAUSED):
                    self._set_state(STOPPING_COMPLETED)
        self._check_done()
        self.persist()
        self.master.kick()

    def _new_trial(self, op: S.Create) -> Trial:
        db = self.master.db
        tid = db.exec("INSERT INTO trials(experiment_id,request_id,state,hparams,seed,start_time,warm_start_checkpoint,searcher_state) VALUES(?,?,?,?,?,?,?,?)",
                      (self.id, op.request_id, ACTIVE, dbmod.J(op.hparams), op.trial_seed, time.time(),
                       op.checkpoint or self.config["searcher"].get("source_checkpoint_uuid"), "{}"))
        db.exec("UPDATE trials SET task_id=? WHERE id=?", (f"trial-{tid}", tid))
        db.exec("INSERT OR REPLACE INTO tasks(id,type,start_time,job_id,state,owner_id) VALUES(?,?,?,?,?,?)",
                (f"trial-{tid}", "TRIAL", time.time(), self.job_id, "ACTIVE", self.owner_id))
        t = Trial(self, {"id": tid, "request_id": op.request_id, "state": ACTIVE, "hparams": op.hparams,
                         "seed": op.trial_seed, "warm_start_checkpoint": op.checkpoint or self.config["searcher"].get("source_checkpoint_uuid")})
        self.trials[op.request_id] = t
        self.trials_by_id[tid] = t
        self.master.trials[tid] = t
        return t

    def _maybe_finish_trial(self, t: Trial, queue: List[Any]) -> None:
        """A trial is closed once Close was requested, it has no pending ops and no allocation."""
        if t.state in TERMINAL or not t.close_requested or t.ops or t.allocation is not None:
            return
        t.state = COMPLETED if not t.early_exit_reason else (ERROR if t.early_exit_reason == S.ExitedReason.ERRORED else COMPLETED)
        t.persist()
        self.master.db.exec("UPDATE tasks SET state='COMPLETED', end_time=? WHERE id=?", (time.time(), t.task_id))
        queue.extend(self.searcher.trial_closed(t.request_id))
        self.master.fire_webhooks("TRIAL_STATE_CHANGE", {"experiment_id": self.id, "trial_id": t.id, "state": t.state})

    def _set_state(self, state: str) -> None:
        if self.state in TERMINAL:
            return
        if self.state != state:
            logger.info("experiment %d: %s -> %s", self.id, self.state, state)
            self.state = state
            self.master.fire_webhooks("EXPERIMENT_STATE_CHANGE", {"experiment_id": self.id, "state": state})

    def _check_done(self) -> None:
        if self.state in STOPPING_TO_TERMINAL:
            running = [t for t in self.trials.values() if t.allocation is not None]
            if self.state in (STOPPING_CANCELED, STOPPING_KILLED, STOPPING_ERROR):
                for t in running:
                    if self.state == STOPPING_KILLED:
                        t.allocation.kill()
                    else:
                        t.allocation.preempt()
            if not running:
                for t in self.trials.values():
                    if t.state not in TERMINAL:
                        t.state = CANCELED if self.state != STOPPING_COMPLETED else COMPLETED
                        t.persist()
                self._set_terminal(STOPPING_TO_TERMINAL[self.state])

    def _set_terminal(self, state: str) -> None:
        self.state = state
        logger.info("experiment %d: finished as %s", self.id, state)
        if getattr(self.master, "telemetry", None) is not None:
            self.master.telemetry.track("experiment_state_changed", experiment_id=self.id, state=state, num_trials=len(self.trials))
        self.master.fire_webhooks("EXPERIMENT_STATE_CHANGE", {"experiment_id": self.id, "state": state})
        self.master.schedule_gc(self)

    # -- events from trials --------------------------------------------------------------------------
    def op_completed(self, t: Trial, length: int, metric: Any) -> None:
        """Harness reported the searcher metric for the op ending at absolute ``length``."""
        if not t.ops or t.ops[0] != length:
            raise Conflict(f"trial {t.id} has no operation of length {length} in flight (ops={t.ops})")
        prev = t.completed_units
        t.ops.pop(0)
        t.completed_units = length
        if isinstance(metric, (int, float)):
            m = float(metric)
            if t.best_validation is None or (m < t.best_validation) == self.smaller_is_better():
                t.best_validation = m
            if self.best_metric is None or (m < self.best_metric) == self.smaller_is_better():
                self.best_metric = m
        t.persist()
        self.searcher.set_trial_progress(t.request_id, float(length))
        ops = self.searcher.validation_completed(t.request_id, metric, S.ValidateAfter(t.request_id, length - prev))
        self.process_operations(ops)

    def trial_progress(self, t: Trial, units: float) -> None:
        self.searcher.set_trial_progress(t.request_id, float(units))
        self.master.db.exec("UPDATE experiments SET progress=? WHERE id=?", (self.searcher.progress(), self.id))

    def trial_early_exit(self, t: Trial, reason: str) -> None:
        if t.early_exit_reason is not None:
            raise Conflict("early exit already reported")
        t.early_exit_reason = reason
        t.ops = []
        t.close_requested = True
        t.persist()
        self.process_operations(self.searcher.trial_exited_early(t.request_id, reason))

    def allocation_exited(self, t: Trial, alloc: Allocation) -> None:
        """Reference ``trial.go:534-673``."""
        t.allocation = None
        code = alloc.exit_code or 0
        queue: List[Any] = []
        if self.state in STOPPING_TO_TERMINAL or self.state in TERMINAL:
            pass
        elif t.early_exit_reason is not None:
            pass
        elif alloc.killed:
            pass
        elif code != 0 and not alloc.preempt_acked:
            t.restarts += 1
            max_restarts = int(self.config.get("max_restarts", 5))
            if t.restarts > max_restarts or t.cancel_retries:
                logger.error("trial %d failed (exit code %d) and exhausted max_restarts=%d", t.id, code, max_restarts)
                t.early_exit_reason = S.ExitedReason.ERRORED
                t.ops = []
                t.close_requested = True
                t.state = ERROR
                t.persist()
                self.master.db.exec("UPDATE tasks SET state='ERROR', end_time=? WHERE id=?", (time.time(), t.task_id))
                queue.extend(self.searcher.trial_exited_early(t.request_id, S.ExitedReason.ERRORED))
                queue.extend(self.searcher.trial_closed(t.request_id))
            else:
                logger.warning("trial %d exited with code %d; restart %d/%d from checkpoint %s", t.id, code,
                               t.restarts, max_restarts, t.latest_checkpoint)
        t.run_id += 1
        t.persist()
        self._maybe_finish_trial(t, queue)
        self.process_operations(queue)

    # -- user actions --------------------------------------------------------------------------------
    def pause(self) -> None:
        if self.state == ACTIVE:
            self._set_state(PAUSED)
            for t in self.trials.values():
                if t.allocation is not None:
                    t.allocation.preempt()
            self.persist()

    def activate(self) -> None:
        if self.state == PAUSED:
            self._set_state(ACTIVE)
            self.persist()
            self.master.kick()

    def cancel(self) -> None:
        if self.state not in TERMINAL:
            self._set_state(STOPPING_CANCELED)
            self._check_done()
            self.persist()

    def kill(self) -> None:
        if self.state not in TERMINAL:
            self._set_state(STOPPING_KILLED)
            self._check_done()
            self.persist()


class Master:
    def __init__(self, db_path: str = ":memory:", host: str = "127.0.0.1", port: int = 8080,
                 scheduler_type: str = "fair_share", fitting_policy: str = "best", preemption: bool = True,
                 num_slots: Optional[int] = None, work_dir: Optional[str] = None,
                 cluster_defaults: Optional[Dict[str, Any]] = None, launch: bool = True,
                 resource_pools: Optional[List[Dict[str, Any]]] = None) -> None:
        self.db = dbmod.DB(db_path)
        self.host, self.port = host, port
        self.lock = threading.RLock()
        self.cluster_id = self.db.get_meta("cluster_id") or "cluster"
        self.agent_id = "b200-box"
        self.launch_enabled = launch
        gpus, _ = gpu.get_gpus() if num_slots is None else ([], "")
        self.gpus = gpus
        self.num_gpus = len(gpus)
        n = num_slots if num_slots is not None else (len(gpus) or 1)
        self.slot_type = "cuda" if (self.num_gpus and num_slots is None) else "cpu"
        # Resource pools (reference master.yaml ``resource_pools``): "default" always exists and owns this box's slots; more
        # pools come from the config and are filled by agents that register into them.  Each pool runs its own scheduler.
        self.pools: Dict[str, scheduler.ResourcePool] = {"default": scheduler.ResourcePool("default", scheduler_type, fitting_policy, preemption)}
        self.pool_descriptions: Dict[str, str] = {"default": "this box"}
        for rp in resource_pools or []:
            name = str(rp.get("pool_name") or rp.get("name") or "")
            if not name:
                raise ValueError("resource_pools entries need a pool_name")
            sch = rp.get("scheduler") or {}
            pre = (sch.get("preemption") or {}).get("enabled", preemption) if isinstance(sch.get("preemption"), dict) else preemption
            self.pools[name] = scheduler.ResourcePool(name, sch.get("type", scheduler_type), sch.get("fitting_policy", fitting_policy), bool(pre))
            self.pool_descriptions[name] = str(rp.get("description") or "")
        self.pools["default"].add_agent(scheduler.Agent(self.agent_id, n))      # num_slots=0: every slot comes from remote agents
        self.agents = agents_mod.AgentRegistry(self)
        from determined_b200.master import portregistry

        self.ports = portregistry.PortRegistry()
        self.work_dir = work_dir or tempfile.mkdtemp(prefix="detb200-master-")
        os.makedirs(self.work_dir, exist_ok=True)
        self.cluster_defaults = cluster_defaults or {}
        if "checkpoint_storage" not in self.cluster_defaults:
            self.cluster_defaults["checkpoint_storage"] = {"type": "shared_fs", "host_path": constants.DEFAULT_CHECKPOINT_PATH}
        self.experiments: Dict[int, Experiment] = {}
        self.trials: Dict[int, Trial] = {}
        self.allocations: Dict[str, Allocation] = {}
        self._kick = threading.Event()
        self._stop = threading.Event()
        self._sched_thread: Optional[threading.Thread] = None
        self.start_time = time.time()
        from determined_b200.master import commands as _commands

        self.commands = _commands.CommandManager(self)
        # pools backed by a workload manager instead of agents (reference dispatcher RM): master/slurm.py
        self.backends: Dict[str, Any] = {}
        for rp in resource_pools or []:
            prov = rp.get("provider") or {}
            kind = str(prov.get("type", "")).lower()
            if kind in ("slurm", "pbs"):
                from determined_b200.master import pbs, slurm

                name = str(rp.get("pool_name") or rp.get("name"))
                self.backends[name] = (slurm.SlurmBackend if kind == "slurm" else pbs.PBSBackend)(self, name, prov)
        self._restore()

    config: Any = None                         # master.config.MasterConfig when started from a configuration (``apply_config``)
    cluster_name = "b200-box"
    tls_cert_file: Optional[str] = None       # set by the API server when it serves HTTPS
    tls_cert_name: Optional[str] = None       # host name in the certificate, when tasks reach the master by another name

    @property
    def url(self) -> str:
        return f"{'https' if self.tls_cert_file else 'http'}://{self.host}:{self.port}"

    def master_env(self) -> Dict[str, str]:
        """How a task finds (and, over TLS, trusts) this master."""
        env = {"DET_MASTER": self.url, "DET_MASTER_HOST": self.host, "DET_MASTER_PORT": str(self.port)}
        if self.tls_cert_file:
            env["DET_MASTER_CERT_FILE"] = self.tls_cert_file
            if self.tls_cert_name:
                env["DET_MASTER_CERT_NAME"] = self.tls_cert_name
        return env

    def apply_config(self, cfg: Any) -> None:
        """The parts of a :class:`master.config.MasterConfig` that are not constructor arguments: cluster name, agent token,
        TLS name, the initial password of the built-in users (only while they still have the blank one), default idle
        timeouts of notebooks / tensorboards, log retention."""
        from determined_b200.common.api import authentication as auth

        self.config = cfg
        self.cluster_name = cfg["cluster_name"]
        self.agent_token = cfg["security.agent_token"]
        self.tls_cert_name = cfg["security.tls.name"]
        self.strict_job_queue_control = bool(cfg["security.authz.strict_job_queue_control"])
        self.authz_type = cfg["security.authz.type"]            # "basic", or "rbac": master/authz.py restricts requests by role
        pw = cfg["security.initial_user_password"]
        if pw:
            blank = auth.salt_and_hash("")
            self.db.exec("UPDATE users SET password_hash=?, modified_at=? WHERE username IN ('admin','determined') AND password_hash=?",
                         (auth.salt_and_hash(pw), time.time(), blank))
        self.commands.default_idle_timeout = {"NOTEBOOK": cfg["notebook_timeout"], "TENSORBOARD": cfg["tensorboard_timeout"]}
        self.log_retention_days = int(cfg["retention_policy.log_retention_days"])
        if cfg["webhooks.signing_key"]:
            self.db.set_meta("webhook_signing_key", cfg["webhooks.signing_key"])
        from determined_b200 import __version__
        from determined_b200.master import telemetry

        self.telemetry = telemetry.Telemetry(self.cluster_id, cfg["telemetry.enabled"], cfg["telemetry.segment_url"], cfg["telemetry.segment_master_key"], __version__)
        self.telemetry.track("master_started", num_pools=len(self.pools), num_slots=sum(a.num_slots for p in self.pools.values() for a in p.agents.values()),
                             scheduler=self.pool.scheduler)
        lg = cfg.data.get("logging") or {}
        if lg.get("type") == "elastic":
            if not lg.get("host"):
                raise ValueError("logging.type elastic needs logging.host")
            from determined_b200.master import elastic

            self.db.log_backend = elastic.ElasticLogs.from_config(lg)
            try:
                self.db.log_backend.ensure_template()
            except Exception as e:  # noqa: BLE001  (a cluster that is down now may be up when the first line arrives)
                logger.warning("could not install the Elasticsearch index template: %s", e)
        else:
            self.db.log_backend = None
        oidc = dict(cfg.data.get("oidc") or {})
        if oidc.get("enabled"):
            missing = [k for k in ("client_id", "idp_sso_url", "token_url", "userinfo_url", "idp_recipient_url") if not oidc.get(k)]
            if missing:
                raise ValueError("oidc.enabled needs: " + ", ".join("oidc." + k for k in missing))
        self.oidc = oidc
        sc = dict(cfg.data.get("scim") or {})
        if sc.get("enabled") and (sc.get("auth") or {}).get("type", "basic") == "basic" and not (sc.get("auth") or {}).get("password"):
            raise ValueError("scim.enabled with basic auth needs scim.auth.username and scim.auth.password")
        self.scim = sc

    log_retention_days = -1
    _last_retention_sweep = 0.0

    def sweep_old_logs(self, now: Optional[float] = None) -> int:
        """Delete task logs older than ``retention_policy.log_retention_days`` (reference ``master/internal/logretention``);
        returns how many lines went.  Called from the scheduler loop at most once an hour."""
        if self.log_retention_days < 0:
            return 0
        now = time.time() if now is None else now
        self._last_retention_sweep = now
        n = self.db.delete_logs_before(now - 86400.0 * self.log_retention_days)
        if n:
            logger.info("log retention: deleted %d task log lines older than %d days", n, self.log_retention_days)
        return n

    @property
    def pool(self) -> scheduler.ResourcePool:
        """The default pool (what every caller meant while there was only one)."""
        return self.pools["default"]

    def pool_for(self, resources: Optional[Dict[str, Any]]) -> scheduler.ResourcePool:
        name = (resources or {}).get("resource_pool") or "default"
        if name not in self.pools:
            raise expconf.ConfigError([f"<config>.resources.resource_pool: unknown resource pool {name!r} (have: {', '.join(sorted(self.pools))})"])
        return self.pools[name]

    def pool_of_agent(self, agent_id: str) -> Optional[scheduler.ResourcePool]:
        for p in self.pools.values():
            if agent_id in p.agents:
                return p
        return None

    def release_request(self, key: str) -> None:
        """Drop a request / allocation from whichever pool holds it (release is idempotent)."""
        for p in self.pools.values():
            p.release(key)

    # -- boot / shutdown -----------------------------------------------------------------------------
    def start(self) -> None:
        self._sched_thread = threading.Thread(target=self._scheduler_loop, daemon=True, name="scheduler")
        self._sched_thread.start()

    def stop(self) -> None:
        self._stop.set()
        self._kick.set()
        self.commands.stop_all()
        with self.lock:
            for a in list(self.allocations.values()):
                a.kill()
        for b in self.backends.values():
            b.stop()

    def _restore(self) -> None:
        """Re-hydrate non-terminal experiments (reference ``core.go:821-873``, ``restore.go``).
        Their processes died with the old master, so trials are simply re-queued and resume from
        their latest checkpoints."""
        for row in self.db.all("SELECT * FROM experiments WHERE state NOT IN ('COMPLETED','CANCELED','ERROR') AND unmanaged=0"):
            try:
                exp = self._hydrate(row)
                self.experiments[exp.id] = exp
                logger.info("restored experiment %d (%s) with %d trials", exp.id, exp.state, len(exp.trials))
                # processes died with the old master: trials that only waited for their allocation
                # to exit before closing can close now
                queue: List[Any] = []
                for t in list(exp.trials.values()):
                    exp._maybe_finish_trial(t, queue)
                exp.process_operations(queue)
            except Exception as e:
                logger.error("failed to restore experiment %s: %s", row["id"], e)
                self.db.exec("UPDATE experiments SET state='ERROR', end_time=? WHERE id=?", (time.time(), row["id"]))

    def _hydrate(self, row: Any) -> Experiment:
        u = self.db.one("SELECT username FROM users WHERE id=?", (row["owner_id"],))
        exp = Experiment(self, {"id": row["id"], "state": row["state"], "config": dbmod.U(row["config"], {}),
                                "owner_id": row["owner_id"], "owner_name": u["username"] if u else "determined",
                                "job_id": row["job_id"]})
        snap = dbmod.U(row["searcher_snapshot"])
        if snap:
            exp.searcher.restore(snap["searcher"])
            exp.shutdown_requested = bool(snap.get("shutdown_requested"))
            exp.best_metric = snap.get("best_metric")
        for tr in self.db.all("SELECT * FROM trials WHERE experiment_id=?", (exp.id,)):
            t = Trial(exp, {**dict(tr), "hparams": dbmod.U(tr["hparams"], {}), "searcher_state": dbmod.U(tr["searcher_state"], {})})
            exp.trials[t.request_id] = t
            exp.trials_by_id[t.id] = t
            self.trials[t.id] = t
        return exp

    # -- experiments -----------------------------------------------------------------------------------
    def create_experiment(self, config_text: Any, model_def: Optional[bytes] = None, owner: str = "determined",
                          activate: bool = True, parent_id: Optional[int] = None, template: Optional[str] = None,
                          validate_only: bool = False, project_id: int = 1, unmanaged: bool = False) -> Tuple[Optional[Experiment], Dict[str, Any]]:
        tmpl = None
        raw = expconf.parse(config_text)
        tname = template or raw.get("template")
        if tname:
            r = self.db.one("SELECT config FROM templates WHERE name=?", (tname,))
            if r is None:
                raise NotFound(f"template {tname!r} not found")
            tmpl = dbmod.U(r["config"], {})
        cfg = expconf.finalize(raw, cluster_defaults=self.cluster_defaults, template=tmpl)
        if (cfg.get("reproducibility") or {}).get("experiment_seed") is None:
            cfg.setdefault("reproducibility", {})["experiment_seed"] = int(time.time()) & 0x7FFFFFFF
        slots = int((cfg.get("resources") or {}).get("slots_per_trial", 1))
        cap = max((a.num_slots for a in self.pool_for(cfg.get("resources")).agents.values()), default=0)      # a trial runs on one agent
        if cap and slots > cap:           # no slots at all yet (agents still connecting): queue it
            raise expconf.ConfigError([f"<config>.resources.slots_per_trial: {slots} exceeds the {cap} slots of the largest agent"])
        S.new_search_method(cfg["searcher"])     # fail early on impossible searcher configs
        if validate_only:
            return None, cfg
        u = self.db.one("SELECT id FROM users WHERE username=?", (owner,))
        owner_id = u["id"] if u else 1
        with self.lock:
            eid = self.db.exec(
                "INSERT INTO experiments(state,config,original_config,model_definition,owner_id,start_time,parent_id,project_id,unmanaged) VALUES(?,?,?,?,?,?,?,?,?)",
                (ACTIVE if activate else PAUSED, dbmod.J(cfg), config_text if isinstance(config_text, str) else dbmod.J(raw),
                 model_def, owner_id, time.time(), parent_id, project_id, int(unmanaged)))
            self.db.exec("UPDATE experiments SET job_id=? WHERE id=?", (f"exp-{eid}", eid))
            exp = Experiment(self, {"id": eid, "state": ACTIVE if activate else PAUSED, "config": cfg,
                                    "owner_id": owner_id, "owner_name": owner})
            self.experiments[eid] = exp
            if getattr(self, "telemetry", None) is not None:
                self.telemetry.track("experiment_created", experiment_id=eid, searcher=(cfg.get("searcher") or {}).get("name"),
                                     slots_per_trial=(cfg.get("resources") or {}).get("slots_per_trial", 1), unmanaged=bool(unmanaged))
            if not unmanaged:
                exp.start()
        return exp, cfg

    def continue_experiment(self, eid: int, override: Any = None) -> Experiment:
        """Re-open a finished single-trial experiment (reference ``api_experiment.go`` ``ContinueExperiment``): the same trial
        trains on from its latest checkpoint, typically with a larger ``searcher.max_length`` from ``override``.  A completed
        experiment must ask for more training than it has done; an errored / cancelled one may simply be retried."""
        from determined_b200.master.searcher import methods as S_methods

        with self.lock:
            live = self.experiments.get(eid)
            if live is not None and live.state not in TERMINAL:
                raise Conflict(f"experiment {eid} is {live.state}; only finished experiments can be continued")
            row = self.db.one("SELECT * FROM experiments WHERE id=?", (eid,))
            if row is None:
                raise NotFound(f"experiment {eid} not found")
            if row["state"] not in TERMINAL:
                raise Conflict(f"experiment {eid} is {row['state']}; only finished experiments can be continued")
            cfg = dbmod.U(row["config"], {})
            if override:
                cfg = expconf.merge(expconf.parse(override), cfg)
                expconf.validate_sanity(cfg)
                expconf.validate_complete(cfg)
            if cfg["searcher"]["name"] != "single":
                raise expconf.ConfigError([f"<config>.searcher.name: only single-trial experiments can be continued, not {cfg['searcher']['name']}"])
            prev_state = row["state"]
            self.db.exec("UPDATE experiments SET state=?, config=?, end_time=NULL WHERE id=?", (ACTIVE, dbmod.J(cfg), eid))
            exp = self._hydrate(self.db.one("SELECT * FROM experiments WHERE id=?", (eid,)))
            if len(exp.trials) != 1:
                self.db.exec("UPDATE experiments SET state=?, config=? WHERE id=?", (prev_state, row["config"], eid))
                raise Conflict(f"experiment {eid} has {len(exp.trials)} trials; exactly one is needed to continue")
            t = next(iter(exp.trials.values()))
            method = exp.searcher.method
            assert isinstance(method, S_methods.RandomSearch)
            if prev_state == COMPLETED and t.completed_units >= method.max_length:
                self.db.exec("UPDATE experiments SET state=?, config=? WHERE id=?", (prev_state, row["config"], eid))
                raise Conflict(f"experiment {eid} already trained {t.completed_units} units; raise searcher.max_length above that to continue")
            # a fresh single-trial search that has already created this trial and still waits for its one operation
            seed = int((cfg.get("reproducibility") or {}).get("experiment_seed") or 0)
            exp.searcher = S.Searcher(seed, method, cfg.get("hyperparameters") or {})
            method.s.update(created_trials=1, pending_trials=1)
            exp.searcher.trials_requested = 1
            exp.searcher.trials_created[t.request_id] = True
            exp.searcher.trial_progress[t.request_id] = float(t.completed_units)
            exp.shutdown_requested = False
            exp.state = ACTIVE
            t.state, t.restarts, t.early_exit_reason, t.cancel_retries = ACTIVE, 0, None, False
            t.target = max(method.max_length, t.completed_units)
            t.ops = [t.target] if t.target > t.completed_units else []
            t.close_requested = True
            self.db.exec("UPDATE trials SET end_time=NULL WHERE id=?", (t.id,))
            self.db.exec("UPDATE tasks SET state='ACTIVE', end_time=NULL WHERE id=?", (t.task_id,))
            t.persist()
            self.experiments[eid] = exp
            self.trials[t.id] = t
            if not t.ops:                         # nothing left to train (an errored run that had in fact finished): close it again
                exp.process_operations([])
                queue: List[Any] = []
                exp._maybe_finish_trial(t, queue)
                exp.process_operations(queue)
            exp.persist()
            logger.info("experiment %d continued: trial %d trains on from %d to %d", eid, t.id, t.completed_units, t.target)
        self.kick()
        return exp

    def move_experiment(self, eid: int, project_id: int) -> None:
        if self.db.one("SELECT id FROM projects WHERE id=?", (project_id,)) is None:
            raise NotFound(f"project {project_id} not found")
        if self.db.one("SELECT id FROM experiments WHERE id=?", (eid,)) is None:
            raise NotFound(f"experiment {eid} not found")
        self.db.exec("UPDATE experiments SET project_id=? WHERE id=?", (project_id, eid))

    def get_experiment(self, eid: int) -> Experiment:
        with self.lock:
            exp = self.experiments.get(eid)
            if exp is None:
                row = self.db.one("SELECT * FROM experiments WHERE id=?", (eid,))
                if row is None:
                    raise NotFound(f"experiment {eid} not found")
                exp = self._hydrate(row)       # terminal experiment: read-only view
            return exp

    def get_trial(self, tid: int) -> Trial:
        with self.lock:
            t = self.trials.get(tid)
            if t is None:
                row = self.db.one("SELECT experiment_id FROM trials WHERE id=?", (tid,))
                if row is None:
                    raise NotFound(f"trial {tid} not found")
                exp = self.get_experiment(row["experiment_id"])
                t = exp.trials_by_id[tid]
            return t

    def get_allocation(self, aid: str) -> Allocation:
        a = self.allocations.get(aid)
        if a is None:
            raise NotFound(f"allocation {aid} not found")
        return a

    # -- scheduling ------------------------------------------------------------------------------------
    def kick(self) -> None:
        self._kick.set()

    def _scheduler_loop(self) -> None:
        while not self._stop.is_set():
            self._kick.wait(timeout=1.0)
            self._kick.clear()
            try:
                self.schedule_once()
                if self.log_retention_days >= 0 and time.time() - self._last_retention_sweep > 3600:
                    self.sweep_old_logs()
            except Exception:
                logger.exception("scheduler tick failed")

    def schedule_once(self) -> None:
        with self.lock:
            # submit / withdraw requests
            for exp in list(self.experiments.values()):
                res = exp.config.get("resources") or {}
                try:
                    pool = self.pool_for(res)
                except expconf.ConfigError:
                    pool = self.pool                 # the pool went away with a config change: fall back rather than strand the trial
                for t in exp.trials.values():
                    rid = f"req-{t.id}-{t.run_id}"
                    if t.wants_resources():
                        if rid not in pool.pending and not any(a.trial is t for a in self.allocations.values()):
                            pool.allocate(scheduler.AllocateRequest(
                                rid, exp.job_id, exp.slots_per_trial(), priority=res.get("priority"),
                                weight=float(res.get("weight", 1) or 1), max_slots=res.get("max_slots"),
                                preemptible=True, name=f"exp {exp.id} trial {t.id}",
                                submit_time=time.time()))
                    elif rid in pool.pending:
                        pool.release(rid)
            self.agents.reap()
            self.commands.reap_idle()
            to_release: List[str] = []
            for pool in self.pools.values():
                d = pool.tick()
                to_release += list(d.to_release)
                self._start_assigned(pool, d.to_allocate)
            for rid in to_release:
                for a in self.allocations.values():
                    if getattr(a, "pool_key", None) == rid:
                        logger.info("preempting allocation %s (scheduler)", a.id)
                        a.preempt()
            now = time.time()
            for a in self.allocations.values():
                if a.preempt_deadline is not None and now > a.preempt_deadline and a.state == "RUNNING":
                    logger.warning("allocation %s ignored preemption for %.0fs; killing", a.id, PREEMPT_KILL_DEADLINE_S)
                    a.kill()

    def _start_assigned(self, pool: scheduler.ResourcePool, assigned: Any) -> None:
        for req, asg in assigned:
            # a task runs on one host: the (single-node) launch layer has no cross-agent rendezvous
            hosts = [a for a, sl in asg.items() if sl] or list(asg)[:1]
            agent_id = hosts[0] if hosts else self.agent_id
            if len(hosts) > 1:
                logger.error("request %s was spread over agents %s; only single-agent placements can run", req.allocation_id, hosts)
                pool.release(req.allocation_id)
                continue
            if req.allocation_id.startswith("cmd-"):          # command / shell / notebook / tensorboard task
                if not self.commands.on_assigned(req.allocation_id, sorted(s for slots in asg.values() for s in slots), agent_id):
                    pool.release(req.allocation_id)
                continue
            tid = int(req.allocation_id.split("-")[1])
            t = self.trials.get(tid)
            if t is None or not t.wants_resources():
                pool.release(req.allocation_id)
                continue
            alloc = Allocation(self, t, req.allocation_id + "-" + uuid.uuid4().hex[:8], req.slots_needed)
            alloc.pool_key = req.allocation_id
            alloc.pool_name = pool.name
            t.allocation = alloc
            self.allocations[alloc.id] = alloc
            slot_ids = sorted(s for slots in asg.values() for s in slots)
            if self.launch_enabled:
                try:
                    alloc.start(slot_ids, agent_id)
                except Exception:
                    logger.exception("failed to start allocation %s", alloc.id)
                    alloc.exit_code = 1
                    alloc.state = "TERMINATED"
                    self.on_allocation_exit(alloc)
            else:
                alloc.slot_ids = slot_ids
                alloc.state = "RUNNING"

    def on_allocation_exit(self, alloc: Allocation) -> None:
        with self.lock:
            self.allocations.pop(alloc.id, None)
            self.release_request(getattr(alloc, "pool_key", alloc.id))
            alloc.trial.experiment.allocation_exited(alloc.trial, alloc)
        self.kick()

    # -- log policies / webhooks / gc ----------------------------------------------------------------
    def log_policies(self, trial: Trial, line: str) -> None:
        """Reference ``master/internal/logpattern``: regex policies with ``cancel_retries`` (and
        ``exclude_node``, meaningless on one box)."""
        pols = trial.experiment.config.get("log_policies") or []
        if not pols:
            return
        import re

        for p in pols:
            pat = p.get("pattern")
            if pat and re.search(pat, line):
                action = (p.get("action") or {}).get("type")
                if action == "cancel_retries":
                    trial.cancel_retries = True

    # webhook task-log triggers: compiled regexes cached per webhook row (rebuilt when the set of webhooks changes)
    _log_hooks: Optional[List[Tuple[Any, Any]]] = None

    def _task_log_hooks(self) -> List[Tuple[Any, Any]]:
        if self._log_hooks is None:
            import re

            hooks = []
            for r in self.db.all("SELECT * FROM webhooks"):
                for t in dbmod.U(r["triggers"], []):
                    if str(t.get("triggerType", "")).endswith("TASK_LOG"):
                        rx = (t.get("condition") or {}).get("regex")
                        if rx:
                            try:
                                hooks.append((re.compile(rx), dict(r)))
                            except re.error:
                                logger.warning("webhook %s: bad regex %r", r["id"], rx)
            self._log_hooks = hooks
        return self._log_hooks

    def webhooks_changed(self) -> None:
        self._log_hooks = None

    def webhook_log_line(self, task_id: str, line: str) -> None:
        """``TRIGGER_TYPE_TASK_LOG`` (reference ``internal/webhooks`` + ``logpattern``): post when a task's output matches."""
        for rx, row in self._task_log_hooks():
            if rx.search(line):
                self._post_webhook(row, "TASK_LOG", {"task_id": task_id, "log": line, "regex": rx.pattern})

    def _post_webhook(self, row: Dict[str, Any], trigger: str, payload: Dict[str, Any]) -> None:
        """One delivery: the default JSON event, or a Slack message for ``WEBHOOK_TYPE_SLACK``; signed with the cluster's
        webhook key (``X-Determined-AuthSignature`` = HMAC-SHA256 over ``timestamp,body``) like the reference."""
        import hashlib
        import hmac

        ts = int(time.time())
        if str(row.get("webhook_type", "")).endswith("SLACK"):
            what = ", ".join(f"{k} {v}" for k, v in payload.items())
            body = {"text": f"Determined: {trigger.replace('_', ' ').lower()} — {what}",
                    "blocks": [{"type": "section", "text": {"type": "mrkdwn", "text": f"*{trigger.replace('_', ' ').title()}*\n{what}"}}]}
        else:
            body = {"event_id": uuid.uuid4().hex, "event_type": trigger, "timestamp": ts, "data": payload, "condition": row.get("_condition")}
        raw = json.dumps(body).encode()
        key = (self.db.get_meta("webhook_signing_key") or "").encode()
        if not key:
            key = secrets.token_hex(16).encode()
            self.db.set_meta("webhook_signing_key", key.decode())
        sig = hmac.new(key, f"{ts},".encode() + raw, hashlib.sha256).hexdigest()
        url = row["url"]

        def post() -> None:
            try:
                import requests

                requests.post(url, data=raw, timeout=5, headers={"Content-Type": "application/json", "X-Determined-AuthTimestamp": str(ts),
                                                                 "X-Determined-AuthSignature": sig})
            except Exception as e:
                logger.debug("webhook %s failed: %s", url, e)

        threading.Thread(target=post, daemon=True).start()

    def fire_webhooks(self, trigger: str, payload: Dict[str, Any]) -> None:
        """State-change triggers.  A trigger may carry ``condition: {state: COMPLETED}`` (or a list of states); without one every
        change is delivered."""
        rows = self.db.all("SELECT * FROM webhooks")
        for r in rows:
            trig = dbmod.U(r["triggers"], [])
            for t in trig or [{}]:
                if trig and not str(t.get("triggerType", "")).endswith(trigger):
                    continue
                want = (t.get("condition") or {}).get("state")
                if want:
                    states = [want] if isinstance(want, str) else list(want)
                    if str(payload.get("state", "")).replace("STATE_", "") not in [str(x).replace("STATE_", "") for x in states]:
                        continue
                row = dict(r)
                row["_condition"] = t.get("condition")
                self._post_webhook(row, trigger, payload)
                break

    def schedule_gc(self, exp: Experiment) -> None:
        threading.Thread(target=self.run_checkpoint_gc, args=(exp.id,), daemon=True, name=f"gc-{exp.id}").start()

    def checkpoints_to_keep(self, exp_id: int, cfg: Dict[str, Any], smaller_is_better: bool) -> Tuple[List[str], List[str]]:
        """Apply ``save_experiment_best / save_trial_best / save_trial_latest`` (reference
        ``checkpoint_gc.go`` + SQL in ``static/srv/gc_checkpoints.sql``)."""
        cs = cfg.get("checkpoint_storage") or {}
        n_exp, n_best, n_latest = int(cs.get("save_experiment_best", 0)), int(cs.get("save_trial_best", 1)), int(cs.get("save_trial_latest", 1))
        metric = cfg["searcher"].get("metric")
        rows = self.db.all("SELECT c.uuid, c.trial_id, c.steps_completed, c.report_time FROM checkpoints c JOIN trials t ON t.id=c.trial_id "
                           "WHERE t.experiment_id=? AND c.state='COMPLETED'", (exp_id,))
        registered = {r["checkpoint_uuid"] for r in self.db.all("SELECT checkpoint_uuid FROM model_versions")}
        scored = []
        for r in rows:
            v = self.db.one("SELECT metrics FROM metrics WHERE trial_id=? AND grp='validation' AND steps_completed=? ORDER BY id DESC LIMIT 1",
                            (r["trial_id"], r["steps_completed"]))
            m = None
            if v is not None and metric:
                val = dbmod.U(v["metrics"], {}).get(metric)
                if isinstance(val, (int, float)):
                    m = float(val) if smaller_is_better else -float(val)
            scored.append((r["uuid"], r["trial_id"], r["steps_completed"], m))
        keep = set(u for u in registered)
        by_trial: Dict[int, List[Tuple[str, int, int, Optional[float]]]] = {}
        for s in scored:
            by_trial.setdefault(s[1], []).append(s)
        for lst in by_trial.values():
            for s in sorted(lst, key=lambda s: -s[2])[:n_latest]:
                keep.add(s[0])
            for s in sorted([s for s in lst if s[3] is not None], key=lambda s: s[3])[:n_best]:
                keep.add(s[0])
        for s in sorted([s for s in scored if s[3] is not None], key=lambda s: s[3])[:n_exp]:
            keep.add(s[0])
        delete = [s[0] for s in scored if s[0] not in keep]
        return sorted(keep), delete

    def run_checkpoint_gc(self, exp_id: int) -> List[str]:
        from determined_b200.common import storage

        row = self.db.one("SELECT config FROM experiments WHERE id=?", (exp_id,))
        if row is None:
            return []
        cfg = dbmod.U(row["config"], {})
        _, delete = self.checkpoints_to_keep(exp_id, cfg, bool(cfg["searcher"].get("smaller_is_better", True)))
        if not delete:
            return []
        try:
            mgr = storage.build(cfg["checkpoint_storage"], None)
        except Exception as e:
            logger.warning("checkpoint GC: cannot build storage manager: %s", e)
            return []
        for u in delete:
            try:
                mgr.delete(u, ["**/*"])
            except Exception as e:
                logger.warning("checkpoint GC: failed to delete %s: %s", u, e)
                continue
            self.db.exec("UPDATE checkpoints SET state='DELETED', resources='{}' WHERE uuid=?", (u,))
        logger.info("checkpoint GC for experiment %d deleted %d checkpoints", exp_id, len(delete))
        return delete
