"""Client-side custom hyperparameter search (``searcher: {name: custom}``).

Parity: reference ``harness/determined/searcher``: ``SearchMethod`` (initial_operations /
on_trial_created / on_validation_completed / on_trial_closed / on_trial_exited_early / progress /
save / load) ``_search_method.py:201-484``; operations ``Create / ValidateAfter / Close / Progress /
Shutdown`` ``:99-199``; ``SearcherState`` ``:16-68``; ``SearchRunner`` event loop
``_search_runner.py:24-241``; ``LocalSearchRunner`` (state files under a local directory)
``:243-362``; ``RemoteSearchRunner`` (state in checkpoint storage via a core.Context)
``_remote_search_runner.py:14-98``.

The master queues events (``GET /experiments/{id}/searcher_events``, long poll) and applies the
operations posted back (``POST /experiments/{id}/searcher_operations``).
"""
import abc
import dataclasses
import enum
import json
import logging
import os
import pathlib
import pickle
import time
import uuid
from typing import Any, Dict, List, Optional, Set, Tuple

import yaml

from determined_b200.common import api, context as det_context

logger = logging.getLogger("determined.searcher")

STATE_FILE = "state"


class SearcherState:
    """Bookkeeping the runner keeps for the method: what failed, was created, closed, and
    per-trial progress."""

    def __init__(self) -> None:
        self.failures: Set[uuid.UUID] = set()
        self.trial_progress: Dict[uuid.UUID, float] = {}
        self.trials_closed: Set[uuid.UUID] = set()
        self.trials_created: Set[uuid.UUID] = set()
        self.last_event_id: int = 0
        self.experiment_completed: bool = False
        self.experiment_failed: bool = False

    def to_dict(self) -> Dict[str, Any]:
        return {"failures": [str(x) for x in self.failures], "trial_progress": {str(k): v for k, v in self.trial_progress.items()},
                "trials_closed": [str(x) for x in self.trials_closed], "trials_created": [str(x) for x in self.trials_created],
                "last_event_id": self.last_event_id, "experiment_completed": self.experiment_completed,
                "experiment_failed": self.experiment_failed}

    def from_dict(self, d: Dict[str, Any]) -> None:
        self.failures = {uuid.UUID(x) for x in d.get("failures", [])}
        self.trial_progress = {uuid.UUID(k): v for k, v in d.get("trial_progress", {}).items()}
        self.trials_closed = {uuid.UUID(x) for x in d.get("trials_closed", [])}
        self.trials_created = {uuid.UUID(x) for x in d.get("trials_created", [])}
        self.last_event_id = int(d.get("last_event_id", 0))
        self.experiment_completed = bool(d.get("experiment_completed", False))
        self.experiment_failed = bool(d.get("experiment_failed", False))


class ExitedReason(enum.Enum):
    ERRORED = "ERRORED"
    USER_CANCELED = "USER_CANCELED"
    INVALID_HP = "INVALID_HP"

    @classmethod
    def _from_wire(cls, s: str) -> "ExitedReason":
        s = s.replace("EXITED_REASON_", "")
        if s == "INIT_INVALID_HP":
            s = "INVALID_HP"
        return cls(s)


class Operation(metaclass=abc.ABCMeta):
    @abc.abstractmethod
    def _to_wire(self) -> Dict[str, Any]:
        ...


class ValidateAfter(Operation):
    """Train the trial ``length`` more units (the experiment's searcher unit), then validate."""

    def __init__(self, request_id: uuid.UUID, length: int) -> None:
        self.request_id, self.length = request_id, length

    def _to_wire(self) -> Dict[str, Any]:
        return {"trialOperation": {"validateAfter": {"requestId": str(self.request_id), "length": str(self.length)}}}


class Close(Operation):
    def __init__(self, request_id: uuid.UUID) -> None:
        self.request_id = request_id

    def _to_wire(self) -> Dict[str, Any]:
        return {"closeTrial": {"requestId": str(self.request_id)}}


class Progress(Operation):
    def __init__(self, progress: float) -> None:
        self.progress = progress

    def _to_wire(self) -> Dict[str, Any]:
        return {"setSearcherProgress": {"progress": self.progress}}


class Shutdown(Operation):
    def __init__(self, cancel: bool = False, failure: bool = False) -> None:
        self.cancel, self.failure = cancel, failure

    def _to_wire(self) -> Dict[str, Any]:
        return {"shutDown": {"cancel": self.cancel, "failure": self.failure}}


class Create(Operation):
    def __init__(self, request_id: uuid.UUID, hparams: Dict[str, Any], checkpoint: Optional[str]) -> None:
        self.request_id, self.hparams, self.checkpoint = request_id, hparams, checkpoint

    def _to_wire(self) -> Dict[str, Any]:
        return {"createTrial": {"requestId": str(self.request_id), "hyperparams": json.dumps(self.hparams), "checkpoint": self.checkpoint}}


class SearchMethod:
    """Implement the event handlers; each returns the operations to apply next."""

    @abc.abstractmethod
    def initial_operations(self, searcher_state: SearcherState) -> List[Operation]:
        ...

    @abc.abstractmethod
    def on_trial_created(self, searcher_state: SearcherState, request_id: uuid.UUID) -> List[Operation]:
        ...

    @abc.abstractmethod
    def on_validation_completed(self, searcher_state: SearcherState, request_id: uuid.UUID, metric: Any, train_length: int) -> List[Operation]:
        ...

    @abc.abstractmethod
    def on_trial_closed(self, searcher_state: SearcherState, request_id: uuid.UUID) -> List[Operation]:
        ...

    @abc.abstractmethod
    def progress(self, searcher_state: SearcherState) -> float:
        ...

    @abc.abstractmethod
    def on_trial_exited_early(self, searcher_state: SearcherState, request_id: uuid.UUID, exited_reason: ExitedReason) -> List[Operation]:
        ...

    def save(self, searcher_state: SearcherState, path: pathlib.Path, *, experiment_id: int) -> None:
        path.mkdir(parents=True, exist_ok=True)
        d = searcher_state.to_dict()
        d["experimentId"] = experiment_id
        (path / STATE_FILE).write_text(json.dumps(d))
        self.save_method_state(path)

    def save_method_state(self, path: pathlib.Path) -> None:
        """Override to persist the method's own state (called by ``save``)."""

    def load(self, path: pathlib.Path) -> Tuple[SearcherState, int]:
        st = SearcherState()
        d = json.loads((path / STATE_FILE).read_text())
        st.from_dict(d)
        self.load_method_state(path)
        return st, int(d["experimentId"])

    def load_method_state(self, path: pathlib.Path) -> None:
        """Override to restore what ``save_method_state`` wrote."""


class _ExperimentInactive(Exception):
    def __init__(self, state: str) -> None:
        self.exp_state = state


class SearchRunner:
    def __init__(self, search_method: SearchMethod) -> None:
        self.search_method = search_method
        self.state = SearcherState()

    def _ops_for_event(self, ev: Dict[str, Any]) -> List[Operation]:
        m, st = self.search_method, self.state
        if "initialOperations" in ev:
            return m.initial_operations(st)
        if "trialCreated" in ev:
            rid = uuid.UUID(ev["trialCreated"]["requestId"])
            st.trials_created.add(rid)
            st.trial_progress[rid] = 0.0
            return m.on_trial_created(st, rid)
        if "trialClosed" in ev:
            rid = uuid.UUID(ev["trialClosed"]["requestId"])
            st.trials_closed.add(rid)
            ops = m.on_trial_closed(st, rid)
            if not ops and st.trials_created and st.trials_closed >= st.trials_created and not st.experiment_completed:
                pass
            return ops
        if "trialExitedEarly" in ev:
            rid = uuid.UUID(ev["trialExitedEarly"]["requestId"])
            reason = ExitedReason._from_wire(ev["trialExitedEarly"].get("exitedReason", "ERRORED"))
            if reason == ExitedReason.INVALID_HP:
                st.trial_progress.pop(rid, None)
            elif reason == ExitedReason.ERRORED:
                st.failures.add(rid)
            return m.on_trial_exited_early(st, rid, reason)
        if "validationCompleted" in ev:
            v = ev["validationCompleted"]
            rid = uuid.UUID(v["requestId"])
            return m.on_validation_completed(st, rid, v.get("metric"), int(v.get("validateAfterLength", 0)))
        if "trialProgress" in ev:
            rid = uuid.UUID(ev["trialProgress"]["requestId"])
            st.trial_progress[rid] = float(ev["trialProgress"].get("partialUnits", 0))
            return [Progress(m.progress(st))]
        if "experimentInactive" in ev:
            raise _ExperimentInactive(ev["experimentInactive"].get("experimentState", ""))
        raise RuntimeError(f"Unsupported searcher event: {ev}")

    def run_experiment(self, experiment_id: int, session: api.Session, prior_operations: Optional[List[Operation]] = None,
                       poll_timeout: int = 30) -> None:
        done = False
        while not done:
            events = self.get_events(session, experiment_id, poll_timeout)
            if not events:
                e = session.get(f"/api/v1/experiments/{experiment_id}").json()["experiment"]
                if e["state"] in ("STATE_COMPLETED", "STATE_CANCELED", "STATE_ERROR"):
                    self.state.experiment_completed = True
                    return
                if e["state"] == "STATE_PAUSED":
                    self._show_experiment_paused_msg()
                    time.sleep(2)
                continue
            for ev in events:
                if ev["id"] <= self.state.last_event_id:
                    ops: List[Operation] = list(prior_operations or [])     # replay after a crash between save and post
                else:
                    ops = self._ops_for_event(ev)
                    self.state.last_event_id = ev["id"]
                    self.save_state(experiment_id, ops)
                if any(isinstance(o, Shutdown) for o in ops):
                    done = True
                    self.state.experiment_completed = True
                self.post_operations(session, experiment_id, ev, ops)
                prior_operations = None

    def post_operations(self, session: api.Session, experiment_id: int, event: Dict[str, Any], operations: List[Operation]) -> None:
        session.post(f"/api/v1/experiments/{experiment_id}/searcher_operations",
                     json={"experimentId": experiment_id, "searcherOperations": [o._to_wire() for o in operations],
                           "triggeredByEvent": {"id": event["id"]}})

    def get_events(self, session: api.Session, experiment_id: int, timeout: int = 30) -> List[Dict[str, Any]]:
        r = session.get(f"/api/v1/experiments/{experiment_id}/searcher_events", params={"timeout_seconds": timeout}, timeout=timeout + 15)
        return r.json().get("searcherEvents", [])

    def save_state(self, experiment_id: int, operations: List[Operation]) -> None:
        pass

    def _show_experiment_paused_msg(self) -> None:
        logger.warning("Experiment is paused; activate it to continue the search.")


class LocalSearchRunner(SearchRunner):
    """Run the search from a user machine / login node; state lives under ``searcher_dir``."""

    def __init__(self, search_method: SearchMethod, searcher_dir: Optional[pathlib.Path] = None, session: Optional[api.Session] = None) -> None:
        super().__init__(search_method)
        self.searcher_dir = pathlib.Path(searcher_dir or pathlib.Path.cwd())
        self.session = session
        self.searcher_dir.mkdir(parents=True, exist_ok=True)

    def run(self, exp_config: Any, model_dir: Optional[str] = None, includes: Any = None) -> int:
        sess = self.session or api.authentication.login_with_cache(os.environ.get("DET_MASTER", "http://127.0.0.1:8080"))
        marker = self.searcher_dir / "experiment_id"
        prior: Optional[List[Operation]] = None
        if marker.exists():
            eid = int(marker.read_text())
            logger.info("Resuming search for experiment %d", eid)
            _, prior = self.load_state(eid)
        else:
            text = exp_config if isinstance(exp_config, str) and not os.path.exists(exp_config) else (
                pathlib.Path(exp_config).read_text() if isinstance(exp_config, (str, pathlib.Path)) else yaml.safe_dump(exp_config))
            files = det_context.read_v1_context(pathlib.Path(model_dir)) if model_dir else []
            r = sess.post("/api/v1/experiments", json={"config": text, "modelDefinition": files, "activate": True}).json()
            eid = r["experiment"]["id"]
            marker.write_text(str(eid))
        self.run_experiment(eid, sess, prior)
        return eid

    def _state_path(self, experiment_id: int) -> pathlib.Path:
        return self.searcher_dir / f"exp_{experiment_id}"

    def load_state(self, experiment_id: int) -> Tuple[int, List[Operation]]:
        p = self._state_path(experiment_id)
        self.state, eid = self.search_method.load(p)
        with (p / "ops").open("rb") as f:
            ops = pickle.load(f)
        return eid, ops

    def save_state(self, experiment_id: int, operations: List[Operation]) -> None:
        p = self._state_path(experiment_id)
        self.search_method.save(self.state, p, experiment_id=experiment_id)
        tmp = p / "ops.tmp"
        with tmp.open("wb") as f:
            pickle.dump(operations, f)
        os.replace(tmp, p / "ops")


class RemoteSearchRunner(SearchRunner):
    """Run the search *as a task on the cluster*: state is checkpointed through
    ``core_context.checkpoint`` so the runner itself can be preempted and resumed."""

    def __init__(self, search_method: SearchMethod, context: Any) -> None:
        super().__init__(search_method)
        self.context = context
        self.latest_checkpoint: Optional[str] = None

    def run(self, exp_config: Any, model_dir: Optional[str] = None, includes: Any = None) -> int:
        import determined_b200 as det

        info = det.get_cluster_info()
        assert info is not None, "RemoteSearchRunner only runs on-cluster"
        sess = api.Session(info.master_url, None, info.session_token, None)
        prior: Optional[List[Operation]] = None
        if info.latest_checkpoint:
            eid, prior = self.load_state(info.latest_checkpoint)
        else:
            text = pathlib.Path(exp_config).read_text() if isinstance(exp_config, (str, pathlib.Path)) and os.path.exists(str(exp_config)) else (
                exp_config if isinstance(exp_config, str) else yaml.safe_dump(exp_config))
            files = det_context.read_v1_context(pathlib.Path(model_dir)) if model_dir else []
            eid = sess.post("/api/v1/experiments", json={"config": text, "modelDefinition": files, "activate": True}).json()["experiment"]["id"]
        self.run_experiment(eid, sess, prior)
        return eid

    def load_state(self, storage_id: str) -> Tuple[int, List[Operation]]:
        with self.context.checkpoint.restore_path(storage_id) as p:
            self.state, eid = self.search_method.load(pathlib.Path(p))
            with (pathlib.Path(p) / "ops").open("rb") as f:
                return eid, pickle.load(f)

    def save_state(self, experiment_id: int, operations: List[Operation]) -> None:
        with self.context.checkpoint.store_path({"steps_completed": self.state.last_event_id}) as (p, sid):
            self.search_method.save(self.state, pathlib.Path(p), experiment_id=experiment_id)
            with (pathlib.Path(p) / "ops").open("wb") as f:
                pickle.dump(operations, f)
        self.latest_checkpoint = sid
