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
from typing import Any, Dict, List, Optional, Sequence, Set, Tuple

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
        self.experiment_id: Optional[int] = None
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

    @classmethod
    def _from_bindings(cls, reason: Any) -> "ExitedReason":
        """``v1TrialExitedEarlyExitedReason`` -> the method-facing reason (an unspecified reason is a crash)."""
        name = getattr(reason, "name", str(reason))
        return {"INVALID_HP": cls.INVALID_HP, "USER_REQUESTED_STOP": cls.USER_CANCELED, "USER_CANCELED": cls.USER_CANCELED}.get(name, cls.ERRORED)


class Operation(metaclass=abc.ABCMeta):
    @abc.abstractmethod
    def _to_wire(self) -> Dict[str, Any]:
        ...

    def _to_searcher_operation(self) -> Any:
        """The typed form (``bindings.v1SearcherOperation``) of this operation."""
        from determined_b200.common.api import bindings

        return bindings.v1SearcherOperation.from_json(self._to_wire())


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
    def __init__(self, state: Any) -> None:
        self.exp_state = state


def _rid(sub: Any) -> uuid.UUID:
    return uuid.UUID(sub.requestId)


class SearchRunner:
    """The event loop between the master's queue of searcher events and a :class:`SearchMethod`.

    Events and operations travel as the typed messages of ``common.api.bindings`` (``v1SearcherEvent`` in,
    ``v1PostSearcherOperationsRequest`` out); ``get_events`` / ``post_operations`` are the two places that touch the
    network and can be replaced (the reference's mock-master tests do exactly that)."""

    poll_timeout = 30          # seconds the master may hold an idle GetSearcherEvents

    def __init__(self, search_method: SearchMethod) -> None:
        self.search_method = search_method
        self.state = SearcherState()

    # ---- one event -> the method's answer ----------------------------------------------------------------------------------
    def _on_initial(self, sub: Any) -> List[Operation]:
        return self.search_method.initial_operations(self.state)

    def _on_created(self, sub: Any) -> List[Operation]:
        rid = _rid(sub)
        self.state.trials_created.add(rid)
        self.state.trial_progress[rid] = 0.0
        return self.search_method.on_trial_created(self.state, rid)

    def _on_closed(self, sub: Any) -> List[Operation]:
        rid = _rid(sub)
        self.state.trials_closed.add(rid)
        return self._with_progress(self.search_method.on_trial_closed(self.state, rid))

    def _on_exited_early(self, sub: Any) -> List[Operation]:
        from determined_b200.common.api import bindings

        if sub.exitedReason is None:
            raise RuntimeError("trialExitedEarly event is invalid without exitedReason")
        rid = _rid(sub)
        if sub.exitedReason == bindings.v1TrialExitedEarlyExitedReason.INVALID_HP:
            self.state.trial_progress.pop(rid, None)          # the trial never counted
        elif sub.exitedReason == bindings.v1TrialExitedEarlyExitedReason.UNSPECIFIED:
            self.state.failures.add(rid)
        return self._with_progress(self.search_method.on_trial_exited_early(self.state, rid, exited_reason=ExitedReason._from_bindings(sub.exitedReason)))

    def _on_validation(self, sub: Any) -> List[Operation]:
        if sub.metric is None:
            raise RuntimeError("validationCompleted event is invalid without a metric")
        ops = self.search_method.on_validation_completed(self.state, _rid(sub), sub.metric, int(sub.validateAfterLength))
        return self._with_progress(ops)

    def _on_progress(self, sub: Any) -> List[Operation]:
        self.state.trial_progress[_rid(sub)] = float(sub.partialUnits)
        return self._with_progress([])

    def _on_inactive(self, sub: Any) -> List[Operation]:
        raise _ExperimentInactive(sub.experimentState)

    def _with_progress(self, ops: List[Operation]) -> List[Operation]:
        return list(ops) + [Progress(self.search_method.progress(self.state))]

    _HANDLERS = (("initialOperations", "_on_initial"), ("trialCreated", "_on_created"), ("trialClosed", "_on_closed"),
                 ("trialExitedEarly", "_on_exited_early"), ("validationCompleted", "_on_validation"),
                 ("experimentInactive", "_on_inactive"), ("trialProgress", "_on_progress"))

    def _get_operations(self, event: Any) -> List[Operation]:
        for member, handler in self._HANDLERS:
            sub = getattr(event, member)
            if sub:
                logger.debug("searcher event %s: %s", event.id, member)
                return getattr(self, handler)(sub)
        raise RuntimeError(f"Unsupported event {event}")

    # ---- the loop -----------------------------------------------------------------------------------------------------------
    def run_experiment(self, experiment_id: int, session: api.Session, prior_operations: Optional[List[Operation]] = None,
                       sleep_time: float = 0.0) -> None:
        """Until the experiment goes inactive: fetch the pending events, answer each one.  The state is saved BEFORE the
        answer is posted; if the process died in between, the first event of the next fetch is one the saved state has seen,
        and the saved ``prior_operations`` are posted again instead of asking the method twice."""
        from determined_b200.common.api import bindings

        self.state.experiment_id = experiment_id
        active = True
        try:
            while active:
                if sleep_time:
                    time.sleep(sleep_time)
                events = self.get_events(session, experiment_id)
                if not events:
                    continue
                seen_before = self.state.last_event_id
                for n, event in enumerate(events):
                    if n == 0 and prior_operations is not None and 0 <= event.id <= seen_before and seen_before != 0:
                        logger.info("re-posting the saved operations for event %d", event.id)
                        ops = prior_operations
                    elif event.experimentInactive:
                        st = event.experimentInactive.experimentState
                        logger.info("experiment %s is inactive: %s", experiment_id, st)
                        self.state.experiment_completed = self.state.experiment_completed or st == bindings.experimentv1State.COMPLETED
                        self.state.experiment_failed = self.state.experiment_failed or st == bindings.experimentv1State.ERROR
                        if st == bindings.experimentv1State.PAUSED:
                            self._show_experiment_paused_msg()
                        else:
                            active = False
                        break
                    else:
                        ops = self._get_operations(event)
                        self.state.last_event_id = event.id
                        self.save_state(experiment_id, ops)
                    self.post_operations(session, experiment_id, event, ops)
                prior_operations = None
        except KeyboardInterrupt:
            print("Runner interrupted")

    def post_operations(self, session: api.Session, experiment_id: int, event: Any, operations: List[Operation]) -> None:
        from determined_b200.common.api import bindings

        body = bindings.v1PostSearcherOperationsRequest(experimentId=experiment_id, triggeredByEvent=event,
                                                        searcherOperations=[o._to_searcher_operation() for o in operations])
        try:
            bindings.post_PostSearcherOperations(session, body=body, experimentId=experiment_id)
        except api.errors.APIException as e:
            # a trial that exits on its own (an autotuning probe) can be gone before its Close arrives: not an error
            if any(isinstance(o, Close) for o in operations) and "could not be found" in str(e):
                logger.warning("ignoring %s", e)
            else:
                raise

    def get_events(self, session: api.Session, experiment_id: int) -> Optional[Sequence[Any]]:
        from determined_b200.common.api import bindings

        return bindings.get_GetSearcherEvents(session, experimentId=experiment_id, timeoutSeconds=self.poll_timeout).searcherEvents or []

    def save_state(self, experiment_id: int, operations: List[Operation]) -> None:
        pass

    def _show_experiment_paused_msg(self) -> None:
        logger.warning("Experiment is paused; activate it to continue the search.")

    @staticmethod
    def _searcher_event_as_dict(event: Any) -> Dict[str, Any]:
        return {k: v for k, v in event.to_json().items() if v is not None}


class LocalSearchRunner(SearchRunner):
    """Run the search from a user machine / login node; state lives under ``searcher_dir``."""

    def __init__(self, search_method: SearchMethod, searcher_dir: Optional[pathlib.Path] = None, session: Optional[api.Session] = None) -> None:
        super().__init__(search_method)
        self.searcher_dir = pathlib.Path(searcher_dir or pathlib.Path.cwd())
        self.session = session
        self.searcher_dir.mkdir(parents=True, exist_ok=True)

    def run(self, exp_config: Any, context_dir: Optional[str] = None, includes: Any = None, model_dir: Optional[str] = None) -> int:
        model_dir = model_dir or context_dir
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

    def _get_state_path(self, experiment_id: int) -> pathlib.Path:
        return self.searcher_dir / f"exp_{experiment_id}"

    def load_state(self, experiment_id: int) -> Tuple[int, List[Operation]]:
        p = self._get_state_path(experiment_id)
        self.state, eid = self.search_method.load(p)
        with (p / "ops").open("rb") as f:
            ops = pickle.load(f)
        return eid, ops

    def save_state(self, experiment_id: int, operations: List[Operation]) -> None:
        p = self._get_state_path(experiment_id)
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

    def run(self, exp_config: Any, context_dir: Optional[str] = None, includes: Any = None, model_dir: Optional[str] = None) -> int:
        import determined_b200 as det

        model_dir = model_dir or context_dir

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
