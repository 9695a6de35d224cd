"""Python SDK: ``from determined.experimental import client``.

Parity: reference ``harness/determined/experimental/client.py:1-646`` (module-level helpers bound
to a default ``Determined`` object) and ``harness/determined/common/experimental/*``
(``Determined``, ``Experiment``, ``Trial``, ``Checkpoint`` with ``download(mode=…)`` /
``write_metadata_file`` ``checkpoint/_checkpoint.py:19-505``, ``Model``/``ModelVersion``,
``Project``/``Workspace``)."""
import enum
import json
import os
import pathlib
import time
from typing import Any, Dict, Iterable, Iterator, List, Optional, Union

import yaml

from determined_b200.common import api, context as det_context, storage
from determined_b200.experimental import _sdk_more
from determined_b200.experimental._sdk_more import Oauth2ScimClient  # noqa: F401


class ExperimentState(enum.Enum):
    ACTIVE = "STATE_ACTIVE"
    PAUSED = "STATE_PAUSED"
    STOPPING_COMPLETED = "STATE_STOPPING_COMPLETED"
    STOPPING_CANCELED = "STATE_STOPPING_CANCELED"
    STOPPING_KILLED = "STATE_STOPPING_KILLED"
    STOPPING_ERROR = "STATE_STOPPING_ERROR"
    COMPLETED = "STATE_COMPLETED"
    CANCELED = "STATE_CANCELED"
    ERROR = "STATE_ERROR"
    DELETED = "STATE_DELETED"
    DELETING = "STATE_DELETING"
    DELETE_FAILED = "STATE_DELETE_FAILED"
    # sub-states of ACTIVE that the job queue reports
    QUEUED = "STATE_QUEUED"
    PULLING = "STATE_PULLING"
    STARTING = "STATE_STARTING"
    RUNNING = "STATE_RUNNING"


TERMINAL = {ExperimentState.COMPLETED, ExperimentState.CANCELED, ExperimentState.ERROR}


class TrialState(enum.Enum):
    """States of a trial (the experiment states that apply to one trial)."""

    UNSPECIFIED = "STATE_UNSPECIFIED"
    ACTIVE = "STATE_ACTIVE"
    PAUSED = "STATE_PAUSED"
    STOPPING_CANCELED = "STATE_STOPPING_CANCELED"
    STOPPING_KILLED = "STATE_STOPPING_KILLED"
    STOPPING_COMPLETED = "STATE_STOPPING_COMPLETED"
    STOPPING_ERROR = "STATE_STOPPING_ERROR"
    CANCELED = "STATE_CANCELED"
    COMPLETED = "STATE_COMPLETED"
    ERROR = "STATE_ERROR"
    QUEUED = "STATE_QUEUED"
    PULLING = "STATE_PULLING"
    STARTING = "STATE_STARTING"
    RUNNING = "STATE_RUNNING"


class DownloadMode(enum.Enum):
    """DIRECT: read checkpoint storage yourself; MASTER: proxy through the master; AUTO: DIRECT
    if the storage is reachable from here else MASTER."""

    DIRECT = "direct"
    MASTER = "master"
    AUTO = "auto"


class CheckpointState(enum.Enum):
    UNSPECIFIED = "STATE_UNSPECIFIED"
    ACTIVE = "STATE_ACTIVE"
    COMPLETED = "STATE_COMPLETED"
    ERROR = "STATE_ERROR"
    DELETED = "STATE_DELETED"
    PARTIALLY_DELETED = "STATE_PARTIALLY_DELETED"


class OrderBy(enum.Enum):
    """Ascending or descending order of a sorted listing."""

    ASCENDING = "ORDER_BY_ASC"
    ASC = "ORDER_BY_ASC"
    DESCENDING = "ORDER_BY_DESC"
    DESC = "ORDER_BY_DESC"


TrialOrderBy = ModelOrderBy = CheckpointOrderBy = OrderBy          # deprecated names of the same enum


class TrialSortBy(enum.Enum):
    """Field a listing of trials is sorted on."""

    ID = "SORT_BY_ID"
    START_TIME = "SORT_BY_START_TIME"
    END_TIME = "SORT_BY_END_TIME"
    STATE = "SORT_BY_STATE"
    BEST_VALIDATION_METRIC = "SORT_BY_BEST_VALIDATION_METRIC"
    LATEST_VALIDATION_METRIC = "SORT_BY_LATEST_VALIDATION_METRIC"
    BATCHES_PROCESSED = "SORT_BY_BATCHES_PROCESSED"
    DURATION = "SORT_BY_DURATION"
    RESTARTS = "SORT_BY_RESTARTS"
    CHECKPOINT_SIZE = "SORT_BY_CHECKPOINT_SIZE"
    LOG_RETENTION_DAYS = "SORT_BY_LOG_RETENTION_DAYS"


class ModelSortBy(enum.Enum):
    NAME = "SORT_BY_NAME"
    DESCRIPTION = "SORT_BY_DESCRIPTION"
    CREATION_TIME = "SORT_BY_CREATION_TIME"
    LAST_UPDATED_TIME = "SORT_BY_LAST_UPDATED_TIME"
    NUM_VERSIONS = "SORT_BY_NUM_VERSIONS"
    WORKSPACE = "SORT_BY_WORKSPACE"


class CheckpointSortBy(enum.Enum):
    """Field a listing of checkpoints is sorted on."""

    UUID = "SORT_BY_UUID"
    TRIAL_ID = "SORT_BY_TRIAL_ID"
    BATCH_NUMBER = "SORT_BY_BATCH_NUMBER"
    END_TIME = "SORT_BY_END_TIME"
    STATE = "SORT_BY_STATE"
    SEARCHER_METRIC = "SORT_BY_SEARCHER_METRIC"


class ExperimentSortBy(enum.Enum):
    """Field a listing of experiments is sorted on."""

    ID = "SORT_BY_ID"
    DESCRIPTION = "SORT_BY_DESCRIPTION"
    START_TIME = "SORT_BY_START_TIME"
    END_TIME = "SORT_BY_END_TIME"
    STATE = "SORT_BY_STATE"
    NUM_TRIALS = "SORT_BY_NUM_TRIALS"
    PROGRESS = "SORT_BY_PROGRESS"
    USER = "SORT_BY_USER"
    NAME = "SORT_BY_NAME"
    FORKED_FROM = "SORT_BY_FORKED_FROM"
    RESOURCE_POOL = "SORT_BY_RESOURCE_POOL"
    PROJECT_ID = "SORT_BY_PROJECT_ID"
    CHECKPOINT_SIZE = "SORT_BY_CHECKPOINT_SIZE"
    CHECKPOINT_COUNT = "SORT_BY_CHECKPOINT_COUNT"
    SEARCHER_METRIC_VAL = "SORT_BY_SEARCHER_METRIC_VAL"


ExperimentOrderBy = OrderBy


class TrialMetrics:
    """One metrics report of a trial (reference ``determined.experimental.TrialMetrics``): ``trial_id``, ``trial_run_id``,
    ``steps_completed``, ``end_time``, ``metrics``, ``group``, ``batch_metrics``.  Also readable like the dict earlier
    versions of this SDK yielded (``m["metrics"]``)."""

    _FIELDS = ("trial_id", "trial_run_id", "steps_completed", "end_time", "metrics", "group", "batch_metrics")
    __slots__ = _FIELDS

    def __init__(self, trial_id: int, trial_run_id: int, steps_completed: int, end_time: Any, metrics: Dict[str, Any], group: str,
                 batch_metrics: Optional[List[Dict[str, Any]]] = None) -> None:
        self.trial_id, self.trial_run_id, self.steps_completed, self.end_time = trial_id, trial_run_id, steps_completed, end_time
        self.metrics, self.group, self.batch_metrics = metrics, group, batch_metrics

    @classmethod
    def _from_json(cls, m: Dict[str, Any]) -> "TrialMetrics":
        kind = {"training": TrainingMetrics, "validation": ValidationMetrics}.get(m.get("group", ""), cls)
        return kind(int(m["trialId"]), int(m.get("trialRunId") or 0), int(m["totalBatches"]), m.get("endTime"), (m.get("metrics") or {}).get("avgMetrics") or {},
                    m.get("group", ""), (m.get("metrics") or {}).get("batchMetrics"))

    def keys(self) -> List[str]:
        return list(self._FIELDS)

    def __getitem__(self, k: str) -> Any:
        if k not in self._FIELDS:
            raise KeyError(k)
        return getattr(self, k)

    def get(self, k: str, default: Any = None) -> Any:
        return getattr(self, k, default) if k in self._FIELDS else default

    def __eq__(self, other: Any) -> bool:
        return isinstance(other, TrialMetrics) and all(getattr(self, k) == getattr(other, k) for k in self._FIELDS)

    def __repr__(self) -> str:
        return f"{type(self).__name__}(" + ", ".join(f"{k}={getattr(self, k)!r}" for k in self._FIELDS) + ")"


class TrainingMetrics(TrialMetrics):
    """A report of the ``training`` group."""

    __slots__ = ()


class ValidationMetrics(TrialMetrics):
    """A report of the ``validation`` group."""

    __slots__ = ()


class Checkpoint(_sdk_more.CheckpointMore):
    def __init__(self, session: api.Session, uuid: str, data: Optional[Dict[str, Any]] = None) -> None:
        self._session, self.uuid = session, uuid
        self._hydrate(data or {})

    def _hydrate(self, d: Dict[str, Any]) -> None:
        self.task_id = d.get("taskId")
        self.allocation_id = d.get("allocationId")
        self.report_time = d.get("reportTime")
        self.resources = {k: int(v) for k, v in (d.get("resources") or {}).items()}
        self.metadata = d.get("metadata") or {}
        self.state = CheckpointState(d["state"]) if d.get("state") in {s.value for s in CheckpointState} else None
        self.training = d.get("training") or {}

    def reload(self) -> None:
        self._hydrate(self._session.get(f"/api/v1/checkpoints/{self.uuid}").json()["checkpoint"])

    def _storage_config(self) -> Dict[str, Any]:
        if not self.training:
            self.reload()
        cfg = (self.training.get("experimentConfig") or {}).get("checkpoint_storage")
        if not cfg:
            raise ValueError("checkpoint has no recorded storage configuration")
        return cfg

    def download(self, path: Optional[str] = None, mode: DownloadMode = DownloadMode.AUTO) -> str:
        """Download to ``path`` (default ``checkpoints/<uuid>``); returns the directory."""
        dst = pathlib.Path(path) if path else pathlib.Path("checkpoints", self.uuid)
        if (dst / "metadata.json").exists() and path is None:
            return str(dst)
        dst.mkdir(parents=True, exist_ok=True)
        mgr = storage.build(self._storage_config(), None)
        mgr.download(self.uuid, str(dst))
        self.write_metadata_file(str(dst / "metadata.json"))
        return str(dst)

    def write_metadata_file(self, path: str) -> None:
        if os.path.exists(path):
            return
        with open(path, "w") as f:
            json.dump(self.metadata, f, indent=2)

    def add_metadata(self, metadata: Dict[str, Any]) -> None:
        self._session.post(f"/api/v1/checkpoints/{self.uuid}/metadata", json={"checkpoint": {"uuid": self.uuid, "metadata": metadata}})
        self.metadata.update(metadata)

    def remove_metadata(self, keys: List[str]) -> None:
        self.reload()
        md = {k: v for k, v in self.metadata.items() if k not in keys}
        self._session.post(f"/api/v1/checkpoints/{self.uuid}/metadata", json={"checkpoint": {"uuid": self.uuid, "metadata": {k: None for k in keys}}})
        self.metadata = md

    def delete(self) -> None:
        self._session.delete("/api/v1/checkpoints", json={"checkpointUuids": [self.uuid]})

    def remove_files(self, globs: List[str]) -> None:
        self._session.post("/api/v1/checkpoints/rm", json={"checkpointUuids": [self.uuid], "checkpointGlobs": globs})

    def __repr__(self) -> str:
        return f"Checkpoint(uuid={self.uuid}, task_id={self.task_id})"


class Trial(_sdk_more.TrialMore):
    def __init__(self, session: api.Session, trial_id: int, data: Optional[Dict[str, Any]] = None) -> None:
        self._session, self.id = session, trial_id
        self._hydrate(data or {})

    def _hydrate(self, d: Dict[str, Any]) -> None:
        self.experiment_id = d.get("experimentId")
        self.hparams = d.get("hparams")
        self.state = d.get("state")
        self.restarts = d.get("restarts")
        self.total_batches = d.get("totalBatchesProcessed")
        self.latest_checkpoint_uuid = d.get("latestCheckpoint")
        self.summary = d

    def reload(self) -> None:
        self._hydrate(self._session.get(f"/api/v1/trials/{self.id}").json()["trial"])

    def logs(self, follow: bool = False, head: Optional[int] = None, tail: Optional[int] = None, container_ids: Optional[List[str]] = None,
             rank_ids: Optional[List[int]] = None, stdtypes: Optional[List[str]] = None, min_level: Any = None, timestamp_before: Any = None,
             timestamp_after: Any = None, sources: Optional[List[str]] = None, search_text: Optional[str] = None) -> Iterator[str]:
        """Log lines of the trial (oldest first); filters as in the reference's ``Trial.logs`` / ``det trial logs``."""
        from determined_b200.common.api import logs as api_logs

        if isinstance(min_level, str):
            min_level = min_level.replace("LOG_LEVEL_", "").upper()
        for e in api_logs.trial_logs(self._session, self.id, head=head, tail=tail, follow=follow, container_ids=container_ids, rank_ids=rank_ids,
                                     sources=sources, stdtypes=stdtypes, min_level=min_level, timestamp_before=timestamp_before,
                                     timestamp_after=timestamp_after):
            line = e.log or e.message or ""
            if search_text is None or search_text in line:
                yield line

    def iter_metrics(self, group: Optional[str] = None) -> Iterator[TrialMetrics]:
        params = {"group": group} if group else {}
        for m in self._session.get(f"/api/v1/trials/{self.id}/metrics", params=params).json()["metrics"]:
            yield TrialMetrics._from_json(m)

    def stream_training_metrics(self) -> Iterator[Dict[str, Any]]:
        return self.iter_metrics("training")

    def stream_validation_metrics(self) -> Iterator[Dict[str, Any]]:
        return self.iter_metrics("validation")

    def list_checkpoints(self, order_desc: bool = False, sort_by: Any = None, order_by: Any = None, max_results: Optional[int] = None) -> List[Checkpoint]:
        """Completed checkpoints of the trial.  ``sort_by``: a :class:`CheckpointSortBy` member, or the name of a validation
        metric; ``order_by``: :class:`OrderBy` (both or neither); ``max_results`` cuts the list."""
        if (sort_by is None) != (order_by is None):
            raise AssertionError("sort_by and order_by must be either both set, or neither.")
        desc = order_desc if order_by is None else getattr(order_by, "value", order_by) == "ORDER_BY_DESC"
        params: Dict[str, Any] = {"orderBy": "ORDER_BY_DESC" if desc else "ORDER_BY_ASC"}
        if isinstance(sort_by, CheckpointSortBy):
            params["sortBy"] = sort_by.value
        r = self._session.get(f"/api/v1/trials/{self.id}/checkpoints", params=params).json()
        cks = [Checkpoint(self._session, c["uuid"], c) for c in r["checkpoints"] if c["state"] == "STATE_COMPLETED"]
        if isinstance(sort_by, str):
            key = lambda c: (c.training.get("validationMetrics", {}).get("avgMetrics") or {}).get(sort_by)  # noqa: E731
            cks = sorted((c for c in cks if key(c) is not None), key=key, reverse=desc)
        elif sort_by is CheckpointSortBy.SEARCHER_METRIC:
            cks = sorted((c for c in cks if c.training.get("searcherMetric") is not None), key=lambda c: c.training["searcherMetric"], reverse=desc)
        return cks[:max_results] if max_results else cks

    get_checkpoints = list_checkpoints

    def select_checkpoint(self, latest: bool = False, best: bool = False, uuid: Optional[str] = None, sort_by: Optional[str] = None,
                          smaller_is_better: Optional[bool] = None) -> Checkpoint:
        if sum([latest, best, uuid is not None]) != 1:
            raise AssertionError("Exactly one of latest, best, or uuid must be set")
        if uuid is not None:
            c = self._session.get(f"/api/v1/checkpoints/{uuid}").json()["checkpoint"]
            return Checkpoint(self._session, uuid, c)
        cks = self.list_checkpoints()
        if not cks:
            raise ValueError(f"trial {self.id} has no checkpoints")
        if latest:
            return max(cks, key=lambda c: c.metadata.get("steps_completed", 0))
        if sort_by is None:
            key = lambda c: c.training.get("searcherMetric")  # noqa: E731
            if smaller_is_better is None:
                smaller_is_better = bool(((cks[0].training.get("experimentConfig") or {}).get("searcher") or {}).get("smaller_is_better", True))
        else:
            key = lambda c: (c.training.get("validationMetrics", {}).get("avgMetrics") or {}).get(sort_by)  # noqa: E731
            if smaller_is_better is None:
                raise AssertionError("smaller_is_better must be set when sort_by is specified")
        cks = [c for c in cks if key(c) is not None]
        if not cks:
            raise ValueError("no checkpoint has the requested validation metric")
        return (min if smaller_is_better else max)(cks, key=key)

    def top_checkpoint(self, sort_by: Optional[str] = None, smaller_is_better: Optional[bool] = None) -> Checkpoint:
        return self.select_checkpoint(best=True, sort_by=sort_by, smaller_is_better=smaller_is_better)

    def kill(self) -> None:
        self._session.post(f"/api/v1/trials/{self.id}/kill")

    def __repr__(self) -> str:
        return f"Trial(id={self.id})"


class Experiment(_sdk_more.ExperimentMore):
    def __init__(self, session: api.Session, experiment_id: int, data: Optional[Dict[str, Any]] = None) -> None:
        self._session, self.id = session, experiment_id
        self._hydrate(data or {})

    def _hydrate(self, d: Dict[str, Any]) -> None:
        self.config = d.get("config")
        self.name = d.get("name")
        self.description = d.get("description")
        self.labels = d.get("labels")
        self.archived = d.get("archived")
        self.progress = d.get("progress")
        self.state = ExperimentState(d["state"]) if d.get("state") in {s.value for s in ExperimentState} else None
        self.start_time, self.end_time = d.get("startTime"), d.get("endTime")

    def reload(self) -> None:
        self._hydrate(self._session.get(f"/api/v1/experiments/{self.id}").json()["experiment"])

    def activate(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/activate")

    def pause(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/pause")

    def cancel(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/cancel")

    def kill(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/kill")

    def archive(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/archive")

    def unarchive(self) -> None:
        self._session.post(f"/api/v1/experiments/{self.id}/unarchive")

    def delete(self) -> None:
        self._session.delete(f"/api/v1/experiments/{self.id}")

    def set_name(self, name: str) -> None:
        self._session.patch(f"/api/v1/experiments/{self.id}", json={"name": name})

    def set_description(self, description: str) -> None:
        self._session.patch(f"/api/v1/experiments/{self.id}", json={"description": description})

    def add_label(self, label: str) -> None:
        self.reload()
        labels = sorted(set((self.labels or []) + [label]))
        self._session.patch(f"/api/v1/experiments/{self.id}", json={"labels": labels})

    def remove_label(self, label: str) -> None:
        self.reload()
        self._session.patch(f"/api/v1/experiments/{self.id}", json={"labels": [l for l in (self.labels or []) if l != label]})

    def move_to_project(self, workspace_name: str, project_name: str) -> None:
        ws = [w for w in self._session.get("/api/v1/workspaces", params={"name": workspace_name}).json()["workspaces"]]
        if not ws:
            raise ValueError(f"workspace {workspace_name!r} not found")
        pr = [p for p in self._session.get(f"/api/v1/workspaces/{ws[0]['id']}/projects").json()["projects"] if p["name"] == project_name]
        if not pr:
            raise ValueError(f"project {project_name!r} not found")
        self._session.post(f"/api/v1/experiments/{self.id}/move", json={"destinationProjectId": pr[0]["id"]})

    def list_trials(self, sort_by: TrialSortBy = TrialSortBy.ID, order_by: OrderBy = OrderBy.ASCENDING) -> List[Trial]:
        r = self._session.get(f"/api/v1/experiments/{self.id}/trials", params={"limit": 10000}).json()
        rows = r["trials"]

        def metric(d: Optional[Dict[str, Any]]) -> Any:
            m = ((d or {}).get("metrics") or {})
            m = m.get("avgMetrics", m)
            name = ((self.config or {}).get("searcher") or {}).get("metric")
            return m.get(name) if isinstance(m, dict) else None

        keys = {
            TrialSortBy.ID: lambda t: t["id"], TrialSortBy.START_TIME: lambda t: t.get("startTime") or "",
            TrialSortBy.END_TIME: lambda t: t.get("endTime") or "", TrialSortBy.STATE: lambda t: t.get("state") or "",
            TrialSortBy.BEST_VALIDATION_METRIC: lambda t: metric(t.get("bestValidation")),
            TrialSortBy.LATEST_VALIDATION_METRIC: lambda t: metric(t.get("latestValidation")),
            TrialSortBy.BATCHES_PROCESSED: lambda t: t.get("totalBatchesProcessed") or 0, TrialSortBy.RESTARTS: lambda t: t.get("restarts") or 0,
            TrialSortBy.DURATION: lambda t: (t.get("endTime") or "") > "" and (t.get("endTime"), t.get("startTime")) or ("", ""),
            TrialSortBy.CHECKPOINT_SIZE: lambda t: int(t.get("totalCheckpointSize") or 0), TrialSortBy.LOG_RETENTION_DAYS: lambda t: t.get("logRetentionDays") or 0,
        }
        key = keys[TrialSortBy(sort_by)]
        have = [t for t in rows if key(t) is not None]
        rest = [t for t in rows if key(t) is None]                        # trials without the value go last in either order
        have.sort(key=key, reverse=OrderBy(order_by) in (OrderBy.DESCENDING, OrderBy.DESC))
        return [Trial(self._session, t["id"], t) for t in have + rest]

    get_trials = list_trials

    def await_first_trial(self, interval: float = 0.1) -> Trial:
        while True:
            tr = self.list_trials()
            if tr:
                return tr[0]
            time.sleep(interval)

    def wait(self, interval: float = 5.0, timeout: Optional[float] = None) -> ExperimentState:
        t0 = time.time()
        while True:
            self.reload()
            if self.state in TERMINAL:
                return self.state
            if self.state == ExperimentState.PAUSED:
                raise ValueError(f"Experiment {self.id} is in paused state. Make sure the experiment is active.")
            if timeout is not None and time.time() - t0 > timeout:
                raise TimeoutError(f"experiment {self.id} still {self.state} after {timeout}s")
            time.sleep(interval)

    def list_checkpoints(self, sort_by: Any = None, smaller_is_better: Optional[bool] = None, limit: Optional[int] = None, order_by: Any = None,
                         max_results: Optional[int] = None) -> List[Checkpoint]:
        """``sort_by``: the name of a validation metric (default: the searcher metric); ``order_by`` (:class:`OrderBy`) is the
        reference's spelling of ``smaller_is_better`` (ascending = smaller first), ``max_results`` of ``limit``."""
        if order_by is not None and smaller_is_better is None:
            smaller_is_better = getattr(order_by, "value", order_by) != "ORDER_BY_DESC"
        limit = limit if limit is not None else max_results
        if isinstance(sort_by, CheckpointSortBy):
            sort_by = None if sort_by is CheckpointSortBy.SEARCHER_METRIC else sort_by
        params: Dict[str, Any] = {"states": "STATE_COMPLETED"}
        if sort_by is not None:
            params["sortByMetric"] = sort_by
        else:
            params["sortBy"] = "SORT_BY_SEARCHER_METRIC"
        r = self._session.get(f"/api/v1/experiments/{self.id}/checkpoints", params=params).json()["checkpoints"]
        cks = [Checkpoint(self._session, c["uuid"], c) for c in r]
        if sort_by is not None and smaller_is_better is False:
            cks.reverse()
        return cks[:limit] if limit else cks

    def top_checkpoint(self, sort_by: Optional[str] = None, smaller_is_better: Optional[bool] = None) -> Checkpoint:
        cks = self.top_n_checkpoints(1, sort_by, smaller_is_better)
        if not cks:
            raise AssertionError(f"No checkpoints found for experiment {self.id}")
        return cks[0]

    def top_n_checkpoints(self, limit: int, sort_by: Optional[str] = None, smaller_is_better: Optional[bool] = None) -> List[Checkpoint]:
        return self.list_checkpoints(sort_by, smaller_is_better, limit)

    def __repr__(self) -> str:
        return f"Experiment(id={self.id})"


class ModelVersion(_sdk_more.ModelVersionMore):
    def __init__(self, session: api.Session, d: Dict[str, Any]) -> None:
        self._session = session
        self.model_version = d.get("version")
        self.model_id = (d.get("model") or {}).get("id")
        self.model_name = (d.get("model") or {}).get("name")
        ck = d.get("checkpoint") or {}
        self.checkpoint = Checkpoint(session, ck.get("uuid"), ck) if ck.get("uuid") else None
        self.name, self.comment, self.notes = d.get("name"), d.get("comment"), d.get("notes")
        self.metadata = d.get("metadata") or {}

    def delete(self) -> None:
        self._session.delete(f"/api/v1/models/{self.model_name}/versions/{self.model_version}")


class Model(_sdk_more.ModelMore):
    def __init__(self, session: api.Session, d: Dict[str, Any]) -> None:
        self._session = session
        self._hydrate(d)

    def _hydrate(self, d: Dict[str, Any]) -> None:
        self.model_id, self.name = d.get("id"), d.get("name")
        self.description, self.metadata, self.labels = d.get("description"), d.get("metadata") or {}, d.get("labels") or []
        self.archived = d.get("archived")
        self.workspace_id, self.notes = d.get("workspaceId"), d.get("notes")

    def reload(self) -> None:
        self._hydrate(self._session.get(f"/api/v1/models/{self.name}").json()["model"])

    def register_version(self, checkpoint_uuid: str) -> ModelVersion:
        r = self._session.post(f"/api/v1/models/{self.name}/versions", json={"checkpointUuid": checkpoint_uuid}).json()
        return ModelVersion(self._session, r["modelVersion"])

    def get_version(self, version: int = -1) -> Optional[ModelVersion]:
        if version == -1:
            vs = self.list_versions(order_desc=True)
            return vs[0] if vs else None
        return ModelVersion(self._session, self._session.get(f"/api/v1/models/{self.name}/versions/{version}").json()["modelVersion"])

    def list_versions(self, order_desc: bool = True, order_by: Any = None) -> List[ModelVersion]:
        if order_by is not None:
            order_desc = getattr(order_by, "value", order_by) == "ORDER_BY_DESC"
        r = self._session.get(f"/api/v1/models/{self.name}/versions", params={"orderBy": "ORDER_BY_DESC" if order_desc else "ORDER_BY_ASC"}).json()
        return [ModelVersion(self._session, v) for v in r["modelVersions"]]

    get_versions = list_versions

    def add_metadata(self, metadata: Dict[str, Any]) -> None:
        self.metadata.update(metadata)
        self._session.patch(f"/api/v1/models/{self.name}", json={"model": {"metadata": self.metadata}})

    def set_labels(self, labels: List[str]) -> None:
        self._session.patch(f"/api/v1/models/{self.name}", json={"model": {"labels": labels}})
        self.labels = labels

    def set_description(self, description: str) -> None:
        self._session.patch(f"/api/v1/models/{self.name}", json={"model": {"description": description}})
        self.description = description

    def archive(self) -> None:
        self._session.post(f"/api/v1/models/{self.name}/archive")

    def unarchive(self) -> None:
        self._session.post(f"/api/v1/models/{self.name}/unarchive")

    def delete(self) -> None:
        self._session.delete(f"/api/v1/models/{self.name}")


class User(_sdk_more.UserMore):
    """Reference ``common/experimental/user.py``."""

    def __init__(self, session: api.Session, data: Dict[str, Any]) -> None:
        self._session = session
        self._hydrate(data)

    def _hydrate(self, d: Dict[str, Any]) -> None:
        self.user_id, self.username = d["id"], d["username"]
        self.admin, self.active, self.display_name = bool(d.get("admin")), bool(d.get("active", True)), d.get("displayName")

    def _patch(self, **fields: Any) -> None:
        self._hydrate(self._session.patch(f"/api/v1/users/{self.user_id}", json={"user": fields}).json()["user"])

    def rename(self, new_username: str) -> None:
        self._patch(username=new_username)

    def activate(self) -> None:
        self._patch(active=True)

    def deactivate(self) -> None:
        self._patch(active=False)

    def change_display_name(self, display_name: str) -> None:
        self._patch(displayName=display_name)

    def change_password(self, new_password: str) -> None:
        self._patch(password=api.authentication.salt_and_hash(new_password), isHashed=True)

    def reload(self) -> None:
        for u in self._session.get("/api/v1/users").json()["users"]:
            if u["id"] == self.user_id:
                self._hydrate(u)

    def __repr__(self) -> str:
        return f"User(id={self.user_id}, username={self.username!r})"


class Project(_sdk_more.ProjectMore):
    """Reference ``common/experimental/project.py``."""

    def __init__(self, session: api.Session, data: Dict[str, Any]) -> None:
        self._session = session
        self.id, self.name, self.description = data["id"], data["name"], data.get("description", "")
        self.workspace_id, self.n_experiments = data.get("workspaceId"), data.get("numExperiments", 0)
        self.archived, self.notes = bool(data.get("archived")), data.get("notes") or []

    def list_experiments(self) -> List["Experiment"]:
        r = self._session.get("/api/v1/experiments", params={"projectId": self.id, "limit": 1000}).json()
        return [Experiment(self._session, e["id"], e) for e in r["experiments"]]

    def delete(self) -> None:
        self._session.delete(f"/api/v1/projects/{self.id}")

    def __repr__(self) -> str:
        return f"Project(id={self.id}, name={self.name!r})"


class Workspace(_sdk_more.WorkspaceMore):
    """Reference ``common/experimental/workspace.py``."""

    def __init__(self, session: api.Session, data: Dict[str, Any]) -> None:
        self._session = session
        self.id, self.name = data["id"], data["name"]
        self.archived, self.immutable, self.n_projects = bool(data.get("archived")), bool(data.get("immutable")), data.get("numProjects", 0)

    def create_project(self, name: str, description: str = "") -> Project:
        r = self._session.post(f"/api/v1/workspaces/{self.id}/projects", json={"name": name, "description": description}).json()
        return Project(self._session, r["project"])

    def list_projects(self) -> List[Project]:
        return [Project(self._session, p) for p in self._session.get(f"/api/v1/workspaces/{self.id}/projects").json()["projects"]]

    def get_project(self, project_name: str) -> Project:
        for p in self.list_projects():
            if p.name == project_name:
                return p
        raise ValueError(f"workspace {self.name!r} has no project named {project_name!r}")

    def delete(self) -> None:
        self._session.delete(f"/api/v1/workspaces/{self.id}")

    def __repr__(self) -> str:
        return f"Workspace(id={self.id}, name={self.name!r})"


class ResourcePool(_sdk_more.ResourcePoolMore):
    """Reference ``common/experimental/resource_pool.py``."""

    def __init__(self, session: api.Session, data: Dict[str, Any]) -> None:
        self._session = session
        self.name = data["name"]
        self.slots_available, self.slots_used, self.num_agents = data.get("slotsAvailable", 0), data.get("slotsUsed", 0), data.get("numAgents", 0)
        self.scheduler_type, self.fitting_policy = data.get("schedulerType"), data.get("schedulerFittingPolicy")

    def __repr__(self) -> str:
        return f"ResourcePool(name={self.name!r}, slots={self.slots_used}/{self.slots_available})"


class Determined(_sdk_more.DeterminedMore):
    """Entry point of the SDK.  ``master`` defaults to ``$DET_MASTER`` or localhost:8080."""

    def __init__(self, master: Optional[str] = None, user: Optional[str] = None, password: Optional[str] = None,
                 cert_path: Optional[str] = None, cert_name: Optional[str] = None, noverify: bool = False) -> None:
        from determined_b200.common.api import certs

        self._master = master or os.environ.get("DET_MASTER") or "http://127.0.0.1:8080"
        cert = certs.default_load(self._master, cert_path, cert_name, noverify) if (cert_path or cert_name or noverify) else None
        self._session = api.authentication.login_with_cache(self._master, user, password, cert=cert)

    @classmethod
    def _from_session(cls, session: api.Session) -> "Determined":
        d = cls.__new__(cls)
        d._master, d._session = session.master, session
        return d

    def create_experiment(self, config: Union[str, pathlib.Path, Dict[str, Any]], model_dir: Union[str, pathlib.Path, None] = None,
                          includes: Optional[Iterable[Union[str, pathlib.Path]]] = None, parent_id: Optional[int] = None,
                          project_id: Optional[int] = None, template: Optional[str] = None, activate: bool = True) -> Experiment:
        if isinstance(config, (str, pathlib.Path)) and os.path.exists(str(config)):
            text = pathlib.Path(config).read_text()
        elif isinstance(config, dict):
            text = yaml.safe_dump(config)
        else:
            text = str(config)
        files: List[Dict[str, Any]] = []
        if model_dir is not None:
            files = det_context.read_v1_context(pathlib.Path(model_dir))
        for inc in includes or []:
            files += det_context.read_v1_context(pathlib.Path(inc))
        r = self._session.post("/api/v1/experiments", json={"config": text, "modelDefinition": files, "activate": activate, "parentId": parent_id,
                                                            "projectId": project_id, "template": template}).json()
        return Experiment(self._session, r["experiment"]["id"], r["experiment"])

    def continue_experiment(self, experiment_id: int, config: Union[str, Dict[str, Any], None] = None) -> Experiment:
        """Re-open a finished single-trial experiment; ``config`` holds overrides such as a larger ``searcher.max_length``."""
        text = yaml.safe_dump(config) if isinstance(config, dict) else config
        r = self._session.post("/api/v1/experiments/continue", json={"id": experiment_id, "overrideConfig": text}).json()
        return Experiment(self._session, experiment_id, r["experiment"])

    def get_experiment(self, experiment_id: int) -> Experiment:
        return Experiment(self._session, experiment_id, self._session.get(f"/api/v1/experiments/{experiment_id}").json()["experiment"])

    def list_experiments(self, **params: Any) -> List[Experiment]:
        r = self._session.get("/api/v1/experiments", params=params).json()
        return [Experiment(self._session, e["id"], e) for e in r["experiments"]]

    def get_trial(self, trial_id: int) -> Trial:
        return Trial(self._session, trial_id, self._session.get(f"/api/v1/trials/{trial_id}").json()["trial"])

    def get_checkpoint(self, uuid: str) -> Checkpoint:
        return Checkpoint(self._session, uuid, self._session.get(f"/api/v1/checkpoints/{uuid}").json()["checkpoint"])

    def create_model(self, name: str, description: str = "", metadata: Optional[Dict[str, Any]] = None, labels: Optional[List[str]] = None,
                     workspace_name: Optional[str] = None) -> Model:
        body: Dict[str, Any] = {"name": name, "description": description, "metadata": metadata or {}, "labels": labels or []}
        if workspace_name:
            body["workspaceName"] = workspace_name
        r = self._session.post("/api/v1/models", json=body).json()
        return Model(self._session, r["model"])

    def get_model(self, identifier: Union[str, int]) -> Model:
        return Model(self._session, self._session.get(f"/api/v1/models/{identifier}").json()["model"])

    def get_models(self, name: Optional[str] = None, **_: Any) -> List[Model]:
        r = self._session.get("/api/v1/models", params={"name": name} if name else {}).json()
        return [Model(self._session, m) for m in r["models"]]

    def get_model_by_id(self, model_id: int) -> Model:
        for m in self.get_models():
            if getattr(m, "model_id", getattr(m, "id", None)) == model_id:
                return m
        raise ValueError(f"no model with id {model_id}")

    def get_model_labels(self) -> List[str]:
        return sorted({label for m in self._session.get("/api/v1/models").json()["models"] for label in (m.get("labels") or [])})

    def get_master_info(self) -> Dict[str, Any]:
        return self._session.get("/api/v1/master").json()

    def whoami(self) -> Dict[str, Any]:
        return self._session.get("/api/v1/me").json()["user"]

    # -- users ---------------------------------------------------------------------------------
    def create_user(self, username: str, admin: bool = False, password: Optional[str] = None, active: bool = True, remote: bool = False) -> User:
        body: Dict[str, Any] = {"user": {"username": username, "admin": admin, "active": active, "remote": remote}}
        if password is not None:
            body.update(password=api.authentication.salt_and_hash(password), isHashed=True)
        return User(self._session, self._session.post("/api/v1/users", json=body).json()["user"])

    def list_users(self, active: Optional[bool] = None) -> List[User]:
        users = [User(self._session, u) for u in self._session.get("/api/v1/users", params={"limit": 10000}).json()["users"]]
        return [u for u in users if active is None or u.active == active]

    def get_user_by_id(self, user_id: int) -> User:
        for u in self.list_users():
            if u.user_id == user_id:
                return u
        raise ValueError(f"no user with id {user_id}")

    def get_user_by_name(self, user_name: str) -> User:
        for u in self.list_users():
            if u.username == user_name:
                return u
        raise ValueError(f"no user named {user_name!r}")

    def get_session_username(self) -> str:
        return str(self.whoami()["username"])

    # -- workspaces / projects / pools -------------------------------------------------------------
    def list_workspaces(self) -> List[Workspace]:
        return [Workspace(self._session, w) for w in self._session.get("/api/v1/workspaces", params={"limit": 10000}).json()["workspaces"]]

    def get_workspace(self, name: str) -> Workspace:
        for w in self.list_workspaces():
            if w.name == name:
                return w
        raise ValueError(f"no workspace named {name!r}")

    def create_workspace(self, name: str) -> Workspace:
        return Workspace(self._session, self._session.post("/api/v1/workspaces", json={"name": name}).json()["workspace"])

    def delete_workspace(self, name: str) -> None:
        self.get_workspace(name).delete()

    def list_resource_pools(self) -> List[ResourcePool]:
        return [ResourcePool(self._session, p) for p in self._session.get("/api/v1/resource-pools").json()["resourcePools"]]

    # -- metrics across trials ---------------------------------------------------------------------
    def iter_trials_metrics(self, trial_ids: List[int], group: str) -> Iterator[TrialMetrics]:
        """Metrics of several trials in one call (reference ``Determined.iter_trials_metrics``)."""
        r = self._session.get("/api/v1/trials/metrics/trial_metrics", params={"trialIds": list(trial_ids), "group": group}).json()
        for m in r["metrics"]:
            yield TrialMetrics._from_json(m)

    def stream_trials_metrics(self, trial_ids: List[int], group: str) -> Iterator[Dict[str, Any]]:
        return self.iter_trials_metrics(trial_ids, group)

    def stream_trials_training_metrics(self, trial_ids: List[int]) -> Iterator[Dict[str, Any]]:
        return self.iter_trials_metrics(trial_ids, "training")

    def stream_trials_validation_metrics(self, trial_ids: List[int]) -> Iterator[Dict[str, Any]]:
        return self.iter_trials_metrics(trial_ids, "validation")

    def preview_search(self, config: Union[str, Dict[str, Any]], seed: int = 0) -> Dict[str, Any]:
        text = yaml.safe_dump(config) if isinstance(config, dict) else config
        return self._session.post("/api/v1/preview-hp-search", json={"config": text, "seed": seed}).json()["simulation"]

    def logout(self) -> None:
        api.authentication.logout(self._master)


_determined: Optional[Determined] = None


def login(master: Optional[str] = None, user: Optional[str] = None, password: Optional[str] = None, **kw: Any) -> None:
    global _determined
    _determined = Determined(master, user, password)


def _d() -> Determined:
    global _determined
    if _determined is None:
        _determined = Determined()
    return _determined


def create_experiment(config: Any, model_dir: Any = None, includes: Any = None, **kw: Any) -> Experiment:
    return _d().create_experiment(config, model_dir, includes, **kw)


def get_experiment(experiment_id: int) -> Experiment:
    return _d().get_experiment(experiment_id)


def get_trial(trial_id: int) -> Trial:
    return _d().get_trial(trial_id)


def get_checkpoint(uuid: str) -> Checkpoint:
    return _d().get_checkpoint(uuid)


def create_model(name: str, description: str = "", metadata: Optional[Dict[str, Any]] = None, labels: Optional[List[str]] = None,
                 workspace_name: Optional[str] = None) -> Model:
    return _d().create_model(name, description, metadata, labels, workspace_name)


def get_model(identifier: Union[str, int]) -> Model:
    return _d().get_model(identifier)


def get_models(**kw: Any) -> List[Model]:
    return _d().get_models(**kw)


def whoami() -> Dict[str, Any]:
    return _d().whoami()


def logout() -> None:
    _d().logout()


def _delegate(name: str) -> Any:
    def fn(*a: Any, **kw: Any) -> Any:
        return getattr(_d(), name)(*a, **kw)
    fn.__name__ = name
    fn.__doc__ = f"``Determined.{name}`` of the default client (``client.login`` first, or the environment's master and user)."
    return fn


for _name in ("list_experiments", "list_models", "get_model_by_id", "get_model_labels", "create_user", "list_users", "get_user_by_id", "get_user_by_name",
              "get_session_username", "list_workspaces", "get_workspace", "create_workspace", "delete_workspace", "iter_trials_metrics",
              "stream_trials_metrics", "stream_trials_training_metrics", "stream_trials_validation_metrics", "add_oauth_client", "list_oauth_clients",
              "remove_oauth_client", "list_resource_pools", "preview_search", "get_master_info"):
    globals()[_name] = _delegate(_name)
del _name


ExperimentReference = Experiment          # names of earlier releases of the reference SDK
TrialReference = Trial
