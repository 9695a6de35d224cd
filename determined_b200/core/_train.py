"""``core.TrainContext`` — metrics, status, early-exit reporting.

Parity: reference ``harness/determined/core/_train.py:18-297``: ``EarlyExitReason`` ``:18-21``,
``set_status`` ``:48-56``, ``report_training_metrics``/``report_validation_metrics``/
``report_metrics(group,…)`` ``:111-167``, ``upload_tensorboard_files`` ``:177-196``,
``report_early_exit`` ``:223-239``, ``get_experiment_best_validation`` ``:241-258``, Dummy ``:261-297``.
"""
import datetime
import enum
import logging
import pathlib
from typing import Any, Callable, Dict, List, Optional, Union

from determined_b200 import util
from determined_b200.common.api import errors as api_errors
from determined_b200.core._tensorboard_mode import TensorboardMode

logger = logging.getLogger("determined.core")


class EarlyExitReason(enum.Enum):
    INVALID_HP = "EXITED_REASON_INVALID_HP"
    # USER_REQUESTED_STOP is deprecated in the reference but still accepted on the wire.
    USER_REQUESTED_STOP = "EXITED_REASON_USER_REQUESTED_STOP"


class TrainContext:
    def __init__(self, session: Any, trial_id: int, run_id: int, exp_id: int, distributed: Any,
                 tensorboard_mode: TensorboardMode, tensorboard_manager: Any = None,
                 tbd_writer: Any = None) -> None:
        self._session = session
        self._trial_id = trial_id
        self._run_id = run_id
        self._exp_id = exp_id
        self._distributed = distributed
        self._tensorboard_mode = tensorboard_mode
        self._tensorboard_manager = tensorboard_manager
        self._tbd_writer = tbd_writer

    def set_status(self, status: str) -> None:
        """``"training" | "validating" | "checkpointing"`` (free-form string on the wire)."""
        self._session.post(f"/api/v1/trials/{self._trial_id}/runner/metadata", json={"state": status})

    def _get_last_validation(self) -> Optional[int]:
        r = self._session.get(f"/api/v1/trials/{self._trial_id}").json()
        lv = (r.get("trial") or {}).get("latestValidation") or {}
        tb = lv.get("totalBatches")
        return None if tb is None else int(tb)

    def report_metrics(self, group: str, steps_completed: int, metrics: Dict[str, Any],
                       batch_metrics: Optional[List[Dict[str, Any]]] = None,
                       report_time: Optional[datetime.datetime] = None) -> None:
        """Report a dict of metrics under an arbitrary ``group`` (``training``, ``validation``,
        or custom e.g. ``inference``).  ``metrics`` must be JSON-serialisable."""
        body = {
            "group": group,
            "metrics": {
                "trialId": self._trial_id, "trialRunId": self._run_id, "stepsCompleted": steps_completed,
                "reportTime": (report_time or datetime.datetime.now(datetime.timezone.utc)).isoformat(),
                "metrics": {"avgMetrics": metrics, **({"batchMetrics": batch_metrics} if batch_metrics else {})},
            },
        }
        self._session.post(f"/api/v1/trials/{self._trial_id}/metrics", data=util.json_encode(body))
        if self._tbd_writer is not None and group in ("training", "validation"):
            w = self._tbd_writer
            if hasattr(w, "on_train_step_end"):          # tensorboard.metric_writers.BatchMetricWriter
                if group == "training":
                    w.on_train_step_end(steps_completed, metrics, batch_metrics)
                else:
                    w.on_validation_step_end(steps_completed, metrics)
            else:                                        # a bare SummaryWriter
                prefix = "Determined/" if group == "training" else "Determined/val_"
                for k, v in metrics.items():
                    if util.is_numerical_scalar(v):
                        w.add_scalar(prefix + k, float(v), steps_completed)
        if self._tensorboard_mode == TensorboardMode.AUTO and self._tensorboard_manager is not None:
            self._tensorboard_manager.sync()

    def report_training_metrics(self, steps_completed: int, metrics: Dict[str, Any],
                                batch_metrics: Optional[List[Dict[str, Any]]] = None) -> None:
        logger.info("report_training_metrics(steps_completed=%s, metrics=%s)", steps_completed, metrics)
        self.report_metrics("training", steps_completed, metrics, batch_metrics)

    def report_validation_metrics(self, steps_completed: int, metrics: Dict[str, Any]) -> None:
        logger.info("report_validation_metrics(steps_completed=%s, metrics=%s)", steps_completed, metrics)
        self.report_metrics("validation", steps_completed, metrics)

    def get_tensorboard_path(self) -> pathlib.Path:
        if self._tensorboard_manager is None:
            raise RuntimeError("tensorboard is not configured for this context")
        return self._tensorboard_manager.base_path

    def upload_tensorboard_files(self, selector: Callable[[pathlib.Path], bool] = lambda _: True,
                                 mangler: Callable[[pathlib.Path, int], pathlib.Path] = lambda p, __: p) -> None:
        if self._tensorboard_mode == TensorboardMode.AUTO:
            raise RuntimeError("upload_tensorboard_files can only be used with TensorboardMode.MANUAL")
        if self._tensorboard_manager is not None:
            self._tensorboard_manager.sync(selector, mangler, self._distributed.rank)

    def report_early_exit(self, reason: EarlyExitReason) -> None:
        """Tell the searcher this trial ended early (InvalidHP → ASHA replaces it)."""
        if self._distributed.rank != 0:
            raise RuntimeError("report_early_exit() may only be called by the chief")
        reason = EarlyExitReason(reason)
        try:
            self._session.post(f"/api/v1/trials/{self._trial_id}/early_exit", json={"reason": reason.value})
        except api_errors.APIException as e:
            if e.status_code != 400:   # already reported
                raise

    def get_experiment_best_validation(self) -> Optional[float]:
        try:
            r = self._session.get(f"/api/v1/experiments/{self._exp_id}/searcher/best_searcher_validation_metric")
        except api_errors.NotFoundException:
            return None
        return float(r.json()["metric"])


class DummyTrainContext(TrainContext):
    def __init__(self) -> None:
        self._metrics_log: List[Dict[str, Any]] = []

    def set_status(self, status: str) -> None:
        logger.info("status: %s", status)

    def _get_last_validation(self) -> Optional[int]:
        return None

    def report_metrics(self, group: str, steps_completed: int, metrics: Dict[str, Any],
                       batch_metrics: Optional[List[Dict[str, Any]]] = None,
                       report_time: Optional[datetime.datetime] = None) -> None:
        self._metrics_log.append({"group": group, "steps_completed": steps_completed, "metrics": metrics})
        logger.info("report_metrics(group=%s, steps_completed=%s, metrics=%s)", group, steps_completed, metrics)

    def report_training_metrics(self, steps_completed: int, metrics: Dict[str, Any],
                                batch_metrics: Optional[List[Dict[str, Any]]] = None) -> None:
        self.report_metrics("training", steps_completed, metrics, batch_metrics)

    def report_validation_metrics(self, steps_completed: int, metrics: Dict[str, Any]) -> None:
        self.report_metrics("validation", steps_completed, metrics)

    def upload_tensorboard_files(self, selector: Any = None, mangler: Any = None) -> None:
        logger.info("upload_tensorboard_files()")

    def get_tensorboard_path(self) -> pathlib.Path:
        return pathlib.Path("/tmp/tensorboard")

    def report_early_exit(self, reason: EarlyExitReason) -> None:
        logger.info("report_early_exit(%s)", reason)

    def get_experiment_best_validation(self) -> Optional[float]:
        return None
