"""Workloads: the legacy unit of work (RUN_STEP / COMPUTE_VALIDATION_METRICS / CHECKPOINT_MODEL)
that the workload sequencer feeds to old-style trial controllers.

Parity: reference ``harness/determined/workload.py:8-212`` (``Workload``, ``Source``,
``WorkloadResponseInterceptor``, ``train_workload``/``validation_workload``/``checkpoint_workload``).
"""
import enum
from typing import Any, Callable, Dict, Iterator, Optional, Tuple

Metrics = Dict[str, Any]
Response = Any
ResponseFunc = Callable[[Response], None]
Stream = Iterator[Tuple["Workload", ResponseFunc]]


class Workload:
    class Kind(enum.Enum):
        RUN_STEP = 1
        COMPUTE_VALIDATION_METRICS = 2
        CHECKPOINT_MODEL = 3

    def __init__(self, kind: "Workload.Kind", e_id: int, t_id: int, s_id: int, num_batches: int, total_batches_processed: int) -> None:
        self.kind = kind
        self.experiment_id, self.trial_id, self.step_id = e_id, t_id, s_id
        self.num_batches = num_batches
        self.total_batches_processed = total_batches_processed

    def __eq__(self, other: object) -> bool:
        return isinstance(other, Workload) and vars(self) == vars(other)

    def __repr__(self) -> str:
        extra = f" ({self.num_batches} Batches)" if self.kind == self.Kind.RUN_STEP else ""
        return f"<{self.kind.name}{extra}: ({self.experiment_id},{self.trial_id},{self.step_id})>"


class InvalidHP:
    """Response meaning "these hyperparameters cannot be trained"."""


class Skipped:
    """Response from non-chief workers."""


class Source:
    def __iter__(self) -> Stream:
        raise NotImplementedError


class WorkloadResponseInterceptor:
    """Feed one workload at a time into a controller-style stream and capture its response."""

    def __init__(self) -> None:
        self._response: Optional[Response] = None

    def _respond(self, resp: Response) -> None:
        self._response = resp

    def send(self, wkld: Workload) -> Stream:
        self._response = None
        yield wkld, self._respond

    def metrics_result(self) -> Metrics:
        if self._response is None:
            raise ValueError("no response was received for the workload")
        if not isinstance(self._response, dict):
            raise ValueError(f"response is not metrics: {self._response!r}")
        return self._response


def ignore_workload_response(*_: Any) -> None:
    pass


def train_workload(step_id: int, exp_id: int = 1, trial_id: int = 1, num_batches: int = 1, total_batches_processed: int = 0) -> Workload:
    return Workload(Workload.Kind.RUN_STEP, exp_id, trial_id, step_id, num_batches, total_batches_processed)


def validation_workload(step_id: int = 1, exp_id: int = 1, trial_id: int = 1, total_batches_processed: int = 0) -> Workload:
    return Workload(Workload.Kind.COMPUTE_VALIDATION_METRICS, exp_id, trial_id, step_id, 0, total_batches_processed)


def checkpoint_workload(step_id: int = 1, exp_id: int = 1, trial_id: int = 1, total_batches_processed: int = 0) -> Workload:
    return Workload(Workload.Kind.CHECKPOINT_MODEL, exp_id, trial_id, step_id, 0, total_batches_processed)
