"""``core.SearcherContext`` — pull searcher operations from the master.

Parity: reference ``harness/determined/core/_searcher.py:12-363``: ``Unit`` ``:12-15``,
``_parse_searcher_units`` ``:18-32``, ``SearcherOperation`` (absolute ``length``,
``report_progress``, ``report_completed`` exactly once by the chief) ``:35-111``,
``SearcherMode`` ``:114-135``, ``operations(searcher_mode, auto_ack)`` ``:209-256``,
``acknowledge_out_of_ops`` ``:258-270``, Dummy variants ``:297-363``.
"""
import enum
import logging
import math
from typing import Any, Iterator, Optional

from determined_b200 import util

logger = logging.getLogger("determined.core")


class Unit(enum.Enum):
    EPOCHS = "UNIT_EPOCHS"
    RECORDS = "UNIT_RECORDS"
    BATCHES = "UNIT_BATCHES"


def _parse_searcher_units(experiment_config: dict) -> Optional[Unit]:
    searcher = experiment_config.get("searcher", {}) or {}
    if searcher.get("name") == "custom":
        u = searcher.get("unit")
        return {"batches": Unit.BATCHES, "records": Unit.RECORDS, "epochs": Unit.EPOCHS}.get(u)
    length = searcher.get("max_length")
    if isinstance(length, dict) and len(length) == 1:
        key = next(iter(length))
        return {"batches": Unit.BATCHES, "records": Unit.RECORDS, "epochs": Unit.EPOCHS}.get(key)
    return None


class SearcherMode(enum.Enum):
    """``WorkersAskChief`` (default): all workers iterate ``operations()`` in step, the chief
    talks to the master and broadcasts.  ``ChiefOnly``: only the chief may iterate."""

    WorkersAskChief = "WORKERS_ASK_CHIEF"
    ChiefOnly = "CHEIF_ONLY"  # (sic) spelling is part of the reference's public enum value


class SearcherOperation:
    """"Train until your *total* training length is ``.length`` then report the searcher metric"."""

    def __init__(self, session: Any, trial_id: int, length: int, is_chief: bool) -> None:
        self._session = session
        self._trial_id = trial_id
        self._length = length
        self._is_chief = is_chief
        self._completed = False

    @property
    def length(self) -> int:
        """Absolute (cumulative) training length, in the searcher's configured unit."""
        return self._length

    def report_progress(self, length: float) -> None:
        if not self._is_chief:
            raise RuntimeError("you must only call op.report_progress() from the chief worker")
        if self._completed and length != self._length:
            raise RuntimeError("you must not call op.report_progress() after op.report_completed()")
        self._session.post(f"/api/v1/trials/{self._trial_id}/progress", data=util.json_encode(float(length)))

    def report_completed(self, searcher_metric: Any) -> None:
        if not self._is_chief:
            raise RuntimeError("you must only call op.report_completed() from the chief worker")
        if self._completed:
            raise RuntimeError("you may only call op.report_completed() once")
        if isinstance(searcher_metric, float) and math.isnan(searcher_metric):
            raise RuntimeError("searcher_metric may not be NaN")
        self._completed = True
        body = {"op": {"length": self._length}, "searcherMetric": searcher_metric}
        self._session.post(f"/api/v1/trials/{self._trial_id}/searcher/completed_operation", data=util.json_encode(body))


class SearcherContext:
    def __init__(self, session: Any, dist: Any, trial_id: int, run_id: int, allocation_id: str,
                 units: Optional[Unit] = None) -> None:
        self._session = session
        self._dist = dist
        self._trial_id = trial_id
        self._run_id = run_id
        self._allocation_id = allocation_id
        self._units = units

    def _get_searcher_op(self) -> Optional[SearcherOperation]:
        body = self._session.get(f"/api/v1/trials/{self._trial_id}/searcher/operation").json()
        if body.get("completed"):
            return None
        length = int(body["op"]["validateAfter"]["length"])   # uint64 arrives as a string
        return SearcherOperation(self._session, self._trial_id, length, self._dist.rank == 0)

    def operations(self, searcher_mode: SearcherMode = SearcherMode.WorkersAskChief,
                   auto_ack: bool = True) -> Iterator[SearcherOperation]:
        searcher_mode = SearcherMode(searcher_mode)
        if self._dist.rank == 0:
            while True:
                op = self._get_searcher_op()
                if searcher_mode == SearcherMode.WorkersAskChief:
                    self._dist.broadcast(op.length if op is not None else None)
                if op is None:
                    if auto_ack:
                        self.acknowledge_out_of_ops()
                    return
                yield op
                if not op._completed:
                    raise RuntimeError("you must call op.report_completed() on each operation")
        else:
            if searcher_mode != SearcherMode.WorkersAskChief:
                raise RuntimeError("you cannot call searcher.operations(searcher_mode=ChiefOnly) from a non-chief worker.")
            while True:
                n = self._dist.broadcast(None)
                if n is None:
                    return
                yield SearcherOperation(self._session, self._trial_id, n, False)

    def acknowledge_out_of_ops(self) -> None:
        """Tell the master we exit because there is nothing to do *right now* (it may restart us
        from the latest checkpoint if ASHA promotes this trial later)."""
        self._session.post(f"/api/v1/allocations/{self._allocation_id}/signals/ack_preemption")

    def get_configured_units(self) -> Optional[Unit]:
        return self._units


class DummySearcherOperation(SearcherOperation):
    def __init__(self, length: int, is_chief: bool) -> None:
        super().__init__(None, 0, length, is_chief)

    def report_progress(self, length: float) -> None:
        if not self._is_chief:
            raise RuntimeError("you must only call op.report_progress() from the chief worker")
        if self._completed and length != self._length:
            raise RuntimeError("you must not call op.report_progress() after op.report_completed()")
        logger.info("progress report: %s/%s", length, self._length)

    def report_completed(self, searcher_metric: Any) -> None:
        if not self._is_chief:
            raise RuntimeError("you must only call op.report_completed() from the chief worker")
        if self._completed:
            raise RuntimeError("you may only call op.report_completed() once")
        if isinstance(searcher_metric, float) and math.isnan(searcher_metric):
            raise RuntimeError("searcher_metric may not be NaN")
        self._completed = True
        logger.info("SearcherOperation Complete: searcher_metric=%s", util.json_encode(searcher_metric))


class DummySearcherContext(SearcherContext):
    """Yields a single operation of ``length`` (default 1)."""

    def __init__(self, dist: Any, length: int = 1) -> None:
        self._dist = dist
        self._length = length
        self._units: Optional[Unit] = None

    def operations(self, searcher_mode: SearcherMode = SearcherMode.WorkersAskChief,
                   auto_ack: bool = True) -> Iterator[SearcherOperation]:
        searcher_mode = SearcherMode(searcher_mode)
        if self._dist.rank > 0 and searcher_mode != SearcherMode.WorkersAskChief:
            raise RuntimeError("you cannot call searcher.operations(searcher_mode=ChiefOnly) from a non-chief worker.")
        op = DummySearcherOperation(self._length, self._dist.rank == 0)
        yield op
        if self._dist.rank == 0 and not op._completed:
            raise RuntimeError("you must call op.report_completed() on each operation")

    def acknowledge_out_of_ops(self) -> None:
        pass

    def get_configured_units(self) -> Optional[Unit]:
        return self._units
