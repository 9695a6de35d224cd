"""Legacy ``WorkloadSequencer``: fuses searcher operations and preemption into one stream of
RUN_STEP / COMPUTE_VALIDATION_METRICS / CHECKPOINT_MODEL workloads for controllers written against
the old workload protocol.

Parity: reference ``harness/determined/layers/_workload_sequencer.py``: ``SavableState`` +
period precomputation ``:55-170``, ``train``/``validate``/``checkpoint`` generators with response
handling (metrics report, searcher completion, ``checkpoint_policy`` best/all/none, InvalidHP)
``:172-360``, the op loop with ``scheduling_unit``-sized steps, ``min_validation_period`` /
``min_checkpoint_period``, initial validation and preemption checks ``:362-438``; chief→worker
broadcast of workloads ``make_compatibility_workloads`` ``:441-482``.  The PyTorch and DeepSpeed
controllers of this build do not need it (they pull searcher ops directly); it is kept for users'
own controllers and for checkpoint compatibility (``workload_sequencer.pkl``).
"""
import logging
import pickle
from typing import Any, Dict, Generator, Iterator, Optional, Tuple

import determined_b200 as det
from determined_b200 import core, errors, workload

logger = logging.getLogger("determined.layers")
WorkloadGenerator = Generator[Tuple[workload.Workload, workload.ResponseFunc], None, None]


class ShouldExit(Exception):
    pass


class WorkloadSequencer(workload.Source):
    class SavableState:
        def __init__(self, trial_id: int, last_ckpt: int = 0, steps_completed: int = 0, step_id: int = 0, last_val: int = 0) -> None:
            self.trial_id, self.last_ckpt, self.steps_completed, self.step_id, self.last_val = trial_id, last_ckpt, steps_completed, step_id, last_val

    def __init__(self, core_context: core.Context, exp_conf: Dict[str, Any], trial_id: int, exp_id: int, global_batch_size: int,
                 steps_completed: int = 0) -> None:
        self.core_context = core_context
        self.cfg = det.ExperimentConfig(exp_conf)
        self._trial_id, self._exp_id = trial_id, exp_id
        unit = core_context.searcher.get_configured_units()
        if unit is None:
            logger.warning("searcher configured without units; assuming epochs")
            unit = core.Unit.EPOCHS
        self._unit = unit
        self.global_batch_size = global_batch_size
        self.records_per_epoch = self.cfg.get_records_per_epoch()
        self.val_from_previous_run = core_context.train._get_last_validation()
        self.want_initial_val = self.cfg.get_perform_initial_validation()
        self.ckpt_policy = self.cfg.get_checkpoint_policy()
        self.scheduling_unit = self.cfg.scheduling_unit()
        self.state = self.SavableState(trial_id=trial_id, steps_completed=steps_completed)
        for name, period in (("min_checkpoint_period", self.cfg.get_min_checkpoint_period()),
                             ("min_validation_period", self.cfg.get_min_validation_period())):
            if period.get("epochs") is not None and self.records_per_epoch is None:
                raise ValueError(f"records_per_epoch must be specified if {name} is configured in epoch units.")
        self.min_val_period_batches = self.as_batches(**self.cfg.get_min_validation_period())
        self.min_ckpt_period_batches = self.as_batches(**self.cfg.get_min_checkpoint_period())
        if self.min_val_period_batches < 1:
            self.min_val_period_batches = None  # type: ignore
        if self.min_ckpt_period_batches < 1:
            self.min_ckpt_period_batches = None  # type: ignore

    # -- unit conversion -------------------------------------------------------------------------------
    def as_batches(self, batches: Optional[int] = None, records: Optional[int] = None, epochs: Optional[int] = None) -> int:
        if batches is None and records is None and epochs is None:
            return 0
        if sum(x is not None for x in (batches, records, epochs)) != 1:
            raise ValueError(f"invalid length: batches={batches} records={records} epochs={epochs}")
        if batches is not None:
            return batches
        if records is not None:
            return max(records // self.global_batch_size, 1) if records > 0 else 0
        assert epochs is not None
        if self.records_per_epoch is None:
            raise ValueError("records_per_epoch is required to use epoch units")
        return max((epochs * self.records_per_epoch) // self.global_batch_size, 1) if epochs > 0 else 0

    def batches_until_op_complete(self, op: core.SearcherOperation) -> int:
        kw = {core.Unit.BATCHES: "batches", core.Unit.RECORDS: "records", core.Unit.EPOCHS: "epochs"}[self._unit]
        return self.as_batches(**{kw: op.length}) - self.state.steps_completed

    def batches_until_val(self) -> int:
        if self.min_val_period_batches is None:
            return 1 << 62
        return self.state.last_val + self.min_val_period_batches - self.state.steps_completed

    def batches_until_ckpt(self) -> int:
        if self.min_ckpt_period_batches is None:
            return 1 << 62
        return self.state.last_ckpt + self.min_ckpt_period_batches - self.state.steps_completed

    # -- state -----------------------------------------------------------------------------------------
    def get_state(self) -> Dict[str, Any]:
        return vars(self.state).copy()

    def load_state(self, state: Dict[str, Any]) -> None:
        if state.get("trial_id") != self._trial_id:
            self.state = self.SavableState(trial_id=self._trial_id)
            return
        self.state = self.SavableState(**{k: state[k] for k in ("trial_id", "last_ckpt", "steps_completed", "step_id", "last_val") if k in state})
        if self.state.steps_completed == self.val_from_previous_run:
            self.state.last_val = self.state.steps_completed

    # -- workload generators -----------------------------------------------------------------------------
    def train(self, num_batches: int, op: core.SearcherOperation) -> WorkloadGenerator:
        self.state.step_id += 1
        wkld = workload.train_workload(self.state.step_id, self._exp_id, self._trial_id, num_batches, self.state.steps_completed)
        box: Dict[str, Any] = {}
        yield wkld, lambda r: box.setdefault("r", r)
        resp = box.get("r")
        if isinstance(resp, workload.InvalidHP):
            self.core_context.train.report_early_exit(core.EarlyExitReason.INVALID_HP)
            raise ShouldExit()
        self.state.steps_completed += num_batches
        if isinstance(resp, dict) and resp:
            m = resp.get("metrics", resp)
            self.core_context.train.report_training_metrics(self.state.steps_completed, m.get("avg_metrics", {}), m.get("batch_metrics"))
            if resp.get("stop_requested"):
                raise ShouldExit()
        unit_done = {core.Unit.BATCHES: self.state.steps_completed, core.Unit.RECORDS: self.state.steps_completed * self.global_batch_size,
                     core.Unit.EPOCHS: (self.state.steps_completed * self.global_batch_size) / (self.records_per_epoch or 1)}[self._unit]
        if not op._completed:
            op.report_progress(unit_done)
        self.check_for_preemption()

    def is_best_validation(self, now: float, before: Optional[float]) -> bool:
        if before is None:
            return True
        return now < before if self.cfg.get_smaller_is_better() else now > before

    def validate(self, op: Optional[core.SearcherOperation]) -> WorkloadGenerator:
        wkld = workload.validation_workload(self.state.step_id, self._exp_id, self._trial_id, self.state.steps_completed)
        box: Dict[str, Any] = {}
        yield wkld, lambda r: box.setdefault("r", r)
        resp = box.get("r")
        if isinstance(resp, workload.InvalidHP):
            self.core_context.train.report_early_exit(core.EarlyExitReason.INVALID_HP)
            raise ShouldExit()
        metrics = (resp or {}).get("metrics", {}).get("validation_metrics", (resp or {}).get("validation_metrics", {}))
        self.state.last_val = self.state.steps_completed
        name = self.cfg.get("searcher", {}).get("metric")
        best_before = self.core_context.train.get_experiment_best_validation()
        self.core_context.train.report_validation_metrics(self.state.steps_completed, metrics)
        sm = metrics.get(name) if name else None
        if name and sm is None:
            raise RuntimeError(f"Search method is configured to use metric '{name}' but model definition returned validation "
                               f"metrics {list(metrics.keys())}.")
        if op is not None and self.batches_until_op_complete(op) < 1 and not op._completed:
            op.report_completed(sm)
        if self.ckpt_policy == "all" or (self.ckpt_policy == "best" and sm is not None and self.is_best_validation(float(sm), best_before)):
            if self.state.last_ckpt != self.state.steps_completed:
                yield from self.checkpoint(already_exiting=False)
        if (resp or {}).get("stop_requested"):
            raise ShouldExit()
        self.check_for_preemption()

    def checkpoint(self, already_exiting: bool) -> WorkloadGenerator:
        self.state.last_ckpt = self.state.steps_completed
        wkld = workload.checkpoint_workload(self.state.step_id, self._exp_id, self._trial_id, self.state.steps_completed)
        box: Dict[str, Any] = {}
        yield wkld, lambda r: box.setdefault("r", r)
        if isinstance(box.get("r"), workload.InvalidHP):
            self.core_context.train.report_early_exit(core.EarlyExitReason.INVALID_HP)
            if not already_exiting:
                raise ShouldExit()
            return
        if not already_exiting:
            self.check_for_preemption()

    def check_for_preemption(self) -> None:
        if self.core_context.preempt.should_preempt():
            raise ShouldExit()

    def __iter__(self) -> workload.Stream:
        try:
            if self.want_initial_val and self.val_from_previous_run is None and self.state.steps_completed == 0:
                yield from self.validate(None)
            for op in self.core_context.searcher.operations(core.SearcherMode.ChiefOnly):
                while self.batches_until_op_complete(op) > 0:
                    n = min(self.batches_until_ckpt(), self.batches_until_val(), self.batches_until_op_complete(op), self.scheduling_unit)
                    yield from self.train(max(n, 1), op)
                    if self.batches_until_val() < 1:
                        yield from self.validate(op)
                    if self.batches_until_ckpt() < 1:
                        yield from self.checkpoint(already_exiting=False)
                if self.state.last_val != self.state.steps_completed or not op._completed:
                    yield from self.validate(op)
                if self.state.last_ckpt != self.state.steps_completed:
                    yield from self.checkpoint(already_exiting=False)
        except ShouldExit:
            if self.state.last_ckpt != self.state.steps_completed:
                yield from self.checkpoint(already_exiting=True)


def make_compatibility_workloads(core_context: core.Context, exp_conf: Dict[str, Any], trial_id: int, exp_id: int,
                                 global_batch_size: int, steps_completed: int = 0) -> Tuple[workload.Stream, Optional[WorkloadSequencer]]:
    """Chief runs the sequencer and broadcasts each workload; workers replay the same stream and
    their responses are discarded."""
    dist = core_context.distributed
    if dist.rank == 0:
        seq = WorkloadSequencer(core_context, exp_conf, trial_id, exp_id, global_batch_size, steps_completed)

        def chief() -> workload.Stream:
            for wkld, fn in seq:
                dist.broadcast(wkld)
                yield wkld, fn
            dist.broadcast(None)

        return chief(), seq

    def worker() -> workload.Stream:
        while True:
            wkld = dist.broadcast(None)
            if wkld is None:
                return
            yield wkld, workload.ignore_workload_response

    return worker(), None
