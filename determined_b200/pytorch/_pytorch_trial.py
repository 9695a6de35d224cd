"""``PyTorchTrial`` and its controller (the training loop).

Parity: reference ``harness/determined/pytorch/_pytorch_trial.py``: ``TrainUnit/Batch/Epoch``
``:43-137``, ``_TrialState`` ``:163-179``, controller validation of arguments ``:182-280``,
``pre_execute_hook`` ``:282-319``, data-loader set-up ``:435-494``, ``run`` ``:548-615``,
searcher-op loop ``:617-679``, hot loop ``_train_with_boundaries`` ``:681-734``, boundary handling
``_train_for_op`` ``:736-804``, LR auto-step ``:829-854``, ``_train_batch`` ``:861-913``,
``_validate`` ``:915-1090``, ``_load`` (+legacy layouts) ``:1092-1285``, ``_save`` ``:1287-1360``,
``PyTorchTrial`` abstract API ``:1391-1648``.  The checkpoint directory layout is byte-for-byte
the reference's contract.

B200-first changes in the hot loop:
  * batches arrive through a pinned-memory prefetcher on a copy stream (K15);
  * per-batch metrics stay on the device and are fetched once per reporting boundary (K16) — the
    reference does ``metric.cpu().detach().numpy()`` (a device sync) on every batch ``:901-907``;
  * optional CUDA-graph replay of the whole ``train_batch`` (forward, backward, overlapped fused
    all-reduce, fused optimizer) — one graph per aggregation phase;
  * step timing uses CUDA events, never a host clock around a sync.
"""
import abc
import collections.abc as col_abc
import contextlib
import copy
import enum
import json
import logging
import pathlib
import pickle
import random
import sys
import time
from typing import Any, Callable, Dict, Iterator, List, Optional, Tuple, Type, Union

import numpy as np
import torch
import torch.distributed as dist

import determined_b200 as det
from determined_b200 import core, errors, tensorboard, util
from determined_b200.errors import ShouldExit
from determined_b200.pytorch import _callback, _data, _lr_scheduler, _metric_utils, _reducer
from determined_b200.pytorch._pytorch_context import PyTorchTrialContext

logger = logging.getLogger("determined.pytorch")


class TrainUnit:
    """A training length or period.  An ``int`` value is a period (``Batch(100)``: every 100
    batches); a container value is a schedule (``Batch([1, 5, 10])``)."""

    def __init__(self, value: Union[int, col_abc.Container]):
        self.value = value

    @staticmethod
    def _from_searcher_unit(length: int, unit: Optional[core.Unit], global_batch_size: Optional[int] = None) -> "TrainUnit":
        if unit == core.Unit.EPOCHS:
            return Epoch(length)
        if unit == core.Unit.RECORDS:
            if global_batch_size is None:
                raise ValueError("global_batch_size required for searcher unit Records.")
            return Batch._from_records(length, global_batch_size)
        if unit == core.Unit.BATCHES:
            return Batch(length)
        raise ValueError(f"unrecognized searcher unit {unit}")

    def _to_searcher_unit(self) -> core.Unit:
        return core.Unit.BATCHES if isinstance(self, Batch) else core.Unit.EPOCHS

    @staticmethod
    def _from_values(batches: Optional[int] = None, records: Optional[int] = None, epochs: Optional[int] = None,
                     global_batch_size: Optional[int] = None) -> "TrainUnit":
        given = [(k, v) for k, v in (("batches", batches), ("records", records), ("epochs", epochs)) if v is not None]
        if len(given) != 1:
            raise ValueError(f"invalid config: batches={batches} records={records} epochs={epochs}")
        kind, v = given[0]
        if v < 1:
            v = sys.maxsize    # "0" means never
        if kind == "batches":
            return Batch(v)
        if kind == "epochs":
            return Epoch(v)
        if not global_batch_size:
            raise ValueError("global_batch_size is required for RECORD units.")
        return Batch._from_records(v, global_batch_size)

    def should_stop(self, step_num: int) -> bool:
        if isinstance(self.value, int):
            return self._divides(step_num)
        return step_num in self.value

    def _divides(self, steps: int) -> bool:
        assert isinstance(self.value, int)
        if self.value < 1:
            return True
        return steps != 0 and steps % self.value == 0

    def __repr__(self) -> str:
        return f"{type(self).__name__}({self.value!r})"


class Epoch(TrainUnit):
    """Epoch(1) = one pass over the (sharded) training loader."""


class Batch(TrainUnit):
    """Batch(1) = one call to ``train_batch``."""

    @staticmethod
    def _from_records(records: int, global_batch_size: int) -> "Batch":
        return Batch(max(records // global_batch_size, 1))


class _Boundary(enum.Enum):
    TRAIN = "TRAIN"
    VALIDATE = "VALIDATE"
    CHECKPOINT = "CHECKPOINT"
    REPORT = "REPORT"


class _TrialState:
    """Pickled as ``trial_state.pkl`` (``vars()`` of this object) — field names are a contract."""

    def __init__(self, trial_id: int = 0, last_ckpt: int = 0, step_id: int = 0, last_val: int = 0,
                 batches_trained: int = 0, epochs_trained: int = 0) -> None:
        self.trial_id = trial_id
        self.last_ckpt = last_ckpt
        self.step_id = step_id
        self.last_val = last_val
        self.batches_trained = batches_trained
        self.epochs_trained = epochs_trained


class _GraphedStep:
    """CUDA-graph capture/replay of ``trial.train_batch`` for one aggregation phase."""

    def __init__(self) -> None:
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.static_in: Any = None
        self.static_out: Any = None
        self.launches = 0     # native kernels recorded in the graph (for launch accounting)


def _copy_into(dst: Any, src: Any) -> None:
    if isinstance(dst, torch.Tensor):
        dst.copy_(src, non_blocking=True)
    elif isinstance(dst, dict):
        for k in dst:
            _copy_into(dst[k], src[k])
    elif isinstance(dst, (list, tuple)):
        for d, s in zip(dst, src):
            _copy_into(d, s)


def _clone_tree(x: Any) -> Any:
    if isinstance(x, torch.Tensor):
        return x.clone()
    if isinstance(x, dict):
        return {k: _clone_tree(v) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_clone_tree(v) for v in x)
    return x


def _graph_outputs(out: Dict[str, Any]) -> Dict[str, Any]:
    """Static graph outputs are overwritten by the next replay.  Scalars are copied to pinned
    memory right after this returns (stream-ordered before the next replay), so they may alias;
    anything bigger is cloned."""
    return {k: (v.clone() if isinstance(v, torch.Tensor) and v.numel() != 1 else v) for k, v in out.items()}


def _same_shapes(a: Any, b: Any) -> bool:
    if isinstance(a, torch.Tensor):
        return isinstance(b, torch.Tensor) and a.shape == b.shape and a.dtype == b.dtype
    if isinstance(a, dict):
        return isinstance(b, dict) and a.keys() == b.keys() and all(_same_shapes(a[k], b[k]) for k in a)
    if isinstance(a, (list, tuple)):
        return isinstance(b, (list, tuple)) and len(a) == len(b) and all(_same_shapes(x, y) for x, y in zip(a, b))
    return True


class _PyTorchTrialController:
    def __init__(self, trial_inst: "PyTorchTrial", context: PyTorchTrialContext,
                 checkpoint_period: TrainUnit, validation_period: TrainUnit, reporting_period: TrainUnit,
                 smaller_is_better: bool, steps_completed: int, latest_checkpoint: Optional[str],
                 local_training: bool, test_mode: bool, searcher_metric_name: Optional[str],
                 checkpoint_policy: str, step_zero_validation: bool, max_length: Optional[TrainUnit],
                 global_batch_size: Optional[int], profiling_enabled: Optional[bool]) -> None:
        if not isinstance(trial_inst, PyTorchTrial):
            raise TypeError("PyTorchTrialController needs a PyTorchTrial")
        self.trial = trial_inst
        self.context = context
        self.core_context = context._core
        self.is_chief = context.distributed.rank == 0
        self.callbacks = self.trial.build_callbacks()
        for name, cb in self.callbacks.items():
            if not isinstance(cb, _callback.PyTorchCallback):
                raise errors.InvalidExperimentException(f"callback {name!r} must be a PyTorchCallback")
        self.checkpoint_period = checkpoint_period
        self.validation_period = validation_period
        self.reporting_period = reporting_period
        self.smaller_is_better = smaller_is_better
        self.local_training = local_training
        self.test_mode = test_mode
        self.searcher_metric_name = searcher_metric_name
        self.ckpt_policy = checkpoint_policy
        self.step_zero_validation = step_zero_validation
        self.max_length = max_length
        self.global_batch_size = global_batch_size
        self.profiling_enabled = profiling_enabled
        self.latest_checkpoint = latest_checkpoint
        self.steps_completed = steps_completed
        self.trial_id = 0 if local_training or context._core.info is None else context.get_trial_id()
        self.state: Optional[_TrialState] = None
        self.start_from_batch = steps_completed
        self.val_from_previous_run = self.core_context.train._get_last_validation()
        self.searcher_unit = None if local_training else self.core_context.searcher.get_configured_units()
        if local_training:
            self.searcher_unit = max_length._to_searcher_unit() if max_length is not None else core.Unit.BATCHES
        self.training_loader: Any = None
        self.validation_loader: Any = None
        self.training_enumerator: Optional[Iterator] = None
        self._graphs: Dict[bool, _GraphedStep] = {}
        self._graph_warmup_left = 3
        self._graphs_failed = False
        self._pending_upload_cbs: List[str] = []
        self.step_timer = core.StepTimer()
        self._metric_writer: Any = None
        self._batch_hooks = [cb for cb in self.callbacks.values()
                             if util.is_overridden(cb.on_training_batch_start, _callback.PyTorchCallback)
                             or util.is_overridden(cb.on_training_batch_end, _callback.PyTorchCallback)]
        # pinned ring for scalar metrics: every batch's scalars are copied D2H asynchronously as soon
        # as they are produced; the host only waits once per reporting window
        self._pin_ring: Optional[torch.Tensor] = None
        self._pin_used = 0
        self._pin_event: Any = None

        if not context.models:
            raise errors.InvalidExperimentException(
                "Must have at least one model. This might be caused by not wrapping your model with wrap_model()")
        if not context.optimizers:
            raise errors.InvalidExperimentException(
                "Must have at least one optimizer. This might be caused by not wrapping your optimizer with wrap_optimizer()")
        self._check_evaluate_implementation()

    # ------------------------------------------------------------------ set-up helpers
    @classmethod
    def pre_execute_hook(cls, trial_seed: int, distributed_backend: Any) -> None:
        """Process-level set-up before the trial class is constructed: process group + seeds."""
        if distributed_backend is not None and distributed_backend.use_torch() and not dist.is_initialized():
            dist.init_process_group(backend="nccl" if torch.cuda.is_available() else "gloo")
        cls._set_random_seeds(trial_seed)

    @staticmethod
    def _set_random_seeds(seed: int) -> None:
        random.seed(seed)
        np.random.seed(seed % (2**32))
        torch.random.manual_seed(seed)

    def _check_evaluate_implementation(self) -> None:
        eb, efd = self._evaluate_batch_defined(), self._evaluate_full_dataset_defined()
        if eb == efd:
            raise errors.InvalidExperimentException(
                "Please define exactly one of: `evaluate_batch()` or `evaluate_full_dataset()`. "
                "For most use cases `evaluate_batch()` is recommended because it can be parallelized across all devices.")

    def _evaluate_batch_defined(self) -> bool:
        return util.is_overridden(self.trial.evaluate_batch, PyTorchTrial)

    def _evaluate_full_dataset_defined(self) -> bool:
        return util.is_overridden(self.trial.evaluate_full_dataset, PyTorchTrial)

    def _set_data_loaders(self) -> None:
        skip = self.start_from_batch
        nreplicas, rank = self.context.distributed.size, self.context.distributed.rank
        train = self.trial.build_training_data_loader()
        if isinstance(train, _data.DataLoader):
            self.training_loader = train.get_data_loader(repeat=True, skip=skip, num_replicas=nreplicas, rank=rank)
            self.context._epoch_len = len(_data.adapt_batch_sampler(train.batch_sampler, num_replicas=nreplicas, rank=rank))
        else:
            if not self.context.experimental._data_repro_checks_disabled:
                raise RuntimeError(
                    "build_training_data_loader() must return a determined.pytorch.DataLoader (which shards, repeats "
                    "and skips reproducibly); you can disable this check by calling "
                    "context.experimental.disable_dataset_reproducibility_checks() to use a plain torch DataLoader and "
                    "own those properties yourself")
            self.training_loader = train
            self.context._epoch_len = len(train)
        if not self.context._epoch_len:
            raise RuntimeError("training data loader is empty for this worker")
        val = self.trial.build_validation_data_loader()
        if self._evaluate_batch_defined():
            if isinstance(val, _data.DataLoader):
                self.validation_loader = val.get_data_loader(repeat=False, skip=0, num_replicas=nreplicas, rank=rank)
            else:
                if not self.context.experimental._data_repro_checks_disabled:
                    raise RuntimeError("build_validation_data_loader() must return a determined.pytorch.DataLoader")
                self.validation_loader = val
        elif self.is_chief:
            self.validation_loader = (val.get_data_loader(repeat=False, skip=0, num_replicas=1, rank=0)
                                      if isinstance(val, _data.DataLoader) else val)

    _restored_rng: Optional[Dict[str, Any]] = None

    def _apply_rng_state(self, rng: Dict[str, Any]) -> None:
        ctx = self.context
        np.random.set_state(rng["np_rng_state"])
        random.setstate(rng["random_rng_state"])
        torch.random.set_rng_state(rng["cpu_rng_state"])
        if torch.cuda.is_available() and ctx.device.type == "cuda":
            if "gpu_rng_state" in rng:
                torch.cuda.set_rng_state(rng["gpu_rng_state"], device=ctx.device)
            else:
                logger.warning("The system has a gpu but no gpu_rng_state exists in the checkpoint.")
        elif "gpu_rng_state" in rng:
            logger.warning("There exists gpu_rng_state in checkpoint but the system has no gpu.")

    def _make_enumerator(self) -> Iterator:
        loader = self.training_loader
        if _repeats(loader):
            it: Iterator = iter(loader)
        else:
            # a plain loader ends after one epoch: chain epochs lazily
            def forever() -> Iterator:
                while True:
                    n = 0
                    for b in loader:
                        n += 1
                        yield b
                    if n == 0:
                        raise RuntimeError("training loader yielded no batches")
            it = forever()
        if self.context.experimental._auto_to_device:
            it = _data.DevicePrefetcher(it, self.context.device)
        return enumerate(it, start=self.start_from_batch)

    # ------------------------------------------------------------------ top level
    def run(self) -> None:
        with contextlib.ExitStack() as stack:
            for cb in self.callbacks.values():
                stack.callback(cb.on_trial_shutdown)
            self._set_data_loaders()
            if self.latest_checkpoint is not None:
                logger.info("Restoring trial from checkpoint %s", self.latest_checkpoint)
                with self.core_context.checkpoint.restore_path(self.latest_checkpoint) as load_path:
                    self._load(pathlib.Path(load_path))
            else:
                self.state = _TrialState(trial_id=self.trial_id)
            assert self.state is not None
            self.start_from_batch = self.state.batches_trained if self.latest_checkpoint else self.steps_completed
            if self.latest_checkpoint is not None and self.state.batches_trained != self.steps_completed:
                # data loader was positioned with steps_completed; rebuild if the checkpoint disagrees
                self.start_from_batch = self.state.batches_trained
                self._set_data_loaders()
            self.state.batches_trained = self.start_from_batch
            self.training_enumerator = self._make_enumerator()
            if self._restored_rng is not None:
                # building the loader iterator draws from the torch generator (worker base seed, shuffling); the
                # reference builds it before it restores the checkpoint's random state (``_pytorch_trial.py:572-579``), so
                # that the first batch after a restart sees exactly the state that was saved: same effect here
                self._apply_rng_state(self._restored_rng)
                self._restored_rng = None
            if self.context.distributed.size > 1:
                self.context._get_engine().sync_params_from_chief(self.context.models)
            for cb in self.callbacks.values():
                cb.on_trial_startup(self.start_from_batch, self.latest_checkpoint)
            if self.context.profiler is not None:
                stack.enter_context(self.context.profiler)
            if self.profiling_enabled:
                self.core_context.profiler.on()
                stack.callback(self.core_context.profiler.off)
            for cb in self.callbacks.values():
                cb.on_training_start()
            stack.callback(self._close_checkpoint_writer)      # runs before on_trial_shutdown: nothing is left in flight
            self._run()

    def _run(self) -> None:
        assert self.state is not None
        try:
            if self.step_zero_validation and self.val_from_previous_run is None and self.state.batches_trained == 0:
                self._validate()
            if self.local_training:
                assert self.max_length is not None and isinstance(self.max_length.value, int)
                ops: Iterator = iter([core.DummySearcherOperation(self.max_length.value, self.is_chief)])
            else:
                ops = self.core_context.searcher.operations()
            for op in ops:
                self._train_for_op(op)
        except ShouldExit as e:
            if not e.skip_exit_checkpoint and not self._checkpoint_is_current():
                self._checkpoint(already_exiting=True)
        except errors.InvalidHP:
            if not self._checkpoint_is_current():
                self._checkpoint(already_exiting=True)
            raise

    # ------------------------------------------------------------------ op / boundary logic
    def _op_length(self, op: core.SearcherOperation) -> TrainUnit:
        if self.test_mode:
            return Batch(1)
        if self.local_training:
            assert self.max_length is not None
            return self.max_length
        return TrainUnit._from_searcher_unit(op.length, self.searcher_unit, self.global_batch_size)

    def _steps_until_complete(self, length: TrainUnit) -> int:
        assert self.state is not None and isinstance(length.value, int)
        if isinstance(length, Batch):
            return length.value - self.state.batches_trained
        return length.value - self.state.epochs_trained

    def _checkpoint_is_current(self) -> bool:
        assert self.state is not None
        return self.state.last_ckpt == self.state.batches_trained

    def _validation_is_current(self) -> bool:
        assert self.state is not None
        return self.state.last_val == self.state.batches_trained

    def _train_for_op(self, op: core.SearcherOperation) -> None:
        length = self._op_length(op)
        periods = {
            _Boundary.TRAIN: length,
            _Boundary.VALIDATE: self.validation_period,
            _Boundary.CHECKPOINT: self.checkpoint_period,
            _Boundary.REPORT: self.reporting_period,
        }
        while self._steps_until_complete(length) > 0:
            reached, per_batch = self._train_until_boundary(periods)
            metrics = self._aggregate_training_metrics(per_batch)
            metrics = self.context.distributed.broadcast(metrics)
            for cb in self.callbacks.values():
                cb.on_training_workload_end(avg_metrics=metrics["avg_metrics"], batch_metrics=metrics["batch_metrics"])
            progressed = False
            for kind in (_Boundary.TRAIN, _Boundary.VALIDATE, _Boundary.CHECKPOINT, _Boundary.REPORT):
                if kind not in reached:
                    continue
                if kind in (_Boundary.TRAIN, _Boundary.REPORT):
                    if self.is_chief and not op._completed and not progressed:
                        self._report_searcher_progress(op)
                        progressed = True
                elif kind == _Boundary.VALIDATE:
                    if not self._validation_is_current():
                        self._validate(op)
                elif kind == _Boundary.CHECKPOINT:
                    if not self._checkpoint_is_current():
                        self._checkpoint(already_exiting=False)
                if self.context.get_enable_tensorboard_logging():
                    self._upload_tb_files()
                self._stop_requested()
        if not self._validation_is_current():
            self._validate(op)
        if not self._checkpoint_is_current():
            self._checkpoint(already_exiting=False)
        if self.is_chief and not self.test_mode and not op._completed:
            raise ShouldExit(skip_exit_checkpoint=True)

    def _report_searcher_progress(self, op: core.SearcherOperation) -> None:
        assert self.state is not None
        if self.searcher_unit == core.Unit.BATCHES:
            op.report_progress(self.state.batches_trained)
        elif self.searcher_unit == core.Unit.RECORDS:
            assert self.global_batch_size
            op.report_progress(self.global_batch_size * self.state.batches_trained)
        elif self.searcher_unit == core.Unit.EPOCHS:
            op.report_progress(self.state.epochs_trained)

    def _stop_requested(self) -> None:
        if self.core_context.preempt.should_preempt():
            raise ShouldExit()
        if self.context.get_stop_requested():
            raise ShouldExit()

    # ------------------------------------------------------------------ hot loop
    def _train_until_boundary(self, periods: Dict[_Boundary, TrainUnit]) -> Tuple[set, List[Dict[str, Any]]]:
        assert self.state is not None and self.training_enumerator is not None
        if self.is_chief:
            self.core_context.train.set_status("training")
        for m in self.context.models:
            m.train()
        self.context.reset_reducers()
        epoch_len = self.context._epoch_len
        assert epoch_len
        per_batch: List[Dict[str, Any]] = []
        num_inputs = 0
        t0 = time.time()
        reached: set = set()
        for batch_idx, batch in self.training_enumerator:
            epoch_idx, in_epoch = divmod(batch_idx, epoch_len)
            self.context._current_batch_idx = batch_idx
            if in_epoch == 0:
                for cb in self.callbacks.values():
                    self._epoch_start(cb, epoch_idx)
            num_inputs += self.trial.get_batch_length(batch)
            for cb in self._batch_hooks:
                cb.on_training_batch_start(batch_idx)
            per_batch.append(self._stage_metrics(self._train_batch(batch, epoch_idx, batch_idx)))
            for cb in self._batch_hooks:
                cb.on_training_batch_end(batch_idx)
            # bookkeeping for the batch that just finished
            self.state.batches_trained += 1
            if in_epoch == epoch_len - 1:
                self.state.epochs_trained += 1
                for cb in self.callbacks.values():
                    cb.on_training_epoch_end(epoch_idx)
            for kind, unit in periods.items():
                if isinstance(unit, Batch) and unit.should_stop(batch_idx + 1):
                    reached.add(kind)
                elif isinstance(unit, Epoch) and in_epoch == epoch_len - 1 and unit.should_stop(epoch_idx + 1):
                    reached.add(kind)
            if self.test_mode:
                reached.add(_Boundary.TRAIN)
            if reached:
                break
        host_metrics = self._collect_metrics(per_batch)   # the one host sync of this window
        dt = time.time() - t0
        if per_batch:
            logger.info(util.make_timing_log("trained", dt, num_inputs, len(per_batch)))
        self._window_inputs = num_inputs
        return reached, host_metrics

    def _train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        ctx = self.context
        boundary = ctx._should_communicate_and_update()
        use_graph = (ctx.experimental._cuda_graphs and ctx.device.type == "cuda" and not self._graphs_failed
                     and ctx.profiler is None)
        self.step_timer.start()
        if use_graph:
            out = self._graphed_train_batch(batch, epoch_idx, batch_idx, boundary)
        else:
            out = self._eager_train_batch(batch, epoch_idx, batch_idx)
        self.step_timer.stop()
        if ctx.profiler is not None:
            ctx.profiler.step()
        if ctx._scaler is not None and ctx.experimental._auto_amp and boundary:
            ctx._scaler.update()
        for sched in ctx.lr_schedulers:
            self._auto_step_lr_scheduler_per_batch(batch_idx, sched)
        return out

    _PIN_SLOTS = 4096

    def _stage_metrics(self, out: Dict[str, Any]) -> Dict[str, Any]:
        """Replace 0-d device tensors by ``("pin", slot)`` markers after queueing an async D2H copy
        into pinned memory (no host synchronisation here)."""
        if self.context.device.type != "cuda":
            return out
        if self._pin_ring is None:
            self._pin_ring = torch.empty(self._PIN_SLOTS, dtype=torch.float32).pin_memory()
        staged: Dict[str, Any] = {}
        for k, v in out.items():
            if isinstance(v, torch.Tensor) and v.is_cuda and v.numel() == 1 and self._pin_used < self._PIN_SLOTS:
                slot = self._pin_used
                self._pin_used += 1
                self._pin_ring[slot:slot + 1].copy_(v.detach().reshape(1).to(torch.float32), non_blocking=True)
                staged[k] = ("pin", slot)
            else:
                staged[k] = v
        return staged

    def _collect_metrics(self, per_batch: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        if self._pin_used:
            torch.cuda.current_stream(self.context.device).synchronize()
            assert self._pin_ring is not None
            vals = self._pin_ring[: self._pin_used].tolist()
            for m in per_batch:
                for k, v in m.items():
                    if isinstance(v, tuple) and len(v) == 2 and v[0] == "pin":
                        m[k] = vals[v[1]]
            self._pin_used = 0
        return _metric_utils.materialize_metrics(per_batch)

    def _eager_train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        out = self.trial.train_batch(batch=batch, epoch_idx=epoch_idx, batch_idx=batch_idx)
        if isinstance(out, torch.Tensor):
            out = {"loss": out}
        if not isinstance(out, dict):
            raise errors.InvalidExperimentException("train_batch() must return a dictionary mapping string names "
                                                    f"to Tensor metrics, got {type(out).__name__}")
        # detach but do NOT move to host: materialised once per reporting window
        return {k: (v.detach() if isinstance(v, torch.Tensor) else v) for k, v in out.items()}

    def _graphed_train_batch(self, batch: Any, epoch_idx: int, batch_idx: int, boundary: bool) -> Dict[str, Any]:
        gs = self._graphs.get(boundary)
        if gs is None or gs.graph is None:
            if self._graph_warmup_left > 0 or (gs is None and self._needs_more_warmup(boundary)):
                self._graph_warmup_left -= 1
                self._graphs.setdefault(boundary, _GraphedStep())
                return self._eager_train_batch(batch, epoch_idx, batch_idx)
            gs = self._graphs.setdefault(boundary, _GraphedStep())
            try:
                return self._capture(gs, batch, epoch_idx, batch_idx)
            except Exception as e:  # capture is best-effort: fall back loudly, once
                logger.warning("CUDA graph capture of train_batch failed (%s: %s); continuing eagerly",
                               type(e).__name__, e)
                self._graphs_failed = True
                torch.cuda.synchronize()
                return self._eager_train_batch(batch, epoch_idx, batch_idx)
        if not _same_shapes(gs.static_in, batch):
            # ragged last batch etc.: run this one eagerly
            return self._eager_train_batch(batch, epoch_idx, batch_idx)
        _copy_into(gs.static_in, batch)
        for opt in self.context.optimizers:
            sync = getattr(opt, "sync_hparams", None)
            if sync is not None:
                sync()
        gs.graph.replay()
        from determined_b200.ops import _lib as _native_lib

        _native_lib.count_launches(gs.launches)          # kernels recorded in the graph ran once more
        return _graph_outputs(gs.static_out)

    def _needs_more_warmup(self, boundary: bool) -> bool:
        # each phase must have run eagerly at least once so lazily created state exists
        return boundary not in self._graphs

    def _capture(self, gs: _GraphedStep, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        torch.cuda.synchronize()
        inval = getattr(self.context._get_engine(), "invalidate_shadows", None)
        if inval is not None:
            inval()      # a captured step refreshes the bf16 weight mirror itself, whatever preceded the capture
        gs.static_in = _clone_tree(batch)
        for opt in self.context.optimizers:
            sync = getattr(opt, "sync_hparams", None)
            if sync is not None:
                sync()
        g = torch.cuda.CUDAGraph()
        from determined_b200.ops import _lib as _native_lib

        n0 = _native_lib.LAUNCHES[0]
        with torch.cuda.graph(g):
            out = self.trial.train_batch(batch=gs.static_in, epoch_idx=epoch_idx, batch_idx=batch_idx)
            if isinstance(out, torch.Tensor):
                out = {"loss": out}
            gs.static_out = {k: (v.detach() if isinstance(v, torch.Tensor) else v) for k, v in out.items()}
        gs.graph = g
        gs.launches = _native_lib.LAUNCHES[0] - n0       # counted once at capture = the single replay below
        # capture does not execute: replay once so this batch is actually trained on
        g.replay()
        return _graph_outputs(gs.static_out)

    def _auto_step_lr_scheduler_per_batch(self, batch_idx: int, sched: _lr_scheduler.LRScheduler) -> None:
        """Step schedulers only on aggregation boundaries, catching up the batches in between."""
        ctx = self.context
        if not ctx._should_communicate_and_update():
            return
        mode, freq, agg = sched._step_mode, sched._frequency, ctx._aggregation_frequency
        SM = _lr_scheduler.LRScheduler.StepMode
        if mode == SM.STEP_EVERY_BATCH:
            for i in range(batch_idx - agg + 1, batch_idx + 1):
                if (i + 1) % freq == 0:
                    sched.step()
        elif mode == SM.STEP_EVERY_OPTIMIZER_STEP:
            if (batch_idx + 1) % freq == 0:
                sched.step()
        elif mode == SM.STEP_EVERY_EPOCH:
            assert ctx._epoch_len
            for e in range(batch_idx // ctx._epoch_len, (batch_idx + agg) // ctx._epoch_len):
                if (e + 1) % freq == 0:
                    sched.step()

    def _aggregate_training_metrics(self, per_batch: List[Dict[str, Any]]) -> Dict[str, Any]:
        assert self.state is not None
        ctx = self.context
        for m in per_batch:
            for k, v in m.items():
                if isinstance(v, np.ndarray) and v.ndim == 0:
                    m[k] = v.item()
        avg_enabled = True
        if ctx._exp_conf is not None:
            avg_enabled = det.ExperimentConfig(ctx._exp_conf).average_training_metrics_enabled()
        if ctx.distributed.size > 1 and avg_enabled:
            per_batch = _metric_utils._average_training_metrics(ctx.distributed, per_batch)
        num_inputs = getattr(self, "_window_inputs", None)
        if ctx.distributed.size > 1 and num_inputs is not None:
            counts = ctx.distributed.gather(num_inputs)
            if self.is_chief and counts is not None:
                num_inputs = sum(counts)
        metrics = util.make_metrics(num_inputs, per_batch)
        reduced = ctx.reduce_metrics(for_training=True)       # collective: all ranks
        metrics["avg_metrics"].update(_metric_utils._convert_metrics_to_numpy(reduced))
        if self.is_chief:
            if ctx.get_enable_tensorboard_logging():
                self._tb_log("train", metrics["avg_metrics"], metrics["batch_metrics"])
            self.core_context.train.report_training_metrics(
                steps_completed=self.state.batches_trained, metrics=metrics["avg_metrics"],
                batch_metrics=metrics["batch_metrics"])
        return metrics

    def _tb_log(self, kind: str, metrics: Dict[str, Any], batch_metrics: Any = None) -> None:
        assert self.state is not None
        try:
            if self._metric_writer is None:
                from determined_b200.tensorboard import metric_writers

                self._metric_writer = metric_writers.BatchMetricWriter.try_build(self.core_context.train.get_tensorboard_path())
            _metric_utils._log_tb_metrics(self._metric_writer, "val" if kind == "val" else "train",
                                          self.state.batches_trained, metrics, batch_metrics)
        except Exception as e:
            logger.debug("tensorboard logging failed: %s", e)

    def _upload_tb_files(self) -> None:
        self.context._maybe_reset_tbd_writer()
        try:
            self.core_context.train.upload_tensorboard_files(
                (lambda _: True) if self.is_chief else (lambda p: "tfevents" not in p.name),
                tensorboard.get_rank_aware_path)
        except RuntimeError:
            pass   # TensorboardMode.AUTO: the core context syncs by itself

    # ------------------------------------------------------------------ validation
    def _check_searcher_metric(self, val_metrics: Dict[str, Any]) -> Any:
        name = self.searcher_metric_name
        if name not in val_metrics:
            raise RuntimeError(
                f"Search method is configured to use metric '{name}' but model definition returned validation "
                f"metrics {list(val_metrics.keys())}. The metric used by the search method must be one of the "
                "validation metrics returned by the model definition.")
        m = val_metrics[name]
        if not util.is_numerical_scalar(m):
            raise RuntimeError(f"Searcher validation metric '{name}' returned a non-scalar value: {m}.")
        return m

    def _is_best_validation(self, now: float, before: Optional[float]) -> bool:
        if before is None:
            return True
        return (now < before) if self.smaller_is_better else (now > before)

    @torch.no_grad()
    def _validate(self, searcher_op: Optional[core.SearcherOperation] = None) -> Dict[str, Any]:
        assert self.state is not None
        ctx = self.context
        if self.is_chief:
            self.core_context.train.set_status("validating")
        ctx.reset_reducers()
        for m in ctx.models:
            m.eval()
        t0 = time.time()
        for cb in self.callbacks.values():
            cb.on_validation_start()
        metrics: Dict[str, Any] = {}
        num_inputs, nb = 0, 0
        input_counts = None
        if self._evaluate_batch_defined():
            for cb in self.callbacks.values():
                cb.on_validation_epoch_start()
            keys = None
            rows: List[Dict[str, Any]] = []
            it: Iterator = iter(self.validation_loader)
            if ctx.experimental._auto_to_device:
                it = _data.DevicePrefetcher(it, ctx.device)
            takes_idx = util.has_param(self.trial.evaluate_batch, "batch_idx", 2)
            for idx, batch in enumerate(it):
                num_inputs += self.trial.get_batch_length(batch)
                out = self.trial.evaluate_batch(batch=batch, batch_idx=idx) if takes_idx else self.trial.evaluate_batch(batch=batch)
                if not isinstance(out, dict):
                    raise TypeError("evaluate_batch() must return a dictionary of string names to Tensor metrics; "
                                    f"got {out}.")
                if keys is None:
                    keys = list(out.keys())
                elif list(out.keys()) != keys:
                    raise ValueError(f"Validation metric names must match across all batches of data: {keys} != {list(out.keys())}.")
                rows.append({k: (v.detach() if isinstance(v, torch.Tensor) else v) for k, v in out.items()})
                nb += 1
                if self.test_mode:
                    break
            if nb == 0:
                raise RuntimeError("validation_loader is empty.")
            host_rows = _metric_utils.materialize_metrics(rows)      # one D2H for the whole pass
            for cb in self.callbacks.values():
                cb.on_validation_epoch_end(host_rows)
            assert keys is not None
            mat = np.array([[np.asarray(r[k], dtype=np.float64) for k in keys] for r in host_rows], dtype=np.float64).reshape(nb, len(keys))
            metrics = _metric_utils._reduce_metrics(
                ctx.distributed, mat, keys,
                _metric_utils._prepare_metrics_reducers(self.trial.evaluation_reducer(), keys))
            input_counts = ctx.distributed.gather((num_inputs, nb))
        else:
            if self.is_chief:
                metrics = self.trial.evaluate_full_dataset(data_loader=self.validation_loader)
                if not isinstance(metrics, dict):
                    raise TypeError(f"eval() must return a dictionary, got {type(metrics).__name__}.")
                metrics = _metric_utils._convert_metrics_to_numpy(metrics)
        metrics.update(_metric_utils._convert_metrics_to_numpy(ctx.reduce_metrics(for_training=False)))
        if ctx.distributed.size > 1 and any(util.is_overridden(c.on_validation_end, _callback.PyTorchCallback)
                                            for c in self.callbacks.values()):
            metrics = ctx.distributed.broadcast(metrics)
        for cb in self.callbacks.values():
            cb.on_validation_end(metrics)
        self.state.last_val = self.state.batches_trained

        best_before = None
        if self.is_chief:
            if input_counts is not None:
                tot_in, tot_b = (sum(x) for x in zip(*input_counts))
                logger.info(util.make_timing_log("validated", time.time() - t0, tot_in, tot_b))
            if ctx.get_enable_tensorboard_logging():
                self._tb_log("val", metrics)
            best_before = self.core_context.train.get_experiment_best_validation()
            self.core_context.train.report_validation_metrics(self.state.batches_trained, metrics)

        searcher_metric = None
        if self.is_chief and searcher_op is not None:
            if self.searcher_metric_name:
                searcher_metric = self._check_searcher_metric(metrics)
            if self._steps_until_complete(self._op_length(searcher_op)) < 1 and not searcher_op._completed:
                searcher_op.report_completed(searcher_metric)

        should_ckpt = False
        if self.is_chief and not self._checkpoint_is_current():
            if self.ckpt_policy == "all":
                should_ckpt = True
            elif self.ckpt_policy == "best" and self.searcher_metric_name:
                sm = searcher_metric if searcher_metric is not None else (
                    self._check_searcher_metric(metrics) if self.searcher_metric_name in metrics else None)
                if sm is not None and self._is_best_validation(float(sm), best_before):
                    should_ckpt = True
        should_ckpt = ctx.distributed.broadcast(should_ckpt)
        if should_ckpt:
            self._checkpoint(already_exiting=False)
        return metrics

    # ------------------------------------------------------------------ checkpointing
    def _checkpoint(self, already_exiting: bool) -> None:
        assert self.state is not None
        if self.is_chief:
            self.core_context.train.set_status("checkpointing")
        self.state.last_ckpt = self.state.batches_trained
        self.state.step_id += 1
        try:
            uuid = ""
            if self.is_chief:
                md = {
                    "determined_version": det.__version__,
                    "steps_completed": self.state.batches_trained,
                    "framework": f"torch-{torch.__version__}",
                    "format": "pickle",
                }
                if self.context.experimental._async_checkpoint and not already_exiting:
                    uuid = self._save_behind(md)
                else:
                    self._drain_checkpoints()
                    with self.core_context.checkpoint.store_path(md) as (path, uuid):
                        self._save(pathlib.Path(path))
            elif not (self.context.experimental._async_checkpoint and not already_exiting):
                self._drain_checkpoints()
            uuid = self.context.distributed.broadcast(uuid)
            if self._ckpt_writer is not None and not already_exiting and self.context.experimental._async_checkpoint:
                self._pending_upload_cbs.append(uuid)          # delivered once the writer is done (next checkpoint / exit)
            else:
                for cb in self.callbacks.values():
                    cb.on_checkpoint_upload_end(uuid=uuid)
        except errors.InvalidHP:
            if not already_exiting:
                self.core_context.train.report_early_exit(core.EarlyExitReason.INVALID_HP)
                raise ShouldExit(skip_exit_checkpoint=True)
            raise
        if not already_exiting:
            self._stop_requested()

    _legacy_epoch_cbs: Optional[set] = None

    def _epoch_start(self, cb: Any, epoch_idx: int) -> None:
        """Callbacks written before ``on_training_epoch_start`` took the epoch index still work (reference
        ``_pytorch_trial.py:368-378`` keeps them alive with a deprecation warning); the arity is looked up once per callback."""
        if self._legacy_epoch_cbs is None:
            import inspect

            self._legacy_epoch_cbs = {id(c) for c in self.callbacks.values() if not inspect.signature(c.on_training_epoch_start).parameters}
            if self._legacy_epoch_cbs:
                logger.warning("on_training_epoch_start() without an epoch_idx parameter is deprecated; please add it")
        if id(cb) in self._legacy_epoch_cbs:
            cb.on_training_epoch_start()
        else:
            cb.on_training_epoch_start(epoch_idx)

    # -- checkpoint fast path (pytorch/_async_checkpoint.py) ----------------------------------------
    _ckpt_writer: Any = None
    _ckpt_pool: Any = None

    def _save_behind(self, md: Dict[str, Any]) -> str:
        """Chief only: snapshot now (non-blocking D2H on the training stream), write and report from the writer thread.
        Returns the storage id as soon as the writer has opened the checkpoint."""
        import queue as _queue

        from determined_b200.pytorch import _async_checkpoint as ac

        if self._ckpt_writer is None:
            self._ckpt_writer, self._ckpt_pool = ac.CheckpointWriter(max_pending=1), ac.PinnedPool()
        self._drain_checkpoints()                 # at most one in flight; also delivers its upload callbacks
        checkpoint, state_vars, load_data = self._collect()
        snap = ac.snapshot(checkpoint, self._ckpt_pool)
        ident: "_queue.Queue[Any]" = _queue.Queue(1)

        def job() -> None:
            try:
                with self.core_context.checkpoint.store_path(md) as (path, uuid):
                    ident.put(uuid)
                    self._write(pathlib.Path(path), snap.wait(), state_vars, load_data)
            except BaseException as e:  # noqa: BLE001
                ident.put(e)               # harmless if the id was already delivered (queue of one, nobody reads twice)
                raise
            finally:
                self._ckpt_pool.give(snap.bufs)

        self._ckpt_writer.submit(job)
        got = ident.get()
        if isinstance(got, BaseException):
            raise got
        return str(got)

    def _drain_checkpoints(self) -> None:
        """Wait for the checkpoint in flight (if any), surface its failure, deliver the deferred upload callbacks."""
        if self._ckpt_writer is not None:
            self._ckpt_writer.join()
        pending, self._pending_upload_cbs = self._pending_upload_cbs, []
        for uuid in pending:
            for cb in self.callbacks.values():
                cb.on_checkpoint_upload_end(uuid=uuid)

    def _close_checkpoint_writer(self) -> None:
        try:
            self._drain_checkpoints()
        finally:
            if self._ckpt_writer is not None:
                w, self._ckpt_writer = self._ckpt_writer, None
                w.close()

    def _save(self, path: pathlib.Path) -> None:
        """Write the reference checkpoint layout: ``state_dict.pth``, ``trial_state.pkl``,
        ``load_data.json``, ``code/`` (``metadata.json`` is added by the CheckpointContext)."""
        checkpoint, state_vars, load_data = self._collect()
        self._write(path, checkpoint, state_vars, load_data)

    def _collect(self) -> Tuple[Dict[str, Any], Dict[str, Any], Dict[str, Any]]:
        """Everything a checkpoint holds, referencing live tensors (main thread, at a step boundary)."""
        assert self.state is not None
        ctx = self.context
        rng_state = {"cpu_rng_state": torch.random.get_rng_state(), "np_rng_state": np.random.get_state(),
                     "random_rng_state": random.getstate()}
        if torch.cuda.is_available() and ctx.device.type == "cuda":
            rng_state["gpu_rng_state"] = torch.cuda.get_rng_state(ctx.device)
        checkpoint: Dict[str, Any] = {
            "models_state_dict": [m.state_dict() for m in ctx.models],
            "optimizers_state_dict": [o.state_dict() for o in ctx.optimizers],
            "lr_schedulers_state_dict": [s.state_dict() for s in ctx.lr_schedulers],
            "callbacks": {name: cb.state_dict() for name, cb in self.callbacks.items()},
            "rng_state": rng_state,
        }
        if ctx._scaler is not None:
            checkpoint["scaler_state_dict"] = ctx._scaler.state_dict()
        for cb in self.callbacks.values():
            cb.on_checkpoint_save_start(checkpoint)
        trial_cls = type(self.trial)
        load_data: Dict[str, Any] = {"trial_type": "PyTorchTrial", "experiment_config": ctx._exp_conf,
                                     "hparams": ctx._hparams,
                                     "trial_cls_spec": f"{trial_cls.__module__}:{trial_cls.__qualname__}",
                                     "is_trainer": True}
        return checkpoint, copy.deepcopy(vars(self.state)), load_data

    def _write(self, path: pathlib.Path, checkpoint: Dict[str, Any], state_vars: Dict[str, Any], load_data: Dict[str, Any]) -> None:
        """Serialise one checkpoint (any thread; ``checkpoint`` must not alias live training state when called off-thread)."""
        path.mkdir(parents=True, exist_ok=True)
        util.write_user_code(path, not self.local_training)
        torch.save(checkpoint, str(path / "state_dict.pth"))
        with (path / "trial_state.pkl").open("wb") as f:
            pickle.dump(state_vars, f)
        (path / "load_data.json").write_text(util.json_encode(load_data))
        for cb in self.callbacks.values():
            if util.is_overridden(cb.on_checkpoint_end, _callback.PyTorchCallback):
                cb.on_checkpoint_end(str(path))
            cb.on_checkpoint_write_end(str(path))

    _LEGACY_LOCATIONS = (("state_dict.pth",), ("determined", "state_dict.pth"), ("pedl", "state_dict.pth"),
                         ("checkpoint.pt",))

    def _load(self, load_path: pathlib.Path) -> None:
        ctx = self.context
        checkpoint = None
        for parts in self._LEGACY_LOCATIONS:
            f = load_path.joinpath(*parts)
            if f.exists():
                checkpoint = torch.load(str(f), map_location="cpu", weights_only=False)
                break
        if not isinstance(checkpoint, dict):
            self.state = self.state or _TrialState(trial_id=self.trial_id)
            return
        for cb in self.callbacks.values():
            cb.on_checkpoint_load_start(checkpoint)

        def one_or_many(single_key: str, multi_key: str, targets: List[Any], what: str) -> None:
            if single_key in checkpoint:
                if multi_key in checkpoint:
                    raise RuntimeError(f"Both {single_key} and {multi_key} in checkpoint.")
                if len(targets) > 1:
                    raise RuntimeError(f"Old-format checkpoint cannot be loaded into a context with more than one {what}.")
                if targets:
                    self._load_one(targets[0], checkpoint[single_key], what)
            elif multi_key in checkpoint:
                for tgt, sd in zip(targets, checkpoint[multi_key]):
                    self._load_one(tgt, sd, what)

        one_or_many("model_state_dict", "models_state_dict", ctx.models, "model")
        one_or_many("optimizer_state_dict", "optimizers_state_dict", ctx.optimizers, "optimizer")
        one_or_many("lr_scheduler", "lr_schedulers_state_dict", ctx.lr_schedulers, "LR scheduler")

        if "scaler_state_dict" in checkpoint:
            if ctx._scaler is not None:
                ctx._scaler.load_state_dict(checkpoint["scaler_state_dict"])
            else:
                logger.warning("There exists scaler_state_dict in checkpoint but the experiment is not using AMP.")
        elif ctx._scaler is not None:
            logger.warning("The experiment is using AMP but scaler_state_dict does not exist in the checkpoint.")
        if "amp_state" in checkpoint:
            logger.warning("checkpoint carries apex amp_state; apex is not part of the B200 stack, ignoring it")

        rng = checkpoint.get("rng_state")
        if rng is not None:
            self._apply_rng_state(rng)
            self._restored_rng = rng
        else:
            logger.warning("The checkpoint has no random state to restore.")

        cb_state = checkpoint.get("callbacks", {})
        for name, cb in self.callbacks.items():
            if name in cb_state:
                cb.load_state_dict(cb_state[name])
            elif util.is_overridden(cb.load_state_dict, _callback.PyTorchCallback):
                logger.warning("Callback '%s' implements load_state_dict(), but no callback state was found for that "
                               "name when restoring from checkpoint. This callback will be initialized from scratch.", name)

        ts, wl = load_path / "trial_state.pkl", load_path / "workload_sequencer.pkl"
        if ts.exists():
            with ts.open("rb") as f:
                self._load_state(pickle.load(f))
        elif wl.exists():
            with wl.open("rb") as f:
                self._load_wlsq_state(pickle.load(f))
        else:
            self.state = _TrialState(trial_id=self.trial_id)

    def _load_one(self, target: Any, sd: Any, what: str) -> None:
        if what != "model":
            target.load_state_dict(sd)
            return
        try:
            target.load_state_dict(sd)
        except Exception:
            # checkpoints written by the reference under DDP carry a "module." prefix
            sd = dict(sd)
            torch.nn.modules.utils.consume_prefix_in_state_dict_if_present(sd, "module.")
            target.load_state_dict(sd)

    def _load_state(self, state: Dict[str, Any]) -> None:
        if state.get("trial_id") != self.trial_id:
            # "continue training from another trial's checkpoint": weights yes, counters no
            self.state = _TrialState(trial_id=self.trial_id)
            return
        known = {k: state[k] for k in ("trial_id", "last_ckpt", "step_id", "last_val", "batches_trained", "epochs_trained") if k in state}
        self.state = _TrialState(**known)
        if self.state.batches_trained == self.val_from_previous_run:
            self.state.last_val = self.state.batches_trained

    def _load_wlsq_state(self, state: Dict[str, Any]) -> None:
        if state.get("trial_id") != self.trial_id:
            self.state = _TrialState(trial_id=self.trial_id)
            return
        done = int(state.get("steps_completed") or 0)
        self.state = _TrialState(trial_id=state.get("trial_id", 0), last_ckpt=state.get("last_ckpt", 0),
                                 last_val=state.get("last_val", 0), step_id=state.get("step_id", 0),
                                 batches_trained=done,
                                 epochs_trained=done // self.context._epoch_len if self.context._epoch_len else 0)
        if self.state.batches_trained == self.val_from_previous_run:
            self.state.last_val = self.state.batches_trained


def _repeats(loader: Any) -> bool:
    from determined_b200.pytorch import samplers

    bs = getattr(loader, "batch_sampler", None)
    seen = 0
    while bs is not None and seen < 8:
        if isinstance(bs, samplers.RepeatBatchSampler):
            return True
        bs = getattr(bs, "batch_sampler", None)
        seen += 1
    return False


class PyTorchTrial(metaclass=abc.ABCMeta):
    """Subclass this to define a model for the harness.  Required: ``__init__(context)``,
    ``train_batch``, ``build_training_data_loader``, ``build_validation_data_loader`` and one of
    ``evaluate_batch`` / ``evaluate_full_dataset``."""

    trial_controller_class = _PyTorchTrialController
    trial_context_class = PyTorchTrialContext

    @abc.abstractmethod
    def __init__(self, context: PyTorchTrialContext) -> None:
        """Build models, optimizers and LR schedulers and wrap them with ``context.wrap_model``,
        ``context.wrap_optimizer`` and ``context.wrap_lr_scheduler`` (in that order)."""

    @abc.abstractmethod
    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Union[torch.Tensor, Dict[str, Any]]:
        """Forward pass, ``context.backward(loss)``, ``context.step_optimizer(opt)``; return a dict
        of metrics (tensors may stay on the device) or the loss tensor."""

    @abc.abstractmethod
    def build_training_data_loader(self) -> Any:
        """Return a ``determined.pytorch.DataLoader`` for training."""

    @abc.abstractmethod
    def build_validation_data_loader(self) -> Any:
        """Return a ``determined.pytorch.DataLoader`` for validation."""

    def build_callbacks(self) -> Dict[str, _callback.PyTorchCallback]:
        return {}

    def evaluate_batch(self, batch: Any, batch_idx: int) -> Dict[str, Any]:
        """Metrics for one validation batch; reduced across batches and slots by
        :meth:`evaluation_reducer`."""
        return {}

    def evaluation_reducer(self) -> Union[_reducer.Reducer, Dict[str, _reducer.Reducer]]:
        return _reducer.Reducer.AVG

    def evaluate_full_dataset(self, data_loader: Any) -> Dict[str, Any]:
        """Alternative to ``evaluate_batch``: chief-only evaluation over the whole loader."""
        return {}

    def get_batch_length(self, batch: Any) -> int:
        """Number of records in ``batch`` (override for exotic batch types)."""
        return _data.data_length(batch)
