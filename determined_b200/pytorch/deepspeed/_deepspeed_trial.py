"""``DeepSpeedTrial`` and its controller.

Parity: reference ``harness/determined/pytorch/deepspeed/_deepspeed_trial.py``: controller
``pre_execute_hook`` (``deepspeed.init_distributed`` unless ``DET_MANUAL_INIT_DISTRIBUTED``)
``:77-99``, loaders sharded by ``mpu.data_parallel_{rank,world_size}`` ``:107-236``, workload loop
``:297-359``, ``_train_for_step`` (``train_batch(iterator, …)`` called gradient-accumulation-steps
times unless pipeline-parallel / manual) ``:364-499``, validation ``:501-600``, per-rank
``det_state_dict_rank{r}.pth`` + ``trial.save`` under ``store_path(shard=True)`` ``:674-723``,
``_load`` ``:608-672``; ``DeepSpeedTrial`` API (iterators in ``train_batch``/``evaluate_batch``,
default save/load = ``engine.save_checkpoint(path, tag=f"model{i}")``) ``:729-945``.

The reference drives this trial type through the legacy workload sequencer
(``layers/_workload_sequencer.py``); here the controller pulls searcher operations directly like
the PyTorchTrial controller does (the sequencer itself lives in
:mod:`determined_b200.layers` for API parity).
"""
import abc
import contextlib
import logging
import os
import pathlib
import pickle
import random
import sys
import time
from typing import Any, Dict, Iterator, List, Optional, Type, Union, cast

import numpy as np
import torch

import determined_b200 as det
from determined_b200 import core, errors, util
from determined_b200.errors import ShouldExit
from determined_b200.pytorch import _callback, _data, _metric_utils, _reducer
from determined_b200.pytorch._pytorch_trial import Batch, Epoch, TrainUnit, _TrialState
from determined_b200.pytorch.deepspeed._deepspeed_context import DeepSpeedTrialContext

logger = logging.getLogger("determined.pytorch.deepspeed")


class _Repeating:
    """Endless iterator over a loader (what DeepSpeed's ``RepeatingLoader`` does)."""

    def __init__(self, loader: Any) -> None:
        self.loader = loader
        self._it = iter(loader)

    def __len__(self) -> int:
        return len(self.loader)

    def __iter__(self) -> "_Repeating":
        return self

    def __next__(self) -> Any:
        try:
            return next(self._it)
        except StopIteration:
            self._it = iter(self.loader)
            return next(self._it)


class DeepSpeedTrialController:
    def __init__(self, trial_inst: "DeepSpeedTrial", context: DeepSpeedTrialContext, *, max_length: Optional[TrainUnit] = None,
                 validation_period: Optional[TrainUnit] = None, checkpoint_period: Optional[TrainUnit] = None,
                 reporting_period: Optional[TrainUnit] = None, checkpoint_policy: str = "best",
                 latest_checkpoint: Optional[str] = None, steps_completed: int = 0, local_training: bool = True,
                 searcher_metric_name: Optional[str] = None, smaller_is_better: bool = True, test_mode: bool = False) -> None:
        if not isinstance(trial_inst, DeepSpeedTrial):
            raise TypeError("DeepSpeedTrialController needs a DeepSpeedTrial")
        self.trial, self.context, self.core_context = trial_inst, context, context._core
        self.is_chief = context.distributed.rank == 0
        # autotuning probe trials (``pytorch.dsat``): one searcher operation, trained batch by batch under the reporting
        # context until the engine's probe exits the process (reference ``_deepspeed_trial.py:46-51,298-308``)
        from determined_b200.pytorch.dsat import defaults as _dsat_defaults

        self._dsat_mode = bool((context.get_hparams() or {}).get(_dsat_defaults.USE_DSAT_MODE_KEY, False))
        if self._dsat_mode and not local_training:
            name = (context.get_experiment_config().get("searcher") or {}).get("name")
            if name != "custom":
                raise errors.InvalidExperimentException("`_use_dsat_mode` can only be set for trials of a custom-searcher experiment")
        self.callbacks = self.trial.build_callbacks()
        self.max_length = max_length
        self.validation_period = validation_period or Batch(sys.maxsize)
        self.checkpoint_period = checkpoint_period or Batch(sys.maxsize)
        self.reporting_period = reporting_period or Batch(100)
        self.ckpt_policy = checkpoint_policy
        self.latest_checkpoint = latest_checkpoint
        self.steps_completed = steps_completed
        self.local_training = local_training
        self.searcher_metric_name = searcher_metric_name
        self.smaller_is_better = smaller_is_better
        self.test_mode = test_mode
        self.trial_id = 0 if local_training or context._core.info is None else context.get_trial_id()
        self.state = _TrialState(trial_id=self.trial_id)
        self.searcher_unit = core.Unit.BATCHES if local_training else self.core_context.searcher.get_configured_units()
        self.training_iterator: Optional[Iterator] = None
        self.validation_loader: Any = None
        self.num_validation_batches: Optional[int] = None
        if not context.models:
            raise errors.InvalidExperimentException(
                "Must have at least one model engine. This might be caused by not wrapping your model with wrap_model_engine()")
        for m in context.models:
            for attr in ("train_micro_batch_size_per_gpu", "gradient_accumulation_steps"):
                if not callable(getattr(m, attr, None)):
                    raise errors.InvalidExperimentException(f"wrapped engine lacks {attr}(): is it a DeepSpeed-compatible engine?")

    @classmethod
    def pre_execute_hook(cls, trial_seed: int, distributed_backend: Any = None) -> None:
        """Process set-up before the trial is built: join the process group (our
        ``zero.init_distributed``, the analogue of ``deepspeed.init_distributed``) unless
        ``DET_MANUAL_INIT_DISTRIBUTED`` is set, then seed the RNGs."""
        if not os.environ.get("DET_MANUAL_INIT_DISTRIBUTED"):
            from determined_b200.parallel import zero

            zero.init_distributed()
        random.seed(trial_seed)
        np.random.seed(trial_seed % (2**32))
        torch.random.manual_seed(trial_seed)

    # ------------------------------------------------------------------ data
    def _set_data_loaders(self) -> None:
        ctx, mpu = self.context, self.context._mpu
        skip = self.steps_completed * (1 if ctx.use_pipeline_parallel else 1)
        nrep, rank = mpu.data_parallel_world_size, mpu.data_parallel_rank
        train = self.trial.build_training_data_loader()
        if mpu.should_build_data_loader:
            if train is None:
                raise errors.InvalidExperimentException("Please provide a training data loader (build_training_data_loader returned None).")
            if isinstance(train, _data.DataLoader):
                # one "batch" of the trial = gradient_accumulation_steps micro batches
                micro_skip = skip * ctx.num_micro_batches_per_slot
                loader = train.get_data_loader(repeat=True, skip=micro_skip, num_replicas=nrep, rank=rank)
                ctx._epoch_len = len(_data.adapt_batch_sampler(train.batch_sampler, num_replicas=nrep, rank=rank)) // ctx.num_micro_batches_per_slot
                self.training_iterator = iter(loader)
            else:
                if not ctx._data_repro_checks_disabled:
                    raise RuntimeError("build_training_data_loader() must return a determined.pytorch.DataLoader unless "
                                       "context.disable_dataset_reproducibility_checks() was called")
                ctx._epoch_len = max(len(train) // ctx.num_micro_batches_per_slot, 1)
                self.training_iterator = _Repeating(train)
        lens = ctx.distributed.allgather(ctx._epoch_len)
        known = [x for x in lens if x]
        if not known:
            raise RuntimeError("no rank built a training data loader")
        ctx._epoch_len = ctx._epoch_len or min(known)
        val = self.trial.build_validation_data_loader()
        if mpu.should_build_data_loader:
            if val is None:
                raise errors.InvalidExperimentException("Please provide a validation data loader.")
            if isinstance(val, _data.DataLoader):
                self.validation_loader = val.get_data_loader(repeat=False, skip=0, num_replicas=nrep, rank=rank)
            else:
                self.validation_loader = val
            n = len(self.validation_loader)
            if ctx.use_pipeline_parallel:
                n //= ctx.num_micro_batches_per_slot
            self.num_validation_batches = n
        nv = ctx.distributed.allgather(self.num_validation_batches)
        self.num_validation_batches = self.num_validation_batches or min(x for x in nv if x)

    # ------------------------------------------------------------------ top level
    def run(self) -> None:
        with contextlib.ExitStack() as stack:
            for cb in self.callbacks.values():
                stack.callback(cb.on_trial_shutdown)
            self._set_data_loaders()
            if self.latest_checkpoint is not None:
                logger.info("Restoring trial from checkpoint %s", self.latest_checkpoint)
                with self.core_context.checkpoint.restore_path(self.latest_checkpoint) as p:
                    self._load(pathlib.Path(p))
                if self.state.batches_trained != self.steps_completed:
                    # the loaders were positioned before the checkpoint told us where we are
                    self.steps_completed = self.state.batches_trained
                    self._set_data_loaders()
            self.state.batches_trained = self.steps_completed = max(self.state.batches_trained, self.steps_completed)
            for cb in self.callbacks.values():
                cb.on_trial_startup(self.steps_completed, self.latest_checkpoint)
            for cb in self.callbacks.values():
                cb.on_training_start()
            try:
                if self.local_training:
                    assert self.max_length is not None and isinstance(self.max_length.value, int)
                    ops: Iterator = iter([core.DummySearcherOperation(self.max_length.value, self.is_chief)])
                else:
                    ops = self.core_context.searcher.operations()
                if self._dsat_mode:
                    self._run_dsat_probe(ops)
                for op in ops:
                    self._train_for_op(op)
            except ShouldExit as e:
                if not e.skip_exit_checkpoint and self.state.last_ckpt != self.state.batches_trained:
                    self._checkpoint(already_exiting=True)

    def _run_dsat_probe(self, ops: Iterator) -> None:
        """Train the one operation of a probe trial a batch at a time, each inside ``dsat_reporting_context``: when the
        engine's probe is done it exits the process and the context reports its result.  A probe that never exits (an
        engine without autotuning support) falls through to the normal loop with the operation still open."""
        from determined_b200.pytorch import dsat

        op = next(ops, None)
        if op is None:
            return
        for m in self.context.models:
            m.train()
        epoch_len = self.context._epoch_len or sys.maxsize
        for _ in range(int(op.length)):
            batch_idx = self.state.batches_trained
            self.context._current_batch_idx = batch_idx
            with dsat.dsat_reporting_context(core_context=self.core_context, op=op):
                self.trial.train_batch(self.training_iterator, batch_idx // epoch_len, batch_idx)
            self.state.batches_trained += 1
        self._train_for_op(op)

    def _op_length(self, op: core.SearcherOperation) -> TrainUnit:
        if self.test_mode:
            return Batch(1)
        if self.local_training:
            assert self.max_length is not None
            return self.max_length
        gbs = self.context.models[0].train_batch_size() if hasattr(self.context.models[0], "train_batch_size") else None
        return TrainUnit._from_searcher_unit(op.length, self.searcher_unit, gbs)

    def _remaining(self, length: TrainUnit) -> int:
        assert isinstance(length.value, int)
        done = self.state.batches_trained if isinstance(length, Batch) else self.state.epochs_trained
        return length.value - done

    def _train_for_op(self, op: core.SearcherOperation) -> None:
        length = self._op_length(op)
        while self._remaining(length) > 0:
            reached = self._train_until_boundary(length)
            if ("train" in reached or "report" in reached) and self.is_chief and not op._completed:
                op.report_progress(self.state.batches_trained if isinstance(length, Batch) else self.state.epochs_trained)
            if "validate" in reached and self.state.last_val != self.state.batches_trained:
                self._validate(op)
            if "checkpoint" in reached and self.state.last_ckpt != self.state.batches_trained:
                self._checkpoint(already_exiting=False)
            if self.core_context.preempt.should_preempt() or self.context.get_stop_requested():
                raise ShouldExit()
        if self.state.last_val != self.state.batches_trained:
            self._validate(op)
        if self.state.last_ckpt != self.state.batches_trained:
            self._checkpoint(already_exiting=False)
        if self.is_chief and not self.test_mode and not op._completed:
            raise ShouldExit(skip_exit_checkpoint=True)

    # ------------------------------------------------------------------ hot loop
    def _train_until_boundary(self, length: TrainUnit) -> set:
        ctx = self.context
        if self.is_chief:
            self.core_context.train.set_status("training")
        for m in ctx.models:
            m.train()
        ctx.reset_reducers()
        epoch_len = ctx._epoch_len
        assert epoch_len
        t0 = time.time()
        per_batch: List[Dict[str, Any]] = []
        num_inputs = 0
        reached: set = set()
        periods = {"train": length, "validate": self.validation_period, "checkpoint": self.checkpoint_period,
                   "report": self.reporting_period}
        while not reached:
            batch_idx = self.state.batches_trained
            epoch_idx, in_epoch = divmod(batch_idx, epoch_len)
            ctx._current_batch_idx = batch_idx
            if in_epoch == 0:
                for cb in self.callbacks.values():
                    cb.on_training_epoch_start(epoch_idx)
            num_inputs += ctx.train_micro_batch_size_per_gpu * ctx.num_micro_batches_per_slot
            calls = 1 if (ctx.use_pipeline_parallel or ctx._manual_grad_accumulation) else ctx.num_micro_batches_per_slot
            for _ in range(calls):
                with contextlib.ExitStack() as es:
                    if ctx.profiler is not None:
                        es.enter_context(ctx.profiler)
                    out = self.trial.train_batch(self.training_iterator, epoch_idx, batch_idx)
                    if ctx.profiler is not None:
                        ctx.profiler.step()
                if ctx._mpu.should_report_metrics:
                    if isinstance(out, torch.Tensor):
                        out = {"loss": out}
                    if not isinstance(out, dict):
                        raise errors.InvalidExperimentException(
                            f"train_batch must return a dictionary mapping string names to Tensor metrics, got {type(out)}")
                    per_batch.append({k: (v.detach() if isinstance(v, torch.Tensor) else v) for k, v in out.items()})
            m0 = ctx.models[0]
            if not ctx.use_pipeline_parallel and hasattr(m0, "micro_steps"):
                if m0.micro_steps % ctx.num_micro_batches_per_slot != 0:
                    raise errors.InvalidExperimentException("did not train for gradient accumulation steps "
                                                            "(call engine.backward() and engine.step() once per train_batch)")
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
                reached.add("train")
        host = _metric_utils.materialize_metrics(per_batch)
        if ctx.distributed.size > 1 and ctx._average_training_metrics:
            host = _metric_utils._average_training_metrics(ctx.distributed, host)
        num_inputs *= ctx._mpu.data_parallel_world_size
        metrics = util.make_metrics(num_inputs, host)
        metrics["avg_metrics"].update(_metric_utils._convert_metrics_to_numpy(ctx.reduce_metrics(for_training=True)))
        if self.is_chief:
            logger.info(util.make_timing_log("trained", time.time() - t0, num_inputs, max(len(host), 1)))
            self.core_context.train.report_training_metrics(self.state.batches_trained, metrics["avg_metrics"], metrics["batch_metrics"])
        bm = ctx.distributed.broadcast(metrics if self.is_chief else None)
        for cb in self.callbacks.values():
            cb.on_training_workload_end(avg_metrics=bm["avg_metrics"], batch_metrics=bm["batch_metrics"])
        return reached

    # ------------------------------------------------------------------ validation
    @torch.no_grad()
    def _validate(self, op: Optional[core.SearcherOperation] = None) -> Dict[str, Any]:
        ctx = self.context
        if self.is_chief:
            self.core_context.train.set_status("validating")
        ctx.reset_reducers()
        for m in ctx.models:
            m.eval()
        for cb in self.callbacks.values():
            cb.on_validation_start()
        t0 = time.time()
        it = iter(self.validation_loader) if self.validation_loader is not None else None
        rows: List[Dict[str, Any]] = []
        keys = None
        n = 1 if self.test_mode else int(self.num_validation_batches or 0)
        for cb in self.callbacks.values():
            cb.on_validation_epoch_start()
        for idx in range(n):
            out = self.trial.evaluate_batch(it, idx)
            if ctx._mpu.should_report_metrics:
                if not isinstance(out, dict):
                    raise errors.InvalidExperimentException(f"evaluate_batch must return a dictionary of metrics, got {type(out)}")
                if keys is None:
                    keys = list(out.keys())
                elif list(out.keys()) != keys:
                    raise ValueError(f"Validation metric names must match across all batches of data: {keys} != {list(out.keys())}.")
                rows.append({k: (v.detach() if isinstance(v, torch.Tensor) else v) for k, v in out.items()})
        host = _metric_utils.materialize_metrics(rows)
        for cb in self.callbacks.values():
            cb.on_validation_epoch_end(host)
        # only ranks that report metrics contribute rows; everyone takes part in the collective
        gathered = ctx.distributed.gather((keys, host))
        metrics: Dict[str, Any] = {}
        if self.is_chief:
            assert gathered is not None
            all_rows = [r for k, hs in gathered if k for r in hs]
            kk = next((k for k, _ in gathered if k), None)
            if kk:
                reducers = _metric_utils._prepare_metrics_reducers(self.trial.evaluation_reducer(), kk)
                for i, name in enumerate(kk):
                    col = np.asarray([np.asarray(r[name], dtype=np.float64) for r in all_rows])
                    metrics[name] = _reducer._simple_reduce_metrics(reducers[name], col)
        metrics.update(_metric_utils._convert_metrics_to_numpy(ctx.reduce_metrics(for_training=False)))
        if ctx.distributed.size > 1 and any(util.is_overridden(c.on_validation_end, _callback.PyTorchCallback) for c in self.callbacks.values()):
            metrics = ctx.distributed.broadcast(metrics)
        for cb in self.callbacks.values():
            cb.on_validation_end(metrics)
        self.state.last_val = self.state.batches_trained
        best_before = None
        if self.is_chief:
            n_in = n * ctx.train_micro_batch_size_per_gpu * ctx._mpu.data_parallel_world_size
            logger.info(util.make_timing_log("validated", time.time() - t0, n_in, n))
            best_before = self.core_context.train.get_experiment_best_validation()
            self.core_context.train.report_validation_metrics(self.state.batches_trained, metrics)
        sm = None
        if self.is_chief and op is not None:
            if self.searcher_metric_name:
                if self.searcher_metric_name not in metrics:
                    raise RuntimeError(f"Search method is configured to use metric '{self.searcher_metric_name}' but model "
                                       f"definition returned validation metrics {list(metrics.keys())}.")
                sm = metrics[self.searcher_metric_name]
            if self._remaining(self._op_length(op)) < 1 and not op._completed:
                op.report_completed(sm)
        should = False
        if self.is_chief and self.state.last_ckpt != self.state.batches_trained:
            if self.ckpt_policy == "all":
                should = True
            elif self.ckpt_policy == "best" and sm is not None:
                should = best_before is None or ((float(sm) < best_before) == self.smaller_is_better)
        if ctx.distributed.broadcast(should):
            self._checkpoint(already_exiting=False)
        return metrics

    # ------------------------------------------------------------------ checkpoints
    def _checkpoint(self, already_exiting: bool) -> None:
        if self.is_chief:
            self.core_context.train.set_status("checkpointing")
        self.state.last_ckpt = self.state.batches_trained
        self.state.step_id += 1
        md = {"steps_completed": self.state.batches_trained, "determined_version": det.__version__,
              "framework": f"torch-{torch.__version__}", "format": "pickle"} if self.is_chief else None
        # every rank writes its shard (optimizer partitions live on all ranks)
        with self.core_context.checkpoint.store_path(md, shard=True) as (path, uuid):
            self._save(pathlib.Path(path))
        for cb in self.callbacks.values():
            cb.on_checkpoint_upload_end(uuid=uuid)
        if not already_exiting and (self.core_context.preempt.should_preempt() or self.context.get_stop_requested()):
            raise ShouldExit()

    def _save(self, path: pathlib.Path) -> None:
        path.mkdir(parents=True, exist_ok=True)
        rank = self.context.distributed.rank
        if self.is_chief:
            util.write_user_code(path, not self.local_training)
            with (path / "trial_state.pkl").open("wb") as f:
                pickle.dump(vars(self.state), f)
            trial_cls = type(self.trial)
            (path / "load_data.json").write_text(util.json_encode({
                "trial_type": "DeepSpeedTrial", "experiment_config": self.context._exp_conf, "hparams": self.context._hparams,
                "trial_cls_spec": f"{trial_cls.__module__}:{trial_cls.__qualname__}"}))
        rng = {"cpu_rng_state": torch.random.get_rng_state(), "np_rng_state": np.random.get_state(), "random_rng_state": random.getstate()}
        if self.context.device.type == "cuda":
            rng["gpu_rng_state"] = torch.cuda.get_rng_state(self.context.device)
        ck = {"rng_state": rng, "callbacks": {n: cb.state_dict() for n, cb in self.callbacks.items()}}
        for cb in self.callbacks.values():
            cb.on_checkpoint_save_start(ck)
        torch.save(ck, str(path / f"det_state_dict_rank{rank}.pth"))
        self.trial.save(self.context, path)
        for cb in self.callbacks.values():
            cb.on_checkpoint_write_end(str(path))

    def _load(self, load_path: pathlib.Path) -> None:
        rank = self.context.distributed.rank
        f = load_path / f"det_state_dict_rank{rank}.pth"
        if not f.exists():
            f = load_path / "det_state_dict_rank0.pth"
        if f.exists():
            ck = torch.load(str(f), map_location="cpu", weights_only=False)
            for cb in self.callbacks.values():
                cb.on_checkpoint_load_start(ck)
            rng = ck.get("rng_state")
            if rng:
                np.random.set_state(rng["np_rng_state"])
                random.setstate(rng["random_rng_state"])
                torch.random.set_rng_state(rng["cpu_rng_state"])
                if self.context.device.type == "cuda" and "gpu_rng_state" in rng:
                    torch.cuda.set_rng_state(rng["gpu_rng_state"], device=self.context.device)
            for name, cb in self.callbacks.items():
                if name in ck.get("callbacks", {}):
                    cb.load_state_dict(ck["callbacks"][name])
        self.trial.load(self.context, load_path)
        ts, wl = load_path / "trial_state.pkl", load_path / "workload_sequencer.pkl"
        if ts.exists():
            st = pickle.loads(ts.read_bytes())
            if st.get("trial_id") == self.trial_id:
                self.state = _TrialState(**{k: st[k] for k in ("trial_id", "last_ckpt", "step_id", "last_val", "batches_trained", "epochs_trained") if k in st})
        elif wl.exists():
            st = pickle.loads(wl.read_bytes())
            if st.get("trial_id") == self.trial_id:
                done = int(st.get("steps_completed") or 0)
                self.state = _TrialState(trial_id=self.trial_id, last_ckpt=st.get("last_ckpt", 0), last_val=st.get("last_val", 0),
                                         step_id=st.get("step_id", 0), batches_trained=done,
                                         epochs_trained=done // self.context._epoch_len if self.context._epoch_len else 0)


class DeepSpeedTrial(metaclass=abc.ABCMeta):
    """Subclass to train with a DeepSpeed-style engine.  ``train_batch`` / ``evaluate_batch``
    receive the data **iterator** (so pipeline engines can pull several micro batches)."""

    trial_controller_class = DeepSpeedTrialController
    trial_context_class = DeepSpeedTrialContext

    @abc.abstractmethod
    def __init__(self, context: DeepSpeedTrialContext) -> None:
        """Build the engine(s) (``determined_b200.parallel.zero.initialize`` or
        ``deepspeed.initialize``) and register them with ``context.wrap_model_engine``."""

    @abc.abstractmethod
    def train_batch(self, dataloader_iter: Optional[Iterator[Any]], epoch_idx: int, batch_idx: int) -> Union[torch.Tensor, Dict[str, Any]]:
        """Process one micro batch: ``batch = next(dataloader_iter)``; forward; ``engine.backward(loss)``;
        ``engine.step()``.  Pipeline engines call ``engine.train_batch(dataloader_iter)`` instead."""

    @abc.abstractmethod
    def build_training_data_loader(self) -> Any:
        ...

    @abc.abstractmethod
    def build_validation_data_loader(self) -> Any:
        ...

    def build_callbacks(self) -> Dict[str, _callback.PyTorchCallback]:
        return {}

    @abc.abstractmethod
    def evaluate_batch(self, dataloader_iter: Optional[Iterator[Any]], batch_idx: int) -> Dict[str, Any]:
        ...

    def evaluation_reducer(self) -> Union[_reducer.Reducer, Dict[str, _reducer.Reducer]]:
        return _reducer.Reducer.AVG

    def save(self, context: DeepSpeedTrialContext, path: pathlib.Path) -> None:
        """Default: every wrapped engine saves under ``path`` with tag ``model{i}``."""
        for i, m in enumerate(context.models):
            m.save_checkpoint(path, tag=f"model{i}")

    def load(self, context: DeepSpeedTrialContext, path: pathlib.Path) -> None:
        for i, m in enumerate(context.models):
            m.load_checkpoint(path, tag=f"model{i}")
