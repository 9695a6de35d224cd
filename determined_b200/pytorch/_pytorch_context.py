"""``PyTorchTrialContext`` — what a ``PyTorchTrial`` uses to wrap its models/optimizers and to run
``backward`` / ``step_optimizer``.

Parity: reference ``harness/determined/pytorch/_pytorch_context.py`` (public surface kept):
``wrap_model`` ``:285-312``, ``wrap_optimizer(optimizer, backward_passes_per_step,
fp16_compression, average_aggregated_gradients)`` ``:314-380``, ``wrap_lr_scheduler`` ``:382-424``,
``set_profiler`` ``:426-461``, ``to_device`` ``:496-503``, ``wrap_scaler`` ``:505-545``,
``configure_apex_amp`` ``:547-689``, ``backward`` ``:708-803``, ``step_optimizer`` ``:814-910``,
``_should_communicate_and_update`` ``:701-706``, epoch/batch accessors ``:912-977``,
tensorboard writer ``:985-1052``.

What is different underneath: there is no DDP/Horovod wrapper.  ``wrap_optimizer`` hands the
optimizer to a gradient engine (:mod:`determined_b200.parallel.ddp`); on CUDA that engine keeps
flat symmetric gradient buffers, overlaps our fused P2P all-reduce with backward, and turns
``optimizer.step()`` into one fused kernel per param group.
"""
import contextlib
import logging
import pathlib
import time
from typing import Any, Callable, Dict, Iterator, List, Optional, Set, Type, Union

import torch
import torch.nn as nn

from determined_b200 import core, errors, util
from determined_b200.parallel import ddp
from determined_b200.pytorch import _data, _experimental, _lr_scheduler, _reducer

logger = logging.getLogger("determined.pytorch")


class PyTorchTrialContext(_reducer._PyTorchReducerContext):
    """Contains runtime information for a ``PyTorchTrial`` and the wrap/step API."""

    def __init__(self, core_context: core.Context, trial_seed: Optional[int], hparams: Optional[Dict],
                 slots_per_trial: int, num_gpus: int, exp_conf: Optional[Dict[str, Any]],
                 aggregation_frequency: int, steps_completed: int, managed_training: bool,
                 debug_enabled: bool, enable_tensorboard_logging: bool = True) -> None:
        self._core = core_context
        self.distributed = core_context.distributed
        _reducer._PyTorchReducerContext.__init__(self, self.distributed.allgather)
        if hparams and hparams.get("global_batch_size") is not None:
            self._per_slot_batch_size, self._global_batch_size = util.calculate_batch_sizes(
                hparams, slots_per_trial, "PyTorchTrial")
        else:
            self._per_slot_batch_size = self._global_batch_size = None
        self._hparams = hparams or {}
        self._num_gpus = num_gpus
        self._debug_enabled = debug_enabled
        self._exp_conf = exp_conf
        self._trial_seed = trial_seed
        self._steps_completed = steps_completed
        self._managed_training = managed_training
        self._aggregation_frequency = max(1, int(aggregation_frequency))
        self.device = self._init_device()

        self.models: List[nn.Module] = []
        self.optimizers: List[torch.optim.Optimizer] = []
        self.lr_schedulers: List[_lr_scheduler.LRScheduler] = []
        self.profiler: Any = None
        self._epoch_len: Optional[int] = None
        self._optimizer_configs: Dict[int, Dict[str, Any]] = {}
        self._scaler: Any = None
        self._use_apex = False
        self._current_batch_idx: Optional[int] = None
        self._last_backward_batch_idx: Optional[int] = None
        self._fp16_compression_default = False
        self._average_aggregated_gradients_default = True
        self._is_pre_trainer = False
        self._stop_requested = False
        self._tbd_writer: Any = None
        self._enable_tensorboard_logging = enable_tensorboard_logging
        self._to_device_warned_types: Set[Type] = set()
        self.experimental = _experimental.PyTorchExperimentalContext(self)
        self._engine: Optional[ddp.GradEngineBase] = None

    # ------------------------------------------------------------------ info accessors
    def get_global_batch_size(self) -> int:
        if self._global_batch_size is None:
            raise ValueError("global_batch_size is undefined: no `global_batch_size` hyperparameter was given")
        return self._global_batch_size

    def get_per_slot_batch_size(self) -> int:
        if self._per_slot_batch_size is None:
            raise ValueError("per_slot_batch_size is undefined: no `global_batch_size` hyperparameter was given")
        return self._per_slot_batch_size

    def get_experiment_config(self) -> Dict[str, Any]:
        if self._exp_conf is None:
            raise ValueError("exp_conf must be set to use context.get_experiment_config")
        return self._exp_conf

    def get_hparam(self, name: str) -> Any:
        if name not in self.get_hparams():
            raise ValueError(f"Could not find name '{name}' in experiment hyperparameters. "
                             "Please check your experiment configuration 'hyperparameters' section.")
        if name == "global_batch_size":
            logger.warning("Please use `context.get_per_slot_batch_size()` and `context.get_global_batch_size()` "
                           "instead of accessing `global_batch_size` directly.")
        return self.get_hparams()[name]

    def get_hparams(self) -> Dict[str, Any]:
        return self._hparams

    def get_stop_requested(self) -> bool:
        return self._stop_requested

    def set_stop_requested(self, stop_requested: bool) -> None:
        """Ask the harness to checkpoint and stop this trial at the next opportunity."""
        if not isinstance(stop_requested, bool):
            raise AssertionError("stop_requested must be a bool")
        logger.info("A trial stoppage has been requested. The trial will be stopped at the end of the current step.")
        self._stop_requested = stop_requested

    def set_enable_tensorboard_logging(self, enable_tensorboard_logging: bool) -> None:
        self._enable_tensorboard_logging = enable_tensorboard_logging

    def get_enable_tensorboard_logging(self) -> bool:
        return self._enable_tensorboard_logging

    def get_trial_seed(self) -> int:
        if self._trial_seed is None:
            raise ValueError("trial_seed must be set to use context.get_trial_seed")
        return self._trial_seed

    def get_initial_batch(self) -> int:
        return self._steps_completed

    def get_data_config(self) -> Dict[str, Any]:
        return self.get_experiment_config().get("data", {})

    def get_experiment_id(self) -> int:
        return int(self._core.info.trial.experiment_id) if self._core.info and self._core.info._trial_info else 0

    def get_trial_id(self) -> int:
        return int(self._core.info.trial.trial_id) if self._core.info and self._core.info._trial_info else 0

    def is_epoch_start(self) -> bool:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        if self._epoch_len is None:
            raise errors.InternalException("Training DataLoader uninitialized.")
        return self._current_batch_idx % self._epoch_len == 0

    def is_epoch_end(self) -> bool:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        if self._epoch_len is None:
            raise errors.InternalException("Training DataLoader uninitialized.")
        return self._current_batch_idx % self._epoch_len == self._epoch_len - 1

    def current_train_epoch(self) -> int:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        if self._epoch_len is None:
            raise errors.InternalException("Training DataLoader uninitialized.")
        return self._current_batch_idx // self._epoch_len

    def current_train_batch(self) -> int:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        return self._current_batch_idx

    # ------------------------------------------------------------------ device / engine
    def _init_device(self) -> torch.device:
        if self._num_gpus > 0 and torch.cuda.is_available():
            idx = self.distributed.local_rank if self.distributed.size > 1 else 0
            idx = idx % max(torch.cuda.device_count(), 1)
            dev = torch.device("cuda", idx)
            torch.cuda.set_device(dev)
            return dev
        return torch.device("cpu")

    def _get_engine(self) -> ddp.GradEngineBase:
        if self._engine is None:
            self._engine = ddp.make_engine(self.distributed, self.device, self._aggregation_frequency,
                                           self.experimental._fused_engine, self.experimental._bucket_cap_mb,
                                           overlap=self.experimental._grad_overlap)
        return self._engine

    def _check_pre_wrap(self, what: str) -> None:
        if self.models or self.optimizers:
            raise errors.InvalidExperimentException(
                f"context.experimental.{what} must be called before wrap_model()/wrap_optimizer()")

    def _set_default_gradient_compression(self, gradient_compression: bool) -> None:
        self._fp16_compression_default = gradient_compression

    def _set_default_average_aggregated_gradients(self, average_aggregated_gradients: bool) -> None:
        self._average_aggregated_gradients_default = average_aggregated_gradients

    def _set_is_pre_trainer(self) -> None:
        self._is_pre_trainer = True

    # ------------------------------------------------------------------ wrap API
    def autocast_forward_pass(self, to_wrap: nn.Module) -> nn.Module:
        """Run ``to_wrap.forward`` under autocast (used by ``experimental.use_amp``)."""
        ctx = self
        orig_forward = to_wrap.forward

        def forward(*args: Any, **kwargs: Any) -> Any:
            dtype = ctx.experimental._amp_dtype or torch.float16
            with torch.autocast(device_type=ctx.device.type, dtype=dtype,
                                enabled=(ctx.device.type == "cuda" or dtype == torch.bfloat16)):
                return orig_forward(*args, **kwargs)

        to_wrap.forward = forward  # type: ignore[method-assign]
        return to_wrap

    def wrap_model(self, model: nn.Module) -> nn.Module:
        """Move the model to this worker's device and register it for checkpointing.  The returned
        object is the same ``nn.Module`` (no DDP wrapper: gradient exchange is owned by the engine
        behind :meth:`wrap_optimizer`)."""
        if self._use_apex:
            raise errors.InvalidExperimentException("Must call wrap_model() before configure_apex_amp.")
        model = model.to(self.device)
        if self.device.type == "cuda" and self.experimental._fused_engine is not False:
            # module-level fusion pass of the B200 engine: hot library ops whose ATen kernels are far
            # off the HBM roofline are swapped for native ones (same parameters / state-dict keys)
            from determined_b200.ops import batchnorm
            from determined_b200.ops import conv as conv_ops

            nc = conv_ops.convert_conv_(model)
            if nc:
                logger.info("fused engine: %d Conv2d modules read bf16 weights from the engine's mirror under bf16 autocast and "
                            "hand their weight gradients to the engine's multi-tensor accumulation", nc)
            n = batchnorm.convert_batchnorm_(model)
            if n:
                nb = batchnorm.fuse_residual_blocks_(model)
                logger.info("fused engine: %d BatchNorm2d modules now run the native channels-last kernels "
                            "(%d residual blocks with ReLU / add folded into them)", n, nb)
        if self.experimental._auto_amp:
            model = self.autocast_forward_pass(model)
        self.models.append(model)
        return model

    def wrap_optimizer(self, optimizer: torch.optim.Optimizer, backward_passes_per_step: int = 1,
                       fp16_compression: Optional[bool] = None,
                       average_aggregated_gradients: Optional[bool] = None) -> torch.optim.Optimizer:
        """Register an optimizer (its parameters must belong to models passed to
        :meth:`wrap_model`).  ``backward_passes_per_step`` multiplies the number of backward
        passes per ``step_optimizer`` call; ``fp16_compression`` selects a 16-bit wire format for
        the gradient exchange; ``average_aggregated_gradients`` divides by
        ``aggregation_frequency``."""
        if self._use_apex:
            raise errors.InvalidExperimentException("Must call wrap_optimizer() before configure_apex_amp.")
        if backward_passes_per_step < 1:
            raise errors.InvalidExperimentException("backward_passes_per_step for local gradient aggregation must be >= 1")
        if fp16_compression is None:
            fp16_compression = self._fp16_compression_default or self.experimental._wire_dtype in (torch.bfloat16, torch.float16)
        if average_aggregated_gradients is None:
            average_aggregated_gradients = self._average_aggregated_gradients_default
        model_params = {id(p) for m in self.models for p in m.parameters()}
        for g in optimizer.param_groups:
            for p in g["params"]:
                if self.models and id(p) not in model_params:
                    raise errors.InvalidExperimentException(
                        "wrap_optimizer() got an optimizer whose parameters do not belong to a wrapped model; "
                        "call wrap_model() first and build the optimizer from the wrapped model's parameters")
        self._optimizer_configs[id(optimizer)] = {
            "backward_passes_per_step": backward_passes_per_step,
            "fp16_compression": bool(fp16_compression),
            "average_aggregated_gradients": bool(average_aggregated_gradients),
        }
        optimizer = self._get_engine().add_optimizer(optimizer, bool(average_aggregated_gradients),
                                                     backward_passes_per_step=backward_passes_per_step,
                                                     fp16_compression=bool(fp16_compression), models=self.models)
        self.optimizers.append(optimizer)
        return optimizer

    def wrap_lr_scheduler(self, lr_scheduler: Any, step_mode: _lr_scheduler.LRScheduler.StepMode,
                          frequency: int = 1) -> Any:
        """Register an LR scheduler built on a wrapped optimizer; the harness steps it according
        to ``step_mode`` and checkpoints it."""
        if isinstance(lr_scheduler, torch.optim.lr_scheduler.ReduceLROnPlateau):
            if step_mode != _lr_scheduler.LRScheduler.StepMode.MANUAL_STEP:
                raise errors.InvalidExperimentException(
                    "torch.optim.lr_scheduler.ReduceLROnPlateau needs the validation metric in step(), so it can "
                    "only be used with step_mode=MANUAL_STEP: call lr_scheduler.step(metric) yourself, e.g. from a "
                    "PyTorchCallback.on_validation_end().")
        opt = getattr(lr_scheduler, "optimizer", None)
        if opt is not None and all(opt is not o for o in self.optimizers):
            raise errors.InvalidExperimentException("Must use an optimizer that is returned by wrap_optimizer()")
        wrapped = _lr_scheduler.LRScheduler(lr_scheduler, step_mode, frequency)
        self.lr_schedulers.append(wrapped)
        return lr_scheduler

    def wrap_scaler(self, scaler: Any) -> Any:
        """Register a ``torch.amp.GradScaler`` so it is checkpointed and updated by the harness."""
        if self._scaler is not None:
            raise errors.InvalidExperimentException("Please only call wrap_scaler or use_amp once.")
        if self._use_apex:
            raise errors.InvalidExperimentException("Do not mix APEX with PyTorch AMP.")
        self._scaler = scaler
        return scaler

    def configure_apex_amp(self, models: Any, optimizers: Any, **kwargs: Any) -> Any:
        """apex.amp is deprecated upstream (reference ``:625-630``) and absent from the B200 stack.
        This shim maps the request onto native autocast: ``opt_level`` O1/O2/O3 → bf16 autocast of
        the models' forward passes; loss scaling is unnecessary for bf16."""
        level = kwargs.get("opt_level", "O1")
        logger.warning("configure_apex_amp(opt_level=%s) is mapped to native bf16 autocast on B200", level)
        if level != "O0":
            self.experimental._auto_amp = True
            self.experimental._amp_dtype = torch.bfloat16
            for m in (models if isinstance(models, (list, tuple)) else [models]):
                self.autocast_forward_pass(m)
        return models, optimizers

    def set_profiler(self, *args: Any, **kwargs: Any) -> None:
        """Configure a ``torch.profiler.profile`` that the harness enters around training and steps
        per batch; traces go to the TensorBoard directory."""
        kwargs.setdefault("on_trace_ready",
                          torch.profiler.tensorboard_trace_handler(str(self.get_tensorboard_path())))
        self.profiler = torch.profiler.profile(*args, **kwargs)

    def to_device(self, data: _data._Data) -> _data.TorchData:
        return _data.to_device(data, self.device, self._to_device_warned_types)

    # ------------------------------------------------------------------ backward / step
    def _should_communicate_and_update(self) -> bool:
        if not self._managed_training:
            return True
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        return (self._current_batch_idx + 1) % self._aggregation_frequency == 0

    def backward(self, loss: torch.Tensor, gradient: Optional[torch.Tensor] = None,
                 retain_graph: bool = False, create_graph: bool = False) -> None:
        """``loss.backward(...)`` with the harness' gradient exchange attached.  On batches that
        are not aggregation boundaries gradients only accumulate locally.  With distributed
        training each parameter's gradient may be computed once per batch."""
        if self._scaler is not None and self.experimental._auto_amp:
            loss = self._scaler.scale(loss)
        engine = self._get_engine()
        engine.backward_begin(self._should_communicate_and_update(), scaler=self._scaler)
        loss.backward(gradient=gradient, retain_graph=retain_graph, create_graph=create_graph)
        engine.backward_end()

    @staticmethod
    def _average_gradients(parameters: Any, divisor: int) -> None:
        if divisor == 1:
            return
        inv = 1.0 / float(divisor)
        for p in parameters:
            if p.grad is not None:
                p.grad.mul_(inv)

    def step_optimizer(self, optimizer: torch.optim.Optimizer,
                       clip_grads: Optional[Callable[[Iterator], None]] = None,
                       auto_zero_grads: bool = True, scaler: Optional[Any] = None) -> None:
        """One optimization step (a no-op except on aggregation boundaries).  ``clip_grads`` is a
        callable taking the parameters (e.g. ``lambda ps: clip_grad_norm_(ps, 1.0)``) or a
        :class:`determined_b200.pytorch.FusedClipGradNorm`."""
        if self._aggregation_frequency > 1 and not auto_zero_grads:
            raise errors.InvalidExperimentException(
                "if optimizations.aggregation_frequency is larger than 1, auto_zero_grads must be set to true.")
        if id(optimizer) not in self._optimizer_configs:
            raise errors.InvalidExperimentException("step_optimizer() needs an optimizer returned by wrap_optimizer()")
        if not self._should_communicate_and_update():
            return
        if scaler is None and self.experimental._auto_amp:
            scaler = self._scaler
        self._get_engine().step(optimizer, clip_grads, auto_zero_grads, scaler,
                                unscale_for_clip=bool(self._scaler is not None and self.experimental._auto_amp))

    # ------------------------------------------------------------------ tensorboard
    def get_tensorboard_path(self) -> pathlib.Path:
        return self._core.train.get_tensorboard_path()

    def get_tensorboard_writer(self) -> Any:
        """A ``torch.utils.tensorboard.SummaryWriter`` on the harness' TensorBoard directory."""
        if self._tbd_writer is None:
            from torch.utils.tensorboard import SummaryWriter

            self._tbd_writer = SummaryWriter(str(self.get_tensorboard_path()))
        return self._tbd_writer

    def _maybe_reset_tbd_writer(self) -> None:
        if self._tbd_writer is not None:
            self._tbd_writer.flush()

    def _close(self) -> None:
        if self._engine is not None:
            self._engine.close()
