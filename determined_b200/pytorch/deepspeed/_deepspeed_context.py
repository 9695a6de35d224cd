"""``DeepSpeedTrialContext`` (reference ``harness/determined/pytorch/deepspeed/_deepspeed_context.py``):
``overwrite_deepspeed_config`` ``:18-42``, ``wrap_model_engine`` ``:178-209``, ``set_mpu``
``:148-176``, ``disable_auto_grad_accumulation`` ``:211-218``, rejection of expconf
``mixed_precision``/``aggregation_frequency`` ``:113-146``, GPU requirement ``:264-273``.

``wrap_model_engine`` accepts any engine with the DeepSpeedEngine surface: our
:class:`determined_b200.parallel.zero.ZeroEngine` (the B200 path) or a real DeepSpeed engine if
the user installed one."""
import json
import logging
import pathlib
from typing import Any, Dict, List, Optional, Set, Type, Union

import torch

import determined_b200 as det
from determined_b200 import core, errors, util
from determined_b200.pytorch import _data, _reducer
from determined_b200.pytorch.deepspeed import _mpu

logger = logging.getLogger("determined.pytorch.deepspeed")


def overwrite_deepspeed_config(base_ds_config: Union[str, Dict], source_ds_dict: Dict[str, Any]) -> Dict[str, Any]:
    """Overlay ``source_ds_dict`` (e.g. the ``overwrite_deepspeed_args`` hyperparameter) on a
    DeepSpeed config given as a dict or a path to a JSON file."""
    if isinstance(base_ds_config, str):
        def no_dups(pairs: Any) -> Dict[str, Any]:
            d: Dict[str, Any] = {}
            for k, v in pairs:
                if k in d:
                    raise ValueError(f"Duplicate key in DeepSpeed config: {k}")
                d[k] = v
            return d

        with open(base_ds_config) as f:
            base_ds_config = json.load(f, object_pairs_hook=no_dups)
    elif not isinstance(base_ds_config, dict):
        raise TypeError("Expected string or dict for base_ds_config argument.")
    return util.merge_dicts(base_ds_config, source_ds_dict)


def _is_pipeline_engine(engine: Any) -> bool:
    return bool(getattr(engine, "is_pipe_parallel", False)) or type(engine).__name__ == "PipelineEngine"


class DeepSpeedTrialContext(_reducer._PyTorchReducerContext):
    def __init__(self, core_context: core.Context, trial_seed: Optional[int], hparams: Optional[Dict], slots_per_trial: int,
                 num_gpus: int, exp_conf: Optional[Dict[str, Any]], steps_completed: int, enable_tensorboard_logging: bool = True,
                 require_gpu: bool = True) -> None:
        self._core = core_context
        self.distributed = core_context.distributed
        _reducer._PyTorchReducerContext.__init__(self, self.distributed.allgather)
        self._hparams = hparams or {}
        self._exp_conf = exp_conf
        self._trial_seed = trial_seed
        self._num_gpus = num_gpus
        self._steps_completed = steps_completed
        self._require_gpu = require_gpu
        self.device = self._init_device()
        self.models: List[Any] = []
        self.profiler: Any = None
        self._epoch_len: Optional[int] = None
        self._mpu = _mpu.make_data_parallel_mpu(self.distributed)
        self._called_set_mpu = False
        self._train_micro_batch_size_per_gpu: Optional[int] = None
        self._num_micro_batches_per_slot: Optional[int] = None
        self._use_pipeline_parallel = False
        self._data_repro_checks_disabled = False
        self._manual_grad_accumulation = False
        self._current_batch_idx: Optional[int] = None
        self._tbd_writer: Any = None
        self._enable_tensorboard_logging = enable_tensorboard_logging
        self._stop_requested = False
        self._to_device_warned_types: Set[Type] = set()
        self._average_training_metrics = True
        self._check_experiment_config_optimizations()

    def _check_experiment_config_optimizations(self) -> None:
        """``optimizations.mixed_precision`` / ``aggregation_frequency`` belong to the engine
        config in DeepSpeed-style trials; reject them in the experiment config."""
        if not self._exp_conf:
            return
        opt = self._exp_conf.get("optimizations") or {}
        self._average_training_metrics = bool(opt.get("average_training_metrics", True))
        if opt.get("mixed_precision", "O0") != "O0":
            raise errors.InvalidExperimentException(
                "Mixed precision must be specified through the engine (DeepSpeed-style) config, not optimizations.mixed_precision")
        if int(opt.get("aggregation_frequency", 1)) > 1:
            raise errors.InvalidExperimentException(
                "Gradient aggregation is specified through gradient_accumulation_steps in the engine config, "
                "not optimizations.aggregation_frequency")

    def _init_device(self) -> torch.device:
        if torch.cuda.is_available() and self._num_gpus > 0:
            idx = self.distributed.local_rank % torch.cuda.device_count()
            torch.cuda.set_device(idx)
            return torch.device("cuda", idx)
        if self._require_gpu:
            raise errors.InvalidExperimentException("GPUs are required for DeepSpeedTrial.")
        return torch.device("cpu")

    # -- public API ------------------------------------------------------------------------------------
    def set_mpu(self, mpu: _mpu.ModelParallelUnit) -> None:
        """Use a custom model-parallel topology (e.g. tensor parallel ranks that are not
        data-parallel peers).  Must precede ``wrap_model_engine``-derived decisions; may be called
        once; ignored with a warning under pipeline parallelism (the engine's grid wins)."""
        if self._called_set_mpu:
            raise errors.InvalidExperimentException("Only one MPU can be passed to DeepSpeedTrialContext. set_mpu was already called.")
        if self._use_pipeline_parallel:
            logger.warning("Pipeline parallelism in use: the engine's own MPU is kept; set_mpu() ignored.")
            return
        self._called_set_mpu = True
        self._mpu = mpu

    def wrap_model_engine(self, model: Any) -> Any:
        """Register a model engine (tracked for checkpointing; the first one defines micro batch
        size and gradient accumulation steps)."""
        if hasattr(model, "to"):
            model = model.to(self.device) or model
        if _is_pipeline_engine(model):
            self._use_pipeline_parallel = True
            if not self.models:
                self._mpu = _mpu.make_deepspeed_mpu(model.mpu)
            else:
                logger.warning("Using the MPU corresponding to the first wrapped model engine.")
        if not self.models:
            self._train_micro_batch_size_per_gpu = int(model.train_micro_batch_size_per_gpu())
            self._num_micro_batches_per_slot = int(model.gradient_accumulation_steps())
        elif int(model.train_micro_batch_size_per_gpu()) != self._train_micro_batch_size_per_gpu:
            logger.warning("Train micro batch size for wrapped model engine %d does not match that of the first engine; "
                           "sample counts are reported for engine 1 only.", len(self.models) + 1)
        self.models.append(model)
        return model

    def disable_auto_grad_accumulation(self) -> None:
        """The controller will call ``train_batch`` once per batch instead of once per micro batch;
        the trial then consumes ``gradient_accumulation_steps`` micro batches itself."""
        self._manual_grad_accumulation = True

    def disable_dataset_reproducibility_checks(self) -> None:
        self._data_repro_checks_disabled = True

    @property
    def use_pipeline_parallel(self) -> bool:
        return self._use_pipeline_parallel

    @property
    def train_micro_batch_size_per_gpu(self) -> int:
        if self._train_micro_batch_size_per_gpu is None:
            raise errors.InvalidExperimentException("Please call wrap_model_engine before accessing train_micro_batch_size_per_gpu.")
        return self._train_micro_batch_size_per_gpu

    @property
    def num_micro_batches_per_slot(self) -> int:
        if self._num_micro_batches_per_slot is None:
            raise errors.InvalidExperimentException("Please call wrap_model_engine before accessing num_micro_batches_per_slot.")
        return self._num_micro_batches_per_slot

    def to_device(self, data: Any) -> Any:
        return _data.to_device(data, self.device, self._to_device_warned_types)

    def get_hparams(self) -> Dict[str, Any]:
        return self._hparams

    def get_hparam(self, name: str) -> Any:
        if name not in self._hparams:
            raise ValueError(f"Could not find name '{name}' in experiment hyperparameters.")
        return self._hparams[name]

    def get_experiment_config(self) -> Dict[str, Any]:
        return self._exp_conf or {}

    def get_data_config(self) -> Dict[str, Any]:
        return (self._exp_conf or {}).get("data", {})

    def get_trial_seed(self) -> int:
        return int(self._trial_seed or 0)

    def get_trial_id(self) -> int:
        return int(self._core.info.trial.trial_id) if self._core.info and self._core.info._trial_info else 0

    def get_experiment_id(self) -> int:
        return int(self._core.info.trial.experiment_id) if self._core.info and self._core.info._trial_info else 0

    def get_initial_batch(self) -> int:
        return self._steps_completed

    def set_stop_requested(self, stop_requested: bool) -> None:
        self._stop_requested = bool(stop_requested)

    def get_stop_requested(self) -> bool:
        return self._stop_requested

    def _need(self) -> None:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        if self._epoch_len is None:
            raise errors.InternalException("Training DataLoader uninitialized.")

    def is_epoch_start(self) -> bool:
        self._need()
        return self._current_batch_idx % self._epoch_len == 0

    def is_epoch_end(self) -> bool:
        self._need()
        return self._current_batch_idx % self._epoch_len == self._epoch_len - 1

    def current_train_epoch(self) -> int:
        self._need()
        return self._current_batch_idx // self._epoch_len

    def current_train_batch(self) -> int:
        if self._current_batch_idx is None:
            raise errors.InternalException("Training hasn't started.")
        return self._current_batch_idx

    def set_profiler(self, *args: Any, **kwargs: Any) -> None:
        kwargs.setdefault("on_trace_ready", torch.profiler.tensorboard_trace_handler(str(self.get_tensorboard_path())))
        self.profiler = torch.profiler.profile(*args, **kwargs)

    def get_tensorboard_path(self) -> pathlib.Path:
        return self._core.train.get_tensorboard_path()

    def get_tensorboard_writer(self) -> Any:
        if self._tbd_writer is None:
            from torch.utils.tensorboard import SummaryWriter

            self._tbd_writer = SummaryWriter(str(self.get_tensorboard_path()))
        return self._tbd_writer

    def _maybe_reset_tbd_writer(self) -> None:
        if self._tbd_writer is not None:
            self._tbd_writer.flush()

    def set_enable_tensorboard_logging(self, enable_tensorboard_logging: bool) -> None:
        self._enable_tensorboard_logging = enable_tensorboard_logging

    def get_enable_tensorboard_logging(self) -> bool:
        return self._enable_tensorboard_logging
