"""``pytorch.init()`` and ``Trainer`` — the user-facing way to run a ``PyTorchTrial``.

Parity: reference ``harness/determined/pytorch/_trainer.py``: ``Trainer.fit(checkpoint_period,
validation_period, max_length, reporting_period, checkpoint_policy, latest_checkpoint,
step_zero_validation, test_mode, profiling_enabled)`` ``:70-203``;
``_initialize_distributed_backend`` ``:206-227``; ``init(hparams, exp_conf, distributed,
aggregation_frequency, enable_tensorboard_logging)`` context manager ``:240-310``.
"""
import contextlib
import logging
import os
import random
import sys
from typing import Any, Dict, Iterator, Optional

import numpy as np
import torch
import torch.distributed as dist

import determined_b200 as det
from determined_b200 import core, errors
from determined_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_b200.pytorch._pytorch_trial import Batch, Epoch, PyTorchTrial, TrainUnit, _PyTorchTrialController

logger = logging.getLogger("determined.pytorch")


class Trainer:
    """Runs a ``PyTorchTrial`` built on a context from :func:`init`, locally or on-cluster."""

    def __init__(self, trial: PyTorchTrial, context: PyTorchTrialContext):
        self._trial = trial
        self._context = context
        self._core = context._core
        self._info = det.get_cluster_info()
        self._local_training = self._info is None or self._info.task_type != "TRIAL"
        self._controller: Optional[_PyTorchTrialController] = None

    def fit(self, checkpoint_period: Optional[TrainUnit] = None, validation_period: Optional[TrainUnit] = None,
            max_length: Optional[TrainUnit] = None, reporting_period: TrainUnit = Batch(100),
            checkpoint_policy: str = "best", latest_checkpoint: Optional[str] = None,
            step_zero_validation: bool = False, test_mode: bool = False,
            profiling_enabled: bool = False) -> None:
        """Train.  Periods default to "never" (a final validation + checkpoint always happen).
        ``max_length`` is required locally and forbidden on-cluster (the searcher decides);
        ``latest_checkpoint`` resumes from a storage id; ``test_mode`` runs one batch of
        train/validate/checkpoint to shake out errors."""
        if checkpoint_policy not in ("all", "best", "none"):
            raise ValueError(f"checkpoint_policy must be all|best|none, got {checkpoint_policy!r}")
        checkpoint_period = checkpoint_period or Batch(sys.maxsize)
        validation_period = validation_period or Batch(sys.maxsize)
        for name, u in (("checkpoint_period", checkpoint_period), ("validation_period", validation_period),
                        ("reporting_period", reporting_period)):
            if not isinstance(u, TrainUnit):
                raise TypeError(f"{name} must be a pytorch.Batch or pytorch.Epoch")

        if self._local_training:
            if max_length is None and not test_mode:
                raise ValueError("max_length must be defined in local training mode.")
            if max_length is None:
                max_length = Batch(1)
            if not isinstance(max_length.value, int):
                raise TypeError("max_length must be configured in TrainUnit(int) types.")
            if profiling_enabled:
                logger.warning("Profiling is not supported in local training mode.")
                profiling_enabled = False
            smaller_is_better, searcher_metric_name = True, None
            steps_completed = 0
            global_batch_size = None
        else:
            if test_mode:
                raise ValueError("test_mode is only supported in local training mode.")
            if max_length is not None:
                logger.warning("max_length is ignored when training on-cluster; the searcher config decides.")
                max_length = None
            assert self._info is not None
            cfg = det.ExperimentConfig(self._info.trial._config)
            smaller_is_better = cfg.get_smaller_is_better()
            searcher_metric_name = cfg.get("searcher", {}).get("metric")
            steps_completed = int(self._info.trial._steps_completed)
            global_batch_size = self._context._global_batch_size
            latest_checkpoint = latest_checkpoint or self._info.latest_checkpoint

        self._controller = _PyTorchTrialController(
            trial_inst=self._trial, context=self._context, checkpoint_period=checkpoint_period,
            validation_period=validation_period, reporting_period=reporting_period,
            smaller_is_better=smaller_is_better, steps_completed=steps_completed,
            latest_checkpoint=latest_checkpoint, local_training=self._local_training, test_mode=test_mode,
            searcher_metric_name=searcher_metric_name, checkpoint_policy=checkpoint_policy,
            step_zero_validation=step_zero_validation, max_length=max_length,
            global_batch_size=global_batch_size, profiling_enabled=profiling_enabled)
        self._controller.run()

    def configure_profiler(self, *args: Any, **kwargs: Any) -> None:
        self._context.set_profiler(*args, **kwargs)


def _initialize_distributed_backend() -> Optional[core.DistributedContext]:
    """If launched by ``torch.distributed.run`` / our launcher (``RANK``+``WORLD_SIZE`` set, or
    ``USE_TORCH_DISTRIBUTED``), join the process group and build the DistributedContext."""
    env = os.environ
    if "RANK" in env and "WORLD_SIZE" in env and int(env["WORLD_SIZE"]) > 1:
        if not dist.is_initialized():
            use_cuda = torch.cuda.is_available()
            if use_cuda:
                torch.cuda.set_device(int(env.get("LOCAL_RANK", "0")) % torch.cuda.device_count())
            dist.init_process_group(backend="nccl" if use_cuda else "gloo")
        return core.DistributedContext.from_torch_distributed()
    info = det.get_cluster_info()
    if info is not None and len(info.container_addrs) > 1 or (info is not None and len(info.slot_ids) > 1):
        raise errors.InvalidExperimentException(
            "this task has more than one slot but no distributed launcher was used; set the experiment entrypoint to "
            "`python3 -m determined.launch.torch_distributed -- python3 train.py`")
    return None


def _generate_local_seed() -> int:
    return random.randint(0, 1 << 31)


def _set_random_seeds(seed: int) -> None:
    random.seed(seed)
    np.random.seed(seed % (2**32))
    torch.random.manual_seed(seed)


@contextlib.contextmanager
def init(*, hparams: Optional[Dict] = None, exp_conf: Optional[Dict[str, Any]] = None,
         distributed: Optional[core.DistributedContext] = None, aggregation_frequency: int = 1,
         enable_tensorboard_logging: bool = True) -> Iterator[PyTorchTrialContext]:
    """Create a ``PyTorchTrialContext``.

    On-cluster, ``hparams``/``exp_conf`` come from the trial info and must not be passed;
    ``aggregation_frequency`` falls back to ``optimizations.aggregation_frequency``.  Locally all
    arguments are optional."""
    info = det.get_cluster_info()
    local = info is None or info.task_type != "TRIAL"
    if distributed is None:
        distributed = _initialize_distributed_backend()
    if local:
        trial_seed = _generate_local_seed()
        steps_completed = 0
        managed = True
        debug = False
        num_gpus = torch.cuda.device_count()
        slots = distributed.size if distributed is not None else 1
        if exp_conf is not None:
            opt = det.ExperimentConfig(exp_conf).get_optimizations_config()
            if aggregation_frequency == 1:
                aggregation_frequency = int(opt.get("aggregation_frequency", 1))
    else:
        assert info is not None
        if hparams is not None or exp_conf is not None:
            logger.warning("hparams/exp_conf passed to pytorch.init() are ignored on-cluster")
        hparams = info.trial.hparams
        exp_conf = info.trial._config
        trial_seed = info.trial.trial_seed
        steps_completed = int(info.trial._steps_completed)
        managed = True
        debug = bool(info.trial._debug)
        num_gpus = len(info.gpu_uuids) if torch.cuda.is_available() else 0
        if torch.cuda.is_available() and num_gpus == 0:
            num_gpus = torch.cuda.device_count()
        cfg = det.ExperimentConfig(exp_conf)
        slots = cfg.slots_per_trial()
        opt = cfg.get_optimizations_config()
        if aggregation_frequency == 1:
            aggregation_frequency = int(opt.get("aggregation_frequency", 1))
    _set_random_seeds(trial_seed)

    with core.init(distributed=distributed, preempt_mode=core.PreemptMode.WorkersAskChief,
                   tensorboard_mode=core.TensorboardMode.MANUAL) as core_context:
        context = PyTorchTrialContext(
            core_context=core_context, trial_seed=trial_seed, hparams=hparams, slots_per_trial=slots,
            num_gpus=num_gpus, exp_conf=exp_conf, aggregation_frequency=aggregation_frequency,
            steps_completed=steps_completed, managed_training=managed, debug_enabled=debug,
            enable_tensorboard_logging=enable_tensorboard_logging)
        if exp_conf is not None:
            opt = det.ExperimentConfig(exp_conf).get_optimizations_config()
            context._set_default_gradient_compression(bool(opt.get("gradient_compression", False)))
            context._set_default_average_aggregated_gradients(bool(opt.get("average_aggregated_gradients", True)))
        try:
            yield context
        finally:
            context._close()
