"""``determined.pytorch.deepspeed`` (reference ``harness/determined/pytorch/deepspeed/__init__.py:1-10``)
plus ``init``/``Trainer`` helpers mirroring ``determined.pytorch`` for local runs."""
import contextlib
import logging
import os
import random
from typing import Any, Dict, Iterator, Optional, Type

import torch

import determined_b200 as det
from determined_b200 import core
from determined_b200.pytorch._pytorch_trial import Batch, Epoch, TrainUnit
from determined_b200.pytorch.deepspeed._mpu import ModelParallelUnit, make_data_parallel_mpu, make_deepspeed_mpu
from determined_b200.pytorch.deepspeed._deepspeed_context import DeepSpeedTrialContext, overwrite_deepspeed_config
from determined_b200.pytorch.deepspeed._deepspeed_trial import DeepSpeedTrial, DeepSpeedTrialController
from determined_b200.parallel.zero import ZeroEngine, initialize, init_distributed  # engine playing DeepSpeed's role

logger = logging.getLogger("determined.pytorch.deepspeed")


@contextlib.contextmanager
def init(*, hparams: Optional[Dict] = None, exp_conf: Optional[Dict[str, Any]] = None,
         distributed: Optional[core.DistributedContext] = None, require_gpu: bool = True) -> Iterator[DeepSpeedTrialContext]:
    """Create a ``DeepSpeedTrialContext`` (local or on-cluster), joining the process group first."""
    info = det.get_cluster_info()
    local = info is None or info.task_type != "TRIAL"
    seed = random.randint(0, 1 << 31) if local else info.trial.trial_seed
    DeepSpeedTrialController.pre_execute_hook(seed)
    if distributed is None and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        distributed = core.DistributedContext.from_deepspeed()
    if not local:
        hparams, exp_conf = info.trial.hparams, info.trial._config
    steps = 0 if local else int(info.trial._steps_completed)
    with core.init(distributed=distributed, preempt_mode=core.PreemptMode.WorkersAskChief,
                   tensorboard_mode=core.TensorboardMode.MANUAL) as cc:
        yield DeepSpeedTrialContext(cc, seed, hparams, cc.distributed.size, torch.cuda.device_count(), exp_conf, steps,
                                    require_gpu=require_gpu)


class Trainer:
    def __init__(self, trial: DeepSpeedTrial, context: DeepSpeedTrialContext) -> None:
        self._trial, self._context = trial, context

    def fit(self, max_length: Optional[TrainUnit] = None, validation_period: Optional[TrainUnit] = None,
            checkpoint_period: Optional[TrainUnit] = None, reporting_period: Optional[TrainUnit] = None,
            checkpoint_policy: str = "best", latest_checkpoint: Optional[str] = None, test_mode: bool = False) -> None:
        info = det.get_cluster_info()
        local = info is None or info.task_type != "TRIAL"
        kw: Dict[str, Any] = {}
        if local:
            if max_length is None and not test_mode:
                raise ValueError("max_length must be defined in local training mode.")
            max_length = max_length or Batch(1)
        else:
            cfg = det.ExperimentConfig(info.trial._config)
            kw = {"searcher_metric_name": cfg.get("searcher", {}).get("metric"), "smaller_is_better": cfg.get_smaller_is_better(),
                  "steps_completed": int(info.trial._steps_completed)}
            latest_checkpoint = latest_checkpoint or info.latest_checkpoint
            max_length = None
        DeepSpeedTrialController(self._trial, self._context, max_length=max_length, validation_period=validation_period,
                                 checkpoint_period=checkpoint_period, reporting_period=reporting_period,
                                 checkpoint_policy=checkpoint_policy, latest_checkpoint=latest_checkpoint,
                                 local_training=local, test_mode=test_mode, **kw).run()


def _run_trial_class(trial_class: Type[DeepSpeedTrial], info: det.ClusterInfo) -> int:
    """``exec.harness`` entry for ``entrypoint: module:DeepSpeedTrialSubclass``."""
    cfg = det.ExperimentConfig(info.trial._config)
    with init() as ctx:
        if ctx.distributed.rank != 0:
            logging.getLogger().setLevel(logging.WARNING)
        trial = trial_class(ctx)

        def unit(d: Any) -> TrainUnit:
            d = d or {"batches": 0}
            if isinstance(d, int):
                d = {"batches": d}
            gbs = ctx.models[0].train_batch_size() if ctx.models and hasattr(ctx.models[0], "train_batch_size") else None
            return TrainUnit._from_values(**d, global_batch_size=gbs)

        Trainer(trial, ctx).fit(validation_period=unit(cfg.get_min_validation_period()),
                                checkpoint_period=unit(cfg.get_min_checkpoint_period()),
                                reporting_period=Batch(cfg.scheduling_unit()), checkpoint_policy=cfg.get_checkpoint_policy())
    return 0
