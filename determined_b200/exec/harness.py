"""``python -m determined.exec.harness module:TrialClass`` — run a legacy class entrypoint.

Parity: reference ``harness/determined/exec/harness.py:26-201``: load the trial class from the
model directory; PyTorchTrial goes through ``pytorch.init`` + ``Trainer.fit`` with periods derived
from the experiment config (``:133-193``: ``min_validation_period``, ``min_checkpoint_period``,
``scheduling_unit`` as reporting period, ``checkpoint_policy``, ``perform_initial_validation``,
``profiling.enabled``); DeepSpeedTrial goes through its controller (``:26-130``)."""
import faulthandler
import logging
import sys
from typing import Any, Type

import determined_b200 as det
from determined_b200 import load

logger = logging.getLogger("determined.exec.harness")


def _run_pytorch_trial(trial_class: Type[Any], info: det.ClusterInfo) -> int:
    from determined_b200 import pytorch

    cfg = det.ExperimentConfig(info.trial._config)
    with pytorch.init() as ctx:
        if ctx.distributed.rank != 0:
            logging.getLogger().setLevel(logging.WARNING)
        opt = cfg.get("optimizations") or {}
        if opt.get("fused_engine") is not None or opt.get("bucket_cap_mb") is not None or opt.get("grad_overlap") is not None:
            ctx.experimental.use_fused_engine(opt.get("fused_engine"), bucket_cap_mb=float(opt.get("bucket_cap_mb") or 25.0),
                                              overlap=opt.get("grad_overlap"))
        if opt.get("cuda_graphs"):
            ctx.experimental.enable_cuda_graphs(True)
        if (cfg.get("optimizations") or {}).get("async_checkpoint"):
            ctx.experimental.enable_async_checkpoint(True)
        trial = trial_class(ctx)
        gbs = ctx._global_batch_size
        def unit(d: Any) -> pytorch.TrainUnit:
            d = d or {"batches": 0}
            if isinstance(d, int):
                d = {"batches": d}
            return pytorch.TrainUnit._from_values(**d, global_batch_size=gbs)
        pytorch.Trainer(trial, ctx).fit(
            checkpoint_period=unit(cfg.get_min_checkpoint_period()),
            validation_period=unit(cfg.get_min_validation_period()),
            reporting_period=pytorch.Batch(cfg.scheduling_unit()),
            checkpoint_policy=cfg.get_checkpoint_policy(),
            step_zero_validation=cfg.get_perform_initial_validation(),
            profiling_enabled=cfg.profiling_enabled())
    return 0


def _run_deepspeed_trial(trial_class: Type[Any], info: det.ClusterInfo) -> int:
    from determined_b200.pytorch import deepspeed as det_ds

    return det_ds._run_trial_class(trial_class, info)


def main(entrypoint: str) -> int:
    logging.basicConfig(level=logging.INFO, format=det.LOG_FORMAT)
    info = det.get_cluster_info()
    assert info is not None and info.task_type == "TRIAL", "exec.harness runs inside trial tasks"
    if info.trial._debug:
        logging.getLogger().setLevel(logging.DEBUG)
        faulthandler.dump_traceback_later(30, repeat=True)
    trial_class = load.trial_class_from_entrypoint(entrypoint)
    from determined_b200 import pytorch

    if issubclass(trial_class, pytorch.PyTorchTrial):
        return _run_pytorch_trial(trial_class, info)
    try:
        from determined_b200.pytorch import deepspeed as det_ds

        if issubclass(trial_class, det_ds.DeepSpeedTrial):
            return _run_deepspeed_trial(trial_class, info)
    except ImportError:
        pass
    raise det.errors.InvalidExperimentException(f"{entrypoint} is not a PyTorchTrial or DeepSpeedTrial")


if __name__ == "__main__":
    if len(sys.argv) != 2:
        print("usage: python -m determined.exec.harness module:TrialClass", file=sys.stderr)
        sys.exit(2)
    sys.exit(main(sys.argv[1]))
