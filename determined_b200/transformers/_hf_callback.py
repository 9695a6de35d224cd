"""``DetCallback`` — plug a HuggingFace ``Trainer`` into the Core API (BASELINE config #3).

Parity: reference ``harness/determined/transformers/_hf_callback.py``: metric reporting ``on_log``
``:50-103``; sharded ``checkpoint.upload(selector=checkpoint-N/)`` ``on_save`` ``:115-146``; resume
through ``checkpoint.download`` ``:158-180``; searcher progress/completion ``:182-265``;
preemption → ``control.should_save`` ``:97-103``; searcher/TrainingArguments compatibility checks
``:267-317``.

HF's own DDP wrapper does the gradient exchange inside ``Trainer``; ``determined_b200.parallel.ddp.install_hf_comm_hook(trainer)``
adds a callback that registers a DDP communication hook on the wrapped model, so that every gradient bucket goes through
this repository's fused all-reduce instead of NCCL (``FusedAllReduceHook``; any ``DistributedDataParallel`` module can use it
directly through ``install_comm_hook(ddp_model)``).
"""
import logging
import os
from typing import Any, Dict, List, Optional, Tuple

import determined_b200 as det

try:  # transformers is a soft dependency
    import transformers
    _Base = transformers.TrainerCallback
except Exception:  # pragma: no cover
    transformers = None  # type: ignore
    _Base = object

logger = logging.getLogger("determined.transformers")

TRAIN, EVAL, TEST, TRAIN_AVG = "train_progress", "eval", "test", "train"


def get_metric_type(d: Dict[str, Any]) -> str:
    """HF logs mix metric families in one dict; classify by key prefix."""
    for k in d:
        if k.startswith("eval"):
            return EVAL
        if k.startswith("test"):
            return TEST
        if k.startswith("train_runtime") or k.startswith("train_samples"):
            return TRAIN_AVG
    return TRAIN


class DetCallback(_Base):  # type: ignore[misc, valid-type]
    def __init__(self, core_context: det.core.Context, args: Any, filter_metrics: Optional[List[str]] = None,
                 user_data: Optional[Dict[str, Any]] = None) -> None:
        super().__init__()
        self.core_context = core_context
        self.filter_metrics = filter_metrics
        self.user_data = user_data
        self.load_last_checkpoint(args)
        self.last_metrics: Dict[str, Any] = {"train_step": -1, "eval_step": -1}
        self.searcher_ops = self.core_context.searcher.operations()
        self.current_op = next(self.searcher_ops)
        self.updating_searcher = False
        info = det.get_cluster_info()
        if info is not None and info.task_type == "TRIAL":
            sc = info.trial._config["searcher"]
            self.searcher_metric = sc.get("metric")
            if sc.get("name") == "custom" or not isinstance(sc.get("max_length"), dict):
                self.searcher_unit, self.searcher_max_length = "batches", self.current_op.length
            else:
                (self.searcher_unit, self.searcher_max_length), = sc["max_length"].items()
                self._check_searcher_compatibility(args)
        else:   # local run against Dummy contexts
            self.searcher_metric, self.searcher_unit, self.searcher_max_length = None, "batches", self.current_op.length

    # -- metrics -----------------------------------------------------------------------------------
    def _get_metrics(self, logs: Dict[str, Any]) -> Tuple[Dict[str, Any], str]:
        kind = get_metric_type(logs)
        if self.filter_metrics:
            logs = {k: v for k, v in logs.items() if any(m in k for m in self.filter_metrics)}
        return dict(logs), kind

    def on_log(self, args: Any, state: Any, control: Any, logs: Optional[Dict[str, Any]] = None, **kwargs: Any) -> None:
        if logs is None:
            logger.warning("on_log called with empty logs")
            return
        metrics, kind = self._get_metrics(logs)
        if kind == TRAIN:
            if self.last_metrics["train_step"] != state.global_step:
                if state.is_world_process_zero:
                    self.core_context.train.report_training_metrics(steps_completed=state.global_step, metrics=metrics)
                metrics["train_step"] = state.global_step
        elif kind == EVAL:
            if self.last_metrics["eval_step"] != state.global_step:
                if state.is_world_process_zero:
                    self.core_context.train.report_validation_metrics(steps_completed=state.global_step, metrics=metrics)
                metrics["eval_step"] = state.global_step
        else:
            logger.debug("metrics of type %s are not reported", kind)
        self.last_metrics.update(metrics)
        if self.updating_searcher:
            self._update_searcher(state, control)
        elif self.core_context.preempt.should_preempt():
            control.should_save = True     # checkpoint first; on_save() then exits

    # -- checkpoints -------------------------------------------------------------------------------
    def on_save(self, args: Any, state: Any, control: Any, **kwargs: Any) -> None:
        info = det.get_cluster_info()
        local_path = os.path.join(args.output_dir, f"checkpoint-{state.global_step}")
        if state.is_world_process_zero and self.user_data is not None:
            self._on_save_user_data(local_path)
        md = {"steps_completed": state.global_step}
        if info is not None and info.task_type == "TRIAL":
            md["trial_id"] = info.trial.trial_id

        def selector(x: str) -> bool:
            return x.startswith((f"checkpoint-{state.global_step}/", "runs/"))

        self.core_context.checkpoint.upload(args.output_dir, metadata=md, shard=True, selector=selector)
        if self.core_context.preempt.should_preempt():
            raise SystemExit(0)

    def _on_save_user_data(self, save_path: str) -> None:
        """Override to persist ``self.user_data`` next to the HF checkpoint."""

    def load_last_checkpoint(self, args: Any) -> None:
        info = det.get_cluster_info()
        latest = info.latest_checkpoint if info is not None else None
        if latest is None:
            return
        if getattr(args, "overwrite_output_dir", False):
            logger.info("Skip load last checkpoint from Determined due to overwrite_output_dir=True.")
            return
        md = self.core_context.checkpoint.get_metadata(latest)
        prev_trial, trial_id = md.get("trial_id"), info.trial.trial_id
        if trial_id != prev_trial:
            logger.info("New trial runs from the checkpoint of trial %s: only weights are restored.", prev_trial)
        cpath = f"checkpoint-{md['steps_completed']}"
        self.core_context.checkpoint.download(latest, args.output_dir)
        ckpt_dir = os.path.join(args.output_dir, cpath)
        if os.path.isdir(ckpt_dir):
            args.resume_from_checkpoint = ckpt_dir

    # -- searcher ----------------------------------------------------------------------------------
    def on_step_end(self, args: Any, state: Any, control: Any, **kwargs: Any) -> None:
        if state.is_world_process_zero and args.logging_steps and state.global_step % int(args.logging_steps) == 0:
            self.current_op.report_progress(state.epoch if self.searcher_unit == "epochs" else state.global_step)
        if self.searcher_unit == "batches" and state.global_step >= self.current_op.length:
            self._update_searcher(state, control)

    def on_epoch_end(self, args: Any, state: Any, control: Any, **kwargs: Any) -> None:
        if self.searcher_unit == "epochs":
            if state.is_world_process_zero:
                self.current_op.report_progress(state.epoch)
            if state.epoch is not None and state.epoch >= self.current_op.length:
                self._update_searcher(state, control)

    def _metrics_reported(self, step: int) -> bool:
        return self.last_metrics["eval_step"] == step and self.last_metrics["train_step"] == step

    def _wait_for_metrics(self, control: Any) -> None:
        # ask the Trainer to log + evaluate + save now; on_log will finish the searcher op
        control.should_evaluate = True
        control.should_log = True
        control.should_save = True
        self.updating_searcher = True

    def _update_searcher(self, state: Any, control: Any) -> None:
        if not self._metrics_reported(state.global_step):
            self._wait_for_metrics(control)
            return
        if state.is_world_process_zero and self.searcher_metric is None:     # local run: nothing is searching
            self.current_op.report_completed(float(self.last_metrics.get("eval_loss", 0.0)))
        elif state.is_world_process_zero:
            if self.searcher_metric not in self.last_metrics:
                raise ValueError(f"Searcher metric {self.searcher_metric} was not reported by the Trainer; available: "
                                 f"{sorted(self.last_metrics)} — make sure an evaluation dataset and compute_metrics are set.")
            self.current_op.report_completed(self.last_metrics[self.searcher_metric])
        self.updating_searcher = False
        try:
            self.current_op = next(self.searcher_ops)
        except StopIteration:
            control.should_training_stop = True

    def _check_searcher_compatibility(self, args: Any) -> None:
        if self.searcher_unit == "batches":
            if getattr(args, "max_steps", -1) == -1:
                self._log_config_mismatch("epochs", getattr(args, "num_train_epochs", None))
            elif args.max_steps != self.searcher_max_length:
                self._log_config_mismatch("batches", args.max_steps)
        elif self.searcher_unit == "epochs":
            if getattr(args, "max_steps", -1) != -1:
                self._log_config_mismatch("batches", args.max_steps)
            elif getattr(args, "num_train_epochs", None) != self.searcher_max_length:
                self._log_config_mismatch("epochs", getattr(args, "num_train_epochs", None))

    def _log_config_mismatch(self, trainer_units: str, trainer_len: Any) -> None:
        logger.warning("Searcher configuration does not match HF Trainer configuration: searcher uses %s=%s, Trainer uses %s=%s. "
                       "The searcher configuration wins: training stops when the searcher's last operation completes.",
                       self.searcher_unit, self.searcher_max_length, trainer_units, trainer_len)
