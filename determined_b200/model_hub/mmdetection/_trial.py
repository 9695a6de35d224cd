"""``MMDetTrial``: train any MMDetection config as a PyTorchTrial.

Parity: reference ``model_hub/model_hub/mmdetection/_trial.py:45-281``.  Hyperparameters:

``config_file``            path of an mmdet config (``$MMDETECTION_CONFIG_DIR`` is prepended to relative paths)
``merge_config``           optional second config merged over it
``override_mmdet_config``  ``{"dotted.key": value}`` overrides (``cfg.merge_from_dict``)
``use_pretrained``         load the checkpoint the model zoo lists for ``config_file``
``data_backend``           ``{"type": "s3" | "gcs" | "fake", ...}``: swap the image file backend

mmdet and mmcv are not part of this image: constructing the trial without them raises a clear ImportError, while the
pieces that do not depend on them (samplers, LR schedules, file backends, config helpers) import and are tested.
"""
import logging
import os
from typing import Any, Dict, List

import torch

from determined_b200 import pytorch
from determined_b200.model_hub.mmdetection import _callbacks, _data, _data_backends

logger = logging.getLogger("determined.model_hub.mmdetection")


def _require_mm() -> Any:
    try:
        import mmcv
        import mmdet  # noqa: F401
    except ImportError as e:
        raise ImportError("MMDetTrial needs the mmcv and mmdet packages, which are not installed in this environment "
                          f"({e}); the helpers in determined_b200.model_hub.mmdetection work without them") from e
    return mmcv


def resolve_config_path(config_file: str) -> str:
    base = os.environ.get("MMDETECTION_CONFIG_DIR")
    if base and not os.path.isabs(config_file) and not os.path.exists(config_file):
        return os.path.join(base, config_file)
    return config_file


def flatten_overrides(overrides: Dict[str, Any], prefix: str = "") -> Dict[str, Any]:
    """``{"optimizer": {"lr": 0.1}}`` and ``{"optimizer.lr": 0.1}`` both become ``{"optimizer.lr": 0.1}``."""
    out: Dict[str, Any] = {}
    for k, v in (overrides or {}).items():
        key = f"{prefix}{k}"
        if isinstance(v, dict) and v and not key.endswith("_delete_"):
            out.update(flatten_overrides(v, key + "."))
        else:
            out[key] = v
    return out


class MMDetTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
        self.mmcv = _require_mm()
        import mmdet.models

        self.context = context
        self.hparams = context.get_hparams()
        self.data_config = context.get_data_config() if hasattr(context, "get_data_config") else {}
        self.cfg = self.build_mmdet_config()
        self.model = mmdet.models.build_detector(self.cfg.model)
        self.model.init_weights()
        if self.hparams.get("use_pretrained"):
            from determined_b200.model_hub.mmdetection import utils

            ckpt, _ = utils.get_pretrained_ckpt_path("/tmp", self.hparams["config_file"])
            if ckpt is not None:
                self.mmcv.runner.load_checkpoint(self.model, ckpt, map_location="cpu")
        self.model = context.wrap_model(self.model)
        self.optimizer = context.wrap_optimizer(self.mmcv.runner.build_optimizer(self.model, self.cfg.optimizer))
        fp16 = self.cfg.get("fp16")
        if fp16 is not None:
            self.setup_torch_amp(fp16)
        gc = (self.cfg.get("optimizer_config") or {}).get("grad_clip")
        self.clip_grads_fn = (lambda params: torch.nn.utils.clip_grad_norm_(params, **gc)) if gc else None

    def build_mmdet_config(self) -> Any:
        cfg = self.mmcv.Config.fromfile(resolve_config_path(self.hparams["config_file"]))
        if self.hparams.get("merge_config"):
            cfg.merge_from_dict(dict(self.mmcv.Config.fromfile(self.hparams["merge_config"])))
        if self.hparams.get("override_mmdet_config"):
            cfg.merge_from_dict(flatten_overrides(self.hparams["override_mmdet_config"]))
        cfg.data.val.test_mode = True
        backend = self.data_config.get("file_client_args") or self.hparams.get("data_backend")
        if backend:
            _data_backends.register()
            args = dict(backend)
            args.setdefault("backend", args.pop("type", None))
            _data_backends.sub_backend(args, cfg)
        logger.info("mmdet config:\n%s", cfg.pretty_text)
        return cfg

    def setup_torch_amp(self, fp16_cfg: Any) -> None:
        """bf16 autocast replaces mmcv's fp16 hook: no loss scaling on B200."""
        self.context.experimental.use_amp() if hasattr(self.context.experimental, "use_amp") else None
        self.model.fp16_enabled = False

    def build_callbacks(self) -> Dict[str, pytorch.PyTorchCallback]:
        cbs: Dict[str, pytorch.PyTorchCallback] = {}
        if self.cfg.get("lr_config"):
            n = len(self.train_loader) if getattr(self, "train_loader", None) is not None else 1
            epochs = int((self.cfg.get("runner") or {}).get("max_epochs", self.cfg.get("total_epochs", 12)))
            cbs["lr_updater"] = _callbacks.LrUpdaterCallback(self.context, dict(self.cfg.lr_config), epochs, n)
        return cbs

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        batch = self.to_device(batch)
        losses = self.model(**batch, return_loss=True)
        loss, log_vars = self.model._parse_losses(losses)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip_grads_fn)
        log_vars["lr"] = self.optimizer.param_groups[0]["lr"]
        return {k: (v if torch.is_tensor(v) else torch.tensor(float(v))) for k, v in log_vars.items()}

    def evaluate_batch(self, batch: Any, batch_idx: int) -> Dict[str, Any]:
        batch = self.to_device(batch)
        idx = batch.pop("idx", None)
        with torch.no_grad():
            results = self.model(return_loss=False, rescale=True, **batch)
        self.reducer.update(([int(i) for i in idx] if idx is not None else [], results))
        return {}

    def build_training_data_loader(self) -> pytorch.DataLoader:
        import mmdet.datasets

        ds = mmdet.datasets.build_dataset(self.cfg.data.train)
        d = self.context.distributed
        self.train_loader = _data.build_dataloader(ds, self.context.get_per_slot_batch_size(), self.cfg.data.workers_per_gpu, d.get_size(),
                                                   d.get_rank(), seed=self.context.get_trial_seed() if hasattr(self.context, "get_trial_seed") else 0)
        self.model.CLASSES = getattr(ds, "CLASSES", None)
        return self.train_loader

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        import mmdet.datasets

        ds = mmdet.datasets.build_dataset(self.cfg.data.val, dict(test_mode=True))
        eval_cfg = dict(self.cfg.get("evaluation") or {})
        for k in ("interval", "tmpdir", "start", "gpu_collect", "save_best", "rule", "dynamic_intervals"):
            eval_cfg.pop(k, None)

        def evaluate_fn(per_rank: List[Any]) -> Any:
            order: List[int] = []
            results: List[Any] = []
            for idx, res in per_rank:
                order.extend(idx)
                results.extend(res)
            ranked = [r for _, r in sorted(zip(order, results), key=lambda t: t[0])] if order else results
            return ds.evaluate(ranked, **eval_cfg)

        self.reducer = self.context.wrap_reducer(evaluate_fn, for_training=False)
        d = self.context.distributed
        return _data.build_dataloader(ds, 1, self.cfg.data.workers_per_gpu, d.get_size(), d.get_rank(), test_mode=True)

    def get_batch_length(self, batch: Any) -> int:
        img = batch["img"]
        data = img.data if hasattr(img, "data") else img
        first = data[0] if isinstance(data, (list, tuple)) else data
        return int(first.shape[0]) if torch.is_tensor(first) else len(first)

    def to_device(self, batch: Any) -> Dict[str, Any]:
        """Unwrap mmcv DataContainers (scatter to this one device) and move tensors."""
        try:
            from mmcv.parallel import scatter

            dev = self.context.device
            return scatter(batch, [dev.index if dev.type == "cuda" else -1])[0]
        except ImportError:
            return self.context.to_device(batch)
