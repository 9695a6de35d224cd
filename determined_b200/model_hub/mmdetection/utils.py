"""Helpers of the MMDetection adapter.

Parity: reference ``model_hub/model_hub/mmdetection/utils.py`` (model-zoo lookup of pretrained checkpoints for a
config file, checkpoint download, fp16 loss-scaler construction).  On B200 mixed precision is bf16 autocast, so
``build_fp16_loss_scaler`` only has to interpret the config (``"dynamic"`` / a number / a dict) into the arguments of
``torch.amp.GradScaler`` for users that still want fp16.
"""
import os
import re
from typing import Any, Dict, Optional, Tuple

import yaml


def get_config_pretrained_url_mapping(config_dir: Optional[str] = None) -> Dict[str, str]:
    """``{config path relative to the mmdet config dir: checkpoint URL}`` from every ``metafile.yml`` of the model zoo
    under ``config_dir`` (default ``$MMDETECTION_CONFIG_DIR``)."""
    config_dir = config_dir or os.environ.get("MMDETECTION_CONFIG_DIR", "/mmdetection/configs")
    out: Dict[str, str] = {}
    if not os.path.isdir(config_dir):
        return out
    for root, _, files in os.walk(config_dir):
        for f in files:
            if f not in ("metafile.yml", "metafile.yaml"):
                continue
            try:
                with open(os.path.join(root, f)) as fh:
                    meta = yaml.safe_load(fh) or {}
            except (OSError, yaml.YAMLError):
                continue
            for m in meta.get("Models") or []:
                cfg, weights = m.get("Config"), m.get("Weights")
                if cfg and weights:
                    out[re.sub(r"^configs/", "", str(cfg))] = str(weights)
    return out


def get_pretrained_ckpt_path(download_directory: str, config_file: str, config_dir: Optional[str] = None) -> Tuple[Optional[str], Optional[Dict[str, Any]]]:
    """Local path of the zoo checkpoint for ``config_file`` (downloaded into ``download_directory`` if needed) and the
    loaded checkpoint, or ``(None, None)`` when the zoo lists none / there is no network."""
    mapping = get_config_pretrained_url_mapping(config_dir)
    key = re.sub(r"^.*configs/", "", config_file)
    url = mapping.get(key)
    if url is None:
        return None, None
    os.makedirs(download_directory, exist_ok=True)
    path = os.path.join(download_directory, os.path.basename(url))
    if not os.path.exists(path):
        try:
            import requests

            with requests.get(url, stream=True, timeout=30) as r:
                r.raise_for_status()
                with open(path + ".part", "wb") as f:
                    for chunk in r.iter_content(1 << 20):
                        f.write(chunk)
            os.replace(path + ".part", path)
        except Exception:  # noqa: BLE001 - offline boxes train from scratch
            return None, None
    import torch

    return path, torch.load(path, map_location="cpu", weights_only=False)


def build_fp16_loss_scaler(loss_scale: Any) -> Dict[str, Any]:
    """Keyword arguments for ``torch.amp.GradScaler`` from an mmcv ``fp16.loss_scale`` setting."""
    if loss_scale == "dynamic" or loss_scale is None:
        return {}
    if isinstance(loss_scale, (int, float)):
        return {"init_scale": float(loss_scale), "growth_interval": 1 << 62}       # static scale: never grows
    if isinstance(loss_scale, dict):
        allowed = {"init_scale", "growth_factor", "backoff_factor", "growth_interval", "enabled"}
        unknown = set(loss_scale) - allowed
        if unknown:
            raise ValueError(f"unsupported loss_scale keys: {sorted(unknown)}")
        return dict(loss_scale)
    raise ValueError(f"loss_scale must be 'dynamic', a number or a dict, got {loss_scale!r}")
