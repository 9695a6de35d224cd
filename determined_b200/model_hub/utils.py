"""Helpers shared by model_hub trials (reference ``model_hub/model_hub/utils.py``)."""
from typing import Any, Dict, List, Union

import numpy as np
import torch


def expand_like(arrays: List[np.ndarray], fill: float = -100) -> np.ndarray:
    """Stack arrays whose trailing dimension differs by right-padding dim 1 with ``fill``."""
    full_shape = list(arrays[0].shape)
    if len(full_shape) == 1:
        return np.concatenate(arrays)
    full_shape[0] = sum(a.shape[0] for a in arrays)
    full_shape[1] = max(a.shape[1] for a in arrays)
    out = np.full(full_shape, fill)
    row = 0
    for a in arrays:
        out[row:row + a.shape[0], : a.shape[1]] = a
        row += a.shape[0]
    return out


def numpify(x: Union[List, np.ndarray, torch.Tensor]) -> np.ndarray:
    if isinstance(x, np.ndarray):
        return x
    if isinstance(x, list):
        return np.array(x)
    if isinstance(x, torch.Tensor):
        return x.detach().float().cpu().numpy() if x.dtype == torch.bfloat16 else x.detach().cpu().numpy()
    raise TypeError(f"expected List, np.ndarray or torch.Tensor, got {type(x)}")


def download_url(download_directory: str, url: str) -> str:
    """Fetch ``url`` into ``download_directory`` (kept under the URL's file name; an existing file is reused)."""
    import os
    import urllib.parse

    import requests

    name = os.path.basename(urllib.parse.urlparse(url).path) or "download"
    os.makedirs(download_directory, exist_ok=True)
    path = os.path.join(download_directory, name)
    if os.path.exists(path):
        return path
    tmp = path + ".part"
    with requests.get(url, stream=True, timeout=60) as r:
        r.raise_for_status()
        with open(tmp, "wb") as f:
            for chunk in r.iter_content(1 << 20):
                f.write(chunk)
    os.replace(tmp, path)
    return path


def compute_num_training_steps(experiment_config: Dict[str, Any], global_batch_size: int) -> int:
    max_length = experiment_config["searcher"]["max_length"]
    if isinstance(max_length, int):
        return max_length
    unit, value = next(iter(max_length.items()))
    if unit == "batches":
        return int(value)
    if unit == "epochs":
        if "records_per_epoch" not in experiment_config:
            raise ValueError("Missing num_training_steps hyperparameter in the experiment configuration, which is needed to configure the "
                             "learning rate scheduler when max_length is given in epochs without records_per_epoch.")
        return int(value) * (int(experiment_config["records_per_epoch"]) // global_batch_size)      # whole batches per epoch
    if unit == "records":
        return int(value / global_batch_size)
    raise ValueError(f"unknown max_length unit {unit}")


class AttrDict(dict):
    """dict with attribute access, applied recursively."""

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        for k, v in list(self.items()):
            if isinstance(v, dict) and not isinstance(v, AttrDict):
                self[k] = AttrDict(v)

    def __getattr__(self, item: str) -> Any:
        try:
            return self[item]
        except KeyError as e:
            raise AttributeError(item) from e

    def __setattr__(self, item: str, value: Any) -> None:
        self[item] = AttrDict(value) if isinstance(value, dict) and not isinstance(value, AttrDict) else value
