"""Cross-rank metric combination for training and validation.

Parity: reference ``harness/determined/pytorch/_metric_utils.py``: ``_combine_and_average_training_metrics``
``:27-62``, ``_prepare_metrics_reducers`` ``:110-140``, ``_reduce_metrics`` ``:143-181``,
``_log_tb_metrics`` ``:184-211``; ``_convert_metrics_to_numpy`` ``:214-231``.

B200 note: the controller keeps per-batch metric tensors on the device and calls
:func:`materialize_metrics` once per *reporting boundary* — one stacked D2H — instead of the
reference's ``metric.cpu().detach().numpy()`` sync on every batch.
"""
from typing import Any, Dict, List, Optional, Tuple, Union

import numpy as np
import torch

from determined_b200 import util
from determined_b200.pytorch import _reducer


def materialize_metrics(per_batch: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
    """Convert a list of per-batch metric dicts (values may be 0-d device tensors) to host
    scalars/arrays with a single synchronising transfer for all scalar tensors."""
    slots: List[Tuple[int, str]] = []
    scalars: List[torch.Tensor] = []
    out: List[Dict[str, Any]] = []
    for i, m in enumerate(per_batch):
        row: Dict[str, Any] = {}
        for k, v in m.items():
            if isinstance(v, torch.Tensor):
                if v.numel() == 1:
                    slots.append((i, k))
                    scalars.append(v.detach().reshape(()).to(torch.float32))
                    row[k] = None
                else:
                    row[k] = v.detach().cpu().numpy()
            else:
                row[k] = v
        out.append(row)
    if scalars:
        host = torch.stack(scalars).cpu().tolist()
        for (i, k), val in zip(slots, host):
            out[i][k] = val
    return out


def _convert_metrics_to_numpy(metrics: Dict[str, Any]) -> Dict[str, Any]:
    return {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in metrics.items()}


def _average_training_metrics(dist: Any, per_batch_metrics: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
    """Gather every rank's per-batch metrics on the chief; batch ``i``'s metrics become the mean
    over ranks of each numeric metric.  Non-chief ranks get their input back."""
    gathered = dist.gather(per_batch_metrics)
    if dist.rank != 0:
        return per_batch_metrics
    assert gathered is not None
    out: List[Dict[str, Any]] = []
    for batch_rows in zip(*gathered):
        row: Dict[str, Any] = {}
        for k in batch_rows[0]:
            vals = [r[k] for r in batch_rows if k in r and r[k] is not None]
            try:
                row[k] = float(np.mean(np.asarray(vals, dtype=np.float64))) if all(np.ndim(v) == 0 for v in vals) \
                    else np.mean(np.stack([np.asarray(v) for v in vals]), axis=0)
            except (TypeError, ValueError):
                row[k] = batch_rows[0][k]
        out.append(row)
    return out


def _combine_and_average_training_metrics(dist: Any, per_batch_metrics: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
    return _average_training_metrics(dist, per_batch_metrics)


def _prepare_metrics_reducers(reducer: Union[_reducer.Reducer, Dict[str, _reducer.Reducer]],
                              keys: Any) -> Dict[str, _reducer.Reducer]:
    if isinstance(reducer, _reducer.Reducer):
        return {k: reducer for k in keys}
    if isinstance(reducer, dict):
        missing = set(keys) - set(reducer)
        if missing:
            raise ValueError(f"evaluation_reducer() is missing reducers for metrics {sorted(missing)}")
        for k, r in reducer.items():
            if not isinstance(r, _reducer.Reducer):
                raise TypeError(f"reducer for {k!r} must be a pytorch.Reducer")
        return {k: reducer[k] for k in keys}
    raise TypeError("evaluation_reducer() must return a pytorch.Reducer or a dict of them")


def _reduce_metrics(dist: Any, batch_metrics: np.ndarray, keys: Any,
                    metrics_reducers: Dict[str, _reducer.Reducer],
                    num_batches: Optional[List[int]] = None) -> Dict[str, Any]:
    """``batch_metrics`` is [n_batches, n_keys] on this rank; gather on the chief and reduce each
    column with its reducer.  Only the chief returns values."""
    if dist.size > 1:
        all_m = dist.gather(batch_metrics)
        all_n = dist.gather(num_batches)
        if dist.rank != 0:
            return {}
        assert all_m is not None and all_n is not None
        batch_metrics = np.concatenate([np.asarray(m).reshape(-1, len(list(keys))) for m in all_m], axis=0)
        num_batches = None if any(n is None for n in all_n) else [x for n in all_n for x in n]
    keys = list(keys)
    return {k: _reducer._simple_reduce_metrics(metrics_reducers[k], np.asarray(batch_metrics)[:, i], num_batches)
            for i, k in enumerate(keys)}


def _log_tb_metrics(writer: Any, metric_type: str, steps_completed: int, metrics: Dict[str, Any],
                    batch_metrics: Optional[List[Dict[str, Any]]] = None) -> None:
    if writer is None:
        return
    if metric_type == "val":
        writer.on_validation_step_end(steps_completed, metrics)
    else:
        writer.on_train_step_end(steps_completed, metrics, batch_metrics)
