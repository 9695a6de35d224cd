"""Dataset helpers of the transformers trials (reference ``model_hub/model_hub/huggingface/_utils.py`` and
``_trial.py:147-204``): column pruning against the model's forward signature, and the default dataset loader.

The ``datasets`` library is used when it is installed.  Without it (this image has none, and no network) data files still
load: :class:`ColumnDataset` reads csv / json / jsonl / txt files into a small column store with the part of the
``datasets.Dataset`` surface the trials use (``column_names``, ``format`` / ``set_format``, ``map``, indexing).
"""
import csv
import inspect
import json
from typing import Any, Callable, Dict, List, Optional, Sequence, Union

from determined_b200.model_hub import utils
from determined_b200.model_hub.huggingface import _config_parser as hf_parse


class ColumnDataset:
    def __init__(self, columns: Dict[str, List[Any]]) -> None:
        n = {len(v) for v in columns.values()}
        if len(n) > 1:
            raise ValueError(f"columns differ in length: { {k: len(v) for k, v in columns.items()} }")
        self._cols = dict(columns)
        self.format: Dict[str, Any] = {"type": None, "columns": list(columns)}

    @classmethod
    def from_rows(cls, rows: Sequence[Dict[str, Any]]) -> "ColumnDataset":
        names: List[str] = []
        for r in rows:
            names += [k for k in r if k not in names]
        return cls({k: [r.get(k) for r in rows] for k in names})

    @classmethod
    def from_file(cls, path: str, kind: Optional[str] = None) -> "ColumnDataset":
        kind = kind or path.rsplit(".", 1)[-1].lower()
        with open(path, newline="" if kind == "csv" else None) as f:
            if kind == "csv":
                return cls.from_rows(list(csv.DictReader(f)))
            if kind in ("text", "txt"):
                return cls({"text": [line.rstrip("\n") for line in f]})
            if kind in ("json", "jsonl"):
                text = f.read().strip()
                if text.startswith("["):
                    return cls.from_rows(json.loads(text))
                if text.startswith("{") and "\n" not in text:
                    obj = json.loads(text)
                    return cls.from_rows(obj["data"] if isinstance(obj.get("data"), list) else [obj])
                return cls.from_rows([json.loads(line) for line in text.splitlines() if line.strip()])
        raise ValueError(f"cannot load {path}: unknown data file type {kind!r} (csv, json, jsonl, txt)")

    @property
    def column_names(self) -> List[str]:
        return list(self._cols)

    def set_format(self, type: Optional[str] = None, columns: Optional[List[str]] = None) -> None:  # noqa: A002
        missing = [c for c in columns or [] if c not in self._cols]
        if missing:
            raise ValueError(f"columns {missing} are not in the dataset ({self.column_names})")
        self.format = {"type": type, "columns": list(columns) if columns is not None else self.column_names}

    def __len__(self) -> int:
        return len(next(iter(self._cols.values()), []))

    def _cast(self, v: Any) -> Any:
        if self.format["type"] in ("torch", "pt"):
            import torch

            try:
                return torch.as_tensor(v)
            except (TypeError, ValueError, RuntimeError):
                return v
        if self.format["type"] in ("numpy", "np"):
            import numpy as np

            return np.asarray(v)
        return v

    def __getitem__(self, i: Union[int, slice, str]) -> Any:
        if isinstance(i, str):
            return self._cols[i]
        return {c: self._cast(self._cols[c][i]) for c in self.format["columns"]}

    def select(self, indices: Sequence[int]) -> "ColumnDataset":
        out = ColumnDataset({k: [v[i] for i in indices] for k, v in self._cols.items()})
        out.format = dict(self.format)
        return out

    def map(self, fn: Callable[..., Dict[str, Any]], batched: bool = False, remove_columns: Optional[List[str]] = None, **_: Any) -> "ColumnDataset":
        cols = {k: list(v) for k, v in self._cols.items()}
        if batched:
            new = fn(dict(cols))
            lens = {len(v) for v in new.values()}
            if lens != {len(self)}:       # the row count changed (e.g. text grouped into blocks): the old columns cannot ride along
                cols = {}
            cols.update({k: list(v) for k, v in new.items()})
        else:
            rows = [fn({k: v[i] for k, v in self._cols.items()}) for i in range(len(self))]
            for k in {k for r in rows for k in r}:
                cols[k] = [r.get(k) for r in rows]
        for k in remove_columns or []:
            cols.pop(k, None)
        return ColumnDataset(cols)


def remove_unused_columns(model: Any, dataset: Any) -> None:
    """Keep only the columns ``model.forward`` has parameters for (plus ``label`` / ``label_ids``, which the default collator
    renames): a tokenised dataset still carries its raw text columns, which cannot be collated."""
    accepted = list(inspect.signature(model.forward).parameters) + ["label", "label_ids"]
    dataset.set_format(type=dataset.format["type"], columns=[c for c in accepted if c in dataset.column_names])


def default_load_dataset(data_config_input: Union[Dict[str, Any], utils.AttrDict]) -> Any:
    """``{"train": ..., "validation": ...}`` from a hub dataset name (needs ``datasets``) or from ``train_file`` /
    ``validation_file``.  A hub dataset without a validation split gives its first ``validation_split_percentage`` percent
    of train to validation; so do local files when only ``train_file`` is given."""
    (cfg,) = hf_parse.parse_dict_to_dataclasses((hf_parse.DatasetKwargs,), data_config_input)
    pct = cfg.validation_split_percentage
    try:
        import datasets as hf_datasets
    except ImportError:
        hf_datasets = None
    if cfg.dataset_name is not None:
        if hf_datasets is None:
            raise ImportError(f"loading the hub dataset {cfg.dataset_name!r} needs the `datasets` package; give train_file / validation_file instead")
        ds = hf_datasets.load_dataset(cfg.dataset_name, cfg.dataset_config_name)
        if "validation" not in ds.keys():
            if pct is None:
                raise ValueError("this dataset has no validation split: set validation_split_percentage in the data config")
            ds["validation"] = hf_datasets.load_dataset(cfg.dataset_name, cfg.dataset_config_name, split=f"train[:{pct}%]")
            ds["train"] = hf_datasets.load_dataset(cfg.dataset_name, cfg.dataset_config_name, split=f"train[{pct}%:]")
        return ds
    if cfg.train_file is None:
        raise ValueError("the data config needs a dataset_name or a train_file")
    files = {k: v for k, v in (("train", cfg.train_file), ("validation", cfg.validation_file)) if v is not None}
    kind = cfg.train_file.rsplit(".", 1)[-1].lower()
    kind = "text" if kind == "txt" else kind
    if hf_datasets is not None:
        out = hf_datasets.load_dataset("json" if kind == "jsonl" else kind, data_files=files)
    else:
        out = {split: ColumnDataset.from_file(path, kind) for split, path in files.items()}
    if "validation" not in out and pct:
        n = len(out["train"]) * int(pct) // 100
        out["validation"], out["train"] = out["train"].select(range(n)), out["train"].select(range(n, len(out["train"])))
    return out
