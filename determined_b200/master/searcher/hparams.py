"""Hyperparameter sampling and grid expansion (reference ``master/pkg/searcher/hyperparameters.go:14-48``,
``grid.go:120-288``).  Hyperparameter specs are the expconf forms ``{type: const|int|double|log|
categorical, ...}``, nested dicts, or bare constants."""
import itertools
import math
from typing import Any, Dict, Iterator, List, Tuple

from determined_b200.master import nprand


def _is_spec(v: Any) -> bool:
    return isinstance(v, dict) and isinstance(v.get("type"), str) and v["type"] in (
        "const", "int", "double", "log", "categorical")


def sample_one(spec: Any, rand: nprand.State) -> Any:
    if _is_spec(spec):
        t = spec["type"]
        if t == "const":
            return spec.get("val")
        if t == "int":
            return int(spec["minval"]) + rand.intn(int(spec["maxval"]) - int(spec["minval"]) + 1)
        if t == "double":
            return rand.uniform(float(spec["minval"]), float(spec["maxval"]))
        if t == "log":
            return math.pow(float(spec.get("base", 10.0)), rand.uniform(float(spec["minval"]), float(spec["maxval"])))
        if t == "categorical":
            vals = spec["vals"]
            return vals[rand.intn(len(vals))]
    if isinstance(spec, dict):
        return {k: sample_one(spec[k], rand) for k in sorted(spec)}
    return spec   # bare constant


def sample_all(hparams: Dict[str, Any], rand: nprand.State) -> Dict[str, Any]:
    """Sample every hyperparameter, visiting names in sorted order (deterministic draw order)."""
    return {name: sample_one(hparams[name], rand) for name in sorted(hparams)}


def _axis(spec: Any) -> List[Any]:
    if _is_spec(spec):
        t = spec["type"]
        if t == "const":
            return [spec.get("val")]
        if t == "categorical":
            return list(spec["vals"])
        count = spec.get("count")
        if count is None:
            raise ValueError("grid search requires `count` on int/double/log hyperparameters")
        count = int(count)
        lo, hi = spec["minval"], spec["maxval"]
        if t == "int":
            lo, hi = int(lo), int(hi)
            n = min(count, hi - lo + 1)
            if n <= 1:
                return [int(round((lo + hi) / 2.0))] if n == 1 and count == 1 else [lo]
            return sorted({int(round(lo + i * (hi - lo) / (n - 1))) for i in range(n)})
        pts = [float((lo + hi) / 2.0)] if count == 1 else [float(lo) + i * (float(hi) - float(lo)) / (count - 1) for i in range(count)]
        if t == "log":
            base = float(spec.get("base", 10.0))
            return [math.pow(base, p) for p in pts]
        return pts
    if isinstance(spec, dict):
        keys = sorted(spec)
        return [dict(zip(keys, combo)) for combo in itertools.product(*[_axis(spec[k]) for k in keys])]
    return [spec]


def grid(hparams: Dict[str, Any]) -> List[Dict[str, Any]]:
    names = sorted(hparams)
    return [dict(zip(names, combo)) for combo in itertools.product(*[_axis(hparams[n]) for n in names])]
