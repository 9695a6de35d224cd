"""Schema-driven experiment-config operations: the two validation phases, defaults and merging (reference
``master/pkg/schemas/{schema,defaults,merge}.go`` + ``expconf/parse.go``), on top of ``schema_engine`` and ``schemas_v0``.

``validate(cfg, url, complete)``   errors of ``cfg`` against the schema ``url`` ("sanity" phase, or "completeness" phase);
``with_defaults(cfg, url)``        ``cfg`` with schema defaults and the run-time defaults filled in;
``merge(cfg, src, url)``           ``cfg`` completed from ``src`` (templates, cluster defaults);
``shim(cfg)``                      rewrites of legacy spellings that the reference still accepts.
"""
import copy
from typing import Any, Callable, Dict, List, Optional

from determined_b200.master import schema_engine as E
from determined_b200.master.schemas_v0 import REGISTRY, url

EXPERIMENT = url("experiment")


def validate(cfg: Any, schema_url: str = EXPERIMENT, complete: bool = False, registry: Optional[E.Registry] = None) -> List[str]:
    reg = registry or REGISTRY
    return E.Validator(reg, complete).errors(reg[schema_url], cfg)


# ---- run-time defaults (reference defaults.go "RuntimeDefaultable" implementations) ------------------------------------
def _image_default(v: Any) -> Any:
    """A string means the same image on every device type; a map falls back to the platform default per device type."""
    if isinstance(v, str):
        return {"cpu": v, "cuda": v, "rocm": v}
    out = dict(v or {})
    if out.get("gpu") is not None and out.get("cuda") is None:
        out["cuda"] = out["gpu"]                      # pre-0.17 spelling
    out.pop("gpu", None)
    for k in ("cpu", "cuda", "rocm"):
        if out.get(k) is None:
            out[k] = f"determined_b200/host-python:{k}"     # tasks are host processes here: the name is informational
    return out


def _envvars_default(v: Any) -> Any:
    if isinstance(v, list):
        return {"cpu": list(v), "cuda": list(v), "rocm": list(v)}
    out = dict(v or {})
    if out.get("gpu") is not None and not out.get("cuda"):
        out["cuda"] = out["gpu"]
    out.pop("gpu", None)
    for k in ("cpu", "cuda", "rocm"):
        if out.get(k) is None:
            out[k] = []
    return out


def _devices_default(v: Any) -> Any:
    out = []
    for d in v or []:
        if isinstance(d, str):                        # docker's --device syntax: host:container[:mode]
            parts = d.split(":")
            d = {"host_path": parts[0], "container_path": parts[1], "mode": parts[2] if len(parts) > 2 else None}
        d = dict(d)
        if d.get("mode") is None:
            d["mode"] = "mrw"
        out.append(d)
    return out


_ADJ = ("amused", "brave", "calm", "deft", "eager", "fair", "glad", "keen", "lucid", "merry", "nimble", "polite", "quick", "steady", "witty")
_ANIMAL = ("badger", "crane", "dingo", "egret", "ferret", "gecko", "heron", "ibis", "jackal", "koala", "lemur", "marten", "newt", "otter", "puffin")


def _experiment_default(v: Any) -> Any:
    """An experiment without a name gets a generated one (reference ``ExperimentConfigV0.RuntimeDefaults``)."""
    if isinstance(v, dict) and v.get("name") is None:
        import random

        v = dict(v, name=f"Experiment ({random.choice(_ADJ)}-{random.choice(_ANIMAL)})")
    return v


def _reproducibility_default(v: Any) -> Any:
    """No seed given: the clock picks one, which is then recorded in the config so that the experiment can be repeated."""
    if isinstance(v, dict) and v.get("experiment_seed") is None:
        import time

        v = dict(v, experiment_seed=int(time.time()))
    return v


DEFAULT_HOOKS: Dict[str, Callable[[Any], Any]] = {
    url("experiment"): _experiment_default,
    url("reproducibility"): _reproducibility_default,
    url("environment-image"): _image_default,
    url("environment-variables"): _envvars_default,
    url("devices"): _devices_default,
}


def with_defaults(cfg: Any, schema_url: str = EXPERIMENT, registry: Optional[E.Registry] = None,
                  hooks: Optional[Dict[str, Callable[[Any], Any]]] = None) -> Any:
    reg = registry or REGISTRY
    return E.fill_defaults(reg, reg[schema_url], copy.deepcopy(cfg), {**DEFAULT_HOOKS, **(hooks or {})})


# ---- merging -----------------------------------------------------------------------------------------------------------
def _append_unique(key: str, normalise: Callable[[Any], Any], fields: List[str]) -> Callable[[Any, Any], Any]:
    def hook(obj: Any, src: Any) -> Any:
        out = [_with_nulls(normalise(x), fields) for x in obj]
        have = {x[key] for x in out}
        for x in src:
            x = _with_nulls(normalise(x), fields)
            if x[key] not in have:
                out.append(x)
                have.add(x[key])
        return out
    return hook


def _with_nulls(d: Dict[str, Any], fields: List[str]) -> Dict[str, Any]:
    return {**{f: None for f in fields}, **d}


def _device_obj(d: Any) -> Dict[str, Any]:
    if isinstance(d, str):
        parts = d.split(":")
        return {"host_path": parts[0], "container_path": parts[1], **({"mode": parts[2]} if len(parts) > 2 else {})}
    return dict(d)


def _hp_norm(v: Any) -> Any:
    """Hyperparameters in canonical form: a bare value is a constant, a dict without "type" is a namespace of hyperparameters."""
    if isinstance(v, dict):
        return v if "type" in v else {k: _hp_norm(x) for k, x in v.items()}
    if isinstance(v, list):
        return [_hp_norm(x) for x in v]
    return {"type": "const", "val": v}


def _hp_merge(obj: Any, src: Any) -> Any:
    """A hyperparameter is atomic: an int range is never completed from another int range.  Only namespaces (nested dicts
    on BOTH sides) merge, key by key.  Inside hyperparameters ``null`` is a value (the constant null), not "absent"."""
    o, s = _hp_norm(obj), _hp_norm(src)
    if isinstance(o, dict) and isinstance(s, dict) and "type" not in o and "type" not in s:
        out = dict(o)
        for k, v in s.items():
            out[k] = _hp_merge(o[k], v) if k in o else v
        return out
    return o


MERGE_HOOKS: Dict[str, Callable[[Any, Any], Any]] = {
    url("hyperparameter"): _hp_merge,
    # mounts and devices from a template are appended unless the experiment already mounts something at that container path
    url("bind-mounts"): _append_unique("container_path", dict, ["host_path", "container_path", "propagation", "read_only"]),
    url("devices"): _append_unique("container_path", _device_obj, ["host_path", "container_path", "mode"]),
}


def merge(cfg: Any, src: Any, schema_url: str = EXPERIMENT, registry: Optional[E.Registry] = None) -> Any:
    reg = registry or REGISTRY
    return E.merge(reg, reg[schema_url], copy.deepcopy(cfg), copy.deepcopy(src), MERGE_HOOKS)
