"""Experiment configuration: parsing, two-phase validation, defaults, merging.

Parity: reference ``schemas/expconf/v0/*.json`` (59 JSON-Schema files with the custom keywords
``union``, ``checks``, ``eventuallyRequired``; ``schemas/README.md:34-60``) and the Go side
``master/pkg/schemas/expconf`` (``parse.go``, ``experiment_config.go``, ``searcher_config.go``,
``hyperparameters_config.go``, ``length.go``).  Semantics kept:

  phase 1 ("sanity")       types / enums / ranges / unknown keys, on the raw user input;
  merge                    user config over templates over cluster defaults (``merge``: user wins,
                           objects merge recursively, arrays replace);
  defaults                 every documented default filled in (``with_defaults``);
  phase 2 ("completeness") ``eventuallyRequired`` fields must now be present (searcher.metric,
                           max_length, max_trials, checkpoint_storage.type/host_path, entrypoint…).

The schema is declared once, below, as data (a compact Python spec instead of JSON-Schema files);
``det preview-search`` and the master both use it.
"""
import copy
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import yaml

Spec = Dict[str, Any]


class ConfigError(ValueError):
    """One or more validation errors; ``errors`` lists them with JSON paths."""

    def __init__(self, errors: List[str]) -> None:
        super().__init__("config is invalid:\n  " + "\n  ".join(errors))
        self.errors = errors


# field spec: t = allowed python types; d = default; enum; min; req (eventually required);
#             obj = nested spec; items = spec for list items; check = callable(value) -> error|None
def F(t: Any, d: Any = None, **kw: Any) -> Dict[str, Any]:
    return {"t": t if isinstance(t, tuple) else (t,), "d": d, **kw}


_NUM = (int, float)
_LENGTH = F((dict, int), None)

_SEARCHER_COMMON = {
    "name": F(str, None, req=True),
    "metric": F(str, None, req=True),
    "smaller_is_better": F(bool, True),
    "source_trial_id": F(int, None),
    "source_checkpoint_uuid": F(str, None),
}
SEARCHERS: Dict[str, Spec] = {
    "single": {**_SEARCHER_COMMON, "max_length": F((dict, int), None, req=True)},
    "random": {**_SEARCHER_COMMON, "max_length": F((dict, int), None, req=True), "max_trials": F(int, None, req=True, min=1),
               "max_concurrent_trials": F(int, 16, min=0)},
    "grid": {**_SEARCHER_COMMON, "max_length": F((dict, int), None, req=True), "max_concurrent_trials": F(int, 16, min=0)},
    "async_halving": {**_SEARCHER_COMMON, "num_rungs": F(int, None, req=True, min=1), "max_length": F((dict, int), None, req=True),
                      "max_trials": F(int, None, req=True, min=1), "divisor": F(_NUM, 4), "max_concurrent_trials": F(int, 16, min=0),
                      "stop_once": F(bool, False)},
    "adaptive_asha": {**_SEARCHER_COMMON, "max_length": F((dict, int), None, req=True), "max_trials": F(int, None, req=True, min=1),
                      "bracket_rungs": F(list, []), "mode": F(str, "standard", enum=["aggressive", "standard", "conservative"]),
                      "divisor": F(_NUM, 4), "max_rungs": F(int, 5, min=1), "max_concurrent_trials": F(int, 16, min=0),
                      "stop_once": F(bool, False)},
    "custom": {"name": F(str, None, req=True), "metric": F(str, None, req=True), "smaller_is_better": F(bool, True),
               "unit": F(str, None, enum=["batches", "records", "epochs"])},
}

_GC = {"save_experiment_best": F(int, 0, min=0), "save_trial_best": F(int, 1, min=0), "save_trial_latest": F(int, 1, min=0)}
STORAGES: Dict[str, Spec] = {
    "shared_fs": {"type": F(str, None, req=True), "host_path": F(str, None, req=True), "storage_path": F(str, None),
                  "propagation": F(str, "rprivate"), "container_path": F(str, None), "checkpoint_path": F(str, None),
                  "tensorboard_path": F(str, None), **_GC},
    "directory": {"type": F(str, None, req=True), "container_path": F(str, None, req=True), **_GC},
    "s3": {"type": F(str, None, req=True), "bucket": F(str, None, req=True), "access_key": F(str, None), "secret_key": F(str, None),
           "endpoint_url": F(str, None), "prefix": F(str, None), **_GC},
    "gcs": {"type": F(str, None, req=True), "bucket": F(str, None, req=True), "prefix": F(str, None), **_GC},
    "azure": {"type": F(str, None, req=True), "container": F(str, None, req=True), "connection_string": F(str, None),
              "account_url": F(str, None), "credential": F(str, None), **_GC},
}

RESOURCES: Spec = {
    "agent_label": F(str, None), "devices": F(list, []), "is_single_node": F(bool, None), "max_slots": F(int, None),
    "native_parallel": F(bool, False), "priority": F(int, None, min=1), "resource_pool": F(str, ""),
    "shm_size": F((int, str), None), "slots": F(int, None), "slots_per_trial": F(int, 1, min=0), "weight": F(_NUM, 1),
}
OPTIMIZATIONS: Spec = {
    "aggregation_frequency": F(int, 1, min=1), "auto_tune_tensor_fusion": F(bool, False),
    "average_aggregated_gradients": F(bool, True), "average_training_metrics": F(bool, True),
    "gradient_compression": F(bool, False), "grad_updates_size_file": F(str, None),
    "mixed_precision": F(str, "O0", enum=["O0", "O1", "O2", "O3"]), "tensor_fusion_cycle_time": F(int, 1, min=0),
    "tensor_fusion_threshold": F(int, 64, min=0),
    # B200 engine extensions (ignored by the reference)
    "fused_engine": F(bool, None), "cuda_graphs": F(bool, False), "bucket_cap_mb": F(_NUM, 25), "async_checkpoint": F(bool, False),
}
PROFILING: Spec = {"enabled": F(bool, False), "begin_on_batch": F(int, 0, min=0), "end_after_batch": F(int, None, min=0),
                   "sync_timings": F(bool, True)}
ENVIRONMENT: Spec = {"image": F((dict, str), {}), "environment_variables": F((dict, list), []), "proxy_ports": F(list, []),
                     "ports": F(dict, {}), "force_pull_image": F(bool, False), "registry_auth": F(dict, None),
                     "add_capabilities": F(list, []), "drop_capabilities": F(list, []), "pod_spec": F(dict, None)}
LOG_POLICY: Spec = {"pattern": F(str, None, req=True), "action": F(dict, None), "name": F(str, None)}

EXPERIMENT: Spec = {
    "bind_mounts": F(list, []), "checkpoint_policy": F(str, "best", enum=["best", "all", "none"]),
    "checkpoint_storage": F(dict, None, req=True), "data": F(dict, {}), "data_layer": F(dict, None), "debug": F(bool, False),
    "description": F(str, None), "entrypoint": F((str, list), None, req=True), "environment": F(dict, {}, obj=ENVIRONMENT),
    "hyperparameters": F(dict, {}), "internal": F(dict, None), "labels": F(list, []), "log_policies": F(list, [], items=LOG_POLICY),
    "retention_policy": F(dict, None), "max_restarts": F(int, 5, min=0), "min_checkpoint_period": F((dict, int), {"batches": 0}),
    "min_validation_period": F((dict, int), {"batches": 0}), "name": F(str, None),
    "optimizations": F(dict, {}, obj=OPTIMIZATIONS), "pbs": F(dict, {}), "perform_initial_validation": F(bool, False),
    "profiling": F(dict, {}, obj=PROFILING), "project": F(str, ""), "records_per_epoch": F(int, 0, min=0),
    "reproducibility": F(dict, {}, obj={"experiment_seed": F(int, None, min=0)}), "resources": F(dict, {}, obj=RESOURCES),
    "scheduling_unit": F(int, 100, min=1), "searcher": F(dict, None, req=True), "security": F(dict, None), "slurm": F(dict, {}),
    "tensorboard_storage": F(dict, None), "workspace": F(str, ""), "template": F(str, None),
    "max_restarts_on_same_node": F(int, None), "preemption_timeout": F(int, None), "integrations": F(dict, None),
}

HP_TYPES = ("const", "int", "double", "log", "categorical")


def parse(text_or_dict: Union[str, bytes, Dict[str, Any]]) -> Dict[str, Any]:
    if isinstance(text_or_dict, (str, bytes)):
        cfg = yaml.safe_load(text_or_dict)
    else:
        cfg = copy.deepcopy(text_or_dict)
    if cfg is None:
        cfg = {}
    if not isinstance(cfg, dict):
        raise ConfigError(["<config>: must be a mapping"])
    return cfg


def merge(override: Any, base: Any) -> Any:
    """``override`` wins; mappings merge key by key; None in override defers to base."""
    if override is None:
        return copy.deepcopy(base)
    if isinstance(override, dict) and isinstance(base, dict):
        out = copy.deepcopy(base)
        for k, v in override.items():
            out[k] = merge(v, base.get(k))
        return out
    return copy.deepcopy(override)


def _type_ok(v: Any, types: Tuple[type, ...]) -> bool:
    if isinstance(v, bool):
        return bool in types
    if isinstance(v, int) and float in types and int not in types:
        return True
    return isinstance(v, types)


def _check_length(v: Any, path: str, errs: List[str]) -> None:
    if isinstance(v, bool):
        errs.append(f"{path}: must be a length")
    elif isinstance(v, int):
        if v < 0:
            errs.append(f"{path}: must be >= 0")
    elif isinstance(v, dict):
        if len(v) != 1 or next(iter(v)) not in ("batches", "records", "epochs"):
            errs.append(f"{path}: must have exactly one of batches|records|epochs")
        else:
            n = next(iter(v.values()))
            if not isinstance(n, int) or isinstance(n, bool) or n < 0:
                errs.append(f"{path}: length value must be a non-negative integer")


def _check_obj(cfg: Dict[str, Any], spec: Spec, path: str, errs: List[str], allow_unknown: bool = False) -> None:
    for k, v in cfg.items():
        if k not in spec:
            if not allow_unknown:
                errs.append(f"{path}.{k}: unknown field")
            continue
        f = spec[k]
        p = f"{path}.{k}"
        if v is None:
            continue
        if not _type_ok(v, f["t"]):
            errs.append(f"{p}: expected {' or '.join(t.__name__ for t in f['t'])}, got {type(v).__name__}")
            continue
        if "enum" in f and v not in f["enum"]:
            errs.append(f"{p}: must be one of {f['enum']}")
        if "min" in f and isinstance(v, (int, float)) and not isinstance(v, bool) and v < f["min"]:
            errs.append(f"{p}: must be >= {f['min']}")
        if "obj" in f and isinstance(v, dict):
            _check_obj(v, f["obj"], p, errs)
        if "items" in f and isinstance(v, list):
            for i, it in enumerate(v):
                if isinstance(it, dict):
                    _check_obj(it, f["items"], f"{p}[{i}]", errs)
                else:
                    errs.append(f"{p}[{i}]: must be a mapping")


def _check_hparam(name: str, v: Any, path: str, errs: List[str], grid: bool) -> None:
    if isinstance(v, dict) and isinstance(v.get("type"), str):
        t = v["type"]
        if t not in HP_TYPES:
            errs.append(f"{path}: unknown hyperparameter type {t!r}")
            return
        if t in ("int", "double", "log"):
            for k in ("minval", "maxval"):
                if not isinstance(v.get(k), (int, float)) or isinstance(v.get(k), bool):
                    errs.append(f"{path}.{k}: required number")
            if all(isinstance(v.get(k), (int, float)) for k in ("minval", "maxval")) and v["minval"] > v["maxval"]:
                errs.append(f"{path}: minval must be <= maxval")
            if t == "int" and any(isinstance(v.get(k), float) and not float(v[k]).is_integer() for k in ("minval", "maxval")):
                errs.append(f"{path}: int hyperparameter bounds must be integers")
            if grid and v.get("count") is None:
                errs.append(f"{path}.count: grid search requires count on {t} hyperparameters")
            if v.get("count") is not None and (not isinstance(v["count"], int) or v["count"] < 1):
                errs.append(f"{path}.count: must be a positive integer")
        if t == "categorical" and (not isinstance(v.get("vals"), list) or not v["vals"]):
            errs.append(f"{path}.vals: must be a non-empty list")
        if t == "const" and "val" not in v:
            errs.append(f"{path}.val: required")
    elif isinstance(v, dict):
        for k, sub in v.items():
            _check_hparam(k, sub, f"{path}.{k}", errs, grid)


def validate_sanity(cfg: Dict[str, Any]) -> None:
    """Phase 1: is what the user wrote well-formed?  (Missing fields are fine here.)"""
    errs: List[str] = []
    _check_obj(cfg, EXPERIMENT, "<config>", errs)
    s = cfg.get("searcher")
    if isinstance(s, dict):
        name = s.get("name")
        if name is not None and name not in SEARCHERS:
            errs.append(f"<config>.searcher.name: must be one of {sorted(SEARCHERS)}")
        elif name is not None:
            _check_obj(s, SEARCHERS[name], "<config>.searcher", errs)
            if "max_length" in s and s["max_length"] is not None:
                _check_length(s["max_length"], "<config>.searcher.max_length", errs)
                ml = s["max_length"]
                n = next(iter(ml.values())) if isinstance(ml, dict) and ml else ml
                if isinstance(n, int) and not isinstance(n, bool) and n < 1:
                    errs.append("<config>.searcher.max_length: must be >= 1")
            if name == "adaptive_asha" and s.get("bracket_rungs"):
                if not all(isinstance(b, int) and b >= 1 for b in s["bracket_rungs"]):
                    errs.append("<config>.searcher.bracket_rungs: must be positive integers")
            d = s.get("divisor")
            if d is not None and isinstance(d, (int, float)) and d <= 1:
                errs.append("<config>.searcher.divisor: must be > 1")
    cs = cfg.get("checkpoint_storage")
    if isinstance(cs, dict):
        t = cs.get("type")
        if t is not None and t not in STORAGES:
            errs.append(f"<config>.checkpoint_storage.type: must be one of {sorted(STORAGES)}")
        elif t is not None:
            _check_obj(cs, STORAGES[t], "<config>.checkpoint_storage", errs)
    for k in ("min_checkpoint_period", "min_validation_period"):
        if cfg.get(k) is not None:
            _check_length(cfg[k], f"<config>.{k}", errs)
    hps = cfg.get("hyperparameters")
    if isinstance(hps, dict):
        grid = isinstance(s, dict) and s.get("name") == "grid"
        for k, v in hps.items():
            _check_hparam(k, v, f"<config>.hyperparameters.{k}", errs, grid)
    ep = cfg.get("entrypoint")
    if isinstance(ep, list) and not all(isinstance(x, str) for x in ep):
        errs.append("<config>.entrypoint: list entrypoint must contain strings")
    if errs:
        raise ConfigError(errs)


def _fill(cfg: Dict[str, Any], spec: Spec) -> Dict[str, Any]:
    out = dict(cfg)
    for k, f in spec.items():
        if out.get(k) is None and f["d"] is not None:
            out[k] = copy.deepcopy(f["d"])
        if "obj" in f and isinstance(out.get(k), dict):
            out[k] = _fill(out[k], f["obj"])
    return out


def with_defaults(cfg: Dict[str, Any]) -> Dict[str, Any]:
    out = _fill(cfg, EXPERIMENT)
    s = out.get("searcher")
    if isinstance(s, dict) and s.get("name") in SEARCHERS:
        out["searcher"] = _fill(s, SEARCHERS[s["name"]])
    cs = out.get("checkpoint_storage")
    if isinstance(cs, dict) and cs.get("type") in STORAGES:
        out["checkpoint_storage"] = _fill(cs, STORAGES[cs["type"]])
    if out.get("name") is None:
        out["name"] = "Experiment (unnamed)"
    return out


def _require(cfg: Dict[str, Any], spec: Spec, path: str, errs: List[str]) -> None:
    for k, f in spec.items():
        if f.get("req") and cfg.get(k) is None:
            errs.append(f"{path}.{k}: is a required property")
        if "obj" in f and isinstance(cfg.get(k), dict):
            _require(cfg[k], f["obj"], f"{path}.{k}", errs)


def validate_complete(cfg: Dict[str, Any]) -> None:
    """Phase 2: after defaults/merging, is everything that must *eventually* be set, set?"""
    errs: List[str] = []
    _require(cfg, EXPERIMENT, "<config>", errs)
    s = cfg.get("searcher")
    if isinstance(s, dict):
        if s.get("name") in SEARCHERS:
            _require(s, SEARCHERS[s["name"]], "<config>.searcher", errs)
        elif s.get("name") is None:
            errs.append("<config>.searcher.name: is a required property")
    cs = cfg.get("checkpoint_storage")
    if isinstance(cs, dict):
        if cs.get("type") in STORAGES:
            _require(cs, STORAGES[cs["type"]], "<config>.checkpoint_storage", errs)
        elif cs.get("type") is None:
            errs.append("<config>.checkpoint_storage.type: is a required property")
    slots = (cfg.get("resources") or {}).get("slots_per_trial", 1)
    gbs = (cfg.get("hyperparameters") or {}).get("global_batch_size")
    if isinstance(gbs, dict) and gbs.get("type") == "const":
        gbs = gbs.get("val")
    if isinstance(gbs, int) and not isinstance(gbs, bool) and slots and gbs < slots:
        errs.append("<config>.hyperparameters.global_batch_size: must be >= resources.slots_per_trial")
    if errs:
        raise ConfigError(errs)


def finalize(raw: Union[str, bytes, Dict[str, Any]], cluster_defaults: Optional[Dict[str, Any]] = None,
             template: Optional[Dict[str, Any]] = None, overrides: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    """The master's pipeline: parse → sanity → merge(template, cluster defaults) → defaults →
    completeness.  Returns the full config dict that trials receive as DET_EXPERIMENT_CONFIG."""
    cfg = parse(raw)
    if overrides:
        cfg = merge(overrides, cfg)
    validate_sanity(cfg)
    if template:
        cfg = merge(cfg, template)
    if cluster_defaults:
        cfg = merge(cfg, cluster_defaults)
    cfg = with_defaults(cfg)
    validate_complete(cfg)
    return cfg


def apply_override(cfg: Dict[str, Any], dotted: str, value: Any) -> None:
    """``--config a.b.c=value`` (reference ``cli/experiment.py:43-53``)."""
    cur = cfg
    parts = dotted.split(".")
    for p in parts[:-1]:
        nxt = cur.get(p)
        if not isinstance(nxt, dict):
            nxt = {}
            cur[p] = nxt
        cur = nxt
    cur[parts[-1]] = value


def length_of(v: Any) -> Tuple[int, Optional[str]]:
    if isinstance(v, dict):
        (u, n), = v.items()
        return int(n), u
    return int(v), None
