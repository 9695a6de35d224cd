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

The schema lives in ``schemas_v0`` (the reference's 59 schema files restated as data under the same ``$id`` URLs) and is
evaluated by ``schema_engine`` (draft-07 subset + the seven custom keywords); ``expconf_schema`` holds the run-time defaults
and merge rules.  The reference's shared test cases (``schemas/test_cases/v0``) run against them in
``tests/test_expconf_schema_cases.py``.  ``det preview-search`` and the master both go through this module.
"""
import copy
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import yaml

from determined_b200.master import expconf_schema as schema

Spec = Dict[str, Any]


class ConfigError(ValueError):
    """One or more validation errors; ``errors`` lists them with JSON paths."""

    def __init__(self, errors: List[str]) -> None:
        super().__init__("config is invalid:\n  " + "\n  ".join(errors))
        self.errors = errors


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


def validate_sanity(cfg: Dict[str, Any]) -> None:
    """Phase 1: is what the user wrote well-formed?  (Missing fields are fine here.)"""
    errs = schema.validate(cfg, schema.EXPERIMENT, complete=False)
    if errs:
        raise ConfigError(errs)


def with_defaults(cfg: Dict[str, Any]) -> Dict[str, Any]:
    """Every default of the schema filled in, plus the run-time ones (generated name, seed, per-device image / variables)."""
    return schema.with_defaults(cfg, schema.EXPERIMENT)


def validate_complete(cfg: Dict[str, Any]) -> None:
    """Phase 2: after defaults/merging, is everything that must *eventually* be set, set?"""
    errs = schema.validate(cfg, schema.EXPERIMENT, complete=True)
    # checks the reference makes in Go next to the schema (master/internal/api_experiment.go, core_experiment.go)
    if cfg.get("entrypoint") is None:
        errs.append("<config>: entrypoint is a required property")
    slots = (cfg.get("resources") or {}).get("slots_per_trial", 1)
    gbs = (cfg.get("hyperparameters") or {}).get("global_batch_size")
    if isinstance(gbs, dict) and gbs.get("type") == "const":
        gbs = gbs.get("val")
    if isinstance(gbs, int) and not isinstance(gbs, bool) and slots and gbs < slots:
        errs.append("<config>.hyperparameters.global_batch_size: must be >= resources.slots_per_trial")
    if errs:
        raise ConfigError(errs)


def schema_merge(cfg: Dict[str, Any], src: Dict[str, Any]) -> Dict[str, Any]:
    """``cfg`` completed from a template or the cluster defaults, with the schema's merge rules (bind mounts and devices
    append, hyperparameters are atomic, a storage / searcher section of another kind only contributes shared fields)."""
    return schema.merge(cfg, src, schema.EXPERIMENT)


def finalize(raw: Union[str, bytes, Dict[str, Any]], cluster_defaults: Optional[Dict[str, Any]] = None,
             template: Optional[Dict[str, Any]] = None, overrides: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
    """The master's pipeline: parse → sanity → merge(template, cluster defaults) → defaults →
    completeness.  Returns the full config dict that trials receive as DET_EXPERIMENT_CONFIG."""
    cfg = parse(raw)
    if overrides:
        cfg = merge(overrides, cfg)
    validate_sanity(cfg)
    if template:
        cfg = schema_merge(cfg, template)
    if cluster_defaults:
        cfg = schema_merge(cfg, cluster_defaults)
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
