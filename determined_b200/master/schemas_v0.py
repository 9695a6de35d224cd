"""Experiment-config schemas, version 0, as data for ``schema_engine``.

Same content and the same ``$id`` URLs as the reference's ``schemas/expconf/v0/*.json`` (so its shared test cases under
``schemas/test_cases/v0`` address them by name), written with a few helpers instead of 59 JSON files: ``n(...)`` is a
nullable property with a default, ``obj(...)`` an object schema that forbids unknown properties, ``union(...)`` the
tagged-alternative keyword.  ``REGISTRY`` maps ``$id`` to schema.
"""
from typing import Any, Dict, Iterable, List, Optional, Tuple

BASE = "http://determined.ai/schemas/expconf/v0/"
REGISTRY: Dict[str, Dict[str, Any]] = {}
_UNSET = object()


def url(name: str) -> str:
    return BASE + name + ".json"


def ref(name: str) -> Dict[str, Any]:
    return {"$ref": url(name)}


def n(types: Any, default: Any = None, **kw: Any) -> Dict[str, Any]:
    """Nullable property of the given JSON type(s) with a default."""
    ts = [types] if isinstance(types, str) else list(types)
    return {"type": ts + ["null"], "default": default, **kw}


def opt(name: str, types: Any, default: Any = None, **kw: Any) -> Dict[str, Any]:
    """Nullable property whose non-null values must satisfy another schema."""
    return n(types, default, optionalRef=url(name), **kw)


def enum(values: Iterable[Any], default: Any = None, **kw: Any) -> Dict[str, Any]:
    return {"enum": [None, *values], "default": default, **kw}


def obj(name: str, props: Dict[str, Any], required: Iterable[str] = (), later: Any = _UNSET, extra: Any = False,
        **kw: Any) -> Dict[str, Any]:
    """Object schema; ``later`` = ``eventuallyRequired`` (needed only once cluster defaults have been merged in)."""
    s: Dict[str, Any] = {"$id": url(name), "type": "object", "additionalProperties": extra, "properties": props}
    req = list(required)
    if req:
        s["required"] = req
    if later is not _UNSET:
        s["eventuallyRequired"] = list(later)
    s.update(kw)
    REGISTRY[s["$id"]] = s
    return s


def schema(name: str, **body: Any) -> Dict[str, Any]:
    s = {"$id": url(name), **body}
    REGISTRY[s["$id"]] = s
    return s


def union(items: List[Tuple[str, Dict[str, Any]]], message: Optional[str] = None) -> Dict[str, Any]:
    u: Dict[str, Any] = {"items": [{"unionKey": k, **v} for k, v in items]}
    if message:
        u["defaultMessage"] = message
    return {"union": u}


def tagged(key: str, names: Dict[str, str]) -> List[Tuple[str, Dict[str, Any]]]:
    """Union members selected by the value of one property: {value: schema name}."""
    return [(f"const:{key}={v}", ref(nm)) for v, nm in names.items()]


# ---------------------------------------------------------------------------------------------------------------------
# building blocks
# ---------------------------------------------------------------------------------------------------------------------
KEEP = {"save_experiment_best": n("integer", 0, minimum=0), "save_trial_best": n("integer", 1, minimum=0),
        "save_trial_latest": n("integer", 1, minimum=0)}
_DOTDOT = {"prefix cannot contain /../": {"not": {"anyOf": [{"type": "string", "pattern": p} for p in
                                                             (r"/\.\./", r"^\.\./", r"/\.\.$", r"^\.\.$")]}}}
SEARCH_COMMON = {"metric": n("string"), "smaller_is_better": n("boolean", True), "source_trial_id": n("integer"),
                 "source_checkpoint_uuid": n("string")}


def _one_key_length(unit: str, minimum: int = 0) -> Dict[str, Any]:
    return {"type": "object", "additionalProperties": False, "required": [unit],
            "properties": {unit: {"type": "integer", "minimum": minimum}}}


schema("length", **union([(f"singleproperty:{u}", _one_key_length(u)) for u in ("batches", "records", "epochs")],
                         'a length object must have one attribute named "batches", "records", or "epochs"'))
schema("check-positive-length", allOf=[ref("length"), {"additionalProperties": {"type": "integer", "minimum": 1}}])
schema("searcher-length", **union([("not:type:object", {"type": "integer", "minimum": 0}), ("always", ref("check-positive-length"))]))

# ---- checkpoint storage ---------------------------------------------------------------------------------------------
obj("shared-fs", {"type": {"const": "shared_fs"}, "host_path": n("string"), "storage_path": n("string"),
                  "propagation": n("string", "rprivate"), "container_path": n("string"), "checkpoint_path": n("string"),
                  "tensorboard_path": n("string"), **KEEP},
    required=["type"], later=["host_path"],
    checks={"storage_path must either be a relative directory or a subdirectory of host_path":
            {"compareProperties": {"type": "a_is_subdir_of_b", "a": "storage_path", "b": "host_path"}}})
obj("directory", {"type": {"const": "directory"}, "container_path": n("string"), **KEEP}, required=["type"], later=["container_path"])
obj("s3", {"type": {"const": "s3"}, "access_key": n("string"), "bucket": n("string"), "secret_key": n("string"),
           "endpoint_url": n("string"), "prefix": n("string", checks=_DOTDOT), **KEEP}, required=["type"], later=["bucket"])
obj("gcs", {"type": {"const": "gcs"}, "bucket": n("string"), "prefix": n("string", checks=_DOTDOT), **KEEP},
    required=["type"], later=["bucket"])
obj("azure", {"type": {"const": "azure"}, "container": n("string"), "connection_string": n("string"), "account_url": n("string"),
              "credential": n("string"), **KEEP}, required=["type"], later=["container"],
    eventually={"checks": {"Exactly one of connection_string or account_url must be set":
                           {"oneOf": [{"eventuallyRequired": ["connection_string"]}, {"eventuallyRequired": ["account_url"]}]}}},
    checks={"credential and connection_string must not both be set":
            {"not": {"required": ["connection_string", "credential"],
                     "properties": {"connection_string": {"type": "string"}, "credential": {"type": "string"}}}}})
_ANY_STORAGE_KEY = ["access_key", "account_url", "bucket", "checkpoint_path", "connection_string", "container", "container_path",
                    "credential", "endpoint_url", "prefix", "host_path", "propagation", "secret_key", "storage_path",
                    "tensorboard_path", "type", "user"]
obj("checkpoint-storage", {**{k: True for k in _ANY_STORAGE_KEY}, **KEEP}, later=["type"],
    **{"if": {"required": ["type"]},
       "then": union(tagged("type", {"shared_fs": "shared-fs", "directory": "directory", "s3": "s3", "gcs": "gcs", "azure": "azure"}),
                     "is not an object where object[\"type\"] is one of 'shared_fs', 'directory', 's3', 'gcs', or 'azure'")})
REGISTRY[url("checkpoint-storage")].pop("type")          # a partially filled storage section (no "type" yet) is legal
_DEPRECATED = "this field is deprecated and will be ignored"
schema("tensorboard-storage", **union([("never", ref("shared-fs")), ("never", ref("s3")), ("never", ref("gcs"))], _DEPRECATED),
       disallowProperties={k: _DEPRECATED for k in KEEP})

# ---- bind mounts, devices, environment ------------------------------------------------------------------------------
obj("bind-mount", {"host_path": {"type": "string", "checks": {"host_path must be an absolute path": {"pattern": "^/"}}},
                   "container_path": {"type": "string", "checks": {'container_path must not be "."': {"not": {"const": "."}}}},
                   "read_only": n("boolean", False), "propagation": n("string", "rprivate")}, required=["host_path", "container_path"])
schema("bind-mounts", type="array", items=ref("bind-mount"))
obj("device", {"host_path": {"type": "string"}, "container_path": {"type": "string"}, "mode": n("string", "mrw")},
    required=["host_path", "container_path"])
schema("devices", type="array", items=union(
    [("never", ref("device")), ("never", {"type": "string", "pattern": "^/[^:]*:/[^:]*(:[rwm]*)?"})],
    "is neither a list of --device strings nor a map containing host_path, container_path, and mode"))
obj("environment-image-map", {k: n("string") for k in ("cpu", "cuda", "rocm", "gpu")}, later=["cpu", "cuda", "rocm"])
schema("environment-image", **union([("never", ref("environment-image-map")), ("never", {"type": "string"})],
                                    "is neither a string nor a map of cpu, cuda, or rocm to strings"))
obj("environment-variables-map", {k: n("array", [], items={"type": "string"}) for k in ("cpu", "cuda", "rocm", "gpu")})
schema("environment-variables", **union([("never", ref("environment-variables-map")), ("never", {"type": "array", "items": {"type": "string"}})],
                                        "is neither a list of strings nor a map of cpu, cuda, or rocm to lists of strings"))
obj("proxy-port", {"proxy_port": {"type": "number"}, "proxy_tcp": n("boolean", False), "unauthenticated": n("boolean", False),
                   "default_service_id": n("boolean", False)}, required=["proxy_port"])
schema("proxy-ports", type="array", items=ref("proxy-port"))
obj("registry-auth", {k: n("string") for k in ("username", "password", "auth", "email", "serveraddress", "identitytoken", "registrytoken")})
_OURS = {"image": "container Image is not configurable, set it in the experiment config", "command": "container Command is not configurable",
         "args": "container Args are not configurable", "working_dir": "container WorkingDir is not configurable",
         "ports": "container Ports are not configurable", "liveness_probe": "container LivenessProbe is not configurable",
         "readiness_probe": "container ReadinessProbe is not configurable", "startup_probe": "container StartupProbe is not configurable",
         "lifecycle": "container Lifecycle is not configurable", "termination_message_path": "container TerminationMessagePath is not configurable",
         "termination_message_policy": "container TerminationMessagePolicy is not configurable",
         "image_pull_policy": "container ImagePullPolicy is not configurable, set it in the experiment config",
         "security_context": "container SecurityContext is not configurable, set it in the experiment config"}
_POD_SPEC = n("object", None, disallowProperties={"name": "pod Name is not a configurable option", "name_space": "pod NameSpace is not a configurable option"},
              properties={"spec": n("object", None, properties={"containers": n("array", None, items={
                  "if": {"properties": {"name": {"const": "determined-container"}}},
                  "then": {"type": "object", "disallowProperties": _OURS}})})})
obj("environment", {"image": opt("environment-image", ["object", "string"], {}),
                    "environment_variables": opt("environment-variables", ["object", "array"], []),
                    "proxy_ports": opt("proxy-ports", "array", []), "ports": n("object", {}, additionalProperties={"type": "integer"}),
                    "force_pull_image": n("boolean", False), "registry_auth": opt("registry-auth", "object"),
                    "add_capabilities": n("array", [], items={"type": "string"}), "drop_capabilities": n("array", [], items={"type": "string"}),
                    "pod_spec": _POD_SPEC}, later=["image"])

# ---- hyperparameters ------------------------------------------------------------------------------------------------
_COUNT = n("integer", None, minimum=1)
_MINMAX = {"compareProperties": {"type": "a<b", "a": "minval", "b": "maxval"}}
obj("hyperparameter-int", {"type": {"const": "int"}, "minval": {"type": "integer"}, "maxval": {"type": "integer"}, "count": _COUNT},
    required=["type", "minval", "maxval"], **_MINMAX)
obj("hyperparameter-double", {"type": {"const": "double"}, "minval": {"type": "number"}, "maxval": {"type": "number"}, "count": _COUNT},
    required=["type", "minval", "maxval"], **_MINMAX)
obj("hyperparameter-log", {"type": {"const": "log"}, "minval": {"type": "number"}, "maxval": {"type": "number"},
                           "base": {"type": "number", "exclusiveMinimum": 0}, "count": _COUNT},
    required=["type", "minval", "maxval", "base"], **_MINMAX)
obj("hyperparameter-const", {"type": {"const": "const"}, "val": True}, required=["type", "val"])
obj("hyperparameter-categorical", {"type": {"const": "categorical"}, "vals": {"type": "array", "minLength": 1}}, required=["type", "vals"])
schema("hyperparameter", **union(
    tagged("type", {k: "hyperparameter-" + k for k in ("int", "double", "log", "const", "categorical")}) + [
        ("always", {"type": "object", "additionalProperties": ref("hyperparameter"),
                    "checks": {'if a hyperparameter object\'s ["type"] is set, it must be one of "int", "double", "log", const", or '
                               '"categorical"': {"properties": {"type": False}}}}),
        ("never", {"not": {"type": "object"}})]))
schema("hyperparameters", type="object", additionalProperties=ref("hyperparameter"))
# with a grid searcher every numeric hyperparameter needs a count, however deeply it is nested
schema("check-grid-hyperparameter", **union([
    ("type:array", {"type": "array", "items": ref("check-grid-hyperparameter")}),
    ("not:hasattr:type", {"type": "object", "properties": {"type": False}, "additionalProperties": ref("check-grid-hyperparameter")}),
    ("never", {"not": {"type": ["object", "array"]}}),
    ("hasattr:type", {"type": "object", "required": ["type"], "properties": {"type": {"type": "string"}},
                      "checks": {"grid search is in use but count was not provided":
                                 {"if": {"properties": {"type": {"enum": ["double", "log", "int"]}}},
                                  "then": {"not": {"properties": {"count": {"type": "null"}}}}}}})]))

# ---- searchers ------------------------------------------------------------------------------------------------------
_SLEN = opt("searcher-length", ["object", "integer"])
_PLEN = opt("check-positive-length", "object")
_MODE = enum(["aggressive", "standard", "conservative"], "standard")
_DIV = n("number", 4, exclusiveMinimum=1)
_CONC = n("integer", 16, minimum=0)
obj("searcher-single", {"name": {"const": "single"}, "max_length": _SLEN, **SEARCH_COMMON}, required=["name"], later=["max_length", "metric"])
obj("searcher-random", {"name": {"const": "random"}, "max_concurrent_trials": _CONC, "max_trials": n("integer", None, minimum=1),
                        "max_length": _SLEN, **SEARCH_COMMON}, required=["name"], later=["max_trials", "max_length", "metric"])
obj("searcher-grid", {"name": {"const": "grid"}, "max_concurrent_trials": _CONC, "max_length": _SLEN, **SEARCH_COMMON},
    required=["name"], later=["max_length", "metric"])
obj("searcher-custom", {"name": {"const": "custom"}, "metric": n("string"), "smaller_is_better": n("boolean", True),
                        "unit": {"enum": ["batches", "records", "epochs", None], "default": None}}, required=["name"], later=["metric"], extra=True)
obj("searcher-async-halving", {"name": {"const": "async_halving"}, "num_rungs": n("integer", None, minimum=1), "max_length": _PLEN,
                               "max_trials": n("integer", None, minimum=1), "divisor": _DIV, "max_concurrent_trials": _CONC,
                               "stop_once": n("boolean", False), **SEARCH_COMMON},
    required=["name"], later=["num_rungs", "max_length", "max_trials", "metric"])
obj("searcher-adaptive-asha", {"name": {"const": "adaptive_asha"}, "bracket_rungs": n("array", [], items={"type": "integer"}),
                               "max_trials": n("integer", None, minimum=1), "mode": _MODE, "divisor": _DIV, "max_rungs": n("integer", 5, minimum=1),
                               "max_concurrent_trials": _CONC, "max_length": _SLEN, "stop_once": n("boolean", False), **SEARCH_COMMON},
    required=["name"], later=["max_length", "max_trials", "metric"])
# end-of-life searchers: still parsed so that old experiments can be read and forked
obj("searcher-adaptive", {"name": {"const": "adaptive"}, "budget": opt("length", "object"), "bracket_rungs": n("array", [], items={"type": "integer"}),
                          "mode": _MODE, "divisor": _DIV, "max_rungs": n("integer", 5, minimum=1), "max_length": _PLEN,
                          "train_stragglers": n("boolean", True), **SEARCH_COMMON}, required=["name"], later=["budget", "max_length", "metric"])
obj("searcher-adaptive-simple", {"name": {"const": "adaptive_simple"}, "max_trials": n("integer", None, minimum=1, maximum=2000), "mode": _MODE,
                                 "divisor": _DIV, "max_rungs": n("integer", 5, minimum=1), "max_length": _PLEN, **SEARCH_COMMON},
    required=["name"], later=["max_trials", "max_length", "metric"])
obj("searcher-sync-halving", {"name": {"const": "sync_halving"}, "budget": _PLEN, "num_rungs": n("integer", None, minimum=1), "max_length": _PLEN,
                              "divisor": _DIV, "train_stragglers": n("boolean", True), **SEARCH_COMMON},
    required=["name"], later=["num_rungs", "max_length", "budget", "metric"])
_SEARCHER_KEYS = ["bracket_rungs", "divisor", "max_concurrent_trials", "max_length", "max_rungs", "max_trials", "mode", "name", "num_rungs",
                  "stop_once", "budget", "train_stragglers", "unit"]
obj("searcher", {**{k: True for k in _SEARCHER_KEYS}, **SEARCH_COMMON}, later=["name", "metric"],
    **{"if": {"required": ["name"]},
       "then": union(tagged("name", {"single": "searcher-single", "random": "searcher-random", "grid": "searcher-grid", "custom": "searcher-custom",
                                     "adaptive_asha": "searcher-adaptive-asha", "async_halving": "searcher-async-halving",
                                     "adaptive": "searcher-adaptive", "adaptive_simple": "searcher-adaptive-simple",
                                     "sync_halving": "searcher-sync-halving"}),
                     "is not an object where object[\"name\"] is one of 'single', 'random', 'grid', 'custom', or 'adaptive_asha'")})
REGISTRY[url("searcher")].pop("type")

# ---- the smaller sections -------------------------------------------------------------------------------------------
obj("optimizations", {"aggregation_frequency": n("integer", 1, minimum=1), "auto_tune_tensor_fusion": n("boolean", False),
                      "average_aggregated_gradients": n("boolean", True), "average_training_metrics": n("boolean", True),
                      "gradient_compression": n("boolean", False), "grad_updates_size_file": n("string"),
                      "mixed_precision": {"enum": [None, "O0", "O1", "O2", "O3"], "default": "O0",
                                          "checks": {"mixed_precision should be a string starting with an uppercase letter 'O'": {"pattern": "^O"}}},
                      "tensor_fusion_cycle_time": n("integer", 1, minimum=0), "tensor_fusion_threshold": n("integer", 64, minimum=0),
                      # this build's own switches (no counterpart in the reference schema): replay train_batch as a CUDA graph,
                      # write periodic checkpoints from a background thread
                      "cuda_graphs": n("boolean"), "async_checkpoint": n("boolean"),
                      # the gradient engine: force the flat-buffer engine on/off (default: on for CUDA), bucket size of the overlapped
                      # exchange, and whether to overlap it with backward at all (default: by gradient volume)
                      "fused_engine": n("boolean"), "bucket_cap_mb": n("number", None, exclusiveMinimum=0), "grad_overlap": n("boolean")})
obj("profiling", {"enabled": n("boolean", False), "begin_on_batch": n("integer", 0, minimum=0), "end_after_batch": n("integer", None, minimum=0),
                  "sync_timings": n("boolean", True)}, compareProperties={"type": "a<=b", "a": "begin_on_batch", "b": "end_after_batch"})
obj("reproducibility", {"experiment_seed": n("integer", None, minimum=0)}, later=["experiment_seed"])
obj("resources", {"agent_label": n("string"), "devices": opt("devices", "array", []), "is_single_node": n("boolean"), "max_slots": n("integer"),
                  "native_parallel": n("boolean", False), "priority": n("integer", None, minimum=1, maximum=99), "resource_pool": n("string", ""),
                  "shm_size": n(["integer", "string"], None, checks={"must be a valid memory size":
                                                                     {"pattern": "^([0-9]*[.])?[0-9]+ ?(([kmgtpKMGTP]([iI]?[bB])?)|[bB])?$"}}),
                  "slots": n("integer"), "slots_per_trial": n("integer", 1, minimum=0), "weight": n("number", 1)})
obj("retention-policy", {"log_retention_days": n("integer", None, minimum=-1, maximum=32767)}, later=["log_retention_days"])
obj("kerberos", {"config_file": {"type": "string"}}, required=["config_file"])
obj("security", {"kerberos": opt("kerberos", "object")})
obj("hpc-cluster-slurm", {"slots_per_node": n("integer", None, minimum=1), "gpu_type": n("string"), "sbatch_args": n("array", None, items={"type": "string"})})
obj("hpc-cluster-pbs", {"slots_per_node": n("integer", None, minimum=1), "pbsbatch_args": n("array", None, items={"type": "string"})})
obj("log-action-cancel-retries", {"type": {"const": "cancel_retries"}}, required=["type"], later=[])
obj("log-action-exclude-node", {"type": {"const": "exclude_node"}}, required=["type"], later=[])
obj("log-action", {"type": True}, later=["type"],
    **{"if": {"required": ["type"]},
       "then": union(tagged("type", {"cancel_retries": "log-action-cancel-retries", "exclude_node": "log-action-exclude-node"}),
                     "is not an object where object[\"type\"] is one of 'cancel_retries' or 'exclude_node'")})
REGISTRY[url("log-action")].pop("type")
obj("log-policy", {"pattern": {"type": ["string"]}, "action": {"type": ["object"], **ref("log-action")}}, required=["pattern", "action"])

# ---- the experiment -------------------------------------------------------------------------------------------------
obj("experiment", {
    "bind_mounts": opt("bind-mounts", "array", []), "checkpoint_policy": enum(["best", "all", "none"], "best"),
    "checkpoint_storage": opt("checkpoint-storage", "object"), "data": n("object", {}), "data_layer": n("object"), "debug": n("boolean", False),
    "description": n("string"), "entrypoint": n(["string", "array"], None, items={"type": "string"}),
    "environment": opt("environment", "object", {}), "hyperparameters": opt("hyperparameters", "object", {}),
    "internal": {"type": "null", "default": None}, "labels": n("array", [], items={"type": "string"}),
    "log_policies": n("array", [], items=ref("log-policy")), "retention_policy": opt("retention-policy", "object"),
    "max_restarts": n("integer", 5, minimum=0), "min_checkpoint_period": opt("length", "object", {"batches": 0}),
    "min_validation_period": opt("length", "object", {"batches": 0}), "name": n("string"), "optimizations": opt("optimizations", "object", {}),
    "pbs": opt("hpc-cluster-pbs", "object", {}), "perform_initial_validation": n("boolean", False), "profiling": opt("profiling", "object", {}),
    "project": n("string", ""), "records_per_epoch": n("integer", 0), "reproducibility": opt("reproducibility", "object", {}),
    "resources": opt("resources", "object", {}), "scheduling_unit": n("integer", 100, minimum=1), "searcher": opt("searcher", "object"),
    "security": opt("security", "object"), "slurm": opt("hpc-cluster-slurm", "object", {}),
    "tensorboard_storage": opt("tensorboard-storage", "object"), "workspace": n("string", "")},
    later=["checkpoint_storage", "name", "hyperparameters", "reproducibility", "searcher"],
    allOf=[{"if": {"properties": {"searcher": {"properties": {"name": {"const": "grid"}}}}},
            "then": {"properties": {"hyperparameters": {"additionalProperties": ref("check-grid-hyperparameter")}}}}])
