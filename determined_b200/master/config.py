"""The master's own configuration: file, environment, command line — one schema, one merged result.

Parity: reference ``master/internal/config/config.go`` (``Config`` with ``db``, ``security`` incl. ``tls`` and
``initial_user_password``, ``checkpoint_storage``, ``task_container_defaults``, ``resource_manager`` / ``resource_pools``,
``telemetry``, ``webhooks``, ``observability``, ``retention_policy``, ``notebook_timeout`` / ``tensorboard_timeout``,
``cluster_name``, ``log``), its viper loading order (defaults < ``master.yaml`` < ``DET_*`` environment < flags) and
``Printable()`` (secrets masked, what ``GET /api/v1/master/config`` serves).

Sources, weakest first:

1. the defaults of :data:`SCHEMA`;
2. the YAML file (``--config-file`` / ``$DET_MASTER_CONFIG``);
3. environment variables ``DET_<PATH>`` where ``<PATH>`` is the dotted key upper-cased with ``_`` for ``.``
   (``DET_PORT``, ``DET_DB_PATH``, ``DET_SECURITY_TLS_CERT``, ``DET_RESOURCE_MANAGER_SCHEDULER_TYPE``); values are parsed as
   YAML scalars, so ``DET_PORT=8443`` is an integer and ``DET_SECURITY_TLS_CERT=/etc/det/cert.pem`` a string;
4. command-line flags (``python -m determined_b200.master --port ...``).

The result is validated against the schema with the same evaluator experiment configs go through
(``master/schema_engine.py``); errors name the offending key.
"""
import copy
import os
from typing import Any, Dict, List, Mapping, Optional

import yaml

from determined_b200.master import schema_engine

_SCHED = {"type": "object", "additionalProperties": False, "properties": {
    "type": {"enum": ["fair_share", "priority", "round_robin"], "default": "fair_share"},
    "fitting_policy": {"enum": ["best", "worst"], "default": "best"},
    "preemption": {"type": "object", "additionalProperties": False, "default": {},
                   "properties": {"enabled": {"type": "boolean", "default": True}}},
    "default_priority": {"type": "integer", "minimum": 1, "maximum": 99, "default": 42},
}}

SCHEMA: Dict[str, Any] = {
    "$id": "http://determined_b200/schemas/master/v0/config.json", "type": "object", "additionalProperties": False,
    "properties": {
        "config_file": {"type": ["string", "null"], "default": None},
        "host": {"type": "string", "default": "127.0.0.1"},
        "port": {"type": "integer", "minimum": 0, "maximum": 65535, "default": 8080},
        "cluster_name": {"type": "string", "default": "b200-box"},
        "root": {"type": ["string", "null"], "default": None, "description": "work directory for task run directories (default: a temporary directory)"},
        "log": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "level": {"enum": ["trace", "debug", "info", "warn", "warning", "error", "fatal"], "default": "info"},
            "color": {"type": "boolean", "default": True}}},
        "db": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "path": {"type": ["string", "null"], "default": None, "description": "sqlite file (default ~/.local/share/determined_b200/master.db)"}}},
        "checkpoint_storage": {"type": ["object", "null"], "default": None},
        "task_container_defaults": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "environment_variables": {"type": "array", "items": {"type": "string", "pattern": "^[^=]+=.*$"}, "default": []},
            "work_dir": {"type": ["string", "null"], "default": None},
            "bind_mounts": {"type": "array", "default": []},
            "shm_size_bytes": {"type": ["integer", "null"], "default": None},
            "network_mode": {"type": "string", "default": "host"}}},
        "resource_manager": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "type": {"enum": ["agent"], "default": "agent"},
            "scheduler": dict(_SCHED, default={}),
            "default_compute_resource_pool": {"type": "string", "default": "default"},
            "default_aux_resource_pool": {"type": "string", "default": "default"}}},
        "resource_pools": {"type": ["array", "null"], "default": None, "items": {"type": "object", "required": ["pool_name"], "properties": {
            "pool_name": {"type": "string", "minLength": 1}, "description": {"type": "string"}, "scheduler": _SCHED,
            "provider": {"type": ["object", "null"]}, "max_aux_containers_per_agent": {"type": "integer", "minimum": 0}}}},
        "security": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "tls": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
                "cert": {"type": ["string", "null"], "default": None}, "key": {"type": ["string", "null"], "default": None},
                "name": {"type": ["string", "null"], "default": None, "description": "host name in the certificate, when tasks reach the master by another name"}}},
            "agent_token": {"type": ["string", "null"], "default": None},
            "initial_user_password": {"type": ["string", "null"], "default": None, "description": "password given to the built-in admin / determined users at first start"},
            "default_task": {"type": "object", "default": {}},
            "authz": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
                "type": {"enum": ["basic", "rbac"], "default": "basic"},
                "strict_job_queue_control": {"type": "boolean", "default": False}}}}},
        "telemetry": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "enabled": {"type": "boolean", "default": False},
            "segment_url": {"type": ["string", "null"], "default": None, "description": "collector that receives the usage events; none are sent without it"},
            "segment_master_key": {"type": ["string", "null"], "default": None},
            "otel_enabled": {"type": "boolean", "default": True},
            "otel_endpoint": {"type": ["string", "null"], "default": None}}},
        "oidc": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "enabled": {"type": "boolean", "default": False}, "provider": {"type": "string", "default": "oidc"},
            "client_id": {"type": ["string", "null"], "default": None}, "client_secret": {"type": ["string", "null"], "default": None},
            "idp_sso_url": {"type": ["string", "null"], "default": None}, "token_url": {"type": ["string", "null"], "default": None},
            "userinfo_url": {"type": ["string", "null"], "default": None}, "idp_recipient_url": {"type": ["string", "null"], "default": None},
            "authentication_claim": {"type": "string", "default": "email"}, "scope": {"type": "string", "default": "openid profile email"},
            "auto_provision_users": {"type": "boolean", "default": False}, "always_redirect": {"type": "boolean", "default": False}}},
        "scim": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "enabled": {"type": "boolean", "default": False},
            "auth": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
                "type": {"enum": ["basic", "oauth"], "default": "basic"},
                "username": {"type": ["string", "null"], "default": None}, "password": {"type": ["string", "null"], "default": None}}}}},
        "logging": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "type": {"enum": ["default", "elastic"], "default": "default", "description": "where task logs live: the master's database, or Elasticsearch"},
            "host": {"type": ["string", "null"], "default": None}, "port": {"type": "integer", "default": 9200},
            "index_prefix": {"type": "string", "default": "determined-tasklogs"},
            "security": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
                "username": {"type": ["string", "null"], "default": None}, "password": {"type": ["string", "null"], "default": None},
                "tls": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
                    "enabled": {"type": "boolean", "default": False}, "skip_verify": {"type": "boolean", "default": False},
                    "certificate": {"type": ["string", "null"], "default": None}}}}}}},
        "observability": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "enable_prometheus": {"type": "boolean", "default": True}}},
        "webhooks": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "signing_key": {"type": ["string", "null"], "default": None},
            "base_url": {"type": ["string", "null"], "default": None}}},
        "retention_policy": {"type": "object", "additionalProperties": False, "default": {}, "properties": {
            "log_retention_days": {"type": "integer", "minimum": -1, "default": -1, "description": "task logs older than this are deleted; -1 keeps them forever"},
            "schedule": {"type": "string", "default": "0 1 * * *", "description": "when the sweep runs (only the hour and minute fields are used)"}}},
        "notebook_timeout": {"type": ["integer", "null"], "minimum": 0, "default": None, "description": "idle seconds after which notebooks are killed"},
        "tensorboard_timeout": {"type": "integer", "minimum": 0, "default": 300},
        "enable_cors": {"type": "boolean", "default": False},
        "feature_switches": {"type": "array", "items": {"type": "string"}, "default": []},
    },
}
REGISTRY = {SCHEMA["$id"]: SCHEMA}

SECRET_KEYS = ("scim.auth.password", "logging.security.password", "oidc.client_secret", "telemetry.segment_master_key", "security.agent_token", "security.initial_user_password", "security.tls.key", "webhooks.signing_key",
               "checkpoint_storage.secret_key", "checkpoint_storage.access_key", "checkpoint_storage.connection_string",
               "checkpoint_storage.credentials")


class MasterConfigError(ValueError):
    def __init__(self, errors: List[str]) -> None:
        super().__init__("invalid master configuration:\n  " + "\n  ".join(errors))
        self.errors = errors


def _leaf_paths(schema: Dict[str, Any], prefix: str = "") -> List[str]:
    out: List[str] = []
    for k, sub in (schema.get("properties") or {}).items():
        p = f"{prefix}.{k}" if prefix else k
        if isinstance(sub, dict) and sub.get("type") == "object" and sub.get("properties"):
            out += _leaf_paths(sub, p)
        else:
            out.append(p)
    return out


def _set(d: Dict[str, Any], dotted: str, value: Any) -> None:
    parts = dotted.split(".")
    for p in parts[:-1]:
        nxt = d.get(p)
        if not isinstance(nxt, dict):
            nxt = d[p] = {}
        d = nxt
    d[parts[-1]] = value


def _get(d: Any, dotted: str) -> Any:
    for p in dotted.split("."):
        if not isinstance(d, dict) or p not in d:
            return None
        d = d[p]
    return d


def _merge(base: Any, over: Any) -> Any:
    if isinstance(base, dict) and isinstance(over, dict):
        out = dict(base)
        for k, v in over.items():
            out[k] = _merge(base.get(k), v) if k in base else copy.deepcopy(v)
        return out
    return copy.deepcopy(over)


def from_env(environ: Mapping[str, str]) -> Dict[str, Any]:
    """``DET_<DOTTED_KEY_IN_UPPER_CASE>`` for every leaf key of the schema."""
    out: Dict[str, Any] = {}
    for path in _leaf_paths(SCHEMA):
        name = "DET_" + path.replace(".", "_").upper()
        if name in environ and path != "config_file":
            try:
                val = yaml.safe_load(environ[name])
            except yaml.YAMLError:
                val = environ[name]
            _set(out, path, val)
    return out


def load(config_file: Optional[str] = None, environ: Optional[Mapping[str, str]] = None, flags: Optional[Dict[str, Any]] = None) -> "MasterConfig":
    """defaults < file < environment < flags (``flags``: dotted key -> value, ``None`` values ignored)."""
    environ = os.environ if environ is None else environ
    config_file = config_file or environ.get("DET_MASTER_CONFIG")
    raw: Dict[str, Any] = {}
    if config_file:
        with open(config_file) as f:
            loaded = yaml.safe_load(f) or {}
        if not isinstance(loaded, dict):
            raise MasterConfigError([f"{config_file}: must contain a mapping at the top level"])
        raw = loaded
        # the pre-schema layout of this file: `db: {path}` was already there, `security.tls` too; nothing to translate
    raw = _merge(raw, from_env(environ))
    over: Dict[str, Any] = {}
    for k, v in (flags or {}).items():
        if v is not None:
            _set(over, k, v)
    raw = _merge(raw, over)
    if config_file:
        raw["config_file"] = config_file
    return MasterConfig(raw)


class MasterConfig:
    def __init__(self, raw: Optional[Dict[str, Any]] = None) -> None:
        raw = copy.deepcopy(raw or {})
        errs = schema_engine.Validator(REGISTRY, complete=True).errors(SCHEMA, raw)
        if errs:
            raise MasterConfigError(errs)
        self.data: Dict[str, Any] = schema_engine.fill_defaults(REGISTRY, SCHEMA, raw)
        tls = self.data["security"]["tls"]
        if bool(tls.get("cert")) != bool(tls.get("key")):
            raise MasterConfigError(["<config>.security.tls: cert and key must be given together"])
        names = [p["pool_name"] for p in (self.data.get("resource_pools") or [])]
        if len(names) != len(set(names)):
            raise MasterConfigError(["<config>.resource_pools: pool names must be unique"])

    def __getitem__(self, dotted: str) -> Any:
        return _get(self.data, dotted)

    def get(self, dotted: str, default: Any = None) -> Any:
        v = _get(self.data, dotted)
        return default if v is None else v

    @property
    def db_path(self) -> str:
        return self["db.path"] or os.path.expanduser("~/.local/share/determined_b200/master.db")

    def cluster_defaults(self) -> Dict[str, Any]:
        """What every experiment config is completed with (reference: ``checkpoint_storage`` and the task-container
        defaults of the master config are merged into experiment configs)."""
        out: Dict[str, Any] = {}
        if self.data.get("checkpoint_storage"):
            out["checkpoint_storage"] = copy.deepcopy(self.data["checkpoint_storage"])
        tcd = self.data["task_container_defaults"]
        if tcd.get("environment_variables"):
            out.setdefault("environment", {})["environment_variables"] = list(tcd["environment_variables"])
        if tcd.get("bind_mounts"):
            out["bind_mounts"] = copy.deepcopy(tcd["bind_mounts"])
        return out

    def printable(self) -> Dict[str, Any]:
        """The configuration with secrets masked (``GET /api/v1/master/config``, the start-up log line)."""
        out = copy.deepcopy(self.data)
        for path in SECRET_KEYS:
            if _get(out, path):
                _set(out, path, "********")
        return out
