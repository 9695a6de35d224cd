"""Checkpoint storage managers (reference ``harness/determined/common/storage/__init__.py:1-111``)."""
import copy
import uuid
from typing import Any, Dict, Optional

from determined_b200.common.storage.base import (  # noqa: F401
    Paths, Selector, StorageManager, apply_globs_to_resources, from_string, list_directory,
)
from determined_b200.common.storage.shared import SharedFSStorageManager, copytree  # noqa: F401
from determined_b200.common.storage.directory import DirectoryStorageManager  # noqa: F401
from determined_b200.common.storage.cloud import (  # noqa: F401
    AzureStorageManager, CloudStorageManager, GCSStorageManager, S3StorageManager,
)

_STORAGE_MANAGERS = {
    "shared_fs": SharedFSStorageManager,
    "directory": DirectoryStorageManager,
    "s3": S3StorageManager,
    "gcs": GCSStorageManager,
    "azure": AzureStorageManager,
}
_GC_KEYS = ("save_experiment_best", "save_trial_best", "save_trial_latest")


def build(config: Dict[str, Any], container_path: Optional[str] = None) -> StorageManager:
    """Instantiate the manager named by ``config["type"]``."""
    if "type" not in config:
        raise ValueError("Missing 'type' parameter of storage configuration")
    cfg = copy.deepcopy(config)
    kind = cfg.pop("type")
    if not isinstance(kind, str):
        raise ValueError("`type` parameter of storage configuration must be a string")
    if kind not in _STORAGE_MANAGERS:
        raise TypeError(f"Unknown storage type: {kind}")
    for k in _GC_KEYS:
        cfg.pop(k, None)
    if kind == "shared_fs" and "storage_path" not in cfg:
        # legacy keys fold into storage_path
        legacy = cfg.get("tensorboard_path", cfg.get("checkpoint_path"))
        if legacy is not None:
            cfg["storage_path"] = legacy
    if kind == "azure":
        if not ("connection_string" in cfg or "account_url" in cfg):
            raise ValueError("one of [connection_string, account_url] is required for azure storage")
        if "container" not in cfg:
            raise ValueError("container is required for azure storage")
    cfg.pop("tensorboard_path", None)
    cfg.pop("checkpoint_path", None)
    cfg = {k: v for k, v in cfg.items() if v is not None}
    try:
        return _STORAGE_MANAGERS[kind].from_config(cfg, container_path)
    except TypeError as e:
        raise TypeError(f"Failed to instantiate {kind} checkpoint storage: {e}")


def validate_manager(manager: StorageManager) -> None:
    """Round-trip a tiny file: store, restore, delete."""
    sid = str(uuid.uuid4())
    with manager.store_path(sid) as p:
        p.mkdir(parents=True, exist_ok=True)
        with (p / "VALIDATE.txt").open("w") as f:
            f.write(sid)
    with manager.restore_path(sid) as p:
        with (p / "VALIDATE.txt").open("r") as f:
            if f.read() != sid:
                raise ValueError("Unable to properly load from storage")
    manager.delete(sid, ["**/*"])


def validate_config(config: Dict[str, Any], container_path: Optional[str] = None) -> None:
    validate_manager(build(config, container_path))


def normalize_prefix(prefix: Optional[str]) -> str:
    """A bucket prefix without leading slashes and with exactly one trailing slash (empty stays empty); ``..`` is refused."""
    new = (prefix or "").strip().lstrip("/")
    if any(part == ".." for part in new.split("/")):
        raise ValueError("prefix must not contain /../")
    return new.rstrip("/") + "/" if new else ""
