"""``directory`` storage: a path that is already mounted in the task (reference
``harness/determined/common/storage/directory.py:13-61``)."""
from typing import Any, Dict, Optional

from determined_b200.common.storage import shared


class DirectoryStorageManager(shared.SharedFSStorageManager):
    @classmethod
    def from_config(cls, config: Dict[str, Any], container_path: Optional[str]) -> "DirectoryStorageManager":
        extra = set(config) - {"container_path"}
        if extra:
            raise TypeError(f"unexpected directory-storage keys: {sorted(extra)}")
        return cls(config["container_path"])
