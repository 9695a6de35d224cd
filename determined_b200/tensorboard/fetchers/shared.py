"""Shared-filesystem / plain-directory fetchers."""
import os
import shutil
from typing import Any, Callable, Dict, Generator, List, Tuple

from determined_b200.tensorboard.fetchers import base


class SharedFSFetcher(base.Fetcher):
    """``storage_paths`` are directories on a filesystem this process can read (absolute, or relative to the configured
    ``host_path`` [+ ``storage_path``])."""

    def __init__(self, storage_config: Dict[str, Any], storage_paths: List[str], local_dir: str) -> None:
        super().__init__(storage_config, storage_paths, local_dir)
        root = storage_config.get("host_path") or storage_config.get("container_path") or ""
        if root and storage_config.get("storage_path"):
            root = os.path.join(root, storage_config["storage_path"])
        self.root = root

    def _abs(self, p: str) -> str:
        return p if os.path.isabs(p) or not self.root else os.path.join(self.root, p)

    def _find(self, storage_path: str) -> Generator[Tuple[str, Any], None, None]:
        top = self._abs(storage_path)
        for dirpath, _, files in os.walk(top):
            for f in files:
                full = os.path.join(dirpath, f)
                try:
                    yield full, os.stat(full).st_mtime_ns
                except FileNotFoundError:
                    continue

    def _fetch(self, filepath: str, new_file_callback: Callable[[], None]) -> None:
        shutil.copyfile(filepath, self._local_path(filepath))
        new_file_callback()


class DirectoryFetcher(SharedFSFetcher):
    """``checkpoint_storage: {type: directory, container_path: ...}``."""
