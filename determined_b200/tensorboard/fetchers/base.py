"""Fetchers bring TensorBoard files from checkpoint storage to a local directory, incrementally (reference
``harness/determined/tensorboard/fetchers``; used by the tensorboard task to follow running trials)."""
import abc
import datetime
import logging
import os
from typing import Any, Callable, Dict, Generator, List, Optional, Tuple

logger = logging.getLogger("determined.tensorboard.fetchers")


class Fetcher(metaclass=abc.ABCMeta):
    """``storage_paths`` are the prefixes to follow (one per trial); files land under ``local_dir`` at the same relative path.
    A file is fetched again only when its modification time in storage moved forward."""

    def __init__(self, storage_config: Dict[str, Any], storage_paths: List[str], local_dir: str) -> None:
        self.storage_config, self.storage_paths, self.local_dir = storage_config, list(storage_paths), local_dir
        self._seen: Dict[str, Any] = {}          # path -> modification stamp at the time it was last offered

    @abc.abstractmethod
    def _find(self, storage_path: str) -> Generator[Tuple[str, Any], None, None]:
        """Yield ``(filepath, modification stamp)`` of everything below ``storage_path``."""

    @abc.abstractmethod
    def _fetch(self, filepath: str, new_file_callback: Callable[[], None]) -> None:
        """Copy one file into ``local_dir`` and call ``new_file_callback()``."""

    def _list(self, storage_path: str) -> Generator[str, None, None]:
        for path, stamp in self._find(storage_path):
            if self._seen.get(path) is None or self._seen[path] < stamp:
                self._seen[path] = stamp
                yield path

    def list_all_generator(self) -> Generator[str, None, None]:
        """New or changed files across every storage path."""
        for sp in self.storage_paths:
            yield from self._list(sp)

    def fetch_new(self, new_file_callback: Callable[[], None] = lambda: None) -> int:
        n = 0
        for path in list(self.list_all_generator()):
            self._fetch(path, new_file_callback)
            n += 1
        return n

    def _local_path(self, filepath: str) -> str:
        dst = os.path.join(self.local_dir, filepath.lstrip("/"))
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        return dst


def _aware(t: Optional[datetime.datetime]) -> datetime.datetime:
    if t is None:
        return datetime.datetime.fromtimestamp(0, datetime.timezone.utc)
    return t if t.tzinfo is not None else t.replace(tzinfo=datetime.timezone.utc)
