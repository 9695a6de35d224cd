"""StorageManager contract.

Parity: reference ``harness/determined/common/storage/base.py:26-245``.  Method names and
semantics (``store_path`` / ``restore_path`` context managers, ``upload`` / ``download`` with
path lists and selectors, ``delete(tgt, globs) -> remaining resources``,
``store_path_is_direct_access``) are the public contract used by ``core.CheckpointContext``.

Implementation notes (ours): glob application for non-filesystem backends is done with a pure
in-memory matcher instead of materialising a temp dir of empty files.
"""
import abc
import contextlib
import fnmatch
import os
import pathlib
import re
import shutil
import tempfile
import uuid
from typing import Any, Callable, Dict, Iterator, List, Optional, Set, Union

Paths = Set[str]
Selector = Callable[[str], bool]


def list_directory(root: Union[str, os.PathLike]) -> Dict[str, int]:
    """``{relative path: size}``; directories end in ``/`` and have size 0."""
    root = os.fspath(root)
    if not os.path.exists(root):
        raise FileNotFoundError(root)
    if not os.path.isdir(root):
        raise NotADirectoryError(root)
    out: Dict[str, int] = {}
    stack = [root]
    while stack:
        cur = stack.pop()
        with os.scandir(cur) as it:
            for e in it:
                rel = os.path.relpath(e.path, root)
                if e.is_dir(follow_symlinks=False):
                    out[rel + "/"] = 0
                    stack.append(e.path)
                else:
                    try:
                        out[rel] = e.stat(follow_symlinks=False).st_size
                    except FileNotFoundError:
                        pass
    return out


def _glob_to_regex(glob: str) -> "re.Pattern[str]":
    """Translate a ``glob.glob(recursive=True)``-style pattern into a regex over ``/`` paths."""
    i, n, out = 0, len(glob), []
    while i < n:
        c = glob[i]
        # ``**`` is recursive only as a whole path component (glob.glob's rule): ``**.txt`` is just ``*.txt``
        at_component_start = i == 0 or glob[i - 1] == "/"
        if at_component_start and glob.startswith("**/", i):
            out.append("(?:[^/]+/)*")
            i += 3
        elif at_component_start and glob.startswith("**", i) and i + 2 == n:
            out.append(".*")
            i += 2
        elif glob.startswith("**", i):
            out.append("[^/]*")
            i += 2
        elif c == "*":
            out.append("[^/]*")
            i += 1
        elif c == "?":
            out.append("[^/]")
            i += 1
        elif c == "[":
            j = glob.find("]", i + 1)
            if j < 0:
                out.append(re.escape(c))
                i += 1
            else:
                out.append(fnmatch.translate(glob[i:j + 1])[4:-3])
                i = j + 1
        else:
            out.append(re.escape(c))
            i += 1
    return re.compile("".join(out) + r"/?\Z")


def apply_globs_to_resources(resources: Dict[str, int], prefix: str, globs: List[str]) -> Dict[str, int]:
    """Return resources that *survive* deleting everything matched by ``globs`` (deleting a
    directory deletes its contents)."""
    pats = [_glob_to_regex(os.path.normpath(os.path.join(prefix, g)) if prefix else g) for g in globs]
    doomed_dirs, doomed = [], set()
    for p in resources:
        bare = p.rstrip("/")
        if any(r.match(bare) for r in pats):
            doomed.add(p)
            if p.endswith("/"):
                doomed_dirs.append(p)
    return {p: s for p, s in resources.items()
            if p not in doomed and not any(p.startswith(d) for d in doomed_dirs)}


class StorageManager(metaclass=abc.ABCMeta):
    """Abstract persistent store addressed by ``storage_id`` (a checkpoint uuid)."""

    def __init__(self, base_path: Union[str, os.PathLike]) -> None:
        base_path = os.fspath(base_path) if isinstance(base_path, os.PathLike) else base_path
        if not isinstance(base_path, str) or not base_path:
            raise ValueError("base_path must be a non-empty string")
        self._base_path = base_path

    @classmethod
    def from_config(cls, config: Dict[str, Any], container_path: Optional[str]) -> "StorageManager":
        return cls(**config)

    def pre_store_path(self, dst: str) -> pathlib.Path:
        """Directory the caller should write into before :meth:`post_store_path`."""
        p = pathlib.Path(self._base_path) / dst
        p.mkdir(parents=True, exist_ok=True)
        return p

    @abc.abstractmethod
    def post_store_path(self, src: Union[str, os.PathLike], dst: str) -> None:
        """Finalize: make what was written under ``src`` durable as ``dst``."""

    @contextlib.contextmanager
    def store_path(self, dst: str) -> Iterator[pathlib.Path]:
        path = self.pre_store_path(dst)
        yield path
        self.post_store_path(path, dst)

    def store_path_is_direct_access(self) -> bool:
        """True if files written in ``store_path`` land in the final location (no upload step)."""
        return False

    @abc.abstractmethod
    def restore_path(self, src: str, selector: Optional[Selector] = None) -> "contextlib.AbstractContextManager[pathlib.Path]":
        ...

    @abc.abstractmethod
    def upload(self, src: Union[str, os.PathLike], dst: str, paths: Optional[Paths] = None) -> None:
        ...

    @abc.abstractmethod
    def download(self, src: str, dst: Union[str, os.PathLike], selector: Optional[Selector] = None) -> None:
        """``selector(path)`` gets relative paths (directories end with ``/``)."""

    @abc.abstractmethod
    def delete(self, tgt: str, globs: List[str]) -> Dict[str, int]:
        """Delete matching files; return the resources that remain."""

    _list_directory = staticmethod(list_directory)
    _apply_globs_to_resources = staticmethod(apply_globs_to_resources)


def from_string(shortcut: str) -> StorageManager:
    """``/abs/path`` | ``file:///p`` | ``s3://bucket/prefix`` | ``gs://bucket/prefix``.
    Parity: reference ``base.py:217-245`` + ``shared.py:101-127``."""
    import urllib.parse

    from determined_b200.common import storage

    u = urllib.parse.urlparse(shortcut)
    if u.params or u.query or u.fragment:
        raise ValueError(f'Malformed checkpoint_storage string "{shortcut}"')
    scheme = u.scheme.lower()
    if scheme in ("", "file"):
        if u.netloc:
            raise ValueError(f'Malformed checkpoint_storage string "{shortcut}": a file path cannot name a host')
        return storage.SharedFSStorageManager(base_path=u.path if scheme else shortcut)
    if scheme == "s3":
        return storage.S3StorageManager(bucket=u.netloc, prefix=u.path.lstrip("/"))
    if scheme == "gs":
        return storage.GCSStorageManager(bucket=u.netloc, prefix=u.path.lstrip("/"))
    raise ValueError(f'Unsupported checkpoint_storage string "{shortcut}": only absolute paths, file://, s3:// and gs:// are understood')
