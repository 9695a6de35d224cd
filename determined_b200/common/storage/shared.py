"""Shared-filesystem storage: checkpoints are plain directories ``<base>/<uuid>/``.

Parity: reference ``harness/determined/common/storage/shared.py:27-278``.  "Direct access": the
trial writes straight into the final directory, so there is no upload step — this is what makes
the ASHA save/restore path (BASELINE config #5) cheap on a single box.
"""
import contextlib
import glob as globlib
import logging
import os
import pathlib
import shutil
from typing import Any, Dict, Iterator, List, Optional, Union

from determined_b200 import errors
from determined_b200.common import check
from determined_b200 import util
from determined_b200.common.storage import base

logger = logging.getLogger("determined.storage")


def copytree(src: str, dst: str, selector: Optional[base.Selector] = None,
             paths: Optional[base.Paths] = None) -> None:
    """Copy the tree under ``src`` into ``dst``, optionally only what ``selector`` / ``paths`` pick.

    Relative names are what both see: files as ``sub/file``, directories as ``sub/``.  Every directory is visited
    whatever the selection says; a directory that is itself selected is created even when empty; a selected file brings
    its parent directories with it; nothing selected means ``dst`` is not created at all.  Symbolic links are followed, so
    a linked directory arrives as a real directory with its contents (the behaviour of the reference's
    ``common/storage/shared.py:copytree``, which its storage conformance tests pin down)."""
    src = os.fspath(src)
    dst = os.fspath(dst)

    def wanted(rel: str) -> bool:
        if paths is not None and rel not in paths:
            return False
        return selector is None or bool(selector(rel))

    unselected = selector is None and paths is None
    if unselected:
        os.makedirs(dst, exist_ok=True)
    touched = []          # (source dir, destination dir) of every directory that now exists under dst
    for cur, dirs, files in os.walk(src, followlinks=True):
        rel_cur = os.path.relpath(cur, src)
        prefix = "" if rel_cur == "." else rel_cur.replace(os.sep, "/") + "/"
        here = dst if rel_cur == "." else os.path.join(dst, rel_cur)
        for d in sorted(dirs):
            if wanted(prefix + d + "/"):
                os.makedirs(os.path.join(here, d), exist_ok=True)
                touched.append((os.path.join(cur, d), os.path.join(here, d)))
        for f in sorted(files):
            if not wanted(prefix + f):
                continue
            if not os.path.isdir(here):
                os.makedirs(here, exist_ok=True)
                touched.append((cur, here))
            shutil.copy2(os.path.join(cur, f), os.path.join(here, f))
    for s_dir, d_dir in reversed(touched):
        with contextlib.suppress(OSError):
            shutil.copystat(s_dir, d_dir)


def full_storage_path(host_path: str, storage_path: Optional[str] = None,
                      container_path: Optional[str] = None) -> str:
    """Where checkpoints live as seen by this process (reference ``common/storage/shared.py:_full_storage_path``):
    ``storage_path`` may be relative to ``host_path`` or an absolute path below it; with ``container_path`` the result is
    re-rooted there (the reference's containers bind ``host_path`` at ``container_path``)."""
    if not os.path.isabs(host_path):
        raise _PathError("`host_path` must be an absolute path.")
    root = host_path if container_path is None else container_path
    if storage_path is None:
        return root
    ab = os.path.normpath(os.path.join(host_path, storage_path))
    if not (ab == host_path or ab.startswith(host_path.rstrip("/") + "/")):
        raise _PathError("storage path must be a subdirectory of host path.")
    return os.path.join(root, os.path.relpath(ab, host_path))


class _PathError(check.CheckFailedError, ValueError):
    """Both what the reference raises (``check.CheckFailedError``) and what a Python caller expects (``ValueError``)."""


_full_storage_path = full_storage_path      # the reference's (private) name, which its own tests import


class SharedFSStorageManager(base.StorageManager):
    def _confined(self, storage_id: str) -> str:
        """Resolve ``storage_id`` under the base path and refuse anything that normalises to the base itself or to a
        location outside it (``../x``, absolute paths): these ids name directories that get copied from and rmtree'd."""
        base_dir = os.path.normpath(self._base_path)
        d = os.path.normpath(os.path.join(base_dir, storage_id))
        if d == base_dir or os.path.commonpath([base_dir, d]) != base_dir:
            raise ValueError(f"storage id {storage_id!r} escapes the checkpoint storage root")
        return d

    @classmethod
    def from_config(cls, config: Dict[str, Any], container_path: Optional[str]) -> "SharedFSStorageManager":
        allowed = {"host_path", "storage_path", "container_path", "propagation"}
        extra = set(config) - allowed
        if extra:
            raise TypeError(f"unexpected shared_fs keys: {sorted(extra)}")
        # Tasks here are host processes, so callers pass container_path=None and the host path is the path we see; a caller
        # that does run inside a mount (the reference's containers bind host_path at container_path) may still say so.
        return cls(full_storage_path(config["host_path"], config.get("storage_path"), container_path))

    def post_store_path(self, src: Union[str, os.PathLike], dst: str) -> None:
        # direct access: nothing to do.
        return None

    def store_path_is_direct_access(self) -> bool:
        return True

    @contextlib.contextmanager
    def restore_path(self, src: str, selector: Optional[base.Selector] = None) -> Iterator[pathlib.Path]:
        # Zero-copy: hand out the real directory.  (selector is irrelevant: nothing is copied.)
        if selector is not None:
            logger.warning("Ignoring partial checkpoint download from shared_fs; all files will be directly accessible from shared_fs.")
        d = self._confined(src)
        if not os.path.isdir(d):
            raise errors.CheckpointNotFound(f"checkpoint directory not found: {d}")
        yield pathlib.Path(d)

    def delete(self, tgt: str, globs: List[str]) -> Dict[str, int]:
        d = self._confined(tgt)
        if not os.path.exists(d):
            logger.info("storage directory does not exist: %s", d)
            return {}
        if not os.path.isdir(d):
            raise NotADirectoryError(d)
        if "**/*" in globs:
            util.rmtree_nfs_safe(d)
            return {}
        files, dirs = [], []
        for g in globs:
            for p in globlib.glob(os.path.join(globlib.escape(d), g), recursive=True):
                if os.path.commonpath([d, os.path.normpath(p)]) != d:
                    logger.warning("refusing to delete path outside checkpoint dir: %s", p)
                    continue
                (files if (os.path.isfile(p) or os.path.islink(p)) else dirs).append(p)
        for p in files:
            with contextlib.suppress(FileNotFoundError):
                os.remove(p)
        for p in dirs:
            if os.path.isdir(p):
                util.rmtree_nfs_safe(p)
        left = base.list_directory(d)
        if not left:
            util.rmtree_nfs_safe(d)
        return left

    def upload(self, src: Union[str, os.PathLike], dst: str, paths: Optional[base.Paths] = None) -> None:
        src = os.fspath(src)
        target = self._confined(dst)
        if os.path.realpath(src) == os.path.realpath(target):
            return
        copytree(src, target, paths=paths)

    def download(self, src: str, dst: Union[str, os.PathLike], selector: Optional[base.Selector] = None) -> None:
        source = self._confined(src)
        if not os.path.isdir(source):
            raise errors.CheckpointNotFound(f"checkpoint directory not found: {source}")
        copytree(source, os.fspath(dst), selector=selector)
