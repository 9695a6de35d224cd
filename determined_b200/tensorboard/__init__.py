"""TensorBoard event-file syncing (reference ``harness/determined/tensorboard/*``).

``TensorboardManager.sync()`` ships new/modified files under the local ``base_path`` to
``<storage>/<cluster_id>/tensorboard/experiment/<e>/trial/<t>/`` (path contract:
reference ``tensorboard/build.py:9-24``).  Uploads happen on a background thread
(reference ``tensorboard/base.py:22-187``); ``close()`` drains it.
"""
import logging
import os
import pathlib
import queue
import shutil
import tempfile
import threading
from typing import Any, Callable, Dict, List, Optional, Tuple

from determined_b200.common import storage
from determined_b200.tensorboard import util  # noqa: E402

logger = logging.getLogger("determined.tensorboard")


def get_sync_path(cluster_id: str, experiment_id: Any, trial_id: Any) -> pathlib.Path:
    return pathlib.Path(cluster_id, "tensorboard", "experiment", str(experiment_id), "trial", str(trial_id))


def get_experiment_sync_path(cluster_id: str, experiment_id: Any) -> pathlib.Path:
    return pathlib.Path(cluster_id, "tensorboard", "experiment", str(experiment_id))


def get_base_path(checkpoint_config: Dict[str, Any]) -> pathlib.Path:
    allocation_id = os.environ.get("DET_ALLOCATION_ID", "")
    rank = os.environ.get("RANK", "0")
    if checkpoint_config.get("base_path"):
        base = pathlib.Path(checkpoint_config["base_path"])
    else:
        base = pathlib.Path(tempfile.gettempdir())
    return base / f"tensorboard-{allocation_id}-{rank}"


def get_rank_aware_path(path: pathlib.Path, rank: int) -> pathlib.Path:
    """Name under which a non-chief rank's file is uploaded (``tensorboard/util.py`` has the rules); rank 0 and event files keep theirs."""
    if rank == 0 or "tfevents" in pathlib.Path(path).name:
        return pathlib.Path(path)
    return util.get_rank_aware_path(pathlib.Path(path), rank)


class TensorboardManager:
    def __init__(self, base_path: pathlib.Path, sync_path: pathlib.Path, async_upload: bool = True) -> None:
        self.base_path = pathlib.Path(base_path)
        self.sync_path = pathlib.Path(sync_path)        # (the directory is created by build(), not here: a manager may point anywhere)
        self._seen: Dict[pathlib.Path, float] = {}
        self._q: "queue.Queue[Any]" = queue.Queue()
        self._thread: Optional[threading.Thread] = None
        self._async = async_upload

    def _changed(self, selector: Callable[[pathlib.Path], bool]) -> List[pathlib.Path]:
        out = []
        for root, _, files in os.walk(self.base_path):
            for f in files:
                p = pathlib.Path(root, f)
                try:
                    mt = p.stat().st_mtime
                except FileNotFoundError:
                    continue
                if self._seen.get(p) == mt or not selector(p):
                    continue
                self._seen[p] = mt
                out.append(p)
        return out

    def list_tb_files(self, rank: int, selector: Callable[[pathlib.Path], bool] = lambda _: True) -> List[pathlib.Path]:
        """Files under ``base_path`` a sync would consider: every rank's event files are theirs alone, so the chief takes all
        files and the other ranks everything except event files (those stay per-rank on the chief's side)."""
        if not self.base_path.exists():
            logger.warning("%s directory does not exist.", self.base_path)
            return []
        out = []
        for root, _, files in os.walk(self.base_path):
            for f in files:
                p = pathlib.Path(root, f)
                if rank != 0 and "tfevents" in f:
                    continue
                if selector(p):
                    out.append(p)
        return out

    def _upload(self, pairs: List[Tuple[pathlib.Path, pathlib.Path]]) -> None:
        raise NotImplementedError

    def sync(self, selector: Callable[[pathlib.Path], bool] = lambda _: True,
             mangler: Callable[[pathlib.Path, int], pathlib.Path] = get_rank_aware_path, rank: int = 0) -> None:
        pairs = []
        for p in self._changed(selector):
            rel = p.relative_to(self.base_path)
            pairs.append((p, self.sync_path / mangler(rel, rank)))
        if not pairs:
            return
        if not self._async:
            self._upload(pairs)
            return
        if self._thread is None:
            self._thread = threading.Thread(target=self._drain, daemon=True, name="tensorboard-sync")
            self._thread.start()
        self._q.put(pairs)

    def _drain(self) -> None:
        while True:
            pairs = self._q.get()
            if pairs is None:
                return
            try:
                self._upload(pairs)
            except Exception as e:
                logger.warning("tensorboard sync failed: %s", e)

    def close(self) -> None:
        if self._thread is not None:
            self._q.put(None)
            self._thread.join(timeout=60)
            self._thread = None

    def delete(self) -> None:
        raise NotImplementedError

    def __enter__(self) -> "TensorboardManager":
        return self

    def __exit__(self, *_: Any) -> None:
        self.close()


class SharedFSTensorboardManager(TensorboardManager):
    def __init__(self, storage_path: str, base_path: pathlib.Path, sync_path: pathlib.Path, **kw: Any) -> None:
        super().__init__(base_path, sync_path, **kw)
        self.storage_path = pathlib.Path(storage_path)

    def _upload(self, pairs: List[Tuple[pathlib.Path, pathlib.Path]]) -> None:
        for src, dst in pairs:
            target = self.storage_path / dst
            target.parent.mkdir(parents=True, exist_ok=True)
            shutil.copy2(src, target)

    def delete(self) -> None:
        shutil.rmtree(self.storage_path / self.sync_path, ignore_errors=True)


class StorageTensorboardManager(TensorboardManager):
    """Cloud backends: reuse the checkpoint StorageManager's ``upload`` with a staged tree."""

    def __init__(self, manager: storage.StorageManager, base_path: pathlib.Path, sync_path: pathlib.Path, **kw: Any) -> None:
        super().__init__(base_path, sync_path, **kw)
        self._manager = manager

    def _upload(self, pairs: List[Tuple[pathlib.Path, pathlib.Path]]) -> None:
        with tempfile.TemporaryDirectory() as td:
            for src, dst in pairs:
                t = pathlib.Path(td) / dst.relative_to(self.sync_path)
                t.parent.mkdir(parents=True, exist_ok=True)
                shutil.copy2(src, t)
            self._manager.upload(td, str(self.sync_path))

    def delete(self) -> None:
        self._manager.delete(str(self.sync_path), ["**/*"])


class S3TensorboardManager(StorageTensorboardManager):
    """TensorBoard files to an S3 bucket (the checkpoint storage's S3 manager does the uploads)."""


class GCSTensorboardManager(StorageTensorboardManager):
    """TensorBoard files to a GCS bucket."""


class AzureTensorboardManager(StorageTensorboardManager):
    """TensorBoard files to an Azure blob container."""


class DirectoryTensorboardManager(SharedFSTensorboardManager):
    """TensorBoard files to a directory mounted into the task."""


def build(cluster_id: str, experiment_id: Any, trial_id: Any, checkpoint_config: Dict[str, Any],
          container_path: Optional[str] = None, async_upload: bool = True) -> TensorboardManager:
    if "type" not in checkpoint_config:
        raise TypeError("Missing 'type' parameter of storage configuration")
    kind = checkpoint_config["type"]
    if not isinstance(kind, str):
        raise TypeError("`type` parameter of storage configuration must be a string")
    if kind not in ("shared_fs", "directory", "s3", "gcs", "azure"):
        raise TypeError(f"Unknown storage type: {kind}")
    base = get_base_path(checkpoint_config)
    base.mkdir(parents=True, exist_ok=True)
    sync = get_sync_path(cluster_id, experiment_id, trial_id)
    storage_config = {k: v for k, v in checkpoint_config.items() if k != "base_path"}      # base_path is ours, not the storage manager's
    mgr = storage.build(storage_config, container_path)
    if kind in ("shared_fs", "directory"):
        cls = SharedFSTensorboardManager if kind == "shared_fs" else DirectoryTensorboardManager
        return cls(mgr._base_path, base, sync, async_upload=async_upload)
    cloud = {"s3": S3TensorboardManager, "gcs": GCSTensorboardManager, "azure": AzureTensorboardManager}[kind]
    return cloud(mgr, base, sync, async_upload=async_upload)


def get_metric_writer() -> Any:
    """A ``torch.utils.tensorboard.SummaryWriter``-backed batch metric writer, if available."""
    from determined_b200.tensorboard.metric_writers import BatchMetricWriter

    # the directory the TensorboardManager built with the default configuration watches and syncs (reference
    # ``tensorboard/__init__.py:get_metric_writer`` -> ``get_base_path({})``)
    return BatchMetricWriter.try_build(get_base_path({}))


from determined_b200.tensorboard.metric_writers import BatchMetricWriter, MetricWriter  # noqa: E402,F401
