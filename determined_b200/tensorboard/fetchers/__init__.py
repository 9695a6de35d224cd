"""``build(checkpoint_storage, paths, local_dir)`` -> the fetcher for that storage type."""
from typing import Any, Dict, List, Type

from determined_b200.tensorboard.fetchers import base, cloud, shared
from determined_b200.tensorboard.fetchers import cloud as s3, cloud as gcs, cloud as azure, shared as directory  # noqa: F401  module aliases of the reference layout
from determined_b200.tensorboard.fetchers.base import Fetcher
from determined_b200.tensorboard.fetchers.cloud import AzureFetcher, GCSFetcher, S3Fetcher
from determined_b200.tensorboard.fetchers.shared import DirectoryFetcher, SharedFSFetcher

__all__ = ["Fetcher", "S3Fetcher", "GCSFetcher", "AzureFetcher", "SharedFSFetcher", "DirectoryFetcher", "build"]

_FETCHERS: Dict[str, Type[Fetcher]] = {"s3": S3Fetcher, "gcs": GCSFetcher, "azure": AzureFetcher, "shared_fs": SharedFSFetcher, "directory": DirectoryFetcher}


def build(storage_config: Dict[str, Any], paths: List[str], local_dir: str) -> Fetcher:
    kind = storage_config.get("type")
    if kind not in _FETCHERS:
        raise ValueError(f"checkpoint_storage type '{kind}' is not supported")
    return _FETCHERS[kind](storage_config, paths, local_dir)
