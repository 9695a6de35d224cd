"""Object-store fetchers; the cloud SDKs are imported when a fetcher is constructed (they are optional dependencies)."""
import urllib.parse
from typing import Any, Callable, Dict, Generator, List, Tuple

from determined_b200.tensorboard.fetchers import base


class S3Fetcher(base.Fetcher):
    def __init__(self, storage_config: Dict[str, Any], storage_paths: List[str], local_dir: str) -> None:
        super().__init__(storage_config, storage_paths, local_dir)
        import boto3

        self.bucket_name = str(storage_config["bucket"])
        self.client = boto3.client("s3", endpoint_url=storage_config.get("endpoint_url"), aws_access_key_id=storage_config.get("access_key"),
                                   aws_secret_access_key=storage_config.get("secret_key"))

    def _find(self, storage_path: str) -> Generator[Tuple[str, Any], None, None]:
        for page in self.client.get_paginator("list_objects_v2").paginate(Bucket=self.bucket_name, Prefix=storage_path.lstrip("/")):
            for obj in page.get("Contents", []):
                yield obj["Key"], base._aware(obj.get("LastModified"))

    def _fetch(self, filepath: str, new_file_callback: Callable[[], None]) -> None:
        self.client.download_file(self.bucket_name, filepath, self._local_path(filepath))
        new_file_callback()


class GCSFetcher(base.Fetcher):
    def __init__(self, storage_config: Dict[str, Any], storage_paths: List[str], local_dir: str) -> None:
        super().__init__(storage_config, storage_paths, local_dir)
        from google.cloud import storage

        self.client = storage.Client()
        self.bucket_name = str(storage_config["bucket"])

    def _find(self, storage_path: str) -> Generator[Tuple[str, Any], None, None]:
        for blob in self.client.list_blobs(self.bucket_name, prefix=storage_path.lstrip("/")):
            yield blob.name, base._aware(blob.updated)

    def _fetch(self, filepath: str, new_file_callback: Callable[[], None]) -> None:
        self.client.bucket(self.bucket_name).blob(filepath).download_to_filename(self._local_path(filepath))
        new_file_callback()


class AzureFetcher(base.Fetcher):
    def __init__(self, storage_config: Dict[str, Any], storage_paths: List[str], local_dir: str) -> None:
        super().__init__(storage_config, storage_paths, local_dir)
        from azure.storage import blob

        if storage_config.get("connection_string"):
            svc = blob.BlobServiceClient.from_connection_string(storage_config["connection_string"])
        else:
            url = storage_config["account_url"]
            if not urllib.parse.urlparse(url).scheme:
                url = "https://" + url
            svc = blob.BlobServiceClient(url, storage_config.get("credential"))
        self.container = svc.get_container_client(str(storage_config["container"]))

    def _find(self, storage_path: str) -> Generator[Tuple[str, Any], None, None]:
        for b in self.container.list_blobs(name_starts_with=storage_path.lstrip("/")):
            yield b["name"], base._aware(b.get("last_modified"))

    def _fetch(self, filepath: str, new_file_callback: Callable[[], None]) -> None:
        with open(self._local_path(filepath), "wb") as f:
            self.container.download_blob(filepath).readinto(f)
        new_file_callback()


