"""Cloud backends share one shape: stage in a local temp dir, then upload / download.

Parity: reference ``common/storage/cloud.py:10-35``, ``s3.py``, ``gcs.py``, ``azure.py``.  The SDKs
(boto3 / google-cloud-storage / azure-storage-blob) are not present in the B200 image; each
manager imports its SDK lazily and raises a clear ImportError if used without it.  The object
layout (``<prefix>/<uuid>/<relative path>``, directories as zero-byte ``name/`` keys) matches.
"""
import contextlib
import os
import pathlib
import shutil
import tempfile
from typing import Any, Dict, Iterator, List, Optional, Union

from determined_b200 import errors
from determined_b200.common.storage import base


class CloudStorageManager(base.StorageManager):
    def __init__(self, temp_dir: Optional[str] = None) -> None:
        super().__init__(temp_dir or tempfile.gettempdir())

    @contextlib.contextmanager
    def restore_path(self, src: str, selector: Optional[base.Selector] = None) -> Iterator[pathlib.Path]:
        dst = os.path.join(self._base_path, src)
        os.makedirs(dst, exist_ok=True)
        self.download(src, dst, selector)
        try:
            yield pathlib.Path(dst)
        finally:
            shutil.rmtree(dst, ignore_errors=True)

    def post_store_path(self, src: Union[str, os.PathLike], dst: str) -> None:
        try:
            self.upload(src, dst)
        finally:
            shutil.rmtree(os.fspath(src), ignore_errors=True)

    def store_path_is_direct_access(self) -> bool:
        return False

    # -- helpers for subclasses --------------------------------------------------------------
    @staticmethod
    def _iter_upload(src: Union[str, os.PathLike], paths: Optional[base.Paths]) -> List[str]:
        rels = sorted(base.list_directory(src))
        if paths is not None:
            rels = [r for r in rels if r in paths]
        return rels


class S3StorageManager(CloudStorageManager):
    def __init__(self, bucket: str, access_key: Optional[str] = None, secret_key: Optional[str] = None,
                 endpoint_url: Optional[str] = None, prefix: Optional[str] = None,
                 temp_dir: Optional[str] = None) -> None:
        super().__init__(temp_dir)
        self.bucket_name = bucket
        self.prefix = (prefix or "").strip("/")
        self._kw = dict(endpoint_url=endpoint_url, aws_access_key_id=access_key, aws_secret_access_key=secret_key)
        self._bucket = None

    @classmethod
    def from_config(cls, config: Dict[str, Any], container_path: Optional[str]) -> "S3StorageManager":
        return cls(**config)

    def _b(self) -> Any:
        if self._bucket is None:
            try:
                import boto3
            except ImportError as e:  # pragma: no cover
                raise ImportError("s3 checkpoint storage requires boto3, which is not installed") from e
            self._bucket = boto3.resource("s3", **self._kw).Bucket(self.bucket_name)
        return self._bucket

    def _key(self, *parts: str) -> str:
        return "/".join([p for p in (self.prefix, *parts) if p])

    def upload(self, src: Union[str, os.PathLike], dst: str, paths: Optional[base.Paths] = None) -> None:
        for rel in self._iter_upload(src, paths):
            key = self._key(dst, rel)
            if rel.endswith("/"):
                self._b().put_object(Key=key, Body=b"")
            else:
                self._b().upload_file(os.path.join(os.fspath(src), rel), key)

    def download(self, src: str, dst: Union[str, os.PathLike], selector: Optional[base.Selector] = None) -> None:
        pfx = self._key(src) + "/"
        found = False
        for obj in self._b().objects.filter(Prefix=pfx):
            found = True
            rel = obj.key[len(pfx):]
            if not rel or (selector is not None and not selector(rel)):
                continue
            target = os.path.join(os.fspath(dst), rel)
            if rel.endswith("/"):
                os.makedirs(target, exist_ok=True)
                continue
            os.makedirs(os.path.dirname(target), exist_ok=True)
            self._b().download_file(obj.key, target)
        if not found:
            raise errors.CheckpointNotFound(f"no objects under s3://{self.bucket_name}/{pfx}")

    def delete(self, tgt: str, globs: List[str]) -> Dict[str, int]:
        pfx = self._key(tgt) + "/"
        objs = {o.key[len(pfx):]: o.size for o in self._b().objects.filter(Prefix=pfx) if o.key != pfx}
        keep = {} if "**/*" in globs else base.apply_globs_to_resources(objs, "", globs)
        doomed = [{"Key": pfx + k} for k in objs if k not in keep]
        for i in range(0, len(doomed), 1000):
            self._b().delete_objects(Delete={"Objects": doomed[i:i + 1000]})
        return keep


class GCSStorageManager(CloudStorageManager):
    def __init__(self, bucket: str, prefix: Optional[str] = None, temp_dir: Optional[str] = None) -> None:
        super().__init__(temp_dir)
        self.bucket_name = bucket
        self.prefix = (prefix or "").strip("/")
        self._bucket = None

    def _b(self) -> Any:
        if self._bucket is None:
            try:
                from google.cloud import storage as gcs
            except ImportError as e:  # pragma: no cover
                raise ImportError("gcs checkpoint storage requires google-cloud-storage, which is not installed") from e
            self._bucket = gcs.Client().bucket(self.bucket_name)
        return self._bucket

    def _key(self, *parts: str) -> str:
        return "/".join([p for p in (self.prefix, *parts) if p])

    def upload(self, src: Union[str, os.PathLike], dst: str, paths: Optional[base.Paths] = None) -> None:
        for rel in self._iter_upload(src, paths):
            blob = self._b().blob(self._key(dst, rel))
            if rel.endswith("/"):
                blob.upload_from_string(b"")
            else:
                blob.upload_from_filename(os.path.join(os.fspath(src), rel))

    def download(self, src: str, dst: Union[str, os.PathLike], selector: Optional[base.Selector] = None) -> None:
        pfx = self._key(src) + "/"
        found = False
        for blob in self._b().list_blobs(prefix=pfx):
            found = True
            rel = blob.name[len(pfx):]
            if not rel or (selector is not None and not selector(rel)):
                continue
            target = os.path.join(os.fspath(dst), rel)
            if rel.endswith("/"):
                os.makedirs(target, exist_ok=True)
                continue
            os.makedirs(os.path.dirname(target), exist_ok=True)
            blob.download_to_filename(target)
        if not found:
            raise errors.CheckpointNotFound(f"no objects under gs://{self.bucket_name}/{pfx}")

    def delete(self, tgt: str, globs: List[str]) -> Dict[str, int]:
        pfx = self._key(tgt) + "/"
        blobs = {b.name[len(pfx):]: b for b in self._b().list_blobs(prefix=pfx) if b.name != pfx}
        sizes = {k: (b.size or 0) for k, b in blobs.items()}
        keep = {} if "**/*" in globs else base.apply_globs_to_resources(sizes, "", globs)
        for k, b in blobs.items():
            if k not in keep:
                b.delete()
        return keep


class AzureStorageManager(CloudStorageManager):
    def __init__(self, container: str, connection_string: Optional[str] = None,
                 account_url: Optional[str] = None, credential: Optional[str] = None,
                 temp_dir: Optional[str] = None) -> None:
        super().__init__(temp_dir)
        self.container = container.rstrip("/")
        self._args = (connection_string, account_url, credential)
        self._client = None

    def _c(self) -> Any:
        if self._client is None:
            try:
                from azure.storage import blob
            except ImportError as e:  # pragma: no cover
                raise ImportError("azure checkpoint storage requires azure-storage-blob, which is not installed") from e
            cs, url, cred = self._args
            svc = (blob.BlobServiceClient.from_connection_string(cs) if cs
                   else blob.BlobServiceClient(url, cred))
            self._client = svc.get_container_client(self.container)
        return self._client

    def upload(self, src: Union[str, os.PathLike], dst: str, paths: Optional[base.Paths] = None) -> None:
        for rel in self._iter_upload(src, paths):
            if rel.endswith("/"):
                # blob stores have no directories: an empty blob whose name ends in "/" marks one (reference azure.py:50-54)
                self._c().upload_blob(f"{dst}/{rel}", b"", overwrite=True)
                continue
            with open(os.path.join(os.fspath(src), rel), "rb") as f:
                self._c().upload_blob(f"{dst}/{rel}", f, overwrite=True)

    def download(self, src: str, dst: Union[str, os.PathLike], selector: Optional[base.Selector] = None) -> None:
        pfx = f"{src}/"
        found = False
        for b in self._c().list_blobs(name_starts_with=pfx):
            found = True
            rel = b["name"][len(pfx):]
            if not rel or (selector is not None and not selector(rel)):
                continue
            target = os.path.join(os.fspath(dst), rel)
            if rel.endswith("/"):
                os.makedirs(target, exist_ok=True)
                continue
            os.makedirs(os.path.dirname(target), exist_ok=True)
            with open(target, "wb") as f:
                self._c().download_blob(b["name"]).readinto(f)
        if not found:
            raise errors.CheckpointNotFound(f"no blobs under {self.container}/{pfx}")

    def delete(self, tgt: str, globs: List[str]) -> Dict[str, int]:
        pfx = f"{tgt}/"
        sizes = {b["name"][len(pfx):]: b["size"] for b in self._c().list_blobs(name_starts_with=pfx)}
        keep = {} if "**/*" in globs else base.apply_globs_to_resources(sizes, "", globs)
        for k in sizes:
            if k not in keep:
                self._c().delete_blob(pfx + k)
        return keep
