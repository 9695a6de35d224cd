"""File backends for mmcv's ``FileClient``: read dataset images straight from cloud buckets, or serve one fake image.

Parity: reference ``model_hub/model_hub/mmdetection/_data_backends.py`` (``S3Backend``, ``GCSBackend``,
``FakeBackend``, ``sub_backend``).  The classes work stand-alone (``get`` / ``get_text`` / ``get_local_path``);
:func:`register` plugs them into mmcv when it is installed.
"""
import contextlib
import io
import os
import tempfile
from typing import Any, Dict, Iterator, Optional


class _Base:
    def get(self, filepath: str) -> bytes:
        raise NotImplementedError

    def get_text(self, filepath: str) -> str:
        return self.get(filepath).decode()

    @contextlib.contextmanager
    def get_local_path(self, filepath: str) -> Iterator[str]:
        """Materialise the object as a temporary local file for APIs that want a path."""
        fd, path = tempfile.mkstemp(suffix=os.path.splitext(filepath)[1])
        try:
            with os.fdopen(fd, "wb") as f:
                f.write(self.get(filepath))
            yield path
        finally:
            with contextlib.suppress(OSError):
                os.remove(path)


class S3Backend(_Base):
    """``FileClient(backend="s3", bucket_name=...)``; needs boto3."""

    def __init__(self, bucket_name: str) -> None:
        import boto3

        self._bucket = boto3.resource("s3").Bucket(bucket_name)

    def get(self, filepath: str) -> bytes:
        buf = io.BytesIO()
        self._bucket.download_fileobj(str(filepath), buf)
        return buf.getvalue()


class GCSBackend(_Base):
    """``FileClient(backend="gcs", bucket_name=...)``; needs google-cloud-storage."""

    def __init__(self, bucket_name: str) -> None:
        from google.cloud import storage

        self._bucket = storage.Client().bucket(bucket_name)

    def get(self, filepath: str) -> bytes:
        return self._bucket.blob(str(filepath)).download_as_bytes()


class FakeBackend(_Base):
    """Every path resolves to the same image: benchmarks the model without a dataset on disk."""

    def __init__(self, fake_img_path: Optional[str] = None) -> None:
        if fake_img_path is not None:
            with open(fake_img_path, "rb") as f:
                self._data = f.read()
        else:
            self._data = _tiny_png()

    def get(self, filepath: str) -> bytes:
        return self._data


def _tiny_png() -> bytes:
    """A valid 32x32 grey PNG, built by hand (no imaging library needed)."""
    import struct
    import zlib

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    raw = b"".join(b"\x00" + bytes([128] * 32 * 3) for _ in range(32))
    return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", 32, 32, 8, 2, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b"")


BACKENDS = {"s3": S3Backend, "gcs": GCSBackend, "fake": FakeBackend}


def register() -> bool:
    """Register the backends with mmcv's FileClient; False when mmcv is not installed."""
    try:
        import mmcv
    except ImportError:
        return False
    for name, cls in BACKENDS.items():
        mmcv.fileio.FileClient.register_backend(name, type(cls.__name__, (cls, mmcv.fileio.BaseStorageBackend), {}), force=True)
    return True


def sub_backend(file_client_args: Dict[str, Any], cfg: Any) -> None:
    """Replace the ``file_client_args`` of every ``LoadImageFromFile``-style pipeline step in an mmdet config (a nested
    dict / list structure), in place."""
    if isinstance(cfg, dict):
        if "type" in cfg and str(cfg["type"]).startswith("LoadImageFrom"):
            cfg["file_client_args"] = dict(file_client_args)
        for v in cfg.values():
            sub_backend(file_client_args, v)
    elif isinstance(cfg, (list, tuple)):
        for v in cfg:
            sub_backend(file_client_args, v)
