"""The reference's storage conformance suite (``harness/tests/storage/util.py``: checkpoint lifecycle with selectors and
path subsets, sharded upload / download from four ranks) run against this package's S3 / GCS / Azure managers.  The cloud
SDKs are not in the image, so each manager is handed an in-memory stand-in for the one SDK object it talks to (an S3
``Bucket`` resource, a GCS ``Bucket``, an Azure ``ContainerClient``); everything above that line — key layout, directory
markers, selectors, globs, CheckpointNotFound — is the manager's own code.

Run through ``tests/fixtures/run_reference_tests.py`` (which aliases ``determined`` to this package and puts the reference's
``tests`` package on the path)."""
import fnmatch  # noqa: F401
import pathlib
import shutil
import threading
from typing import Any, Dict, Iterator, List

import pytest

from determined.common import storage
from determined.common.storage import cloud
from tests import parallel
from tests.storage import util


class _Store:
    def __init__(self) -> None:
        self.objs: Dict[str, bytes] = {}
        self.lock = threading.Lock()


# ---- boto3.resource("s3").Bucket ------------------------------------------------------------------------------------
class _S3Obj:
    def __init__(self, key: str, size: int) -> None:
        self.key, self.size = key, size


class _S3Objects:
    def __init__(self, st: _Store) -> None:
        self.st = st

    def filter(self, Prefix: str = "") -> Iterator[_S3Obj]:  # noqa: N803
        with self.st.lock:
            items = sorted((k, len(v)) for k, v in self.st.objs.items() if k.startswith(Prefix))
        return iter([_S3Obj(k, n) for k, n in items])


class FakeS3Bucket:
    def __init__(self) -> None:
        self.st = _Store()
        self.objects = _S3Objects(self.st)

    def put_object(self, Key: str, Body: bytes = b"") -> None:  # noqa: N803
        with self.st.lock:
            self.st.objs[Key] = bytes(Body)

    def upload_file(self, filename: str, key: str) -> None:
        data = pathlib.Path(filename).read_bytes()
        with self.st.lock:
            self.st.objs[key] = data

    def download_file(self, key: str, filename: str) -> None:
        with self.st.lock:
            data = self.st.objs[key]
        pathlib.Path(filename).write_bytes(data)

    def delete_objects(self, Delete: Dict[str, Any]) -> None:  # noqa: N803
        with self.st.lock:
            for o in Delete["Objects"]:
                self.st.objs.pop(o["Key"], None)


# ---- google.cloud.storage Bucket ------------------------------------------------------------------------------------
class _GcsBlob:
    def __init__(self, st: _Store, name: str) -> None:
        self.st, self.name = st, name

    @property
    def size(self) -> int:
        return len(self.st.objs.get(self.name, b""))

    def upload_from_string(self, data: Any) -> None:
        with self.st.lock:
            self.st.objs[self.name] = data if isinstance(data, bytes) else str(data).encode()

    def upload_from_filename(self, filename: str) -> None:
        data = pathlib.Path(filename).read_bytes()
        with self.st.lock:
            self.st.objs[self.name] = data

    def download_to_filename(self, filename: str) -> None:
        pathlib.Path(filename).write_bytes(self.st.objs[self.name])

    def delete(self) -> None:
        with self.st.lock:
            self.st.objs.pop(self.name, None)


class FakeGcsBucket:
    def __init__(self) -> None:
        self.st = _Store()

    def blob(self, name: str) -> _GcsBlob:
        return _GcsBlob(self.st, name)

    def list_blobs(self, prefix: str = "") -> List[_GcsBlob]:
        with self.st.lock:
            names = sorted(k for k in self.st.objs if k.startswith(prefix))
        return [_GcsBlob(self.st, n) for n in names]


# ---- azure.storage.blob ContainerClient -----------------------------------------------------------------------------
class _AzDownload:
    def __init__(self, data: bytes) -> None:
        self.data = data

    def readinto(self, f: Any) -> int:
        f.write(self.data)
        return len(self.data)


class FakeAzureContainer:
    def __init__(self) -> None:
        self.st = _Store()

    def upload_blob(self, name: str, data: Any, overwrite: bool = False) -> None:
        body = data.read() if hasattr(data, "read") else bytes(data)
        with self.st.lock:
            if name in self.st.objs and not overwrite:
                raise RuntimeError("blob exists")
            self.st.objs[name] = body

    def list_blobs(self, name_starts_with: str = "") -> List[Dict[str, Any]]:
        with self.st.lock:
            return [{"name": k, "size": len(v)} for k, v in sorted(self.st.objs.items()) if k.startswith(name_starts_with)]

    def download_blob(self, name: str) -> _AzDownload:
        return _AzDownload(self.st.objs[name])

    def delete_blob(self, name: str) -> None:
        with self.st.lock:
            self.st.objs.pop(name, None)


def _remotes() -> Dict[str, Any]:
    return {"s3": FakeS3Bucket(), "gcs": FakeGcsBucket(), "azure": FakeAzureContainer()}


def _managers(tmp: pathlib.Path, remotes: Any = None):
    """One manager per backend with its local staging directory under ``tmp``; ``remotes`` lets several managers (one per
    rank, as in the reference's live tests) share one remote store."""
    remotes = remotes or _remotes()
    s3 = cloud.S3StorageManager(bucket="b", prefix="some/prefix", temp_dir=str(tmp / "s3"))
    s3._bucket = remotes["s3"]
    gcs = cloud.GCSStorageManager(bucket="b", prefix="p", temp_dir=str(tmp / "gcs"))
    gcs._bucket = remotes["gcs"]
    az = cloud.AzureStorageManager(container="c", connection_string="unused", temp_dir=str(tmp / "az"))
    az._client = remotes["azure"]
    return {"s3": s3, "gcs": gcs, "azure": az}


def _store_of(m: Any) -> _Store:
    return (m._bucket if hasattr(m, "_bucket") and m._bucket is not None else m._client).st


@pytest.mark.parametrize("kind", ["s3", "gcs", "azure"])
def test_lifecycle(kind: str, tmp_path: pathlib.Path, caplog: Any) -> None:
    m = _managers(tmp_path)[kind]

    def post_delete_cb(storage_id: str) -> None:
        assert not [k for k in _store_of(m).objs if storage_id in k], "objects left behind after delete"

    util.run_storage_lifecycle_test(m, post_delete_cb, caplog)


@pytest.mark.parametrize("kind", ["s3", "gcs", "azure"])
def test_sharded_upload_download(kind: str, tmp_path: pathlib.Path) -> None:
    m = _managers(tmp_path)[kind]

    def clean_up(storage_id: str, manager: Any) -> None:
        manager.delete(storage_id, ["**/*"])

    with parallel.Execution(4, local_size=2) as pex:

        @pex.run
        def do_test() -> None:
            util.run_storage_upload_download_sharded_test(pex, m, tmp_path, clean_up)


@pytest.mark.parametrize("kind", ["s3", "gcs", "azure"])
def test_sharded_store_restore(kind: str, tmp_path: pathlib.Path) -> None:
    remotes = _remotes()

    def clean_up(storage_id: str, manager: Any) -> None:
        manager.delete(storage_id, ["**/*"])

    with parallel.Execution(4, local_size=2) as pex:

        @pex.run
        def do_test() -> None:
            # as in the reference's live tests: one manager (own staging directory) per rank, one remote store
            m = _managers(tmp_path / f"storage_{pex.distributed.rank}", remotes)[kind]
            util.run_storage_store_restore_sharded_test(pex, m, clean_up)


def test_validate_manager_round_trip(tmp_path: pathlib.Path) -> None:
    for m in _managers(tmp_path).values():
        storage.validate_manager(m)
        assert not _store_of(m).objs
    shutil.rmtree(tmp_path, ignore_errors=True)
