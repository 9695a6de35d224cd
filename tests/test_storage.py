"""Storage manager conformance (common conformance suite, reference ``tests/storage/util.py``)."""
import os
import pathlib

import pytest

from determined_b200.common import storage


def make_tree(root: pathlib.Path) -> None:
    (root / "sub" / "deep").mkdir(parents=True)
    (root / "empty").mkdir()
    (root / "root.txt").write_text("r")
    (root / "sub" / "a.bin").write_text("aa")
    (root / "sub" / "deep" / "b.bin").write_text("bbb")


@pytest.mark.parametrize("kind", ["shared_fs", "directory"])
def test_conformance(tmp_path, kind):
    base = tmp_path / "store"
    base.mkdir()
    cfg = {"type": "shared_fs", "host_path": str(base)} if kind == "shared_fs" else {"type": "directory", "container_path": str(base)}
    mgr = storage.build(cfg, None)
    storage.validate_manager(mgr)
    assert mgr.store_path_is_direct_access()
    with mgr.store_path("u1") as p:
        make_tree(p)
    assert storage.list_directory(base / "u1") == {"sub/": 0, "sub/deep/": 0, "empty/": 0, "root.txt": 1, "sub/a.bin": 2, "sub/deep/b.bin": 3}
    with mgr.restore_path("u1") as p:
        assert (p / "sub" / "deep" / "b.bin").read_text() == "bbb"
    dst = tmp_path / "dl"
    mgr.download("u1", dst, selector=lambda n: n.startswith("sub/") and not n.endswith("a.bin"))
    assert sorted(str(x.relative_to(dst)) for x in dst.rglob("*") if x.is_file()) == ["sub/deep/b.bin"]
    src = tmp_path / "up"
    src.mkdir()
    make_tree(src)
    mgr.upload(src, "u2", paths={"root.txt", "sub/", "sub/a.bin"})
    assert sorted(str(x.relative_to(base / "u2")) for x in (base / "u2").rglob("*") if x.is_file()) == ["root.txt", "sub/a.bin"]
    left = mgr.delete("u1", ["sub/deep", "*.txt"])
    assert left == {"sub/": 0, "empty/": 0, "sub/a.bin": 2}
    assert mgr.delete("u1", ["**/*"]) == {} and not (base / "u1").exists()
    with pytest.raises(FileNotFoundError):
        with mgr.restore_path("nope"):
            pass


def test_glob_application_in_memory():
    res = {"a.txt": 1, "d/": 0, "d/x.bin": 2, "d/e/": 0, "d/e/y.bin": 3, "z.bin": 4}
    assert storage.apply_globs_to_resources(res, "", ["d"]) == {"a.txt": 1, "z.bin": 4}
    assert storage.apply_globs_to_resources(res, "", ["**/*.bin"]) == {"a.txt": 1, "d/": 0, "d/e/": 0}
    assert storage.apply_globs_to_resources(res, "", ["*.bin"]) == {"a.txt": 1, "d/": 0, "d/x.bin": 2, "d/e/": 0, "d/e/y.bin": 3}


def test_build_and_shortcuts(tmp_path):
    with pytest.raises(TypeError):
        storage.build({"type": "nope"}, None)
    with pytest.raises(ValueError):
        storage.build({}, None)
    m = storage.build({"type": "shared_fs", "host_path": str(tmp_path), "storage_path": "ck", "save_trial_best": 3}, None)
    assert m._base_path == str(tmp_path / "ck")
    with pytest.raises(ValueError):
        storage.build({"type": "shared_fs", "host_path": str(tmp_path), "storage_path": "../escape"}, None)
    assert isinstance(storage.from_string(str(tmp_path)), storage.SharedFSStorageManager)
    assert isinstance(storage.from_string("s3://bucket/prefix"), storage.S3StorageManager)
    assert isinstance(storage.from_string("gs://bucket/p"), storage.GCSStorageManager)
    with pytest.raises(ValueError):
        storage.from_string("ftp://x/y")
    with pytest.raises(ValueError):
        storage.build({"type": "azure", "container": "c"}, None)


def test_s3_manager_against_fake_bucket(tmp_path):
    """Cloud managers are exercised against an in-memory fake of the SDK surface they use."""
    objs = {}

    class Obj:
        def __init__(self, key):
            self.key, self.size = key, len(objs[key])

    class Bucket:
        class objects:  # noqa
            @staticmethod
            def filter(Prefix):
                return [Obj(k) for k in sorted(objs) if k.startswith(Prefix)]

        def put_object(self, Key, Body):
            objs[Key] = Body

        def upload_file(self, path, key):
            objs[key] = open(path, "rb").read()

        def download_file(self, key, path):
            open(path, "wb").write(objs[key])

        def delete_objects(self, Delete):
            for o in Delete["Objects"]:
                objs.pop(o["Key"], None)

    m = storage.S3StorageManager(bucket="b", prefix="pre", temp_dir=str(tmp_path / "tmp"))
    m._bucket = Bucket()
    src = tmp_path / "src"
    src.mkdir()
    make_tree(src)
    m.upload(src, "u1")
    assert "pre/u1/sub/deep/b.bin" in objs and not m.store_path_is_direct_access()
    with m.restore_path("u1") as p:
        assert (p / "sub" / "a.bin").read_text() == "aa"
    assert not (tmp_path / "tmp" / "u1").exists(), "temp copy must be cleaned up"
    left = m.delete("u1", ["sub/deep"])
    assert "sub/deep/b.bin" not in left and "sub/a.bin" in left
    assert m.delete("u1", ["**/*"]) == {} and not [k for k in objs if k.startswith("pre/u1/")]
