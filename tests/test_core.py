"""Core API tests: distributed object collectives, sharded checkpoints, searcher/preempt/train
contexts against fake sessions (SURVEY §4 "fake master": mock sessions + simulated long-poll)."""
import json
import os
import pathlib
import threading
import time
from typing import Any, Dict, List
from unittest import mock

import pytest

from determined_b200 import core
from determined_b200.common import storage
from tests import parallel_util


@pytest.mark.parametrize("size,local_size", [(1, 1), (3, 3), (4, 2), (6, 3)])
def test_collectives(size, local_size):
    ex = parallel_util.Execution(size, local_size)

    def fn():
        d = ex.distributed
        r = d.rank
        out = {}
        out["gather"] = d.gather(f"g{r}")
        out["allgather"] = d.allgather({"r": r})
        out["bcast"] = d.broadcast("chief-says" if r == 0 else None)
        out["gather_local"] = d.gather_local(r)
        out["allgather_local"] = d.allgather_local(r)
        out["bcast_local"] = d.broadcast_local(f"node{d.cross_rank}" if d.local_rank == 0 else None)
        return out

    res = ex.run(fn)
    for r, o in enumerate(res):
        node = r // local_size
        members = list(range(node * local_size, (node + 1) * local_size))
        assert o["allgather"] == [{"r": i} for i in range(size)]
        assert o["bcast"] == "chief-says"
        assert o["gather"] == ([f"g{i}" for i in range(size)] if r == 0 else None)
        assert o["allgather_local"] == members
        assert o["gather_local"] == (members if r % local_size == 0 else None)
        assert o["bcast_local"] == f"node{node}"


def test_collective_large_payload_and_many_rounds():
    ex = parallel_util.Execution(3)

    def fn():
        d = ex.distributed
        blob = bytes([d.rank]) * (3 << 20)
        got = d.allgather(blob)
        assert [len(g) for g in got] == [3 << 20] * 3 and got[2][:1] == b"\x02"
        for i in range(200):
            assert d.broadcast(i if d.rank == 0 else None) == i
        return True

    assert ex.run(fn) == [True] * 3


def test_merge_metadata_and_resources():
    m, c = core.merge_metadata([{"a": 1, "d": {"x": 1}}, {"b": 2, "d": {"y": 2}}, {"a": 1}])
    assert m == {"a": 1, "b": 2, "d": {"x": 1, "y": 2}} and c == {}
    m, c = core.merge_metadata([{"a": [1]}, {"a": [2]}, {"a": [3]}])
    assert c == {"/a": [0, 1, 2]}
    m, c = core.merge_metadata([{"a": {"c": 1}}, {"a": 1}])
    assert "/a" in c
    r, c = core.merge_resources([{"d/": 0, "d/f0": 3}, {"d/": 0, "d/f1": 4}])
    assert r == {"d/": 0, "d/f0": 3, "d/f1": 4} and c == {}
    r, c = core.merge_resources([{"f": 1}, {"f": 1}, {"g/": 0}, {"g": 2}])
    assert set(c) == {"f", "g"} and c["f"] == [0, 1]


class FakeSession:
    def __init__(self) -> None:
        self.posts: List[Any] = []
        self.lock = threading.Lock()

    def post(self, path: str, json: Any = None, data: Any = None, **kw: Any) -> Any:
        with self.lock:
            self.posts.append((path, json if json is not None else data))
        return mock.MagicMock()

    def get(self, path: str, **kw: Any) -> Any:
        raise AssertionError(f"unexpected GET {path}")


def make_ckpt_ctx(dist, base, sess):
    mgr = storage.SharedFSStorageManager(str(base))
    return core.CheckpointContext(dist, mgr, sess, "task-1", "alloc-1", core.TensorboardMode.MANUAL)


class UploadOnlyStorage(storage.SharedFSStorageManager):
    """Behaves like a cloud backend: store_path stages in a private dir and must upload."""

    def __init__(self, base: str, stage: str) -> None:
        super().__init__(base)
        self._stage = stage

    def store_path_is_direct_access(self) -> bool:
        return False

    def pre_store_path(self, dst: str) -> pathlib.Path:
        p = pathlib.Path(self._stage) / f"{dst}-{threading.get_ident()}"
        p.mkdir(parents=True, exist_ok=True)
        return p

    def post_store_path(self, src, dst) -> None:
        self.upload(src, dst)


def test_sharded_upload_union_and_report(tmp_path):
    sess = FakeSession()
    ex = parallel_util.Execution(3)

    def fn():
        d = ex.distributed
        ctx = make_ckpt_ctx(d, tmp_path / "store", sess)
        src = tmp_path / f"src{d.rank}"
        (src / "sub").mkdir(parents=True)
        (src / f"shard{d.rank}.bin").write_bytes(b"x" * (d.rank + 1))
        (src / "sub" / f"s{d.rank}").write_text("hi")
        (src / "common.txt").write_text("same on all ranks")
        return ctx.upload(str(src), {"steps_completed": 7, f"rank{d.rank}": True, "shared": {"k" + str(d.rank): d.rank}}, shard=True)

    ids = ex.run(fn)
    assert len(set(ids)) == 1
    out = tmp_path / "store" / ids[0]
    names = {str(p.relative_to(out)) for p in out.rglob("*") if p.is_file()}
    assert names == {"shard0.bin", "shard1.bin", "shard2.bin", "sub/s0", "sub/s1", "sub/s2", "common.txt", "metadata.json"}
    md = json.loads((out / "metadata.json").read_text())
    assert md["steps_completed"] == 7 and md["rank1"] is True and md["shared"] == {"k0": 0, "k1": 1, "k2": 2}
    reports = [p for p in sess.posts if p[0] == "/api/v1/checkpoints"]
    assert len(reports) == 1, "only the chief reports"
    body = reports[0][1]
    assert body["uuid"] == ids[0] and body["state"] == "STATE_COMPLETED" and body["resources"]["shard2.bin"] == "3"
    assert body["metadata"]["steps_completed"] == 7 and "metadata.json" in body["resources"]


def test_sharded_upload_conflicts_are_refused(tmp_path):
    sess = FakeSession()
    ex = parallel_util.Execution(2)
    ex._expect_errors = True

    def fn():
        d = ex.distributed
        ctx = make_ckpt_ctx(d, tmp_path / "store", sess)
        src = tmp_path / f"c{d.rank}"
        src.mkdir()
        (src / "same_name").write_text(f"different content {d.rank}")
        ctx.upload(str(src), {"steps_completed": 1}, shard=True)

    ex.run(fn)
    assert all(e is not None and "conflict" in str(e[0]) for e in ex.errors)
    assert not [p for p in sess.posts if p[0] == "/api/v1/checkpoints"]


def test_sharded_upload_shared_dir_uploaded_once_and_none_dirs(tmp_path):
    sess = FakeSession()
    ex = parallel_util.Execution(3)
    shared = tmp_path / "shared"
    shared.mkdir()
    (shared / "model.bin").write_bytes(b"m" * 10)

    def fn():
        d = ex.distributed
        ctx = make_ckpt_ctx(d, tmp_path / "store", sess)
        return ctx.upload(str(shared) if d.rank < 2 else None, {"steps_completed": 1}, shard=True)

    ids = ex.run(fn)
    assert (tmp_path / "store" / ids[0] / "model.bin").read_bytes() == b"m" * 10


def test_upload_requires_chief_and_steps_completed(tmp_path):
    sess = FakeSession()
    ctx = make_ckpt_ctx(core.DummyDistributedContext(), tmp_path / "store", sess)
    (tmp_path / "a").mkdir()
    with pytest.raises(ValueError, match="steps_completed"):
        ctx.upload(str(tmp_path / "a"), {"no": "steps"})
    sid = ctx.upload(str(tmp_path / "a"), {"steps_completed": 3}, selector=lambda p: p != "skipme")
    assert (tmp_path / "store" / sid / "metadata.json").exists()


def test_store_path_sharded_direct_and_uploading(tmp_path):
    for direct in (True, False):
        sess = FakeSession()
        ex = parallel_util.Execution(2)
        base = tmp_path / f"store{direct}"
        base.mkdir()

        def fn():
            d = ex.distributed
            mgr = storage.SharedFSStorageManager(str(base)) if direct else UploadOnlyStorage(str(base), str(tmp_path / "stage"))
            ctx = core.CheckpointContext(d, mgr, sess, "t", "a", core.TensorboardMode.MANUAL)
            with ctx.store_path({"steps_completed": 5, f"r{d.rank}": 1}, shard=True) as (path, sid):
                (path / f"rank{d.rank}.pt").write_text(str(d.rank))
            return sid

        ids = ex.run(fn)
        out = base / ids[0]
        assert {p.name for p in out.iterdir()} == {"rank0.pt", "rank1.pt", "metadata.json"}
        md = json.loads((out / "metadata.json").read_text())
        assert md == {"steps_completed": 5, "r0": 1, "r1": 1}
        assert len([p for p in sess.posts if p[0] == "/api/v1/checkpoints"]) == 1


def test_restore_path_shared_download_with_selectors(tmp_path):
    base = tmp_path / "store"
    (base / "ck" / "d").mkdir(parents=True)
    for n in ("a.bin", "b.bin", "d/c.bin"):
        (base / "ck" / n).write_text(n)
    ex = parallel_util.Execution(4, 2)
    seen = {}

    def fn():
        d = ex.distributed
        mgr = UploadOnlyStorage(str(base), str(tmp_path / "stage"))
        mgr._base_path = str(base)
        # force the cloud-style restore (temp copy) through download()
        ctx = core.CheckpointContext(d, storage.SharedFSStorageManager(str(base)), FakeSession(), "t", "a", core.TensorboardMode.MANUAL)
        dst = tmp_path / f"dl-node{d.cross_rank}"
        sel = (lambda p: p.startswith("a")) if d.local_rank == 0 else (lambda p: p.startswith("d"))
        ctx.download("ck", str(dst), selector=sel)
        with ctx.restore_path("ck") as p:
            assert (pathlib.Path(p) / "b.bin").exists()
        return sorted(str(x.relative_to(dst)) for x in dst.rglob("*") if x.is_file())

    res = ex.run(fn)
    # one download per node containing the UNION of what its workers selected
    assert res[0] == res[1] == ["a.bin", "d/c.bin"] and res[2] == res[3] == ["a.bin", "d/c.bin"]


def test_searcher_operations_chief_and_workers():
    class Sess(FakeSession):
        def __init__(self) -> None:
            super().__init__()
            self.ops = [5, 12]

        def get(self, path, **kw):
            assert path.endswith("/searcher/operation")
            r = mock.MagicMock()
            r.json.return_value = {"completed": False, "op": {"validateAfter": {"length": str(self.ops[0])}}} if self.ops else {"completed": True}
            return r

        def post(self, path, json=None, data=None, **kw):
            if path.endswith("completed_operation"):
                self.ops.pop(0)
            return super().post(path, json, data)

    sess = Sess()
    ex = parallel_util.Execution(3)

    def fn():
        d = ex.distributed
        sc = core.SearcherContext(sess, d, 7, 0, "alloc", core.Unit.BATCHES)
        lens = []
        for op in sc.operations():
            lens.append(op.length)
            if d.rank == 0:
                op.report_progress(op.length)
                op.report_completed(0.1)
            else:
                with pytest.raises(RuntimeError):
                    op.report_completed(0.1)
        return lens

    assert ex.run(fn) == [[5, 12]] * 3
    assert any(p[0].endswith("/signals/ack_preemption") for p in sess.posts), "out-of-ops must be acknowledged"


def test_searcher_must_complete_each_op():
    sc = core.DummySearcherContext(core.DummyDistributedContext(), length=3)
    with pytest.raises(RuntimeError, match="report_completed"):
        for op in sc.operations():
            pass
    for op in sc.operations():
        assert op.length == 3
        op.report_completed(1.0)
        with pytest.raises(RuntimeError):
            op.report_completed(1.0)


def test_preempt_workers_ask_chief_all_ranks_agree():
    class Sess(FakeSession):
        def __init__(self) -> None:
            super().__init__()
            self.flag = threading.Event()

        def get(self, path, params=None, timeout=None, **kw):
            r = mock.MagicMock()
            t = float((params or {}).get("timeout_seconds", 0))
            r.json.return_value = {"preempt": self.flag.wait(min(t, 0.2))}
            return r

    sess = Sess()
    ex = parallel_util.Execution(3)
    barrier = threading.Barrier(3)

    def fn():
        d = ex.distributed
        pc = core.PreemptContext(sess, "alloc", d, core.PreemptMode.WorkersAskChief)
        with pc:
            first = pc.should_preempt()
            barrier.wait()
            if d.rank == 0:
                sess.flag.set()
                time.sleep(0.6)
            barrier.wait()
            seq = [pc.should_preempt() for _ in range(3)]
        return first, seq

    res = ex.run(fn)
    assert all(r == (False, [True, True, True]) for r in res)
    assert len([p for p in sess.posts if p[0].endswith("ack_preemption")]) == 1


def test_preempt_chief_only_rejects_workers():
    ex = parallel_util.Execution(2)
    ex._expect_errors = True

    def fn():
        d = ex.distributed
        pc = core.DummyPreemptContext(d, core.PreemptMode.ChiefOnly).start()
        return pc.should_preempt()

    ex.run(fn)
    assert ex.errors[0] is None and "ChiefOnly" in str(ex.errors[1][0])


def test_train_context_wire_format():
    sess = FakeSession()
    tc = core.TrainContext(sess, 3, 2, 1, core.DummyDistributedContext(), core.TensorboardMode.MANUAL)
    tc.report_training_metrics(10, {"loss": 0.5}, [{"loss": 0.6}, {"loss": 0.4}])
    tc.report_validation_metrics(10, {"acc": 0.9})
    tc.set_status("validating")
    tc.report_early_exit(core.EarlyExitReason.INVALID_HP)
    paths = [p[0] for p in sess.posts]
    assert paths == ["/api/v1/trials/3/metrics", "/api/v1/trials/3/metrics", "/api/v1/trials/3/runner/metadata", "/api/v1/trials/3/early_exit"]
    body = json.loads(sess.posts[0][1])
    assert body["group"] == "training" and body["metrics"]["stepsCompleted"] == 10 and body["metrics"]["trialRunId"] == 2
    assert body["metrics"]["metrics"] == {"avgMetrics": {"loss": 0.5}, "batchMetrics": [{"loss": 0.6}, {"loss": 0.4}]}


def test_context_invalid_hp_reports_early_exit(tmp_path):
    from determined_b200 import InvalidHP

    ctx = core._dummy_init(checkpoint_storage=str(tmp_path))
    ctx.train = mock.MagicMock()
    with pytest.raises(SystemExit) as e:
        with ctx:
            raise InvalidHP("nope")
    assert e.value.code == 0
    ctx.train.report_early_exit.assert_called_once_with(core.EarlyExitReason.INVALID_HP)


def test_profiler_collects_and_ships():
    got = []

    class Train:
        def report_metrics(self, group, steps_completed, metrics, report_time=None):
            got.append((group, metrics))

    p = core.ProfilerContext("agent-x", core.DummyDistributedContext(), Train())
    p.on(sampling_interval=0.1, samples_per_report=2)
    time.sleep(0.8)
    p.off()
    groups = {g for g, _ in got}
    assert {"cpu", "memory"} <= groups
    assert all(list(m) == ["agent-x"] for _, m in got)
