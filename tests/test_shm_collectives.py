"""Shared-memory object collectives (csrc/host/shm_collective.cpp): semantics of allgather / gather / broadcast across
threads and processes, messages larger than a slot, many rounds, stale segments."""
import multiprocessing as mp
import os
import threading
import time
import uuid

import pytest

from determined_b200 import core
from determined_b200.core import _shm_transport as shm

pytestmark = pytest.mark.skipif(not shm.available(), reason="libdetb200_host.so not built")


def _ranks(n, fn, **kw):
    tag = uuid.uuid4().hex
    out, err = [None] * n, [None] * n

    def w(r):
        try:
            t = shm.ShmTransport(r, n, tag, timeout_s=30, **kw)
            try:
                out[r] = fn(t, r)
            finally:
                t.close()
        except BaseException as e:  # noqa: BLE001
            err[r] = e

    th = [threading.Thread(target=w, args=(r,), daemon=True) for r in range(n)]
    [t.start() for t in th]
    [t.join(60) for t in th]
    assert not any(t.is_alive() for t in th), "deadlocked collective"
    for e in err:
        if e is not None:
            raise e
    return out


def test_allgather_gather_broadcast_semantics():
    def fn(t, r):
        a = t.allgather({"rank": r, "blob": b"x" * r})
        g = t.gather(("g", r))
        b = t.broadcast("from-chief" if r == 0 else None)
        e = t.allgather(None)
        return a, g, b, e

    res = _ranks(4, fn)
    for r, (a, g, b, e) in enumerate(res):
        assert a == [{"rank": j, "blob": b"x" * j} for j in range(4)]
        assert g == ([("g", j) for j in range(4)] if r == 0 else None)
        assert b == "from-chief" and e == [None] * 4


def test_messages_larger_than_a_slot_and_many_rounds():
    def fn(t, r):
        big = os.urandom(1) * (300_000 + 7 * r)                  # 74+ rounds through a 4 KiB slot, different length per rank
        got = t.allgather(big)
        assert [len(x) for x in got] == [300_000 + 7 * j for j in range(3)] and got[r] == big
        acc = 0
        for i in range(3000):                                     # sequence numbers just keep counting
            acc += sum(t.allgather(i + r))
        return acc

    res = _ranks(3, fn, slot_bytes=4096)
    assert len(set(res)) == 1 and res[0] == sum(3 * i + 3 for i in range(3000))


def _proc(rank, n, tag, q):
    t = shm.ShmTransport(rank, n, tag, timeout_s=30)
    vals = t.allgather((rank, os.getpid()))
    word = t.broadcast("hello" if rank == 0 else None)
    t.close()
    q.put((rank, vals, word))


def test_across_processes_and_stale_segment_is_ignored():
    tag = uuid.uuid4().hex
    # a segment left behind by a dead creator must not capture late attachers
    stale = f"/dev/shm/detb200-{tag}"
    with open(stale, "wb") as f:
        f.write(b"\0" * (1 << 16))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_proc, args=(r, 3, tag, q)) for r in (1, 2)]
    [p.start() for p in ps]
    time.sleep(0.5)                                               # ranks 1 and 2 are already polling for the segment
    p0 = ctx.Process(target=_proc, args=(0, 3, tag, q))
    p0.start()
    got = sorted(q.get(timeout=60) for _ in range(3))
    [p.join(30) for p in ps + [p0]]
    pids = [v[1] for v in got[0][1]]
    assert len(set(pids)) == 3
    for r, vals, word in got:
        assert [v[0] for v in vals] == [0, 1, 2] and word == "hello"
    assert not os.path.exists(stale), "rank 0 unlinks the segment on close"


def test_distributed_context_uses_it_on_one_node():
    tag_port = 20000 + os.getpid() % 20000
    out = [None] * 2

    def w(r):
        d = core.DistributedContext(rank=r, size=2, local_rank=r, local_size=2, cross_rank=0, cross_size=1, chief_ip="127.0.0.1", pub_port=tag_port)
        try:
            assert type(d._transport).__name__ == "ShmTransport"
            out[r] = (d.allgather(r * 10), d.broadcast("cfg" if r == 0 else None), d.gather_local(r), d.allgather_local(r))
        finally:
            d.close()

    th = [threading.Thread(target=w, args=(r,), daemon=True) for r in range(2)]
    [t.start() for t in th]
    [t.join(60) for t in th]
    assert out[0] == ([0, 10], "cfg", [0, 1], [0, 1]) and out[1] == ([0, 10], "cfg", None, [0, 1])


def test_protocol_is_race_free_under_thread_sanitizer(tmp_path):
    """Builds the collective with -fsanitize=thread and runs 4 ranks as threads on one mapping; also proves the check has teeth:
    the same run with the arrive store weakened to relaxed must be reported."""
    import pathlib
    import subprocess

    src = pathlib.Path(__file__).resolve().parent.parent / "csrc" / "host"

    def build(lib_src, out):
        r = subprocess.run(["g++", "-std=c++17", "-O1", "-g", "-fsanitize=thread", "-pthread", str(lib_src), str(src / "shm_collective_test.cpp"), "-lrt", "-o", str(out)],
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        return r

    good = tmp_path / "shm_tsan"
    r = build(src / "shm_collective.cpp", good)
    if r.returncode != 0:
        pytest.skip("no ThreadSanitizer runtime: " + r.stdout[-200:])
    env = dict(os.environ, TSAN_OPTIONS="exitcode=66")
    r = subprocess.run([str(good), "4", "200"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and "ThreadSanitizer" not in r.stdout, r.stdout[-3000:]
    broken_src = tmp_path / "broken.cpp"
    text = (src / "shm_collective.cpp").read_text()
    weak = text.replace("h->arrive()[h->rank].v.store(s, std::memory_order_release)", "h->arrive()[h->rank].v.store(s, std::memory_order_relaxed)")
    assert weak != text
    broken_src.write_text(weak)
    bad = tmp_path / "shm_broken"
    assert build(broken_src, bad).returncode == 0
    r = subprocess.run([str(bad), "4", "50"], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert "ThreadSanitizer: data race" in r.stdout
