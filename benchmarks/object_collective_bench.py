"""Latency of the control-plane allgather (pickled small dict) between N local processes: shared memory vs the socket star."""
import json
import multiprocessing as mp
import os
import sys
import time
import uuid

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def worker(rank, n, kind, tag, port, iters, q):
    from determined_b200 import core

    d = core.DistributedContext(rank=rank, size=n, local_rank=rank, local_size=n, cross_rank=0, cross_size=1, chief_ip="127.0.0.1",
                                pub_port=port, _transport=kind)
    obj = {"rank": rank, "loss": 0.123, "preempt": False}
    for _ in range(50):
        d.allgather(obj)
    t0 = time.perf_counter()
    for _ in range(iters):
        d.allgather(obj)
    dt = (time.perf_counter() - t0) / iters
    big = os.urandom(4 << 20)
    t1 = time.perf_counter()
    for _ in range(5):
        d.allgather(big)
    dtb = (time.perf_counter() - t1) / 5
    d.close()
    if rank == 0:
        q.put((dt, dtb))


def main():
    n, iters = int(os.environ.get("N", 4)), 2000
    out = {"ranks": n}
    ctx = mp.get_context("spawn")
    for kind in ("shm", "socket"):
        os.environ["DET_ALLOCATION_ID"] = uuid.uuid4().hex
        q = ctx.Queue()
        port = 21000 + os.getpid() % 1000 + (0 if kind == "shm" else 7)
        ps = [ctx.Process(target=worker, args=(r, n, kind, "", port, iters, q)) for r in range(n)]
        [p.start() for p in ps]
        dt, dtb = q.get(timeout=300)
        [p.join(60) for p in ps]
        out[kind] = {"small_allgather_us": dt * 1e6, "allgather_4MiB_ms": dtb * 1e3}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
