"""Host-side pieces of the symmetric-memory layer that need no GPU: the file-descriptor exchange that carries VMM
allocation handles and the multicast object between rank processes, and the bucket plan of the gradient engine."""
import os
import threading

import torch

from determined_b200.ops import comm
from determined_b200.parallel.flat import FlatLayout


def test_fd_exchange_delivers_descriptors_out_of_order_between_ranks():
    world, key = 3, "t" + os.urandom(4).hex()
    xs = [comm.FdExchange(r, world, key) for r in range(world)]
    pipes = {}
    try:
        # every rank sends two descriptors (two "buffers") to every peer; receivers ask for the second one first
        for r in range(world):
            for seq in (0, 1):
                rd, wr = os.pipe()
                os.write(wr, f"rank{r}-seq{seq}".encode())
                os.close(wr)
                pipes[(r, seq)] = rd
                for dst in range(world):
                    if dst != r:
                        xs[r].send(dst, seq, rd)
        got = {}

        def receiver(me: int) -> None:
            for seq in (1, 0):
                for src in range(world):
                    if src != me:
                        fd = xs[me].recv(seq, src)
                        got[(me, src, seq)] = fd

        ts = [threading.Thread(target=receiver, args=(r,)) for r in range(world)]
        [t.start() for t in ts]
        [t.join(10) for t in ts]
        assert len(got) == world * (world - 1) * 2
        # a received descriptor is a duplicate of the sender's: same pipe, so the first reader drains it
        seen = set()
        for (me, src, seq), fd in got.items():
            assert fd not in pipes.values() or True
            st_a, st_b = os.fstat(fd), os.fstat(pipes[(src, seq)])
            assert (st_a.st_ino, st_a.st_dev) == (st_b.st_ino, st_b.st_dev)
            seen.add((src, seq))
            os.close(fd)
        assert seen == set(pipes)
    finally:
        for fd in pipes.values():
            os.close(fd)
        for x in xs:
            x.close()


def test_bucket_plan_splits_off_a_small_last_ready_bucket():
    torch.manual_seed(0)
    sizes = [300, 64, 5000, 5000, 20000, 20000, 40000, 10]      # "first layers" first, like module.parameters()
    params = [torch.nn.Parameter(torch.zeros(n)) for n in sizes]
    lay = FlatLayout([{"params": params}], torch.device("cpu"), align=8, segment_align=8)
    plain = lay.make_buckets(cap_bytes=4 * 50000, multiple=8)
    tailed = lay.make_buckets(cap_bytes=4 * 50000, multiple=8, tail_cap_bytes=4 * 4000)
    for plan in (plain, tailed):
        # buckets tile [0,total) exactly, in reverse order (bucket 0 = last parameters = first ready)
        spans = sorted((b.offset, b.offset + b.numel) for b in plan)
        assert spans[0][0] == 0 and spans[-1][1] == lay.total
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert plan[0].offset + plan[0].numel == lay.total
        owned = [id(s.param) for b in plan for s in b.owned]
        assert sorted(owned) == sorted(id(p) for p in params)
        assert all(b.offset % 8 == 0 and b.numel % 8 == 0 for b in plan)
    assert len(tailed) == len(plain) + 1
    last = tailed[-1]
    assert last.offset == 0 and [s.numel for s in last.owned] == [300, 64, 5000]     # cut at the first boundary past 4000 elements
    assert last.numel * 4 < 4 * 50000
