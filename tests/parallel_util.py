"""Thread-per-rank harness: run a function once per simulated rank, each thread owning a REAL
``core.DistributedContext`` (socket transport on localhost).  Test strategy from SURVEY §4
("multi-rank without a cluster"), mirroring the reference's ``harness/tests/parallel.py:14-153``."""
import contextlib
import socket
import threading
import traceback
from typing import Any, Callable, List, Optional

from determined_b200 import core


def free_port() -> int:
    with contextlib.closing(socket.socket()) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class Execution:
    def __init__(self, size: int, local_size: Optional[int] = None, force_tcp: bool = True, timeout: float = 60) -> None:
        self.size = size
        self.local_size = local_size or size
        assert size % self.local_size == 0
        self.force_tcp = force_tcp
        self.timeout = timeout
        self._tls = threading.local()
        self.port = free_port()

    @property
    def rank(self) -> int:
        return self._tls.rank

    @property
    def distributed(self) -> core.DistributedContext:
        return self._tls.dist

    def run(self, fn: Callable[[], Any]) -> List[Any]:
        results: List[Any] = [None] * self.size
        errors: List[Any] = [None] * self.size

        def worker(rank: int) -> None:
            try:
                self._tls.rank = rank
                self._tls.dist = core.DistributedContext(
                    rank=rank, size=self.size, local_rank=rank % self.local_size, local_size=self.local_size,
                    cross_rank=rank // self.local_size, cross_size=self.size // self.local_size,
                    chief_ip="127.0.0.1", pub_port=self.port, force_tcp=self.force_tcp, _transport="socket")
                try:
                    results[rank] = fn()
                finally:
                    self._tls.dist.close()
            except BaseException as e:  # noqa
                errors[rank] = (e, traceback.format_exc())

        threads = [threading.Thread(target=worker, args=(r,), daemon=True) for r in range(self.size)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(self.timeout)
            if t.is_alive():
                raise TimeoutError("rank thread did not finish (deadlocked collective?)")
        self.errors = errors
        for r, e in enumerate(errors):
            if e is not None and not getattr(self, "_expect_errors", False):
                raise AssertionError(f"rank {r} failed:\n{e[1]}")
        return results
