"""Port registry: hands out TCP ports for task services (tensorboard / notebook servers, rendezvous)
so concurrent tasks on one host never collide.

Parity: reference ``master/internal/portregistry`` (``GetPort`` / ``ReleasePort`` over a configured
range, used for the per-task proxy and DeepSpeed / torch rendezvous ports).
"""
import socket
import threading
from typing import Optional, Set


class PortRegistry:
    def __init__(self, lo: int = 22000, hi: int = 23000, host: str = "127.0.0.1") -> None:
        if not (0 < lo < hi <= 65536):
            raise ValueError("bad port range")
        self.lo, self.hi, self.host = lo, hi, host
        self._used: Set[int] = set()
        self._next = lo
        self._lock = threading.Lock()

    def _bindable(self, port: int) -> bool:
        s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        try:
            s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            s.bind((self.host, port))
            return True
        except OSError:
            return False
        finally:
            s.close()

    def acquire(self) -> int:
        """A port nobody holds and that binds right now; raises when the range is exhausted."""
        with self._lock:
            for _ in range(self.hi - self.lo):
                p = self._next
                self._next = self.lo + (self._next + 1 - self.lo) % (self.hi - self.lo)
                if p not in self._used and self._bindable(p):
                    self._used.add(p)
                    return p
        raise RuntimeError(f"no free port in [{self.lo}, {self.hi})")

    def release(self, port: Optional[int]) -> None:
        if port is not None:
            with self._lock:
                self._used.discard(port)

    def in_use(self) -> Set[int]:
        with self._lock:
            return set(self._used)
