"""ctypes front end of ``csrc/host/shm_collective.cpp``: pickled-object allgather / broadcast / gather between the ranks
of one box through a POSIX shared-memory segment (no sockets, no hub).  Selected by ``DistributedContext`` for single-node
trials when the host library is built; the socket star stays the fallback and the multi-node path.
"""
import ctypes
import os
import pathlib
import pickle
from typing import Any, List, Optional

_LIB: Optional[ctypes.CDLL] = None
_LIB_PATH = pathlib.Path(__file__).resolve().parents[1] / "_native" / "libdetb200_host.so"


def load() -> Optional[ctypes.CDLL]:
    global _LIB
    if _LIB is None and _LIB_PATH.exists():
        lib = ctypes.CDLL(str(_LIB_PATH))
        u64p = ctypes.POINTER(ctypes.c_uint64)
        lib.detb200_shm_last_error.restype = ctypes.c_char_p
        lib.detb200_shm_open.restype = ctypes.c_void_p
        lib.detb200_shm_open.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_size_t, ctypes.c_double]
        lib.detb200_shm_close.argtypes = [ctypes.c_void_p]
        lib.detb200_shm_unlink.argtypes = [ctypes.c_void_p]
        lib.detb200_shm_exchange_sizes.argtypes = [ctypes.c_void_p, ctypes.c_uint64, u64p, ctypes.c_double]
        lib.detb200_shm_allgatherv.argtypes = [ctypes.c_void_p, ctypes.c_char_p, u64p, u64p, ctypes.c_void_p, ctypes.c_double]
        lib.detb200_shm_barrier.argtypes = [ctypes.c_void_p, ctypes.c_double]
        _LIB = lib
    return _LIB


def available() -> bool:
    return load() is not None and os.path.isdir("/dev/shm")


class ShmTransport:
    """Same three verbs as the socket / torch transports of ``core/_distributed.py``."""

    def __init__(self, rank: int, size: int, tag: str, slot_bytes: int = 1 << 20, timeout_s: float = 600.0) -> None:
        lib = load()
        if lib is None:
            raise RuntimeError("libdetb200_host.so is not built (python -m determined_b200._native.build)")
        self.rank, self.size, self.timeout_s = rank, size, timeout_s
        self._lib = lib
        self._name = ("/detb200-" + "".join(c if c.isalnum() else "-" for c in tag))[:200].encode()
        self._h = lib.detb200_shm_open(self._name, rank, size, slot_bytes, min(timeout_s, 120.0))
        if not self._h:
            raise RuntimeError(f"shared-memory transport: {lib.detb200_shm_last_error().decode()}")
        if size > 1:
            # everyone is attached after this barrier; rank 0 then drops the name, so close() is a plain unmap and ranks may
            # close in any order (or not at all, if the process is killed) without leaving a segment behind
            self.barrier()
            if rank == 0:
                lib.detb200_shm_unlink(self._h)

    def _check(self, rc: int) -> None:
        if rc != 0:
            raise RuntimeError(f"shared-memory collective failed (rc={rc}): {self._lib.detb200_shm_last_error().decode()}")

    def _allgather_bytes(self, data: bytes) -> List[bytes]:
        n = self.size
        sizes = (ctypes.c_uint64 * n)()
        self._check(self._lib.detb200_shm_exchange_sizes(self._h, len(data), sizes, self.timeout_s))
        offsets = (ctypes.c_uint64 * n)()
        total = 0
        for j in range(n):
            offsets[j] = total
            total += sizes[j]
        out = ctypes.create_string_buffer(max(total, 1))
        if max(sizes) > 0:
            self._check(self._lib.detb200_shm_allgatherv(self._h, data, sizes, offsets, out, self.timeout_s))
        raw = out.raw
        return [raw[offsets[j]:offsets[j] + sizes[j]] for j in range(n)]

    def allgather(self, obj: Any) -> List[Any]:
        return [pickle.loads(b) for b in self._allgather_bytes(pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL))]

    def gather(self, obj: Any) -> Optional[List[Any]]:
        out = self.allgather(obj)
        return out if self.rank == 0 else None

    def broadcast(self, obj: Any) -> Any:
        parts = self._allgather_bytes(pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL) if self.rank == 0 else b"")
        return pickle.loads(parts[0])

    def barrier(self) -> None:
        self._check(self._lib.detb200_shm_barrier(self._h, self.timeout_s))

    def close(self) -> None:
        if self._h:
            self._lib.detb200_shm_close(self._h)
            self._h = None
