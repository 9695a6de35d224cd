"""Symmetric memory + in-kernel collectives (native side: ``csrc/comm.cu``).

``SymmetricBuffer``  device memory allocated outside torch's caching allocator so it can be
                     exported with CUDA IPC and mapped by the other ranks of the trial.
``PeerGroup``        rank/world + the bootstrap channel (``torch.distributed`` object all-gather)
                     used once to exchange IPC handles; owns the per-rank control block (flags,
                     epoch counter, per-rank statistics).
``allreduce_fused``  two-shot P2P all-reduce over NVLink fused with scaling (1/(world*agg),
                     optional 1/loss_scale), non-finite detection and squared-norm accumulation.

Replaces the NCCL all-reduce + foreach unscale/div/norm kernels the reference triggers through
DDP/Horovod (``_pytorch_context.py:296-299,357-368,805-812,885-888``).
"""
import ctypes
import logging
from ctypes import c_void_p
from typing import Any, List, Optional

import torch

from determined_b200.ops import _lib

logger = logging.getLogger("determined.ops")

DEFAULT_TIMEOUT_S = 120.0
MAX_RANKS = 16


class _CudaArray:
    def __init__(self, ptr: int, nbytes: int, owner: Any) -> None:
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }


class SymmetricBuffer:
    """``nbytes`` of zero-initialised device memory on the current device, exportable to peers."""

    def __init__(self, nbytes: int, device: Optional[torch.device] = None) -> None:
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.nbytes = int((nbytes + 255) // 256 * 256)
        lib = _lib.load()
        out = c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(lib.detb200_symm_alloc(ctypes.byref(out), self.nbytes), "symm_alloc")
        self.ptr = int(out.value)
        self._bytes = torch.as_tensor(_CudaArray(self.ptr, self.nbytes, self), device=self.device)
        self._freed = False

    def tensor(self, dtype: torch.dtype, numel: int, offset_bytes: int = 0) -> torch.Tensor:
        item = torch.empty((), dtype=dtype).element_size()
        assert offset_bytes + numel * item <= self.nbytes
        return self._bytes[offset_bytes: offset_bytes + numel * item].view(dtype)

    def export_handle(self) -> bytes:
        lib = _lib.load()
        n = lib.detb200_ipc_handle_bytes()
        buf = ctypes.create_string_buffer(n)
        _lib.check(lib.detb200_ipc_export(self.ptr, buf), "ipc_export")
        return bytes(buf.raw)

    def free(self) -> None:
        if not self._freed:
            self._freed = True
            try:
                _lib.load().detb200_symm_free(self.ptr)
            except Exception:
                pass


class PeerGroup:
    """One per trial process.  ``dist_group`` is a torch.distributed group used only at set-up."""

    def __init__(self, rank: int = 0, world: int = 1, dist_group: Any = None,
                 device: Optional[torch.device] = None) -> None:
        self.rank, self.world = rank, world
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self._dist_group = dist_group
        self._opened: List[int] = []
        lib = _lib.load()
        self.ctrl = SymmetricBuffer(int(lib.detb200_ctrl_bytes()), self.device)
        self.ctrl_ptrs = self.map(self.ctrl)
        self.found_inf = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._first_of_step = True

    @classmethod
    def from_torch_distributed(cls, device: Optional[torch.device] = None) -> "PeerGroup":
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            return cls(dist.get_rank(), dist.get_world_size(), dist.group.WORLD, device)
        return cls(0, 1, None, device)

    def map(self, buf: SymmetricBuffer) -> Any:
        """Return a ctypes ``void*[world]`` with our pointer at ``rank`` and peers' mapped pointers."""
        arr = (c_void_p * self.world)()
        if self.world == 1:
            arr[0] = buf.ptr
            return arr
        import torch.distributed as dist

        handles: List[Any] = [None] * self.world
        dist.all_gather_object(handles, (self.device.index, buf.export_handle()), group=self._dist_group)
        lib = _lib.load()
        with torch.cuda.device(self.device):
            for r, (dev_index, h) in enumerate(handles):
                if r == self.rank:
                    arr[r] = buf.ptr
                    continue
                out = c_void_p()
                _lib.check(lib.detb200_ipc_open(h, ctypes.byref(out)), f"ipc_open(rank {r})")
                arr[r] = out.value
                self._opened.append(int(out.value))
        dist.barrier(group=self._dist_group)
        return arr

    def stats_tensor(self) -> torch.Tensor:
        """Device view ``float[2, MAX_RANKS]``: row 0 = per-rank sum of squares of the reduced
        gradients this step, row 1 = per-rank non-finite counts."""
        lib = _lib.load()
        p = lib.detb200_comm_stats_ptr(self.ctrl.ptr)
        off = int(p) - self.ctrl.ptr
        return self.ctrl.tensor(torch.float32, 2 * MAX_RANKS, off).view(2, MAX_RANKS)

    def begin_step(self) -> None:
        self._first_of_step = True

    def close(self) -> None:
        lib = _lib.load()
        for p in self._opened:
            try:
                lib.detb200_ipc_close(p)
            except Exception:
                pass
        self._opened.clear()


def allreduce_fused(group: PeerGroup, peer_bufs: Any, offset: int, numel: int, scale: float,
                    inv_loss_scale: Optional[torch.Tensor] = None, found_inf: Optional[torch.Tensor] = None,
                    stream: Optional[torch.cuda.Stream] = None, phases: int = 7, rank: Optional[int] = None,
                    timeout_s: float = DEFAULT_TIMEOUT_S) -> None:
    """In-place all-reduce (sum * scale) of ``[offset, offset+numel)`` fp32 elements of the
    symmetric buffers ``peer_bufs``.  Statistics accumulate in ``group.stats_tensor()`` until
    ``group.begin_step()``."""
    lib = _lib.load()
    first = group._first_of_step
    group._first_of_step = False
    _lib.check(lib.detb200_allreduce_fused(peer_bufs, group.ctrl_ptrs, group.rank if rank is None else rank,
                                           group.world, offset, numel, float(scale), _lib.ptr(inv_loss_scale),
                                           _lib.ptr(found_inf), int(first), phases, float(timeout_s),
                                           _lib.stream_ptr(stream)), "allreduce_fused")


def broadcast(group: PeerGroup, peer_bufs: Any, offset: int, numel: int, root: int = 0,
              stream: Optional[torch.cuda.Stream] = None, phases: int = 7,
              timeout_s: float = DEFAULT_TIMEOUT_S) -> None:
    lib = _lib.load()
    _lib.check(lib.detb200_broadcast(peer_bufs, group.ctrl_ptrs, group.rank, group.world, root, offset, numel,
                                     phases, float(timeout_s), _lib.stream_ptr(stream)), "broadcast")
