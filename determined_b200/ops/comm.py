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
import array
import ctypes
import logging
import os
import secrets
import socket
import struct
from ctypes import c_void_p
from typing import Any, Dict, List, Optional, Tuple

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


def vmm_caps(device: Optional[torch.device] = None) -> Tuple[bool, bool]:
    """(exportable VMM allocations available, multicast objects available) on ``device``."""
    if os.environ.get("DETB200_SYMM", "").lower() == "ipc":
        return False, False
    lib = _lib.load()
    caps = (ctypes.c_int * 2)()
    with torch.cuda.device(device if device is not None else torch.cuda.current_device()):
        try:
            _lib.check0(lib.detb200_vmm_query(caps), "vmm_query")
        except _lib.NativeError:
            return False, False
    mc = bool(caps[1]) and os.environ.get("DETB200_NVLS", "1") != "0"
    return bool(caps[0]), mc


class SymmetricBuffer:
    """``nbytes`` of zero-initialised device memory on the current device that the other ranks of the trial can map.

    ``vmm=True`` allocates through the CUDA virtual-memory API (handle exported as a file descriptor; can be bound to an
    NVSwitch multicast object, see ``PeerGroup.multicast``); ``vmm=False`` is cudaMalloc + legacy CUDA IPC.  ``world`` sizes
    the allocation granularity for a multicast object of that many devices."""

    def __init__(self, nbytes: int, device: Optional[torch.device] = None, vmm: bool = False, world: int = 1) -> None:
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        lib = _lib.load()
        self.vmm = bool(vmm)
        self.fd = -1
        out = c_void_p()
        with torch.cuda.device(self.device):
            if self.vmm:
                gran = ctypes.c_int64()
                _lib.check0(lib.detb200_vmm_granularity(int(world), ctypes.byref(gran)), "vmm_granularity")
                g = max(int(gran.value), 256)
                self.nbytes = int((nbytes + g - 1) // g * g)
                fd = ctypes.c_int(-1)
                _lib.check0(lib.detb200_vmm_alloc(self.nbytes, ctypes.byref(out), ctypes.byref(fd)), "vmm_alloc")
                self.fd = int(fd.value)
            else:
                self.nbytes = int((nbytes + 255) // 256 * 256)
                _lib.check0(lib.detb200_symm_alloc(ctypes.byref(out), self.nbytes), "symm_alloc")
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
        _lib.check0(lib.detb200_ipc_export(self.ptr, buf), "ipc_export")
        return bytes(buf.raw)

    def close_fd(self) -> None:
        if self.fd >= 0:
            os.close(self.fd)
            self.fd = -1

    def free(self) -> None:
        if not self._freed:
            self._freed = True
            self.close_fd()
            try:
                lib = _lib.load()
                (lib.detb200_vmm_free if self.vmm else lib.detb200_symm_free)(self.ptr)
            except Exception:
                pass


class FdExchange:
    """Pass file descriptors between the rank processes of one box: one unix datagram socket per rank in the abstract
    namespace, SCM_RIGHTS messages tagged ``(sequence, source rank)``.  This is how VMM allocation handles and the multicast
    object reach the peers (NCCL does the same for its own cuMem handles)."""

    def __init__(self, rank: int, world: int, key: str) -> None:
        self.rank, self.world, self.key = rank, world, key
        self.sock = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
        self.sock.bind(self._addr(rank))
        self.sock.settimeout(120.0)
        self._stash: Dict[Tuple[int, int], int] = {}

    def _addr(self, r: int) -> str:
        return f"\0detb200-fd-{self.key}-{r}"

    def send(self, dst: int, seq: int, fd: int) -> None:
        # (socket.send_fds drops its `address` argument: call sendmsg directly)
        self.sock.sendmsg([struct.pack("<qq", seq, self.rank)], [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", [fd]))],
                          0, self._addr(dst))

    def recv(self, seq: int, src: int) -> int:
        """The descriptor ``src`` sent with tag ``seq`` (messages that arrive early for other tags are kept)."""
        while (seq, src) not in self._stash:
            msg, fds, _flags, _addr = socket.recv_fds(self.sock, 16, 4)
            s, r = struct.unpack("<qq", msg)
            if fds:
                self._stash[(s, r)] = fds[0]
                for extra in fds[1:]:
                    os.close(extra)
        return self._stash.pop((seq, src))

    def close(self) -> None:
        for fd in self._stash.values():
            os.close(fd)
        self._stash.clear()
        self.sock.close()


class PeerGroup:
    """One per trial process.  ``dist_group`` is a torch.distributed group used only at set-up."""

    def __init__(self, rank: int = 0, world: int = 1, dist_group: Any = None,
                 device: Optional[torch.device] = None) -> None:
        self.rank, self.world = rank, world
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self._dist_group = dist_group
        self._opened: List[int] = []
        self._opened_vmm: List[int] = []
        self._mcs: List[Tuple[int, int, int]] = []
        self._seq = 0
        self._fdx: Optional[FdExchange] = None
        lib = _lib.load()
        self.vmm_ok, self.nvls_ok = vmm_caps(self.device) if world > 1 else (False, False)
        if world > 1:
            import torch.distributed as dist

            # every rank must take the same path
            flags = [None] * world
            dist.all_gather_object(flags, (self.vmm_ok, self.nvls_ok), group=dist_group)
            self.vmm_ok = all(f[0] for f in flags)
            self.nvls_ok = self.vmm_ok and all(f[1] for f in flags)
            if self.vmm_ok:
                key = [secrets.token_hex(8) if rank == 0 else None]
                dist.broadcast_object_list(key, src=dist.get_global_rank(dist_group, 0) if dist_group is not None else 0, group=dist_group)
                self._fdx = FdExchange(rank, world, key[0])
                dist.barrier(group=dist_group)     # every socket is bound before anyone sends
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

    def alloc(self, nbytes: int) -> SymmetricBuffer:
        """A symmetric buffer of the best kind this group can map: VMM (multicast-capable) when every rank supports it."""
        with torch.cuda.device(self.device):
            return SymmetricBuffer(nbytes, self.device, vmm=self.vmm_ok, world=self.world)

    def map(self, buf: SymmetricBuffer) -> Any:
        """Return a ctypes ``void*[world]`` with our pointer at ``rank`` and peers' mapped pointers."""
        arr = (c_void_p * self.world)()
        if self.world == 1:
            arr[0] = buf.ptr
            return arr
        import torch.distributed as dist

        lib = _lib.load()
        if buf.vmm:
            assert self._fdx is not None, "VMM buffers need the fd exchange (vmm_ok)"
            seq = self._seq
            self._seq += 1
            sizes: List[Any] = [None] * self.world
            dist.all_gather_object(sizes, buf.nbytes, group=self._dist_group)
            for r in range(self.world):
                if r != self.rank:
                    self._fdx.send(r, seq, buf.fd)
            with torch.cuda.device(self.device):
                for r in range(self.world):
                    if r == self.rank:
                        arr[r] = buf.ptr
                        continue
                    fd = self._fdx.recv(seq, r)
                    out = c_void_p()
                    try:
                        _lib.check0(lib.detb200_vmm_import(fd, int(sizes[r]), ctypes.byref(out)), f"vmm_import(rank {r})")
                    finally:
                        os.close(fd)
                    arr[r] = out.value
                    self._opened_vmm.append(int(out.value))
            dist.barrier(group=self._dist_group)
            return arr
        handles: List[Any] = [None] * self.world
        dist.all_gather_object(handles, (self.device.index, buf.export_handle()), group=self._dist_group)
        with torch.cuda.device(self.device):
            for r, (dev_index, h) in enumerate(handles):
                if r == self.rank:
                    arr[r] = buf.ptr
                    continue
                out = c_void_p()
                _lib.check0(lib.detb200_ipc_open(h, ctypes.byref(out)), f"ipc_open(rank {r})")
                arr[r] = out.value
                self._opened.append(int(out.value))
        dist.barrier(group=self._dist_group)
        return arr

    def multicast(self, buf: SymmetricBuffer) -> Optional[int]:
        """Bind ``buf`` (same size on every rank, VMM) to a new multicast object and return this rank's mapping of it: the
        address ``multimem.ld_reduce`` / ``multimem.st`` take.  None when the box has no NVLS or the set-up fails on any
        rank (the caller then stays on peer-to-peer loads and stores).  Collective."""
        if self.world == 1 or not (self.nvls_ok and buf.vmm):
            return None
        import torch.distributed as dist

        lib = _lib.load()
        seq = self._seq
        self._seq += 1
        mc = ctypes.c_int64(0)
        ok = True
        err = ""
        with torch.cuda.device(self.device):
            try:
                if self.rank == 0:
                    fd = ctypes.c_int(-1)
                    _lib.check0(lib.detb200_mc_create(self.world, buf.nbytes, ctypes.byref(mc), ctypes.byref(fd)), "mc_create")
                    for r in range(1, self.world):
                        self._fdx.send(r, seq, int(fd.value))
                    os.close(int(fd.value))
            except _lib.NativeError as e:
                ok, err = False, str(e)
            # rank 0's outcome decides whether anybody waits for a descriptor
            flag = [ok if self.rank == 0 else None]
            dist.broadcast_object_list(flag, src=dist.get_global_rank(self._dist_group, 0) if self._dist_group is not None else 0,
                                       group=self._dist_group)
            if not flag[0]:
                logger.info("multicast object could not be created (%s); using peer-to-peer collectives", err)
                return None
            stages = ("import", "add_device", "bind", "map")
            mc_ptr = c_void_p()
            for stage in stages:
                try:
                    if ok and stage == "import" and self.rank != 0:
                        fd_in = self._fdx.recv(seq, 0)
                        try:
                            _lib.check0(lib.detb200_mc_import(fd_in, ctypes.byref(mc)), "mc_import")
                        finally:
                            os.close(fd_in)
                    elif ok and stage == "add_device":
                        _lib.check0(lib.detb200_mc_add_device(mc.value), "mc_add_device")
                    elif ok and stage == "bind":
                        _lib.check0(lib.detb200_mc_bind(mc.value, buf.ptr, buf.nbytes), "mc_bind")
                    elif ok and stage == "map":
                        _lib.check0(lib.detb200_mc_map(mc.value, buf.nbytes, ctypes.byref(mc_ptr)), "mc_map")
                except (_lib.NativeError, OSError) as e:
                    ok, err = False, f"{stage}: {e}"
                # every device must be added before any bind, and everybody must agree on the outcome of each stage
                oks: List[Any] = [None] * self.world
                dist.all_gather_object(oks, ok, group=self._dist_group)
                if not all(oks):
                    if mc.value:
                        lib.detb200_mc_free(mc.value, mc_ptr.value, 0)
                    logger.info("multicast set-up failed at %s on some rank (%s); using peer-to-peer collectives", stage, err)
                    return None
        self._mcs.append((int(mc.value), int(mc_ptr.value), buf.nbytes))
        return int(mc_ptr.value)

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
        for mc, mc_ptr, nbytes in self._mcs:
            try:
                lib.detb200_mc_free(mc, mc_ptr, nbytes)
            except Exception:
                pass
        self._mcs.clear()
        for p in self._opened_vmm:
            try:
                lib.detb200_vmm_free(p)
            except Exception:
                pass
        self._opened_vmm.clear()
        for p in self._opened:
            try:
                lib.detb200_ipc_close(p)
            except Exception:
                pass
        self._opened.clear()
        if self._fdx is not None:
            self._fdx.close()
            self._fdx = None


def default_comm_ctas() -> int:
    """CTA budget of the in-kernel collectives (0 = the library default, 32): ``DETB200_COMM_CTAS`` overrides."""
    try:
        return max(0, int(os.environ.get("DETB200_COMM_CTAS", "0")))
    except ValueError:
        return 0


def allreduce_fused(group: PeerGroup, peer_bufs: Any, offset: int, numel: int, scale: float,
                    inv_loss_scale: Optional[torch.Tensor] = None, found_inf: Optional[torch.Tensor] = None,
                    stream: Optional[torch.cuda.Stream] = None, phases: int = 7, rank: Optional[int] = None,
                    timeout_s: float = DEFAULT_TIMEOUT_S, mc_ptr: Optional[int] = None, max_ctas: Optional[int] = None) -> None:
    """In-place all-reduce (sum * scale) of ``[offset, offset+numel)`` fp32 elements of the
    symmetric buffers ``peer_bufs``.  Statistics accumulate in ``group.stats_tensor()`` until
    ``group.begin_step()``.  ``mc_ptr`` (from ``group.multicast``) selects the in-switch reduction."""
    lib = _lib.load()
    first = group._first_of_step
    group._first_of_step = False
    _lib.check(lib.detb200_allreduce_fused_ex(peer_bufs, group.ctrl_ptrs, mc_ptr, group.rank if rank is None else rank,
                                              group.world, offset, numel, float(scale), _lib.ptr(inv_loss_scale),
                                              _lib.ptr(found_inf), int(first), phases,
                                              default_comm_ctas() if max_ctas is None else int(max_ctas), float(timeout_s),
                                              _lib.stream_ptr(stream)), "allreduce_fused")


def allreduce_bf16(group: PeerGroup, peer_stage: Any, offset: int, numel: int, scale: float,
                   inv_loss_scale: Optional[torch.Tensor] = None, found_inf: Optional[torch.Tensor] = None,
                   stream: Optional[torch.cuda.Stream] = None, timeout_s: float = DEFAULT_TIMEOUT_S,
                   mc_ptr: Optional[int] = None, max_ctas: Optional[int] = None) -> None:
    """All-reduce of ``[offset, offset+numel)`` bf16 elements of the symmetric *staging* buffers: bf16 on the wire, fp32
    accumulation, statistics from the fp32 sums (reference ``hvd.Compression.fp16``, ``_pytorch_context.py:367``)."""
    lib = _lib.load()
    first = group._first_of_step
    group._first_of_step = False
    _lib.check(lib.detb200_allreduce_bf16(peer_stage, group.ctrl_ptrs, mc_ptr, group.rank, group.world, offset, numel,
                                          float(scale), _lib.ptr(inv_loss_scale), _lib.ptr(found_inf), int(first), 7,
                                          default_comm_ctas() if max_ctas is None else int(max_ctas), float(timeout_s),
                                          _lib.stream_ptr(stream)), "allreduce_bf16")


def narrow_bf16(src: torch.Tensor, dst: torch.Tensor, stream: Optional[torch.cuda.Stream] = None) -> None:
    lib = _lib.load()
    _lib.check(lib.detb200_narrow_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _lib.stream_ptr(stream)), "narrow_bf16")


def widen_bf16(src: torch.Tensor, dst: torch.Tensor, stream: Optional[torch.cuda.Stream] = None) -> None:
    lib = _lib.load()
    _lib.check(lib.detb200_widen_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _lib.stream_ptr(stream)), "widen_bf16")


def broadcast(group: PeerGroup, peer_bufs: Any, offset: int, numel: int, root: int = 0,
              stream: Optional[torch.cuda.Stream] = None, phases: int = 7,
              timeout_s: float = DEFAULT_TIMEOUT_S) -> None:
    lib = _lib.load()
    _lib.check(lib.detb200_broadcast(peer_bufs, group.ctrl_ptrs, group.rank, group.world, root, offset, numel,
                                     phases, float(timeout_s), _lib.stream_ptr(stream)), "broadcast")
