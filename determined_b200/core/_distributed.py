"""``core.DistributedContext``: rank identity + object collectives for the control plane.

Parity: reference ``harness/determined/core/_distributed.py:35-437`` — same constructor
arguments, ``from_horovod/from_deepspeed/from_torch_distributed`` env contracts (``:166-233``),
``gather/gather_local/allgather/allgather_local/broadcast/broadcast_local`` (``:287-407``) and
``DummyDistributedContext`` (``:410-419``).

B200-first redesign: the reference ships pickles over ZeroMQ PUB/SUB+PUSH/PULL.  Here a
``DistributedContext`` owns a *transport*:

``_TorchTransport``
    used when ``torch.distributed`` is initialised.  Objects are pickled into a uint8 tensor and
    moved with ONE size all-gather + ONE padded payload all-gather on a dedicated process group;
    with the NCCL backend both run on a private CUDA side stream so control-plane traffic
    (metrics at report boundaries, preemption flags, searcher op lengths) overlaps training
    kernels instead of serialising with them.
``_SocketTransport``
    a plain-socket star (see :mod:`determined_b200.ipc`) for processes/threads with no
    ``torch.distributed`` (launch layers, CPU tests with one thread per rank).

Both keep the reference's serial-number guard against mismatched collectives.
"""
import logging
import os
import pickle
import socket
import tempfile
import threading
from typing import Any, Callable, Dict, List, Optional, Tuple

from determined_b200 import constants, errors, ipc

logger = logging.getLogger("determined.core")


class _Transport:
    def allgather(self, obj: Any) -> List[Any]:
        raise NotImplementedError

    def gather(self, obj: Any) -> Optional[List[Any]]:
        out = self.allgather(obj)
        return out

    def broadcast(self, obj: Any) -> Any:
        raise NotImplementedError

    def close(self) -> None:
        pass


class _SocketTransport(_Transport):
    def __init__(self, rank: int, size: int, chief_ip: str, port: int, unix_path: Optional[str]) -> None:
        self.rank, self.size = rank, size
        self._hub: Optional[ipc.ObjectHub] = None
        self._spoke: Optional[ipc.ObjectSpoke] = None
        addr: Any = unix_path if unix_path else (chief_ip, port)
        if rank == 0:
            self._hub = ipc.ObjectHub(unix_path if unix_path else port, size - 1)
            self._hub.accept_all()
        else:
            self._spoke = ipc.ObjectSpoke(addr, rank)

    def gather(self, obj: Any) -> Optional[List[Any]]:
        if self._hub is not None:
            got = self._hub.gather()
            got[0] = obj
            return [got[r] for r in range(self.size)]
        assert self._spoke is not None
        self._spoke.send(obj)
        return None

    def broadcast(self, obj: Any) -> Any:
        if self._hub is not None:
            self._hub.broadcast(obj)
            return obj
        assert self._spoke is not None
        return self._spoke.recv()

    def allgather(self, obj: Any) -> List[Any]:
        return self.broadcast(self.gather(obj))

    def close(self) -> None:
        if self._hub is not None:
            self._hub.close()
        if self._spoke is not None:
            self._spoke.close()


class _TorchTransport(_Transport):
    """Object collectives as byte-tensor collectives on a private process group / CUDA stream."""

    def __init__(self) -> None:
        import torch
        import torch.distributed as dist

        self._torch, self._dist = torch, dist
        self.rank, self.size = dist.get_rank(), dist.get_world_size()
        backend = dist.get_backend()
        self._cuda = "nccl" in str(backend) and torch.cuda.is_available()
        # A private group: our collectives can never interleave with the user's on the same
        # communicator (DDP hooks, user all_reduce), which is what deadlocks shared groups.
        self._group = dist.new_group(backend="nccl" if self._cuda else "gloo")
        self._device = torch.device("cuda", torch.cuda.current_device()) if self._cuda else torch.device("cpu")
        self._stream = torch.cuda.Stream(device=self._device) if self._cuda else None
        self._serial = 0

    def _run(self, fn: Callable[[], Any]) -> Any:
        if self._stream is None:
            return fn()
        with self._torch.cuda.stream(self._stream):
            out = fn()
        self._stream.synchronize()  # host needs the bytes; compute streams keep running
        return out

    def allgather(self, obj: Any) -> List[Any]:
        torch, dist = self._torch, self._dist
        self._serial += 1
        payload = pickle.dumps((self._serial, obj), protocol=pickle.HIGHEST_PROTOCOL)

        def go() -> List[Any]:
            n = torch.tensor([len(payload)], dtype=torch.int64, device=self._device)
            sizes = torch.empty(self.size, dtype=torch.int64, device=self._device)
            dist.all_gather_into_tensor(sizes, n, group=self._group)
            sizes_h = sizes.cpu().tolist()
            cap = max(sizes_h)
            buf = torch.zeros(cap, dtype=torch.uint8, device=self._device)
            src = torch.frombuffer(bytearray(payload), dtype=torch.uint8)
            buf[: len(payload)].copy_(src, non_blocking=False)
            out = torch.empty(cap * self.size, dtype=torch.uint8, device=self._device)
            dist.all_gather_into_tensor(out, buf, group=self._group)
            host = out.cpu().numpy().tobytes()
            objs = []
            for r, sz in enumerate(sizes_h):
                serial, o = pickle.loads(host[r * cap: r * cap + sz])
                if serial != self._serial:
                    raise errors.InternalException(
                        f"object-collective serial mismatch: rank {r} is at {serial}, we are at {self._serial}")
                objs.append(o)
            return objs

        return self._run(go)

    def gather(self, obj: Any) -> Optional[List[Any]]:
        out = self.allgather(obj)
        return out if self.rank == 0 else None

    def broadcast(self, obj: Any) -> Any:
        torch, dist = self._torch, self._dist
        self._serial += 1

        def go() -> Any:
            if self.rank == 0:
                payload = pickle.dumps((self._serial, obj), protocol=pickle.HIGHEST_PROTOCOL)
                n = torch.tensor([len(payload)], dtype=torch.int64, device=self._device)
            else:
                payload = b""
                n = torch.zeros(1, dtype=torch.int64, device=self._device)
            dist.broadcast(n, src=dist.get_global_rank(self._group, 0), group=self._group)
            cap = int(n.item())
            if self.rank == 0:
                buf = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(self._device)
            else:
                buf = torch.empty(cap, dtype=torch.uint8, device=self._device)
            dist.broadcast(buf, src=dist.get_global_rank(self._group, 0), group=self._group)
            serial, o = pickle.loads(buf.cpu().numpy().tobytes())
            if serial != self._serial:
                raise errors.InternalException(
                    f"object-broadcast serial mismatch: chief is at {serial}, we are at {self._serial}")
            return o

        return self._run(go)

    def close(self) -> None:
        # Groups die with destroy_process_group(); nothing to tear down eagerly.
        pass


class DistributedContext:
    """Rank identity of this worker plus pickled-object collectives among all workers.

    All collectives must be called by every worker in the same order."""

    def __init__(self, *, rank: int, size: int, local_rank: int, local_size: int, cross_rank: int,
                 cross_size: int, chief_ip: Optional[str] = None,
                 pub_port: int = constants.INTER_TRAIN_PROCESS_COMM_PORT_1,
                 pull_port: int = constants.INTER_TRAIN_PROCESS_COMM_PORT_2,
                 port_offset: int = 0, force_tcp: bool = False,
                 _transport: Optional[str] = None) -> None:
        if size < 1 or not (0 <= rank < size):
            raise ValueError(f"invalid rank/size: {rank}/{size}")
        if local_size < 1 or not (0 <= local_rank < local_size):
            raise ValueError(f"invalid local_rank/local_size: {local_rank}/{local_size}")
        if cross_size < 1 or not (0 <= cross_rank < cross_size):
            raise ValueError(f"invalid cross_rank/cross_size: {cross_rank}/{cross_size}")
        self.rank, self.size = rank, size
        self.local_rank, self.local_size = local_rank, local_size
        self.cross_rank, self.cross_size = cross_rank, cross_size
        self._is_chief = rank == 0
        self._is_local_chief = local_rank == 0
        self._closed = False
        self._transport: Optional[_Transport] = None
        if size == 1:
            return
        kind = _transport
        if kind is None:
            kind = "torch" if _torch_dist_ready(size) else "socket"
            if kind == "socket" and cross_size == 1 and not force_tcp and os.environ.get("DET_SHM_COLLECTIVES", "1") != "0":
                from determined_b200.core import _shm_transport

                if _shm_transport.available():
                    kind = "shm"
        if kind == "torch":
            self._transport = _TorchTransport()
        elif kind == "shm":
            # all ranks on this box: one shared-memory segment instead of a socket star through the chief
            from determined_b200.core import _shm_transport

            tag = os.environ.get("DET_ALLOCATION_ID", "") or str(os.getppid())
            self._transport = _shm_transport.ShmTransport(rank, size, f"{tag}-{pub_port + port_offset}")  # type: ignore[assignment]
        else:
            if chief_ip is None:
                raise ValueError("chief_ip is required when size > 1 and torch.distributed is not initialized")
            port = pub_port + port_offset
            unix_path = None
            if cross_size == 1 and not force_tcp:
                tag = os.environ.get("DET_ALLOCATION_ID", "") or str(os.getppid())
                unix_path = os.path.join(tempfile.gettempdir(), f"detb200-{tag}-{port}.sock")
            self._transport = _SocketTransport(rank, size, chief_ip, port, unix_path)

    # -- alternate constructors (env contracts identical to the reference) ---------------------
    @classmethod
    def from_horovod(cls, hvd: Any, chief_ip: Optional[str] = None) -> "DistributedContext":
        """Horovod is not part of the B200 stack; accepted for API compatibility with any object
        exposing ``rank()/size()/local_rank()/local_size()/cross_rank()/cross_size()``."""
        return cls(rank=hvd.rank(), size=hvd.size(), local_rank=hvd.local_rank(), local_size=hvd.local_size(),
                   cross_rank=hvd.cross_rank(), cross_size=hvd.cross_size(),
                   chief_ip=chief_ip or os.environ.get("DET_CHIEF_IP"))

    @classmethod
    def from_deepspeed(cls, chief_ip: Optional[str] = None) -> "DistributedContext":
        """Reads ``RANK, WORLD_SIZE, LOCAL_RANK, LOCAL_SIZE, CROSS_RANK, CROSS_SIZE``
        (reference ``:196-213``)."""
        env = os.environ
        need = ["RANK", "WORLD_SIZE", "LOCAL_RANK"]
        missing = [k for k in need if k not in env]
        if missing:
            raise ValueError(f"from_deepspeed requires env {missing}; was this started by the deepspeed launcher?")
        rank, size, lrank = int(env["RANK"]), int(env["WORLD_SIZE"]), int(env["LOCAL_RANK"])
        lsize = int(env.get("LOCAL_SIZE", env.get("LOCAL_WORLD_SIZE", size)))
        return cls(rank=rank, size=size, local_rank=lrank, local_size=lsize,
                   cross_rank=int(env.get("CROSS_RANK", rank // lsize)),
                   cross_size=int(env.get("CROSS_SIZE", max(size // lsize, 1))),
                   chief_ip=chief_ip or env.get("DET_CHIEF_IP") or env.get("MASTER_ADDR"))

    @classmethod
    def from_torch_distributed(cls, chief_ip: Optional[str] = None) -> "DistributedContext":
        """Reads ``RANK, WORLD_SIZE, LOCAL_RANK, LOCAL_WORLD_SIZE, GROUP_RANK, GROUP_WORLD_SIZE``
        set by ``torch.distributed.run`` (reference ``:216-233``)."""
        env = os.environ
        missing = [k for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK") if k not in env]
        if missing:
            raise ValueError(f"from_torch_distributed requires env {missing}; was this started by torchrun?")
        rank, size, lrank = int(env["RANK"]), int(env["WORLD_SIZE"]), int(env["LOCAL_RANK"])
        lsize = int(env.get("LOCAL_WORLD_SIZE", size))
        return cls(rank=rank, size=size, local_rank=lrank, local_size=lsize,
                   cross_rank=int(env.get("GROUP_RANK", rank // lsize)),
                   cross_size=int(env.get("GROUP_WORLD_SIZE", max(size // lsize, 1))),
                   chief_ip=chief_ip or env.get("DET_CHIEF_IP") or env.get("MASTER_ADDR"))

    # -- lifecycle ---------------------------------------------------------------------------------
    def close(self) -> None:
        if self._closed:
            return
        self._closed = True
        if self._transport is not None:
            self._transport.close()
            self._transport = None

    def __enter__(self) -> "DistributedContext":
        return self

    def __exit__(self, *_: Any) -> None:
        self.close()

    # -- identity ----------------------------------------------------------------------------------
    def get_rank(self) -> int:
        return self.rank

    def get_local_rank(self) -> int:
        return self.local_rank

    def get_size(self) -> int:
        return self.size

    def get_num_agents(self) -> int:
        return self.cross_size

    def get_cross_rank(self) -> int:
        return self.cross_rank

    def get_cross_size(self) -> int:
        return self.cross_size

    def get_local_size(self) -> int:
        return self.local_size

    # -- collectives -------------------------------------------------------------------------------
    def _node_of(self, rank: int) -> int:
        return rank // self.local_size

    def gather(self, stuff: Any) -> Optional[List]:
        """Chief gets ``[obj_rank0, obj_rank1, …]``; everyone else gets None."""
        if self.size < 2:
            return [stuff]
        assert self._transport is not None, "DistributedContext is closed"
        return self._transport.gather(stuff)

    def gather_local(self, stuff: Any) -> Optional[List]:
        """Like gather, but among the workers of this node; the local chief gets the list."""
        if self.local_size < 2:
            return [stuff]
        if self.cross_size == 1:
            return self.gather(stuff)
        everything = self.allgather(stuff)
        if not self._is_local_chief:
            return None
        return [o for r, o in enumerate(everything) if self._node_of(r) == self.cross_rank]

    def allgather(self, stuff: Any) -> List:
        if self.size < 2:
            return [stuff]
        assert self._transport is not None, "DistributedContext is closed"
        return self._transport.allgather(stuff)

    def allgather_local(self, stuff: Any) -> List:
        if self.local_size < 2:
            return [stuff]
        everything = self.allgather(stuff)
        return [o for r, o in enumerate(everything) if self._node_of(r) == self.cross_rank]

    def broadcast(self, stuff: Any) -> Any:
        """Every worker returns the chief's ``stuff``."""
        if self.size < 2:
            return stuff
        assert self._transport is not None, "DistributedContext is closed"
        return self._transport.broadcast(stuff)

    def broadcast_local(self, stuff: Any = None) -> Any:
        """Every worker returns its *local chief's* ``stuff``."""
        if self.local_size < 2:
            return stuff
        if self.cross_size == 1:
            return self.broadcast(stuff)
        everything = self.allgather(stuff if self._is_local_chief else None)
        return everything[self.cross_rank * self.local_size]


def _torch_dist_ready(size: int) -> bool:
    try:
        import torch.distributed as dist
    except ImportError:
        return False
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() == size


class DummyDistributedContext(DistributedContext):
    """Single-worker context for off-cluster / single-slot runs."""

    def __init__(self) -> None:
        super().__init__(rank=0, size=1, local_rank=0, local_size=1, cross_rank=0, cross_size=1)
