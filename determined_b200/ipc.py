"""Host-side IPC: socket object-collectives (the CPU fallback transport of
``core.DistributedContext``) and the worker-liveness watchdog (PID server/client).

Parity: reference ``harness/determined/ipc.py`` — ZMQ broadcast server/client ``:34-245`` and
``PIDServer``/``PIDClient`` ``:272-562``.  This build has no ZeroMQ dependency on the product path:
  * tensor-adjacent control-plane collectives ride NCCL byte tensors (``core/_distributed.py``);
  * this module is the plain-socket star topology used when ``torch.distributed`` is not
    initialised (launch layers, CPU tests with one *thread* per rank).
The protocol keeps the reference's protective measure: every message carries a monotonically
increasing serial number checked on both sides (reference ``ipc.py:23-31,157-172``).
The PID protocol is byte-compatible with the reference: ``"<pid>\\n"`` then any number of ``k``
keep-alives then ``q`` on graceful exit.
"""
import contextlib
import logging
import os
import pickle
import selectors
import signal
import socket
import struct
import subprocess
import sys
import threading
import time
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

from determined_b200 import errors

logger = logging.getLogger("determined.ipc")

_HDR = struct.Struct("!QQ")  # serial, payload length
Addr = Union[str, int, Tuple[str, int]]


def _send_frame(sock: socket.socket, serial: int, payload: bytes) -> None:
    sock.sendall(_HDR.pack(serial, len(payload)) + payload)


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray()
    while len(buf) < n:
        chunk = sock.recv(min(n - len(buf), 1 << 20))
        if not chunk:
            raise ConnectionError("peer closed the connection mid-message")
        buf += chunk
    return bytes(buf)


def _recv_frame(sock: socket.socket) -> Tuple[int, bytes]:
    serial, n = _HDR.unpack(_recv_exact(sock, _HDR.size))
    return serial, _recv_exact(sock, n)


def _listen(addr: Addr, backlog: int) -> socket.socket:
    if isinstance(addr, str):
        s = socket.socket(socket.AF_UNIX)
        with contextlib.suppress(FileNotFoundError):
            os.remove(addr)
        s.bind(addr)
    else:
        s = socket.socket()
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        s.bind(("", addr) if isinstance(addr, int) else addr)
    s.listen(max(backlog, 1))
    return s


def _connect(addr: Addr, timeout: float = 120.0) -> socket.socket:
    deadline = time.monotonic() + timeout
    delay = 0.01
    while True:
        try:
            if isinstance(addr, str):
                s = socket.socket(socket.AF_UNIX)
                s.connect(addr)
            else:
                s = socket.create_connection(("127.0.0.1", addr) if isinstance(addr, int) else addr, timeout=10)
                s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
                s.settimeout(None)
            return s
        except (ConnectionRefusedError, FileNotFoundError, socket.timeout, OSError):
            if time.monotonic() > deadline:
                raise
            time.sleep(delay)
            delay = min(delay * 2, 0.5)


class ObjectHub:
    """Chief side of the star: accepts ``num_spokes`` connections, then ``gather``/``broadcast``."""

    def __init__(self, addr: Addr, num_spokes: int) -> None:
        self._listener = _listen(addr, num_spokes)
        self.num_spokes = num_spokes
        self._conns: Dict[int, socket.socket] = {}
        self._send_serial = 0
        self._recv_serial = 0

    @property
    def port(self) -> Optional[int]:
        name = self._listener.getsockname()
        return name[1] if isinstance(name, tuple) else None

    def accept_all(self, timeout: float = 600.0) -> None:
        self._listener.settimeout(timeout)
        while len(self._conns) < self.num_spokes:
            conn, _ = self._listener.accept()
            if conn.family != socket.AF_UNIX:
                conn.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            _, payload = _recv_frame(conn)
            rank = pickle.loads(payload)
            if rank in self._conns:
                raise errors.InternalException(f"duplicate spoke rank {rank}")
            self._conns[rank] = conn
        # handshake so that no spoke races ahead of a peer that has not connected yet
        for c in self._conns.values():
            _send_frame(c, 0, b"ok")

    def gather(self) -> Dict[int, Any]:
        self._recv_serial += 1
        out = {}
        for rank, c in self._conns.items():
            serial, payload = _recv_frame(c)
            if serial != self._recv_serial:
                raise errors.InternalException(
                    f"out-of-order gather from rank {rank}: got serial {serial}, expected {self._recv_serial}")
            out[rank] = pickle.loads(payload)
        return out

    def broadcast(self, obj: Any) -> None:
        self._send_serial += 1
        payload = pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)
        for c in self._conns.values():
            _send_frame(c, self._send_serial, payload)

    def close(self) -> None:
        for c in self._conns.values():
            with contextlib.suppress(OSError):
                c.close()
        self._conns.clear()
        with contextlib.suppress(OSError):
            name = self._listener.getsockname()
            self._listener.close()
            if isinstance(name, str) and name:
                with contextlib.suppress(FileNotFoundError):
                    os.remove(name)


class ObjectSpoke:
    """Worker side of the star."""

    def __init__(self, addr: Addr, rank: int, timeout: float = 600.0) -> None:
        self._sock = _connect(addr, timeout)
        self._send_serial = 0
        self._recv_serial = 0
        _send_frame(self._sock, 0, pickle.dumps(rank))
        _recv_frame(self._sock)  # "ok" once everybody is connected

    def send(self, obj: Any) -> None:
        self._send_serial += 1
        _send_frame(self._sock, self._send_serial, pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL))

    def recv(self) -> Any:
        self._recv_serial += 1
        serial, payload = _recv_frame(self._sock)
        if serial != self._recv_serial:
            raise errors.InternalException(
                f"out-of-order broadcast: got serial {serial}, expected {self._recv_serial}")
        return pickle.loads(payload)

    def close(self) -> None:
        with contextlib.suppress(OSError):
            self._sock.close()


# ---------------------------------------------------------------------------------------------
# Worker-liveness watchdog
# ---------------------------------------------------------------------------------------------

def _pid_alive(pid: int) -> bool:
    try:
        import psutil

        try:
            return psutil.Process(pid).status() not in (psutil.STATUS_DEAD, psutil.STATUS_ZOMBIE, psutil.STATUS_STOPPED)
        except psutil.NoSuchProcess:
            return False
    except ImportError:
        try:
            os.kill(pid, 0)
            return True
        except ProcessLookupError:
            return False
        except PermissionError:
            return True


class PIDServer:
    """Tracks PIDs reported by :class:`PIDClient`\\ s.  ``run()`` returns when all clients reported
    a graceful shutdown, raises :class:`WorkerError` if any pid vanishes without one.

    On Linux ≥5.3 the server additionally holds a ``pidfd`` per worker (``os.pidfd_open``) so a
    death is observed as a readable fd in the same selector loop rather than by polling."""

    def __init__(self, addr: Addr, num_clients: int) -> None:
        self.addr = addr
        self.num_clients = num_clients
        self.sel: Optional[selectors.BaseSelector] = None
        self.listener: Optional[socket.socket] = None
        self.pids: List[int] = []
        self.graceful_shutdowns: List[int] = []
        self.conns: Dict[socket.socket, int] = {}
        self._pidfds: Dict[int, int] = {}
        self.started = False

    def start(self) -> "PIDServer":
        if self.started:
            return self
        self.started = True
        try:
            self.sel = selectors.DefaultSelector()
            self.listener = _listen(self.addr, self.num_clients)
            self.listener.setblocking(False)
            self.sel.register(self.listener, selectors.EVENT_READ, "listener")
        except Exception:
            self.close()
            raise
        return self

    @property
    def port(self) -> Optional[int]:
        assert self.listener is not None
        name = self.listener.getsockname()
        return name[1] if isinstance(name, tuple) else None

    def close(self) -> None:
        self.started = False
        for fd in self._pidfds.values():
            with contextlib.suppress(OSError):
                os.close(fd)
        self._pidfds.clear()
        if self.listener:
            self.listener.close()
            self.listener = None
        if self.sel:
            self.sel.close()
            self.sel = None

    def __enter__(self) -> "PIDServer":
        return self.start()

    def __exit__(self, *_: Any) -> None:
        self.close()

    def _accept(self) -> None:
        assert self.sel and self.listener
        conn, _ = self.listener.accept()
        conn.setblocking(True)
        buf = b""
        while b"\n" not in buf:
            data = conn.recv(4096)
            if not data:
                raise ValueError("pid_client did not deliver a PID!")
            buf += data
        head, rest = buf.split(b"\n", 1)
        pid = int(head)
        self.pids.append(pid)
        self.conns[conn] = pid
        if hasattr(os, "pidfd_open"):
            with contextlib.suppress(OSError):
                self._pidfds[pid] = os.pidfd_open(pid)
        conn.setblocking(False)
        self.sel.register(conn, selectors.EVENT_READ, "conn")
        if len(self.pids) == self.num_clients:
            self.sel.unregister(self.listener)
            self.listener.close()
            self.listener = None
        if rest:
            self._on_conn(conn, rest)

    def _on_conn(self, conn: socket.socket, data: Optional[bytes] = None) -> None:
        assert self.sel
        pid = self.conns[conn]
        if data is None:
            try:
                data = conn.recv(4096)
            except (BlockingIOError, ConnectionResetError):
                data = b""
        if data:
            last = data[-1:]
            if last == b"k":
                return
            if last == b"q":
                self.graceful_shutdowns.append(pid)
            else:
                raise ValueError(f"invalid message from pid_client: {data!r}")
        if self.listener is not None:
            raise errors.WorkerError("worker died before all workers connected")
        self.sel.unregister(conn)
        conn.close()
        del self.conns[conn]
        if pid not in self.graceful_shutdowns:
            # the connection is gone without the goodbye byte: the worker crashed (its pid may linger a moment as a zombie,
            # so do not leave this to the pid poll, which stops once every connection has closed)
            raise errors.WorkerError(f"Detected that worker process (pid {pid}) exited without a graceful exit")

    def check_pids(self) -> None:
        for pid in self.pids:
            if pid in self.graceful_shutdowns:
                continue
            if not _pid_alive(pid):
                raise errors.WorkerError(f"Detected that worker process (pid {pid}) died without a graceful exit")

    def run(self, health_check: Optional[Callable[[], None]] = None, poll_period: float = 1.0) -> None:
        self.start()
        assert self.sel
        while self.listener is not None or self.conns:
            for key, _ in self.sel.select(timeout=poll_period):
                if key.data == "listener":
                    self._accept()
                else:
                    self._on_conn(key.fileobj)  # type: ignore
            self.check_pids()
            if health_check is not None:
                health_check()

    def run_subprocess(self, cmd: List[str], on_fail: Optional[signal.Signals] = None,
                       on_exit: Optional[signal.Signals] = None, grace_period: int = 3,
                       signal_children: bool = False, env: Optional[Dict[str, str]] = None) -> int:
        """Run ``cmd`` (the launcher) while watching the workers and return an exit code with the reference's meaning
        (``ipc.py:451-502``): the launcher's own code if it ends first (77 if that was 0: it must outlive its workers); after
        a worker failure the launcher is given ``grace_period`` seconds, then ``on_fail`` (78 if it then exits 0), else
        waited for (79 if it exits 0); after a clean run ``on_exit`` is delivered the same way."""
        from determined_b200 import util

        p = subprocess.Popen(cmd, env=env)

        class HealthCheckFail(Exception):
            def __init__(self, code: int):
                self.code = code

        def health_check() -> None:
            ret = p.poll()
            if ret is not None:
                raise HealthCheckFail(ret)

        def send(sig: signal.Signals) -> None:
            if signal_children:
                util.signal_process_tree(p, sig)
            with contextlib.suppress(ProcessLookupError):
                p.send_signal(sig)

        def deliver(sig: signal.Signals, zero_means: int) -> Optional[int]:
            time.sleep(grace_period)            # let things finish logging and exit on their own first
            send(sig)
            if sig != signal.SIGKILL:
                try:
                    return p.wait(timeout=10) or zero_means
                except subprocess.TimeoutExpired:
                    logger.error("launcher ignored %s; sending SIGKILL", sig.name)
                    send(signal.SIGKILL)
            return None

        with util.forward_signals(p):
            try:
                self.run(health_check)
            except HealthCheckFail as e:
                return e.code or 77
            except errors.WorkerError as e:
                logger.error("%s", e)
                if on_fail is not None:
                    rc = deliver(on_fail, 78)
                    if rc is not None:
                        return rc
                return p.wait() or 79
            if on_exit is not None:
                rc = deliver(on_exit, 0)
                if rc is not None:
                    return rc
            return p.wait()


class PIDClient:
    def __init__(self, addr: Addr) -> None:
        self.addr = addr
        self.sock: Optional[socket.socket] = None

    def start(self) -> "PIDClient":
        self.sock = _connect(self.addr, timeout=60)
        self.sock.sendall(b"%d\n" % os.getpid())
        return self

    def close(self, graceful: bool = False) -> None:
        if self.sock is None:
            return
        with contextlib.suppress(OSError):
            if graceful:
                self.sock.sendall(b"q")
            self.sock.close()
        self.sock = None

    def keep_alive(self) -> None:
        if self.sock is None:
            raise ValueError("PIDClient must be started first")
        self.sock.sendall(b"k")

    def __enter__(self) -> "PIDClient":
        return self.start()

    def __exit__(self, e_type: Any, e_val: Any, _: Any) -> None:
        graceful = e_type is None or (isinstance(e_val, SystemExit) and e_val.code in (0, None))
        self.close(graceful=graceful)

    def run_subprocess(self, cmd: List[str]) -> int:
        from determined_b200 import util

        p = subprocess.Popen(cmd)
        with util.forward_signals(p):
            return p.wait()
