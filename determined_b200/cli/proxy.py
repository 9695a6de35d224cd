"""TCP through the master: a websocket to ``/tunnel/<task id>[/<port>]`` carries the bytes of one TCP connection to a port of
a running task (its service port, or one it declared in ``proxy_ports``).

Parity: reference ``harness/determined/cli/proxy.py`` (``http_connect_tunnel`` for ``ssh -o ProxyCommand``,
``http_tunnel_listener`` + ``ListenerConfig`` for ``-p LOCAL:REMOTE`` port forwarding).  Binary frames both ways; closing
either side closes the other.
"""
import contextlib
import dataclasses
import socket
import sys
import threading
from typing import Any, Iterator, List, Optional

from determined_b200.common import websocket


@dataclasses.dataclass
class ListenerConfig:
    service_id: str
    local_port: int
    local_addr: str = "127.0.0.1"
    remote_port: Optional[int] = None          # default: the task's service port


def _open(sess: Any, service_id: str, remote_port: Optional[int]) -> websocket.WebSocket:
    url = sess.master.rstrip("/") + f"/tunnel/{service_id}" + (f"/{int(remote_port)}" if remote_port else "")
    headers = {"Authorization": f"Bearer {sess.token}"} if getattr(sess, "token", None) else {}
    cert = getattr(sess, "cert", None)
    ctx = cert.ssl_context() if cert is not None and url.startswith("https") else None
    return websocket.connect(url, headers=headers, ssl_context=ctx, server_hostname=getattr(cert, "name", None))


def _ws_to(ws: websocket.WebSocket, write: Any, done: threading.Event) -> None:
    try:
        while not done.is_set():
            try:
                _, data = ws.recv(timeout=1.0)
            except (socket.timeout, TimeoutError):
                continue
            if data:
                write(data)
    except (websocket.WebSocketError, OSError):
        pass
    finally:
        done.set()


def copy_socket(ws: websocket.WebSocket, conn: socket.socket) -> None:
    """Shuttle one accepted TCP connection over ``ws`` until either side ends."""
    done = threading.Event()
    lock = threading.Lock()
    t = threading.Thread(target=_ws_to, args=(ws, conn.sendall, done), daemon=True)
    t.start()
    conn.settimeout(1.0)
    try:
        while not done.is_set():
            try:
                data = conn.recv(65536)
            except socket.timeout:
                continue
            if not data:
                break
            with lock:
                ws.send_binary(data)
    except (OSError, websocket.WebSocketError):
        pass
    finally:
        done.set()
        with contextlib.suppress(Exception):
            ws.close()
        with contextlib.suppress(OSError):
            conn.close()
        t.join(timeout=2)


def http_connect_tunnel(sess: Any, service_id: str, remote_port: Optional[int] = None) -> None:
    """stdin / stdout are the two directions of the connection (what ``ssh -o ProxyCommand=...`` wants)."""
    ws = _open(sess, service_id, remote_port)
    done = threading.Event()

    def out(data: bytes) -> None:
        sys.stdout.buffer.write(data)
        sys.stdout.buffer.flush()

    t = threading.Thread(target=_ws_to, args=(ws, out, done), daemon=True)
    t.start()
    try:
        while not done.is_set():
            data = sys.stdin.buffer.read1(65536) if hasattr(sys.stdin.buffer, "read1") else sys.stdin.buffer.read(65536)
            if not data:
                break
            ws.send_binary(data)
    except (OSError, websocket.WebSocketError):
        pass
    finally:
        done.set()
        with contextlib.suppress(Exception):
            ws.close()


@contextlib.contextmanager
def http_tunnel_listener(sess: Any, tunnels: List[ListenerConfig]) -> Iterator[List[int]]:
    """Listen on each ``local_addr:local_port`` (0 picks a free port) and forward every connection to its task; yields the
    ports actually bound."""
    stop = threading.Event()
    socks: List[socket.socket] = []

    def serve(ls: socket.socket, cfg: ListenerConfig) -> None:
        ls.settimeout(0.5)
        while not stop.is_set():
            try:
                conn, _ = ls.accept()
            except socket.timeout:
                continue
            except OSError:
                return
            try:
                ws = _open(sess, cfg.service_id, cfg.remote_port)
            except Exception as e:  # noqa: BLE001
                print(f"tunnel to {cfg.service_id} failed: {e}", file=sys.stderr)
                conn.close()
                continue
            threading.Thread(target=copy_socket, args=(ws, conn), daemon=True).start()

    try:
        for cfg in tunnels:
            ls = socket.socket()
            ls.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
            ls.bind((cfg.local_addr, cfg.local_port))
            ls.listen(16)
            socks.append(ls)
            threading.Thread(target=serve, args=(ls, cfg), daemon=True).start()
        yield [s.getsockname()[1] for s in socks]
    finally:
        stop.set()
        for s in socks:
            with contextlib.suppress(OSError):
                s.close()
