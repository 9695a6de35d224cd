"""RFC 6455 websockets on plain sockets, both ends, no third-party package.

Used by the streaming-updates channel (``GET /stream`` upgraded; reference ``master/internal/stream`` serves a websocket and
``harness/determined/common/streams/_client.py`` reads it with ``lomond``) and by the master's reverse proxy, which tunnels
upgraded connections to task services (notebook kernels, tensorboard).

What is implemented: the opening handshake (client and server side, ``Sec-WebSocket-Accept``), text / binary / close / ping /
pong frames, fragmentation on receive, client-side masking, 7 / 16 / 64-bit lengths.  No extensions (no per-message
compression), no subprotocol negotiation.
"""
import base64
import hashlib
import json
import os
import socket
import ssl
import struct
import threading
import urllib.parse
from typing import Any, Dict, Optional, Tuple

GUID = "258EAFA5-E914-47DA-95CA-C5AB0DC85B11"
OP_CONT, OP_TEXT, OP_BINARY, OP_CLOSE, OP_PING, OP_PONG = 0x0, 0x1, 0x2, 0x8, 0x9, 0xA
MAX_MESSAGE = 64 << 20


class WebSocketError(Exception):
    pass


class ConnectionClosed(WebSocketError):
    def __init__(self, code: int = 1006, reason: str = "") -> None:
        super().__init__(f"websocket closed ({code}{': ' + reason if reason else ''})")
        self.code, self.reason = code, reason


def accept_key(key: str) -> str:
    return base64.b64encode(hashlib.sha1((key + GUID).encode()).digest()).decode()


def is_upgrade(headers: Any) -> bool:
    """Does this request (an ``http.server`` / ``email.message`` header object, or a dict) ask for a websocket?"""
    return "upgrade" in str(headers.get("Connection", "")).lower() and str(headers.get("Upgrade", "")).lower() == "websocket"


class WebSocket:
    """One established connection.  ``send_*`` may be called from several threads; ``recv`` from one."""

    def __init__(self, sock: Any, is_client: bool, rfile: Any = None) -> None:
        self.sock = sock
        self.is_client = is_client
        self._rfile = rfile if rfile is not None else sock.makefile("rb")
        self._wlock = threading.Lock()
        self.closed = False

    # -- sending ---------------------------------------------------------------------------------------------------------------
    def _send_frame(self, opcode: int, payload: bytes) -> None:
        head = bytes([0x80 | opcode])
        n = len(payload)
        mask_bit = 0x80 if self.is_client else 0
        if n < 126:
            head += bytes([mask_bit | n])
        elif n < 1 << 16:
            head += bytes([mask_bit | 126]) + struct.pack(">H", n)
        else:
            head += bytes([mask_bit | 127]) + struct.pack(">Q", n)
        if self.is_client:
            mask = os.urandom(4)
            payload = bytes(b ^ mask[i & 3] for i, b in enumerate(payload)) if n < 4096 else _mask_fast(payload, mask)
            head += mask
        with self._wlock:
            if self.closed and opcode != OP_CLOSE:
                raise ConnectionClosed()
            self.sock.sendall(head + payload)

    def send_text(self, text: str) -> None:
        self._send_frame(OP_TEXT, text.encode())

    def send_json(self, obj: Any) -> None:
        self.send_text(json.dumps(obj, default=str))

    def send_binary(self, data: bytes) -> None:
        self._send_frame(OP_BINARY, data)

    def ping(self, data: bytes = b"") -> None:
        self._send_frame(OP_PING, data)

    def close(self, code: int = 1000, reason: str = "") -> None:
        if self.closed:
            return
        try:
            self._send_frame(OP_CLOSE, struct.pack(">H", code) + reason.encode()[:120])
        except OSError:
            pass
        self.closed = True
        try:
            self.sock.shutdown(socket.SHUT_RDWR)
        except OSError:
            pass
        try:
            self.sock.close()
        except OSError:
            pass

    # -- receiving -------------------------------------------------------------------------------------------------------------
    def _read(self, n: int) -> bytes:
        data = self._rfile.read(n)
        if data is None or len(data) < n:
            self.closed = True
            raise ConnectionClosed(1006, "connection lost")
        return data

    def _recv_frame(self) -> Tuple[bool, int, bytes]:
        b0, b1 = self._read(2)
        fin, opcode = bool(b0 & 0x80), b0 & 0x0F
        if b0 & 0x70:
            raise WebSocketError("reserved bits set (extensions are not supported)")
        masked, n = bool(b1 & 0x80), b1 & 0x7F
        if n == 126:
            n = struct.unpack(">H", self._read(2))[0]
        elif n == 127:
            n = struct.unpack(">Q", self._read(8))[0]
        if n > MAX_MESSAGE:
            raise WebSocketError(f"frame of {n} bytes exceeds the limit")
        if not self.is_client and not masked:
            raise WebSocketError("client frames must be masked")
        mask = self._read(4) if masked else None
        payload = self._read(n) if n else b""
        if mask is not None:
            payload = _mask_fast(payload, mask)
        return fin, opcode, payload

    def recv(self, timeout: Optional[float] = None) -> Tuple[int, bytes]:
        """Next data message as ``(opcode, payload)`` (``OP_TEXT`` / ``OP_BINARY``); pings are answered, pongs dropped.
        Raises ``ConnectionClosed`` on a close frame or a lost connection, ``socket.timeout`` / ``TimeoutError`` when
        ``timeout`` seconds pass without a complete message."""
        self.sock.settimeout(timeout)
        buf, first = b"", None
        while True:
            fin, opcode, payload = self._recv_frame()
            if opcode == OP_PING:
                self._send_frame(OP_PONG, payload)
                continue
            if opcode == OP_PONG:
                continue
            if opcode == OP_CLOSE:
                code = struct.unpack(">H", payload[:2])[0] if len(payload) >= 2 else 1005
                reason = payload[2:].decode(errors="replace")
                if not self.closed:
                    try:
                        self._send_frame(OP_CLOSE, payload[:2])
                    except OSError:
                        pass
                self.closed = True
                raise ConnectionClosed(code, reason)
            if opcode in (OP_TEXT, OP_BINARY):
                first, buf = opcode, payload
            elif opcode == OP_CONT and first is not None:
                buf += payload
                if len(buf) > MAX_MESSAGE:
                    raise WebSocketError("message exceeds the limit")
            else:
                raise WebSocketError(f"unexpected opcode {opcode}")
            if fin:
                return first, buf      # type: ignore[return-value]

    def recv_json(self, timeout: Optional[float] = None) -> Any:
        _, data = self.recv(timeout)
        return json.loads(data.decode())


def _mask_fast(data: bytes, mask: bytes) -> bytes:
    n = len(data)
    if n == 0:
        return data
    reps = (n + 3) // 4
    return (int.from_bytes(data, "big") ^ int.from_bytes((mask * reps)[:n], "big")).to_bytes(n, "big")


# ---------------------------------------------------------------------------------------------------------------------------
def connect(url: str, headers: Optional[Dict[str, str]] = None, ssl_context: Optional[ssl.SSLContext] = None,
            server_hostname: Optional[str] = None, timeout: float = 30.0) -> WebSocket:
    """Client side: ``ws://`` / ``wss://`` (``http`` / ``https`` are accepted as synonyms)."""
    u = urllib.parse.urlparse(url)
    secure = u.scheme in ("wss", "https")
    host, port = u.hostname, u.port or (443 if secure else 80)
    assert host
    raw = socket.create_connection((host, port), timeout=timeout)
    sock: Any = raw
    if secure:
        ctx = ssl_context or ssl.create_default_context()
        sock = ctx.wrap_socket(raw, server_hostname=server_hostname or host)
    key = base64.b64encode(os.urandom(16)).decode()
    path = (u.path or "/") + ("?" + u.query if u.query else "")
    lines = [f"GET {path} HTTP/1.1", f"Host: {host}:{port}", "Upgrade: websocket", "Connection: Upgrade",
             f"Sec-WebSocket-Key: {key}", "Sec-WebSocket-Version: 13"]
    for k, v in (headers or {}).items():
        lines.append(f"{k}: {v}")
    sock.sendall(("\r\n".join(lines) + "\r\n\r\n").encode())
    rfile = sock.makefile("rb")
    status = rfile.readline().decode(errors="replace").strip()
    got: Dict[str, str] = {}
    while True:
        line = rfile.readline().decode(errors="replace")
        if line in ("\r\n", "\n", ""):
            break
        k, _, v = line.partition(":")
        got[k.strip().lower()] = v.strip()
    parts = status.split(" ", 2)
    if len(parts) < 2 or parts[1] != "101":
        body = b""
        try:
            n = int(got.get("content-length") or 0)
            body = rfile.read(n) if n else b""
        except (ValueError, OSError):
            pass
        sock.close()
        raise WebSocketError(f"server refused the upgrade: {status} {body[:200].decode(errors='replace')}")
    if got.get("sec-websocket-accept") != accept_key(key):
        sock.close()
        raise WebSocketError("bad Sec-WebSocket-Accept in the server's answer")
    return WebSocket(sock, is_client=True, rfile=rfile)


def server_handshake(handler: Any) -> WebSocket:
    """Server side, inside an ``http.server.BaseHTTPRequestHandler`` method: answer 101 and take the connection over."""
    key = handler.headers.get("Sec-WebSocket-Key")
    if not key or not is_upgrade(handler.headers):
        raise WebSocketError("not a websocket upgrade request")
    handler.send_response(101, "Switching Protocols")
    handler.send_header("Upgrade", "websocket")
    handler.send_header("Connection", "Upgrade")
    handler.send_header("Sec-WebSocket-Accept", accept_key(key))
    handler.end_headers()
    handler.wfile.flush()
    handler.close_connection = True
    return WebSocket(handler.connection, is_client=False, rfile=handler.rfile)
