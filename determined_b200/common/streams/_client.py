"""Streaming-updates client.

Parity: reference ``harness/determined/common/streams/_client.py``: ``Stream`` (subscribe / iterate,
``Sync(complete=False)`` ... offline messages ... ``Sync(complete=True)`` ... online messages, automatic
resubscribe from a primary-key cache), ``ProjectSpec`` / ``ModelSpec`` / ``ModelVersionSpec``,
``KeyCache``, ``StreamWebSocket`` transport ABC.  Two transports: ``WebSocketStreamTransport`` speaks a websocket to
``GET /stream`` like the reference's lomond client (own RFC 6455 code in ``common/websocket.py``; TLS through the session's
``Cert``; reconnects and re-sends the last startup message, which the ``Stream`` builds from its key cache so that nothing is
missed and nothing is repeated), and ``PollingStreamTransport`` runs the same messages over ``POST /stream`` for places a
websocket cannot go.  ``Stream.connect(session)`` picks the websocket.
"""
import abc
import collections
import time
import uuid
from typing import Any, Deque, Dict, Iterator, List, Optional, Sequence, Set, Tuple, Union

from determined_b200.common import api
from determined_b200.common.streams import _util, wire


class StreamWebSocket(metaclass=abc.ABCMeta):
    @abc.abstractmethod
    def send(self, msg: Dict[str, Any]) -> None:
        ...

    @abc.abstractmethod
    def recv(self, timeout: Optional[float] = None) -> List[Dict[str, Any]]:
        """Return zero or more server messages (blocking up to ``timeout``)."""

    def close(self) -> None:
        pass


class PollingStreamTransport(StreamWebSocket):
    def __init__(self, session: api.Session, poll_interval: float = 1.0) -> None:
        self._session, self._interval = session, poll_interval
        self._startup: Optional[Dict[str, Any]] = None
        self._since = 0
        self._buffer: List[Dict[str, Any]] = []

    def send(self, msg: Dict[str, Any]) -> None:
        self._startup = msg
        r = self._session.post("/stream", json=msg).json()
        self._since = int(r.get("seq", 0))
        self._buffer += r["messages"]

    def recv(self, timeout: Optional[float] = None) -> List[Dict[str, Any]]:
        if self._buffer:
            out, self._buffer = self._buffer, []
            return out
        if self._startup is None:
            return []
        deadline = time.time() + (timeout if timeout is not None else self._interval)
        while True:
            r = self._session.post("/stream", json={"subscribe": self._startup.get("subscribe", {}), "poll_since": self._since}).json()
            self._since = max(self._since, int(r.get("seq", 0)))
            if r["messages"] or time.time() >= deadline:
                return r["messages"]
            time.sleep(min(self._interval, max(deadline - time.time(), 0)))


class WebSocketStreamTransport(StreamWebSocket):
    """A reader thread turns frames into a queue of messages; ``recv`` drains the queue.  When the connection is lost the
    next ``recv`` reconnects (with back-off) and replays the last startup message."""

    def __init__(self, session: api.Session, reconnect: bool = True, max_backoff: float = 10.0) -> None:
        import queue

        self._session = session
        self._reconnect = reconnect
        self._max_backoff = max_backoff
        self._ws: Any = None
        self._q: "queue.Queue[Any]" = queue.Queue()
        self._last_startup: Optional[Dict[str, Any]] = None
        self._closed = False
        self.reconnects = 0
        # called with the transport before a lost connection is re-opened; returns the startup message to send (the Stream
        # sets this to rebuild `known` / `since` from its caches); default: the last one sent, unchanged
        self.on_reconnect: Any = None

    def _connect(self) -> None:
        import threading

        from determined_b200.common import websocket

        master = self._session.master
        cert = self._session._effective_cert()
        ctx = cert.ssl_context() if (cert is not None and master.startswith("https://")) else None
        headers = {"Authorization": f"Bearer {self._session.token}"} if self._session.token else {}
        ws = websocket.connect(master + "/stream", headers=headers, ssl_context=ctx, server_hostname=getattr(cert, "name", None))
        self._ws = ws
        q = self._q

        def reader() -> None:
            try:
                while True:
                    q.put(ws.recv_json())
            except Exception as e:  # noqa: BLE001
                q.put(e)

        threading.Thread(target=reader, daemon=True, name="stream-ws-reader").start()

    def send(self, msg: Dict[str, Any]) -> None:
        self._last_startup = msg
        if self._ws is None or self._ws.closed:
            self._connect()
        self._ws.send_json(msg)

    def _recover(self) -> None:
        delay = 0.2
        while not self._closed:
            try:
                self._connect()
                startup = self.on_reconnect(self) if self.on_reconnect is not None else self._last_startup
                if startup is not None:
                    self._last_startup = startup
                    self._ws.send_json(startup)
                self.reconnects += 1
                return
            except (OSError, Exception) as e:  # noqa: BLE001
                if not self._reconnect:
                    raise
                time.sleep(delay)
                delay = min(delay * 2, self._max_backoff)
                last = e  # noqa: F841

    def recv(self, timeout: Optional[float] = None) -> List[Dict[str, Any]]:
        import queue

        out: List[Dict[str, Any]] = []
        try:
            item = self._q.get(timeout=timeout) if timeout is not None else self._q.get()
        except queue.Empty:
            return out
        while True:
            if isinstance(item, Exception):
                if self._closed:
                    return out
                if not self._reconnect:
                    raise item
                self._q = queue.Queue()
                self._recover()
                return out
            out.append(item)
            try:
                item = self._q.get_nowait()
            except queue.Empty:
                return out

    def close(self) -> None:
        self._closed = True
        if self._ws is not None:
            self._ws.close()


def int_or_list(x: Optional[Union[int, Sequence[int]]]) -> Optional[List[int]]:
    if x is None:
        return None
    return [x] if isinstance(x, int) else list(x)


class ProjectSpec:
    def __init__(self, workspace_id: Any = None, project_id: Any = None) -> None:
        self.workspace_id, self.project_id = int_or_list(workspace_id), int_or_list(project_id)

    def to_wire(self) -> wire.ProjectSubscriptionSpec:
        return wire.ProjectSubscriptionSpec(workspace_ids=self.workspace_id, project_ids=self.project_id)


class ModelSpec:
    def __init__(self, workspace_id: Any = None, model_id: Any = None, user_id: Any = None) -> None:
        self.workspace_id, self.model_id, self.user_id = int_or_list(workspace_id), int_or_list(model_id), int_or_list(user_id)

    def to_wire(self) -> wire.ModelSubscriptionSpec:
        return wire.ModelSubscriptionSpec(workspace_ids=self.workspace_id, model_ids=self.model_id, user_ids=self.user_id)


class ModelVersionSpec:
    def __init__(self, model_id: Any = None, model_version_id: Any = None, user_id: Any = None) -> None:
        self.model_id, self.model_version_id, self.user_id = int_or_list(model_id), int_or_list(model_version_id), int_or_list(user_id)

    def to_wire(self) -> wire.ModelVersionSubscriptionSpec:
        return wire.ModelVersionSubscriptionSpec(model_ids=self.model_id, model_version_ids=self.model_version_id, user_ids=self.user_id)


class Sync:
    def __init__(self, sync_id: Any, complete: bool) -> None:
        self.sync_id, self.complete = sync_id, complete

    def __repr__(self) -> str:
        return f"Sync(sync_id={self.sync_id}, complete={self.complete})"

    def __eq__(self, other: object) -> bool:
        return isinstance(other, Sync) and other.sync_id == self.sync_id and other.complete == self.complete


class KeyCache:
    """Primary keys only: just enough state to resubscribe after a reconnect."""

    def __init__(self, keys: Optional[Set[int]] = None) -> None:
        self.keys = keys or set()
        self.maxseq = 0

    def upsert(self, id_: int, seq: int) -> None:
        self.keys.add(id_)
        self.maxseq = max(self.maxseq, seq or 0)

    def delete_msg(self, deleted: str) -> None:
        for i in _util.range_encoded_keys(deleted):
            self.keys.discard(i)

    def known(self) -> str:
        return _util.encode_keys(self.keys)


_KINDS = {"project": (wire.ProjectMsg, wire.ProjectsDeleted, "projects"), "model": (wire.ModelMsg, wire.ModelsDeleted, "models"),
          "modelversion": (wire.ModelVersionMsg, wire.ModelVersionsDeleted, "model_versions")}


class Stream:
    """Iterate to read events of the current subscription; ``subscribe`` replaces it."""

    def __init__(self, ws: StreamWebSocket) -> None:
        self._ws = ws
        self._caches = {k: KeyCache() for k in ("projects", "models", "model_versions")}
        self._pending: Deque[Any] = collections.deque()
        self._user_sync_id: Any = None
        self._sync_sent: Optional[str] = None
        self._closed = False
        self._specs: Dict[str, Any] = {}
        if hasattr(ws, "on_reconnect"):
            ws.on_reconnect = lambda _t: self._startup_message()      # after a lost connection: resume from the caches

    @classmethod
    def connect(cls, session: api.Session, websocket: bool = True) -> "Stream":
        """A stream on the master behind ``session``: over a websocket, or over long polls with ``websocket=False``."""
        return cls(WebSocketStreamTransport(session) if websocket else PollingStreamTransport(session))

    def _startup_message(self) -> Dict[str, Any]:
        spec: Dict[str, Any] = {}
        known: Dict[str, str] = {}
        for name, s in self._specs.items():
            w = s.to_wire()
            w.since = self._caches[name].maxseq
            spec[name] = w.to_json()
            known[name] = self._caches[name].known()
        return {"sync_id": self._sync_sent, "known": known, "subscribe": spec}

    def subscribe(self, sync_id: Any = None, projects: Optional[ProjectSpec] = None, models: Optional[ModelSpec] = None,
                  model_versions: Optional[ModelVersionSpec] = None) -> "Stream":
        self._specs = {name: s for name, s in (("projects", projects), ("models", models), ("model_versions", model_versions)) if s is not None}
        self._user_sync_id = sync_id
        self._sync_sent = str(uuid.uuid4())
        self._pending.clear()
        self._ws.send(self._startup_message())
        return self

    def _parse(self, msg: Dict[str, Any]) -> None:
        for key, val in msg.items():
            if key == "sync_msg":
                if val.get("sync_id") == self._sync_sent:
                    self._pending.append(Sync(self._user_sync_id, bool(val.get("complete"))))
                continue
            for kind, (up, dele, cache) in _KINDS.items():
                if key == kind:
                    m = up.from_json(val)
                    self._caches[cache].upsert(m.id, getattr(m, "seq", 0))
                    self._pending.append(m)
                elif key == cache + "_deleted":
                    self._caches[cache].delete_msg(val)
                    self._pending.append(dele.from_json(val))

    def __iter__(self) -> Iterator[Any]:
        return self

    def __next__(self) -> Any:
        while not self._pending:
            if self._closed:
                raise StopIteration
            for m in self._ws.recv():
                self._parse(m)
        return self._pending.popleft()

    def next_nowait(self, timeout: float = 0.0) -> Optional[Any]:
        if not self._pending:
            for m in self._ws.recv(timeout=timeout):
                self._parse(m)
        return self._pending.popleft() if self._pending else None

    def close(self) -> None:
        self._closed = True
        self._ws.close()
