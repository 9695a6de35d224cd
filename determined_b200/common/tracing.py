"""Minimal OpenTelemetry-compatible tracing (no SDK dependency).

Parity: reference ``master/pkg/opentelemetry/opentelemetry.go:19-45`` (OTLP exporter wired into the
echo / gRPC servers when ``telemetry.otel_enabled`` is set).  Spans carry W3C ``traceparent``
context, are batched by a background thread and exported either as OTLP/HTTP JSON to
``OTEL_EXPORTER_OTLP_ENDPOINT`` (``/v1/traces`` is appended) or, with a ``file://`` endpoint, as
one OTLP JSON document per line.  Disabled (every call a no-op) unless :func:`configure` ran.
"""
import contextlib
import contextvars
import json
import os
import queue
import secrets
import threading
import time
from typing import Any, Dict, Iterator, List, Optional

_current: "contextvars.ContextVar[Optional[Span]]" = contextvars.ContextVar("detb200_span", default=None)


class Span:
    __slots__ = ("name", "trace_id", "span_id", "parent_id", "start_ns", "end_ns", "attrs", "status_error")

    def __init__(self, name: str, trace_id: str, parent_id: str = "") -> None:
        self.name, self.trace_id, self.parent_id = name, trace_id, parent_id
        self.span_id = secrets.token_hex(8)
        self.start_ns, self.end_ns = time.time_ns(), 0
        self.attrs: Dict[str, Any] = {}
        self.status_error: Optional[str] = None

    def set(self, key: str, value: Any) -> None:
        self.attrs[key] = value

    @property
    def traceparent(self) -> str:
        return f"00-{self.trace_id}-{self.span_id}-01"

    def to_otlp(self) -> Dict[str, Any]:
        def val(v: Any) -> Dict[str, Any]:
            if isinstance(v, bool):
                return {"boolValue": v}
            if isinstance(v, int):
                return {"intValue": str(v)}
            if isinstance(v, float):
                return {"doubleValue": v}
            return {"stringValue": str(v)}

        d: Dict[str, Any] = {"traceId": self.trace_id, "spanId": self.span_id, "name": self.name, "kind": 2,
                             "startTimeUnixNano": str(self.start_ns), "endTimeUnixNano": str(self.end_ns),
                             "attributes": [{"key": k, "value": val(v)} for k, v in self.attrs.items()]}
        if self.parent_id:
            d["parentSpanId"] = self.parent_id
        if self.status_error is not None:
            d["status"] = {"code": 2, "message": self.status_error}
        return d


class Tracer:
    def __init__(self, service: str, endpoint: str, flush_s: float = 2.0) -> None:
        self.service, self.endpoint, self.flush_s = service, endpoint, flush_s
        self._q: "queue.Queue[Span]" = queue.Queue(maxsize=10000)
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True, name="otlp-export")
        self._thread.start()
        self.exported = 0
        self.dropped = 0

    def submit(self, span: Span) -> None:
        try:
            self._q.put_nowait(span)
        except queue.Full:
            self.dropped += 1

    def _drain(self) -> List[Span]:
        out: List[Span] = []
        while len(out) < 512:
            try:
                out.append(self._q.get_nowait())
            except queue.Empty:
                break
        return out

    def _payload(self, spans: List[Span]) -> Dict[str, Any]:
        return {"resourceSpans": [{"resource": {"attributes": [{"key": "service.name", "value": {"stringValue": self.service}}]},
                                   "scopeSpans": [{"scope": {"name": "determined_b200"}, "spans": [s.to_otlp() for s in spans]}]}]}

    def flush(self) -> None:
        while True:
            spans = self._drain()
            if not spans:
                return
            body = json.dumps(self._payload(spans))
            try:
                if self.endpoint.startswith("file://"):
                    with open(self.endpoint[7:], "a") as f:
                        f.write(body + "\n")
                else:
                    import requests

                    requests.post(self.endpoint.rstrip("/") + "/v1/traces", data=body, headers={"Content-Type": "application/json"}, timeout=5)
                self.exported += len(spans)
            except Exception:  # noqa: BLE001 - tracing must never take the service down
                self.dropped += len(spans)

    def _run(self) -> None:
        while not self._stop.wait(self.flush_s):
            self.flush()

    def shutdown(self) -> None:
        self._stop.set()
        self.flush()


_tracer: Optional[Tracer] = None


def configure(service: str, endpoint: Optional[str] = None, flush_s: float = 2.0) -> Optional[Tracer]:
    """Enable tracing for this process.  ``endpoint`` defaults to ``$OTEL_EXPORTER_OTLP_ENDPOINT``; without either tracing stays off."""
    global _tracer
    endpoint = endpoint or os.environ.get("OTEL_EXPORTER_OTLP_ENDPOINT")
    if not endpoint:
        return None
    if _tracer is not None:
        _tracer.shutdown()
    _tracer = Tracer(service, endpoint, flush_s)
    return _tracer


def shutdown() -> None:
    global _tracer
    if _tracer is not None:
        _tracer.shutdown()
        _tracer = None


def enabled() -> bool:
    return _tracer is not None


def parse_traceparent(header: Optional[str]) -> Optional[tuple]:
    if not header:
        return None
    parts = header.strip().split("-")
    if len(parts) != 4 or len(parts[1]) != 32 or len(parts[2]) != 16:
        return None
    return parts[1], parts[2]


@contextlib.contextmanager
def span(name: str, traceparent: Optional[str] = None, **attrs: Any) -> Iterator[Optional[Span]]:
    """``with tracing.span("GET /api/v1/experiments", traceparent=request_header) as s: ...``; nests through a context variable."""
    if _tracer is None:
        yield None
        return
    parent = _current.get()
    remote = parse_traceparent(traceparent)
    if remote is not None:
        s = Span(name, remote[0], remote[1])
    elif parent is not None:
        s = Span(name, parent.trace_id, parent.span_id)
    else:
        s = Span(name, secrets.token_hex(16))
    s.attrs.update(attrs)
    tok = _current.set(s)
    try:
        yield s
    except BaseException as e:
        s.status_error = f"{type(e).__name__}: {e}"
        raise
    finally:
        _current.reset(tok)
        s.end_ns = time.time_ns()
        tr = _tracer
        if tr is not None:
            tr.submit(s)


def current_traceparent() -> Optional[str]:
    s = _current.get()
    return s.traceparent if s is not None else None
