"""Product telemetry: anonymous usage events (reference ``master/internal/telemetry``: Segment ``track`` calls for master
start, experiment creation and state changes, periodic cluster size), off unless ``telemetry.enabled`` and a
``telemetry.segment_url`` are set — there is no default collector.

Events are queued and posted by one background thread, so a slow or dead collector never touches the request path; the last
events are kept in memory for ``GET /api/v1/master/telemetry?events=1`` (admins), which is also how tests read them.
"""
import base64
import collections
import json
import logging
import queue
import threading
import time
from typing import Any, Deque, Dict, Optional

import requests

logger = logging.getLogger("determined.master.telemetry")


class Telemetry:
    def __init__(self, cluster_id: str, enabled: bool = False, url: Optional[str] = None, key: Optional[str] = None, version: str = "") -> None:
        self.cluster_id, self.enabled, self.url, self.key, self.version = cluster_id, bool(enabled), url, key or "", version
        self.recent: Deque[Dict[str, Any]] = collections.deque(maxlen=200)
        self._q: "queue.Queue[Optional[Dict[str, Any]]]" = queue.Queue(maxsize=1000)
        self._thread: Optional[threading.Thread] = None
        self.sent = 0
        self.dropped = 0

    def track(self, event: str, **properties: Any) -> None:
        if not self.enabled:
            return
        msg = {"type": "track", "event": event, "userId": self.cluster_id, "timestamp": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()),
               "properties": dict(properties, master_version=self.version), "context": {"library": {"name": "determined_b200"}}}
        self.recent.append(msg)
        if not self.url:
            return
        if self._thread is None:
            self._thread = threading.Thread(target=self._pump, daemon=True, name="telemetry")
            self._thread.start()
        try:
            self._q.put_nowait(msg)
        except queue.Full:
            self.dropped += 1

    def _pump(self) -> None:
        headers = {"Content-Type": "application/json"}
        if self.key:
            headers["Authorization"] = "Basic " + base64.b64encode(f"{self.key}:".encode()).decode()
        while True:
            msg = self._q.get()
            if msg is None:
                return
            try:
                requests.post(str(self.url), data=json.dumps(msg), headers=headers, timeout=10).raise_for_status()
                self.sent += 1
            except Exception as e:  # noqa: BLE001
                self.dropped += 1
                logger.debug("telemetry event %s not delivered: %s", msg["event"], e)

    def flush(self, timeout: float = 5.0) -> None:
        t0 = time.time()
        while not self._q.empty() and time.time() - t0 < timeout:
            time.sleep(0.02)

    def stop(self) -> None:
        if self._thread is not None:
            self._q.put(None)
