"""In-process stdout/stderr interception + batched shipping to ``POST /task-logs``.

Parity: reference ``harness/determined/core/_log_shipper.py:60-200`` (used by unmanaged
trials, where no container-level ``ship_logs`` wraps the process)."""
import datetime
import io
import logging
import queue
import sys
import threading
from typing import Any, List, Optional

logger = logging.getLogger("determined.core")


class _Tee(io.TextIOBase):
    def __init__(self, orig: Any, q: "queue.Queue[Any]", stdtype: str) -> None:
        self._orig, self._q, self._stdtype, self._buf = orig, q, stdtype, ""

    def write(self, s: str) -> int:
        self._orig.write(s)
        self._buf += s
        while "\n" in self._buf:
            line, self._buf = self._buf.split("\n", 1)
            self._q.put((datetime.datetime.now(datetime.timezone.utc).isoformat(), self._stdtype, line + "\n"))
        return len(s)

    def flush(self) -> None:
        self._orig.flush()


class _LogShipper:
    def __init__(self, session: Any, trial_id: int, task_id: str, distributed: Any = None,
                 batch: int = 1000, period: float = 1.0) -> None:
        self._session, self._trial_id, self._task_id = session, trial_id, task_id
        self._rank = None if distributed is None else distributed.rank
        self._q: "queue.Queue[Any]" = queue.Queue()
        self._batch, self._period = batch, period
        self._thread: Optional[threading.Thread] = None
        self._orig = (sys.stdout, sys.stderr)

    def start(self) -> "_LogShipper":
        sys.stdout = _Tee(self._orig[0], self._q, "stdout")  # type: ignore
        sys.stderr = _Tee(self._orig[1], self._q, "stderr")  # type: ignore
        self._thread = threading.Thread(target=self._run, daemon=True, name="log-shipper")
        self._thread.start()
        return self

    def _ship(self, items: List[Any]) -> None:
        body = [{"task_id": self._task_id, "trial_id": self._trial_id, "timestamp": ts, "stdtype": st,
                 "rank_id": self._rank, "log": line} for ts, st, line in items]
        try:
            self._session.post("/task-logs", json=body)
        except Exception as e:
            self._orig[1].write(f"log shipping failed: {e}\n")

    def _run(self) -> None:
        done = False
        while not done:
            items: List[Any] = []
            try:
                first = self._q.get(timeout=self._period)
                if first is None:
                    done = True
                else:
                    items.append(first)
                while len(items) < self._batch:
                    it = self._q.get_nowait()
                    if it is None:
                        done = True
                        break
                    items.append(it)
            except queue.Empty:
                pass
            if items:
                self._ship(items)

    def close(self) -> None:
        sys.stdout, sys.stderr = self._orig
        self._q.put(None)
        if self._thread is not None:
            self._thread.join(timeout=10)

    def __enter__(self) -> "_LogShipper":
        return self.start()

    def __exit__(self, *_: Any) -> None:
        self.close()
