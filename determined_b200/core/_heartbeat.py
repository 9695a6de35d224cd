"""Unmanaged-trial heartbeat (reference ``harness/determined/core/_heartbeat.py:1-154``): a
daemon thread PATCHes ``/api/v1/trials/{id}`` periodically so the master can tell a detached
(``core_v2`` unmanaged) run is alive, and sets the terminal state on exit."""
import logging
import threading
from typing import Any, Optional

logger = logging.getLogger("determined.core")


class _Heartbeat(threading.Thread):
    def __init__(self, session: Any, trial_id: int, interval: float = 60.0) -> None:
        super().__init__(daemon=True, name="unmanaged-heartbeat")
        self._session, self._trial_id, self._interval = session, trial_id, interval
        self._quit = threading.Event()

    def _beat(self, state: Optional[str] = None) -> None:
        body = {"trialId": self._trial_id}
        if state:
            body["state"] = state
        try:
            self._session.patch(f"/api/v1/trials/{self._trial_id}", json=body)
        except Exception as e:
            logger.debug("heartbeat failed: %s", e)

    def run(self) -> None:
        self._beat("STATE_RUNNING")
        while not self._quit.wait(self._interval):
            self._beat()

    def close(self, exit_ok: bool = True) -> None:
        self._quit.set()
        self._beat("STATE_COMPLETED" if exit_ok else "STATE_ERROR")
