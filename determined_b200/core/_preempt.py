"""``core.PreemptContext`` — cooperative preemption.

Parity: reference ``harness/determined/core/_preempt.py:15-312``: long-poll watcher thread
``:15-121``, ``PreemptMode`` ``:124-145``, ``should_preempt(auto_ack)`` ``:189-241``,
``acknowledge_preemption_signal`` ``:243-257``, Dummy ``:260-312``.

Difference: ``should_preempt()`` on the chief never blocks on the network — the watcher thread
keeps a flag — so calling it *every batch* is cheap; only the broadcast to workers costs, and
that rides the control-plane transport (NCCL side stream / unix socket).
"""
import enum
import logging
import threading
import time
from typing import Any, Optional

from determined_b200.common.api import errors as api_errors

logger = logging.getLogger("determined.core")


class _PreemptionWatcher(threading.Thread):
    """Long-polls ``GET /allocations/{id}/signals/preemption`` until it says ``preempt: true``."""

    def __init__(self, session: Any, allocation_id: str, poll_timeout: int = 60) -> None:
        super().__init__(daemon=True, name="preemption-watcher")
        self._session = session
        self._allocation_id = allocation_id
        self._poll_timeout = poll_timeout
        self._should_preempt = False
        self._should_quit = threading.Event()

    def _poll(self, longpoll_time: int) -> bool:
        r = self._session.get(f"/api/v1/allocations/{self._allocation_id}/signals/preemption",
                              params={"timeout_seconds": str(longpoll_time)},
                              timeout=longpoll_time + 10)
        return bool(r.json().get("preempt", False))

    def run(self) -> None:
        backoff = 0.5
        first = True
        while not self._should_quit.is_set():
            try:
                # First poll returns immediately so a task started already-preempted sees it at once.
                got = self._poll(0 if first else self._poll_timeout)
                first = False
                backoff = 0.5
                if got:
                    self._should_preempt = True
                    return
            except (api_errors.BadRequestException, OSError, Exception) as e:  # noqa: B014
                if self._should_quit.is_set():
                    return
                logger.debug("preemption poll failed (%s); retrying in %.1fs", e, backoff)
                self._should_quit.wait(backoff)
                backoff = min(backoff * 2, 30.0)

    def close(self) -> None:
        self._should_quit.set()

    def should_preempt(self) -> bool:
        return self._should_preempt


class PreemptMode(enum.Enum):
    """``WorkersAskChief``: all workers call ``should_preempt()`` in step; the chief decides and
    broadcasts, so everybody stops on the same batch.  ``ChiefOnly``: only the chief may call.
    ``WorkersAskMaster``: every worker polls the master independently (no synchronisation)."""

    WorkersAskChief = "WORKERS_ASK_CHIEF"
    ChiefOnly = "CHIEF_ONLY"
    WorkersAskMaster = "WORKERS_ASK_MASTER"


class PreemptContext:
    def __init__(self, session: Any, allocation_id: str, dist: Any,
                 preempt_mode: PreemptMode = PreemptMode.WorkersAskChief) -> None:
        self._session = session
        self._allocation_id = allocation_id
        self._dist = dist
        self._preempt_mode = PreemptMode(preempt_mode)
        self._watcher: Optional[_PreemptionWatcher] = None
        self._started = False
        self._ack_sent = False
        if dist.get_rank() == 0 or self._preempt_mode == PreemptMode.WorkersAskMaster:
            self._watcher = _PreemptionWatcher(session, allocation_id)

    def start(self) -> "PreemptContext":
        if self._started:
            raise RuntimeError("you cannot call PreemptContext.start() multiple times")
        self._started = True
        if self._watcher is not None:
            self._watcher.start()
        return self

    def close(self) -> None:
        if self._watcher is not None:
            self._watcher.close()

    def __enter__(self) -> "PreemptContext":
        return self.start()

    def __exit__(self, *_: Any) -> None:
        self.close()

    def should_preempt(self, auto_ack: bool = True) -> bool:
        if not self._started:
            raise RuntimeError("you cannot call PreemptContext.should_preempt() before PreemptContext.start()")
        if self._watcher is not None:
            out = self._watcher.should_preempt()
            if auto_ack and out and not self._ack_sent:
                self.acknowledge_preemption_signal()
                self._ack_sent = True
            if self._preempt_mode == PreemptMode.WorkersAskChief:
                self._dist.broadcast(out)
            return out
        if self._preempt_mode == PreemptMode.ChiefOnly:
            raise RuntimeError("PreemptContext was configured with preempt_mode=ChiefOnly but .should_preempt() "
                               f"was called from non-chief worker of rank={self._dist.get_rank()}")
        out = self._dist.broadcast(None)
        assert isinstance(out, bool)
        return out

    def acknowledge_preemption_signal(self) -> None:
        """Tell the master "I am exiting because you asked; reschedule me later"."""
        self._session.post(f"/api/v1/allocations/{self._allocation_id}/signals/ack_preemption")


class DummyPreemptContext(PreemptContext):
    def __init__(self, dist: Any, preempt_mode: PreemptMode = PreemptMode.WorkersAskChief) -> None:
        self._dist = dist
        self._preempt_mode = PreemptMode(preempt_mode)
        self._started = False
        self._force = False

    def start(self) -> "DummyPreemptContext":
        self._started = True
        return self

    def close(self) -> None:
        pass

    def _force_preemption(self) -> None:
        """Test hook: make the next ``should_preempt`` return True."""
        self._force = True

    def should_preempt(self, auto_ack: bool = True) -> bool:
        if not self._started:
            raise RuntimeError("cannot call .should_preempt() before .start(); use the PreemptContext as a context manager")
        if self._dist.rank == 0:
            out = self._force
            if self._preempt_mode == PreemptMode.WorkersAskChief:
                self._dist.broadcast(out)
            return out
        if self._preempt_mode == PreemptMode.ChiefOnly:
            raise RuntimeError("PreemptContext was configured with preempt_mode=ChiefOnly but .should_preempt() "
                               f"was called from non-chief worker of rank={self._dist.get_rank()}")
        if self._preempt_mode == PreemptMode.WorkersAskChief:
            return bool(self._dist.broadcast(None))
        return False

    def acknowledge_preemption_signal(self) -> None:
        pass
