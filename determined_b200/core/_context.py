"""``core.init()`` and ``core.Context``.

Parity: reference ``harness/determined/core/_context.py``: ``Context`` composition ``:21-106``,
InvalidHP handling on exit ``:102-106``, SIGUSR1 stack dump ``:109-127``, ``_dummy_init``
``:147-183``, ``init(distributed, checkpoint_storage, preempt_mode, tensorboard_mode)`` ``:190-343``,
``_run_prepare`` ``:346-364``.
"""
import faulthandler
import logging
import os
import pathlib
import signal
import sys
import tempfile
import threading
import traceback
from typing import Any, Dict, Optional, Union

from determined_b200 import _info, constants, errors, tensorboard
from determined_b200.common import api, storage
from determined_b200.core import _checkpoint, _distributed, _experimental, _preempt, _profiler, _searcher, _train
from determined_b200.core._tensorboard_mode import TensorboardMode

logger = logging.getLogger("determined.core")


class Context:
    """Bundle of the per-concern Core API contexts; use as a context manager."""

    def __init__(self, checkpoint: _checkpoint.CheckpointContext,
                 distributed: Optional[_distributed.DistributedContext] = None,
                 preempt: Optional[_preempt.PreemptContext] = None,
                 train: Optional[_train.TrainContext] = None,
                 searcher: Optional[_searcher.SearcherContext] = None,
                 info: Optional[_info.ClusterInfo] = None,
                 experimental: Optional[_experimental.ExperimentalCoreContext] = None,
                 profiler: Optional[_profiler.ProfilerContext] = None,
                 _metrics: Any = None, _tensorboard_manager: Any = None, _heartbeat: Any = None,
                 _log_shipper: Any = None) -> None:
        self.checkpoint = checkpoint
        self.distributed = distributed or _distributed.DummyDistributedContext()
        self.preempt = preempt or _preempt.DummyPreemptContext(self.distributed)
        self.train = train or _train.DummyTrainContext()
        self.searcher = searcher or _searcher.DummySearcherContext(self.distributed)
        self.info = info
        self.experimental = experimental or _experimental.DummyExperimentalCoreContext()
        self.profiler = profiler or _profiler.DummyProfilerContext()
        self._tensorboard_manager = _tensorboard_manager
        self._heartbeat = _heartbeat
        self._log_shipper = _log_shipper

    def start(self) -> None:
        self.preempt.start()
        if self._log_shipper is not None:
            self._log_shipper.start()
        if self._heartbeat is not None:
            self._heartbeat.start()

    def __enter__(self) -> "Context":
        self.start()
        return self

    def close(self, exc_type: Optional[type] = None, exc_val: Optional[BaseException] = None,
              exc_tb: Any = None) -> None:
        self.preempt.close()
        self.profiler._close()
        if self._tensorboard_manager is not None:
            self._tensorboard_manager.close()
        if self._heartbeat is not None:
            self._heartbeat.close(exit_ok=exc_type is None)
        if self._log_shipper is not None:
            self._log_shipper.close()
        # Detect InvalidHP and tell the searcher before the process dies.
        if isinstance(exc_val, errors.InvalidHP):
            if self.distributed.rank == 0:
                self.train.report_early_exit(_train.EarlyExitReason.INVALID_HP)
            logger.info("InvalidHP detected during Trial init, converting InvalidHP to exit(0)")
        self.distributed.close()

    def __exit__(self, exc_type: Optional[type], exc_val: Optional[BaseException], exc_tb: Any) -> None:
        self.close(exc_type, exc_val, exc_tb)
        if isinstance(exc_val, errors.InvalidHP):
            # The early-exit report already told the master; the process itself ends cleanly.
            sys.exit(0)


def _install_stacktrace_on_sigusr1() -> None:
    """``kill -USR1 <pid>`` dumps every thread's stack to stderr (debug aid)."""
    if not hasattr(signal, "SIGUSR1") or threading.current_thread() is not threading.main_thread():
        return
    try:
        faulthandler.register(signal.SIGUSR1, all_threads=True, chain=False)
    except (ValueError, RuntimeError, AttributeError):
        pass


def _get_storage_manager(checkpoint_storage: Optional[Union[str, Dict[str, Any]]]) -> Optional[storage.StorageManager]:
    if checkpoint_storage is None:
        return None
    if isinstance(checkpoint_storage, str):
        return storage.from_string(checkpoint_storage)
    if isinstance(checkpoint_storage, dict):
        return storage.build(checkpoint_storage, container_path=None)
    raise TypeError("checkpoint_storage must be a string, dictionary, or None")


def _dummy_init(*, distributed: Optional[_distributed.DistributedContext] = None,
                checkpoint_storage: Optional[Union[str, Dict[str, Any]]] = None,
                preempt_mode: _preempt.PreemptMode = _preempt.PreemptMode.WorkersAskChief,
                tensorboard_path: Optional[pathlib.Path] = None,
                searcher_length: int = 1) -> Context:
    """Build a core.Context for off-cluster use (no master)."""
    distributed = distributed or _distributed.DummyDistributedContext()
    mgr = _get_storage_manager(checkpoint_storage)
    if mgr is None:
        base = os.path.join(tempfile.gettempdir(), "determined_b200-checkpoints")
        os.makedirs(base, exist_ok=True)
        mgr = storage.SharedFSStorageManager(base)
    checkpoint = _checkpoint.DummyCheckpointContext(distributed, mgr)
    _install_stacktrace_on_sigusr1()
    return Context(
        distributed=distributed, checkpoint=checkpoint,
        preempt=_preempt.DummyPreemptContext(distributed, preempt_mode),
        train=_train.DummyTrainContext(), searcher=_searcher.DummySearcherContext(distributed, searcher_length),
        experimental=_experimental.DummyExperimentalCoreContext(), profiler=_profiler.DummyProfilerContext(),
    )


def _run_prepare(distributed: _distributed.DistributedContext, session: Any, trial_id: int,
                 checkpoint_storage: Optional[Dict[str, Any]]) -> Optional[int]:
    """Tell the master which storage backend this run writes to; returns its storage id."""
    sid = None
    if distributed.rank == 0:
        try:
            r = session.post("/api/v1/runs/start", json={"runId": trial_id, "checkpointStorage": checkpoint_storage})
            sid = r.json().get("storageId")
        except api.APIException as e:
            logger.debug("runs/start not supported by master: %s", e)
    return distributed.broadcast(sid)


def init(*, distributed: Optional[_distributed.DistributedContext] = None,
         checkpoint_storage: Optional[Union[str, Dict[str, Any]]] = None,
         preempt_mode: _preempt.PreemptMode = _preempt.PreemptMode.WorkersAskChief,
         tensorboard_mode: TensorboardMode = TensorboardMode.AUTO) -> Context:
    """Build a ``core.Context``.  On-cluster (``det.get_cluster_info()`` is not None) every
    sub-context talks to the master; off-cluster the Dummy twins are used so the same script runs
    anywhere."""
    info = _info.get_cluster_info()
    if info is None:
        return _dummy_init(distributed=distributed, checkpoint_storage=checkpoint_storage, preempt_mode=preempt_mode)

    distributed = distributed or _distributed.DummyDistributedContext()
    preempt_mode = _preempt.PreemptMode(preempt_mode)
    tensorboard_mode = TensorboardMode(tensorboard_mode)
    session = api.Session(info.master_url, username=None, token=info.session_token, cert=None, max_retries=6)

    train = searcher = tbd_mgr = None
    experimental = None
    storage_backend_id = None
    mgr = _get_storage_manager(checkpoint_storage)

    if info.task_type == "TRIAL":
        t = info.trial
        cfg = t._config
        if mgr is None:
            cs = cfg.get("checkpoint_storage")
            if not cs:
                raise ValueError("experiment config has no checkpoint_storage and none was passed to core.init()")
            mgr = storage.build(cs, container_path=None)
            try:
                tbd_mgr = tensorboard.build(info.cluster_id, t.experiment_id, t.trial_id, cs,
                                            async_upload=True)
            except Exception as e:
                logger.debug("tensorboard manager unavailable: %s", e)
        if isinstance(checkpoint_storage, dict):
            storage_backend_id = _run_prepare(distributed, session, t.trial_id, checkpoint_storage)
        train = _train.TrainContext(session, t.trial_id, t._trial_run_id, t.experiment_id, distributed,
                                    tensorboard_mode, tbd_mgr, None)
        searcher = _searcher.SearcherContext(session, distributed, t.trial_id, t._trial_run_id,
                                             info.allocation_id, _searcher._parse_searcher_units(cfg))
        experimental = _experimental.ExperimentalCoreContext(session, t.trial_id)
        checkpoint: _checkpoint.CheckpointContext = _checkpoint.CheckpointContext(
            distributed, mgr, session, info.task_id, info.allocation_id, tensorboard_mode, tbd_mgr,
            storage_backend_id)
        preempt = _preempt.PreemptContext(session, info.allocation_id, distributed, preempt_mode)
        profiler = _profiler.ProfilerContext(info.agent_id, distributed, train)
    else:
        # Non-trial tasks (commands, notebooks): checkpointing allowed if storage given.
        if mgr is None:
            base = os.path.join(tempfile.gettempdir(), "determined_b200-checkpoints")
            os.makedirs(base, exist_ok=True)
            mgr = storage.SharedFSStorageManager(base)
        checkpoint = _checkpoint.DummyCheckpointContext(distributed, mgr)
        preempt = _preempt.PreemptContext(session, info.allocation_id, distributed, preempt_mode)
        profiler = _profiler.DummyProfilerContext()

    _install_stacktrace_on_sigusr1()
    return Context(distributed=distributed, checkpoint=checkpoint, preempt=preempt, train=train,
                   searcher=searcher, info=info, experimental=experimental, profiler=profiler,
                   _tensorboard_manager=tbd_mgr)
