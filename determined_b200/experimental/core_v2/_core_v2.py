"""``core_v2`` — "unmanaged"/detached experiments: a plain script reports to the master without
having been launched by it.

Parity: reference ``harness/determined/experimental/core_v2``: ``DefaultConfig``/``UnmanagedConfig``
``_core_v2.py:23-101``, ``init(defaults, unmanaged, client, distributed, checkpoint_storage,
preempt_mode, tensorboard_mode)`` ``:221-262`` populating module-level singletons (``train``,
``checkpoint``, ``distributed``, ``preempt``, ``searcher``, ``info``), ``close()`` ``:264-273``;
experiment+trial creation over REST ``_unmanaged.py:242-308``; heartbeat ``core/_heartbeat.py``.
In managed mode (launched by the master) it simply wraps ``core.init``.
"""
import atexit
import dataclasses
import logging
import os
import uuid
from typing import Any, Dict, List, Optional, Union

import yaml

import determined_b200 as det
from determined_b200 import core
from determined_b200.common import api, storage
from determined_b200.core import DistributedContext, PreemptMode, TensorboardMode  # noqa: F401
from determined_b200.experimental.core_v2 import _core_context_v2, _unmanaged

logger = logging.getLogger("determined.core_v2")


@dataclasses.dataclass
class DefaultConfig:
    name: Optional[str] = None
    hparams: Optional[Dict[str, Any]] = None
    data: Optional[Dict] = None
    description: Optional[str] = None
    labels: Optional[List[str]] = None
    checkpoint_storage: Optional[Union[str, Dict[str, Any]]] = None
    searcher: Optional[Dict[str, Any]] = None


@dataclasses.dataclass
class UnmanagedConfig:
    workspace: Optional[str] = None
    project: Optional[str] = None
    external_experiment_id: Optional[str] = None
    external_trial_id: Optional[str] = None


train: Optional[core.TrainContext] = None
distributed: Optional[core.DistributedContext] = None
preempt: Optional[core.PreemptContext] = None
checkpoint: Optional[core.CheckpointContext] = None
searcher: Optional[core.SearcherContext] = None
info: Optional[det.ClusterInfo] = None
_context: Optional[core.Context] = None
_session: Optional[api.Session] = None
_exp_id: Optional[int] = None


def _set_globals() -> None:
    global train, distributed, preempt, checkpoint, searcher, info
    assert _context is not None
    train, distributed, preempt = _context.train, _context.distributed, _context.preempt
    checkpoint, searcher, info = _context.checkpoint, _context.searcher, _context.info


def init_context(*, defaults: Optional[DefaultConfig] = None, unmanaged: Optional[UnmanagedConfig] = None, client: Any = None,
                 distributed: Optional[core.DistributedContext] = None,
                 checkpoint_storage: Optional[Union[str, Dict[str, Any]]] = None,
                 preempt_mode: core.PreemptMode = core.PreemptMode.WorkersAskChief,
                 tensorboard_mode: core.TensorboardMode = core.TensorboardMode.AUTO) -> core.Context:
    global _session, _exp_id
    if det.get_cluster_info() is not None:
        return core.init(distributed=distributed, checkpoint_storage=checkpoint_storage, preempt_mode=preempt_mode, tensorboard_mode=tensorboard_mode)
    defaults, unmanaged = defaults or DefaultConfig(), unmanaged or UnmanagedConfig()
    dctx = distributed or core.DummyDistributedContext()
    session = client._session if client is not None else api.authentication.login_with_cache(os.environ.get("DET_MASTER", "http://127.0.0.1:8080"))
    _session = session
    cs = checkpoint_storage or defaults.checkpoint_storage
    if cs is not None and defaults.checkpoint_storage is None:
        defaults = dataclasses.replace(defaults, checkpoint_storage=cs)
    cinfo = _unmanaged._get_or_create_experiment_and_trial(session, defaults, unmanaged, dctx)
    _exp_id = cinfo.trial.experiment_id
    return _core_context_v2.make_v2_context(session=session, info=cinfo, distributed=dctx, checkpoint_storage=checkpoint_storage if isinstance(checkpoint_storage, dict) else None,
                                            preempt_mode=preempt_mode, tensorboard_mode=tensorboard_mode,
                                            ship_logs=os.environ.get("DET_CORE_V2_SHIP_LOGS", "1") != "0")


def init(*, defaults: Optional[DefaultConfig] = None, unmanaged: Optional[UnmanagedConfig] = None, client: Any = None,
         distributed: Optional[core.DistributedContext] = None, checkpoint_storage: Optional[Union[str, Dict[str, Any]]] = None,
         preempt_mode: core.PreemptMode = core.PreemptMode.WorkersAskChief,
         tensorboard_mode: core.TensorboardMode = core.TensorboardMode.AUTO) -> None:
    """Initialise the module-level singletons (``core_v2.train.report_training_metrics(...)`` …)."""
    global _context
    if _context is not None:
        raise RuntimeError("core_v2.init() was already called; call core_v2.close() first")
    _context = init_context(defaults=defaults, unmanaged=unmanaged, client=client, distributed=distributed,
                            checkpoint_storage=checkpoint_storage, preempt_mode=preempt_mode, tensorboard_mode=tensorboard_mode)
    _context.start()
    _set_globals()
    atexit.register(close)


def close() -> None:
    global _context, train, distributed, preempt, checkpoint, searcher, info
    if _context is None:
        return
    atexit.unregister(close)
    _context.close()
    _context = None
    train = distributed = preempt = checkpoint = searcher = info = None


def url_reverse_webui_exp_view() -> str:
    assert _session is not None and _exp_id is not None, "call init() first"
    return f"{_session.master}/det/experiments/{_exp_id}"
