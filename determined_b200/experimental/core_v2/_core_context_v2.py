"""The ``core.Context`` of an unmanaged run.

Parity: reference ``harness/determined/experimental/core_v2/_core_context_v2.py:25-184``: train / searcher / checkpoint
contexts against the rows ``_unmanaged`` made, a TensorBoard manager when there is checkpoint storage to sync event files to
(``TensorboardMode.AUTO`` also writes reported metrics as TensorBoard scalars), a log shipper so that what the script prints
shows up as the trial's logs on the master, a heartbeat so that the master notices when the script dies without saying so,
and no preemption (nobody can preempt a process the master did not start).
"""
import logging
import os
from typing import Any, Dict, Optional, Union

import determined_b200 as det
from determined_b200 import core, tensorboard
from determined_b200.common import api, storage
from determined_b200.core import _log_shipper

logger = logging.getLogger("determined.core_v2")


def _default_storage_manager() -> storage.StorageManager:
    return storage.from_string(os.path.expanduser("~/.local/share/determined_b200/checkpoints"))


def make_v2_context(*, session: api.Session, info: det.ClusterInfo, distributed: Optional[core.DistributedContext] = None,
                    checkpoint_storage: Optional[Union[str, Dict[str, Any]]] = None,
                    preempt_mode: core.PreemptMode = core.PreemptMode.WorkersAskChief,
                    tensorboard_mode: core.TensorboardMode = core.TensorboardMode.AUTO, ship_logs: bool = True) -> core.Context:
    dctx = distributed or core.DummyDistributedContext()
    t = info.trial
    cs_cfg = t._config.get("checkpoint_storage")
    if isinstance(checkpoint_storage, dict):
        cs_cfg = checkpoint_storage
    mgr = storage.build(cs_cfg, None) if cs_cfg else _default_storage_manager()
    tensorboard_mode = core.TensorboardMode(tensorboard_mode)
    tbd_mgr = tbd_writer = None
    if cs_cfg:
        try:
            tbd_mgr = tensorboard.build(info.cluster_id, t.experiment_id, t.trial_id, cs_cfg, async_upload=True)
            if tensorboard_mode == core.TensorboardMode.AUTO:
                tbd_writer = tensorboard.get_metric_writer()
        except Exception as e:  # noqa: BLE001  (no tensorboard package, unwritable path: metrics still reach the master)
            logger.debug("tensorboard manager unavailable: %s", e)
            tbd_mgr = tbd_writer = None
    if tbd_mgr is None:
        tensorboard_mode = core.TensorboardMode.MANUAL
    train = core.TrainContext(session, t.trial_id, t._trial_run_id, t.experiment_id, dctx, tensorboard_mode, tbd_mgr, tbd_writer)
    ck = core.CheckpointContext(dctx, mgr, session, info.task_id, None, tensorboard_mode, tbd_mgr)
    max_len = (t._config.get("searcher") or {}).get("max_length", 1)
    if isinstance(max_len, dict):
        max_len = next(iter(max_len.values()), 1)
    shipper = _log_shipper._LogShipper(session, t.trial_id, info.task_id, dctx) if ship_logs else None
    hb = core._Heartbeat(session, t.trial_id) if dctx.rank == 0 else None
    return core.Context(checkpoint=ck, distributed=dctx, preempt=core.DummyPreemptContext(dctx, preempt_mode), train=train,
                        searcher=core.DummySearcherContext(dctx, length=int(max_len or 1)), info=info,
                        _tensorboard_manager=tbd_mgr, _heartbeat=hb, _log_shipper=shipper)
