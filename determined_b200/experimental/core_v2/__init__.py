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


def _exp_config(defaults: DefaultConfig, hparams: Dict[str, Any]) -> Dict[str, Any]:
    cs = defaults.checkpoint_storage
    if isinstance(cs, str):
        cs = {"type": "shared_fs", "host_path": os.path.abspath(cs)}
    cfg: Dict[str, Any] = {
        "name": defaults.name or f"unmanaged-{uuid.uuid4().hex[:8]}", "entrypoint": "python3 (unmanaged)",
        "hyperparameters": {k: {"type": "const", "val": v} for k, v in hparams.items()},
        "searcher": defaults.searcher or {"name": "single", "metric": "loss", "max_length": 100000000},
        "data": defaults.data or {}, "labels": defaults.labels or [],
    }
    if defaults.description:
        cfg["description"] = defaults.description
    if cs:
        cfg["checkpoint_storage"] = cs
    return cfg


def _get_or_create_experiment_and_trial(session: api.Session, defaults: DefaultConfig, unmanaged: UnmanagedConfig,
                                        distributed_ctx: core.DistributedContext) -> det.ClusterInfo:
    """The chief creates (or finds, by external ids) the experiment and trial; ids are broadcast."""
    ids = None
    if distributed_ctx.rank == 0:
        hparams = defaults.hparams or {}
        cfg = _exp_config(defaults, hparams)
        if unmanaged.external_experiment_id:
            cfg.setdefault("data", {})["external_experiment_id"] = unmanaged.external_experiment_id
        eid = None
        if unmanaged.external_experiment_id:
            for e in session.get("/api/v1/experiments").json()["experiments"]:
                if (e["config"].get("data") or {}).get("external_experiment_id") == unmanaged.external_experiment_id and e["unmanaged"]:
                    eid = e["id"]
                    break
        if eid is None:
            r = session.post("/api/v1/experiments", json={"config": yaml.safe_dump(cfg), "unmanaged": True, "activate": True}).json()
            eid, full_cfg = r["experiment"]["id"], r["config"]
        else:
            full_cfg = session.get(f"/api/v1/experiments/{eid}").json()["experiment"]["config"]
        tid, latest = None, None
        if unmanaged.external_trial_id:
            for t in session.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]:
                if (t["hparams"] or {}).get("__external_trial_id") == unmanaged.external_trial_id:
                    tid, latest = t["id"], t["latestCheckpoint"]
                    break
        if tid is None:
            hp = dict(hparams)
            if unmanaged.external_trial_id:
                hp["__external_trial_id"] = unmanaged.external_trial_id
            tid = session.post(f"/api/v1/experiments/{eid}/trials", json={"hparams": hp, "unmanaged": True}).json()["trial"]["id"]
        ids = (eid, tid, latest, full_cfg, session.get("/api/v1/master").json()["clusterId"])
    eid, tid, latest, full_cfg, cluster_id = distributed_ctx.broadcast(ids)
    trial = det.TrialInfo(trial_id=tid, experiment_id=eid, trial_seed=0, hparams=defaults.hparams or {}, config=full_cfg,
                          steps_completed=0, trial_run_id=0)
    return det.ClusterInfo(master_url=session.master, cluster_id=cluster_id, agent_id="unmanaged", slot_ids=[], task_id=f"trial-{tid}",
                           allocation_id=f"unmanaged-{tid}-{uuid.uuid4().hex[:6]}", session_token=session.token or "", task_type="TRIAL",
                           latest_checkpoint=latest, trial_info=trial, rendezvous_info=det.RendezvousInfo(["127.0.0.1"], 0, [0]))


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
    cinfo = _get_or_create_experiment_and_trial(session, defaults, unmanaged, dctx)
    _exp_id = cinfo.trial.experiment_id
    t = cinfo.trial
    cs_cfg = t._config.get("checkpoint_storage")
    mgr = storage.build(cs_cfg, None) if cs_cfg else storage.from_string(os.path.expanduser("~/.local/share/determined_b200/checkpoints"))
    trainc = core.TrainContext(session, t.trial_id, 0, t.experiment_id, dctx, core.TensorboardMode.MANUAL, None, None)
    ck = core.CheckpointContext(dctx, mgr, session, cinfo.task_id, None, core.TensorboardMode.MANUAL, None)
    hb = core._Heartbeat(session, t.trial_id) if dctx.rank == 0 else None
    return core.Context(checkpoint=ck, distributed=dctx, preempt=core.DummyPreemptContext(dctx, preempt_mode), train=trainc,
                        searcher=core.DummySearcherContext(dctx, length=int((t._config.get("searcher") or {}).get("max_length", 1) if not isinstance((t._config.get("searcher") or {}).get("max_length"), dict) else 1)),
                        info=cinfo, _heartbeat=hb)


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
