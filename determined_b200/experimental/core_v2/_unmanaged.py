"""Unmanaged experiments and trials: rows on the master for work the master did not start.

Parity: reference ``harness/determined/experimental/core_v2/_unmanaged.py`` — create the experiment
(``_create_unmanaged_experiment`` ``:12-34``) or find it again by its external id (``_put_unmanaged_experiment`` ``:36-63``),
the same for the trial (``:74-138``), and build the ``ClusterInfo`` an unmanaged process runs with (``:165-240``).
"""
import os
import uuid
from typing import Any, Dict, Optional

import yaml

import determined_b200 as det
from determined_b200 import core
from determined_b200.common import api


def resolve_project(session: api.Session, workspace: Optional[str], project: Optional[str]) -> Optional[int]:
    """Project id for ``UnmanagedConfig.workspace`` / ``.project`` (both by name).  A workspace without a project means its
    first project; names that do not exist are an error rather than a silent fall-back to "Uncategorized"."""
    if not workspace and not project:
        return None
    workspace = workspace or "Uncategorized"
    ws = session.get("/api/v1/workspaces", params={"name": workspace}).json()["workspaces"]
    if not ws:
        raise ValueError(f"workspace {workspace!r} does not exist")
    projects = session.get(f"/api/v1/workspaces/{ws[0]['id']}/projects").json()["projects"]
    if project is None:
        if not projects:
            raise ValueError(f"workspace {workspace!r} has no project")
        return int(projects[0]["id"])
    hit = [p for p in projects if p["name"] == project]
    if not hit:
        raise ValueError(f"workspace {workspace!r} has no project named {project!r}")
    return int(hit[0]["id"])


def _exp_config(defaults: Any, hparams: Dict[str, Any]) -> Dict[str, Any]:
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


def _get_or_create_experiment_and_trial(session: api.Session, defaults: Any, unmanaged: Any,
                                        distributed_ctx: core.DistributedContext) -> det.ClusterInfo:
    """The chief creates (or finds, by external ids) the experiment and trial; ids are broadcast."""
    if unmanaged.external_trial_id and not unmanaged.external_experiment_id:
        raise ValueError("external_trial_id needs external_experiment_id: a trial is found again inside its experiment")
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
            body: Dict[str, Any] = {"config": yaml.safe_dump(cfg), "unmanaged": True, "activate": True}
            pid = resolve_project(session, unmanaged.workspace, unmanaged.project)
            if pid is not None:
                body["projectId"] = pid
            r = session.post("/api/v1/experiments", json=body).json()
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
