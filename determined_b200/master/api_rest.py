"""The rest of the reference's REST surface: bulk experiment actions, labels, file tree, validation history, unmanaged
trials by external id, log fields, typed metric reports and streams, profiler series, accelerator data, task lookups,
user settings / activity, project notes / columns / metric ranges / moves, workspace pins, role lookups, experiment and run
search.  Found by diffing ``proto/src/determined/api/v1/api.proto`` (230 rpcs) against the route table.

Routes live on the same table as ``api.py`` and ``api_more.py``; free-form per-object facts (user settings, pins, notes,
activity) are rows of the ``attrs`` table.
"""
import base64
import io
import json
import tarfile
import time
from typing import Any, Dict, List, Optional, Tuple

from determined_b200.master import api as A
from determined_b200.master import api_more as M
from determined_b200.master import db as dbmod

route, Ctx, HTTPError = A.route, A.Ctx, A.HTTPError


def _ids(ctx: Ctx, key: str = "experimentIds") -> List[int]:
    b = ctx.body or {}
    ids = [int(x) for x in b.get(key) or []]
    flt = b.get("filters") or {}
    if not ids and flt:
        ids = [e["id"] for e in _filter_experiments(ctx, flt)]
    return ids


def _filter_experiments(ctx: Ctx, flt: Dict[str, Any]) -> List[Dict[str, Any]]:
    """``BulkExperimentFilters`` / search filters: description, name, labels, archived, states, userIds, projectId, excludedExperimentIds."""
    out = []
    for r in ctx.m.db.all("SELECT id, state, config, original_config, owner_id, start_time, end_time, progress, archived, parent_id, notes, project_id, unmanaged, job_id FROM experiments ORDER BY id"):
        e = A.exp_json(ctx.m, r)
        if flt.get("name") and str(flt["name"]).lower() not in str(e["name"] or "").lower():
            continue
        if flt.get("description") and str(flt["description"]).lower() not in str(e["description"] or "").lower():
            continue
        if flt.get("labels") and not set(flt["labels"]) <= set(e["labels"]):
            continue
        if flt.get("archived") is not None and bool(flt["archived"]) != e["archived"]:
            continue
        if flt.get("states") and e["state"] not in [s if str(s).startswith("STATE_") else f"STATE_{s}" for s in flt["states"]]:
            continue
        if flt.get("userIds") and e["userId"] not in [int(u) for u in flt["userIds"]]:
            continue
        if flt.get("projectId") and int(flt["projectId"]) != e["projectId"]:
            continue
        if e["id"] in [int(x) for x in flt.get("excludedExperimentIds") or []]:
            continue
        out.append(e)
    return out


def _bulk(action: str) -> Any:
    """One result row per experiment: ``{"id": ..., "error": ""}`` — a failure on one does not stop the others."""
    single = {"activate": A._exp_action("activate"), "pause": A._exp_action("pause"), "cancel": A._exp_action("cancel"), "kill": A._exp_action("kill"),
              "archive": A.archive_experiment, "unarchive": A.unarchive_experiment}[action]

    def handler(ctx: Ctx) -> Any:
        results = []
        for eid in _ids(ctx):
            try:
                single(ctx, str(eid))
                results.append({"id": eid, "error": ""})
            except HTTPError as e:
                results.append({"id": eid, "error": e.message if hasattr(e, "message") else str(e)})
        return {"results": results}
    return handler


for _a in ("activate", "pause", "cancel", "kill", "archive", "unarchive"):
    route("POST", f"/api/v1/experiments/{_a}")(_bulk(_a))


@route("POST", "/api/v1/experiments/move")
def move_experiments(ctx: Ctx) -> Any:
    dst = int((ctx.body or {}).get("destinationProjectId") or 0)
    if ctx.m.db.one("SELECT id FROM projects WHERE id=?", (dst,)) is None:
        raise HTTPError(404, f"project {dst} not found")
    results = []
    for eid in _ids(ctx):
        if ctx.m.db.one("SELECT id FROM experiments WHERE id=?", (eid,)) is None:
            results.append({"id": eid, "error": "not found"})
            continue
        ctx.m.db.exec("UPDATE experiments SET project_id=? WHERE id=?", (dst, eid))
        results.append({"id": eid, "error": ""})
    return {"results": results}


@route("DELETE", "/api/v1/experiments/delete")
def delete_experiments(ctx: Ctx) -> Any:
    if not ctx.body and ctx.query.get("experimentIds"):
        ctx.body = {"experimentIds": ctx.query["experimentIds"]}
    results = []
    for eid in _ids(ctx):
        try:
            A.delete_experiment(ctx, str(eid))
            results.append({"id": eid, "error": ""})
        except HTTPError as e:
            results.append({"id": eid, "error": e.message})
    return {"results": results}


@route("PUT", "/api/v1/experiments/retain_logs")
def retain_logs_bulk(ctx: Ctx) -> Any:
    days = int((ctx.body or {}).get("numDays", 0))
    results = []
    for eid in _ids(ctx):
        try:
            M.set_attr(ctx.m, "experiment", eid, "logRetentionDays", days)
            for t in ctx.m.db.all("SELECT id FROM trials WHERE experiment_id=?", (eid,)):
                M.set_attr(ctx.m, "trial", t["id"], "logRetentionDays", days)
            results.append({"id": eid, "error": ""})
        except Exception as e:  # noqa: BLE001
            results.append({"id": eid, "error": str(e)})
    return {"results": results}


# ------------------------------------------------------------------------------------------------------------------- labels
@route("GET", "/api/v1/experiment/labels")
def experiment_labels(ctx: Ctx) -> Any:
    """Every label in use, most used first."""
    count: Dict[str, int] = {}
    pid = ctx.qi("projectId", 0)
    for r in ctx.m.db.all("SELECT config, project_id FROM experiments"):
        if pid and r["project_id"] != pid:
            continue
        for lab in dbmod.U(r["config"], {}).get("labels") or []:
            count[lab] = count.get(lab, 0) + 1
    return {"labels": sorted(count, key=lambda k: (-count[k], k))}


def _set_labels(ctx: Ctx, eid: str, label: str, add: bool) -> Any:
    row = A._exp_row(ctx, eid)
    cfg = dbmod.U(row["config"], {})
    labels = [x for x in cfg.get("labels") or [] if x != label] + ([label] if add else [])
    cfg["labels"] = labels
    ctx.m.db.exec("UPDATE experiments SET config=? WHERE id=?", (dbmod.J(cfg), int(eid)))
    with ctx.m.lock:
        live = ctx.m.experiments.get(int(eid))
        if live is not None:
            live.config["labels"] = labels
    return {"labels": labels}


route("PUT", "/api/v1/experiments/{eid}/labels/{label}")(lambda ctx, eid, label: _set_labels(ctx, eid, label, True))
route("DELETE", "/api/v1/experiments/{eid}/labels/{label}")(lambda ctx, eid, label: _set_labels(ctx, eid, label, False))


@route("GET", "/api/v1/model/labels")
def model_labels(ctx: Ctx) -> Any:
    count: Dict[str, int] = {}
    wid = ctx.qi("workspaceId", 0)
    for r in ctx.m.db.all("SELECT labels, workspace_id FROM models"):
        if wid and r["workspace_id"] != wid:
            continue
        for lab in dbmod.U(r["labels"], []):
            count[lab] = count.get(lab, 0) + 1
    return {"labels": sorted(count, key=lambda k: (-count[k], k))}


@route("POST", "/api/v1/models/{name}/move")
def move_model(ctx: Ctx, name: str) -> Any:
    dst = int((ctx.body or {}).get("destinationWorkspaceId") or 0)
    if ctx.m.db.one("SELECT id FROM workspaces WHERE id=?", (dst,)) is None:
        raise HTTPError(404, f"workspace {dst} not found")
    r = A._model(ctx, name)
    ctx.m.db.exec("UPDATE models SET workspace_id=?, last_updated_time=? WHERE id=?", (dst, time.time(), r["id"]))
    return {}


# ------------------------------------------------------------------------------------------------------- model definition
def _model_def_members(ctx: Ctx, eid: str) -> List[Tuple[tarfile.TarInfo, bytes]]:
    row = ctx.m.db.one("SELECT model_definition FROM experiments WHERE id=?", (int(eid),))
    if row is None:
        raise HTTPError(404, f"experiment {eid} not found")
    out: List[Tuple[tarfile.TarInfo, bytes]] = []
    if row["model_definition"]:
        with tarfile.open(fileobj=io.BytesIO(row["model_definition"]), mode="r:*") as tf:
            for ti in tf.getmembers():
                f = tf.extractfile(ti) if ti.isfile() else None
                out.append((ti, f.read() if f is not None else b""))
    return out


@route("GET", "/api/v1/experiments/{eid}/file_tree")
def model_def_tree(ctx: Ctx, eid: str) -> Any:
    """The experiment's code as a tree of ``FileNode``s (directories first, then by name)."""
    root: Dict[str, Any] = {"files": {}}
    for ti, data in _model_def_members(ctx, eid):
        parts = [p for p in ti.name.strip("/").split("/") if p and p != "."]
        cur = root
        for i, p in enumerate(parts):
            leaf = i == len(parts) - 1
            node = cur["files"].setdefault(p, {"path": "/".join(parts[:i + 1]), "name": p, "isDir": not leaf or ti.isdir(), "files": {},
                                               "modifiedTime": A._ts(ti.mtime), "contentLength": 0, "contentType": ""})
            if leaf and not ti.isdir():
                node.update(isDir=False, contentLength=len(data), contentType="text/plain" if _is_text(data) else "application/octet-stream")
            cur = node

    def listed(node: Dict[str, Any]) -> List[Dict[str, Any]]:
        kids = sorted(node["files"].values(), key=lambda n: (not n["isDir"], n["name"]))
        return [dict({k: v for k, v in n.items() if k != "files"}, files=listed(n)) for n in kids]
    return {"files": listed(root)}


def _is_text(data: bytes) -> bool:
    try:
        data[:4096].decode()
        return b"\0" not in data[:4096]
    except UnicodeDecodeError:
        return False


@route("POST", "/api/v1/experiments/{eid}/file")
def model_def_file(ctx: Ctx, eid: str) -> Any:
    want = str((ctx.body or {}).get("path") or "").strip("/")
    for ti, data in _model_def_members(ctx, eid):
        if ti.name.strip("/") == want and ti.isfile():
            return {"file": base64.b64encode(data).decode()}
    raise HTTPError(404, f"no file {want!r} in the code of experiment {eid}")


# ------------------------------------------------------------------------------------------------------ metrics and history
@route("GET", "/api/v1/experiments/{eid}/validation-history")
def validation_history(ctx: Ctx, eid: str) -> Any:
    """The validations that improved on everything before them (the "best so far" staircase of the searcher metric)."""
    row = A._exp_row(ctx, eid)
    s = dbmod.U(row["config"], {}).get("searcher") or {}
    name, smaller = s.get("metric"), bool(s.get("smaller_is_better", True))
    best: Optional[float] = None
    out = []
    rows = ctx.m.db.all("SELECT m.* FROM metrics m JOIN trials t ON t.id=m.trial_id WHERE t.experiment_id=? AND m.grp='validation' ORDER BY m.report_time, m.id", (int(eid),))
    for r in rows:
        v = dbmod.U(r["metrics"], {}).get(name)
        if not isinstance(v, (int, float)) or isinstance(v, bool):
            continue
        if best is None or (v < best if smaller else v > best):
            best = float(v)
            out.append({"trialId": r["trial_id"], "endTime": A._ts(r["report_time"]), "searcherMetric": float(v)})
    return {"validationHistory": out}


@route("GET", "/api/v1/experiments/{eid}/metrics-stream/batches")
def metric_batches(ctx: Ctx, eid: str) -> Any:
    name, grp = ctx.q("metricName"), (ctx.q("group") or {"METRIC_TYPE_TRAINING": "training", "METRIC_TYPE_VALIDATION": "validation"}.get(ctx.q("metricType") or "", "validation"))
    if not name:
        raise HTTPError(400, "metricName is required")
    steps = set()
    for r in ctx.m.db.all("SELECT m.steps_completed s, m.metrics mm FROM metrics m JOIN trials t ON t.id=m.trial_id WHERE t.experiment_id=? AND m.grp=?", (int(eid), grp)):
        if name in dbmod.U(r["mm"], {}):
            steps.add(int(r["s"]))
    return [{"result": {"batches": sorted(steps)}}]


def _typed_metrics(group: str) -> Any:
    def handler(ctx: Ctx) -> Any:
        ids = [int(x) for x in ctx.query.get("trialIds", [])]
        out = []
        for i in ids:
            for r in ctx.m.db.all("SELECT * FROM metrics WHERE trial_id=? AND grp=? ORDER BY steps_completed, id", (i, group)):
                out.append({"trialId": r["trial_id"], "group": r["grp"], "totalBatches": r["steps_completed"], "trialRunId": r["trial_run_id"],
                            "endTime": A._ts(r["report_time"]), "metrics": {"avgMetrics": dbmod.U(r["metrics"], {}), "batchMetrics": dbmod.U(r["batch_metrics"])}})
        return [{"result": {"metrics": out}}]
    return handler


route("GET", "/api/v1/trials/time-series")(A.compare_trials)
route("GET", "/api/v1/trials/metrics/training_metrics")(_typed_metrics("training"))
route("GET", "/api/v1/trials/metrics/validation_metrics")(_typed_metrics("validation"))


def _typed_report(group: str, key: str) -> Any:
    def handler(ctx: Ctx, tid: str) -> Any:
        b = ctx.body or {}
        mm = b.get(key) or b
        ctx.body = {"group": group, "metrics": mm}
        return A.report_metrics(ctx, tid)
    return handler


route("POST", "/api/v1/trials/{tid}/training_metrics")(_typed_report("training", "trainingMetrics"))
route("POST", "/api/v1/trials/{tid}/validation_metrics")(_typed_report("validation", "validationMetrics"))


def _ckpt_metrics(ctx: Ctx, trial_id: int, steps: int) -> Any:
    grp = ctx.q("group")
    typ = {"TRIAL_SOURCE_INFO_TYPE_INFERENCE": "inference", "TRIAL_SOURCE_INFO_TYPE_FINE_TUNING": "fine_tuning"}.get(ctx.q("trialSourceInfoType") or "")
    rows = ctx.m.db.all("SELECT * FROM metrics WHERE trial_id=? AND steps_completed<=?" + (" AND grp=?" if grp else "") + " ORDER BY steps_completed, id",
                        (trial_id, steps) + ((grp,) if grp else ()))
    mets = [{"trialId": r["trial_id"], "group": r["grp"], "totalBatches": r["steps_completed"], "endTime": A._ts(r["report_time"]),
             "metrics": {"avgMetrics": dbmod.U(r["metrics"], {})}} for r in rows]
    return {"metrics": [{"checkpointUuid": None, "trialId": trial_id, "trialSourceInfoType": typ or "unspecified", "metricReports": mets}]}


@route("GET", "/api/v1/checkpoints/{uuid_}/metrics")
def metrics_by_checkpoint(ctx: Ctx, uuid_: str) -> Any:
    r = A._ckpt_row(ctx, uuid_)
    out = _ckpt_metrics(ctx, r["trial_id"], r["steps_completed"])
    out["metrics"][0]["checkpointUuid"] = uuid_
    return out


@route("GET", "/api/v1/models/{name}/versions/{version}/metrics")
def metrics_by_model_version(ctx: Ctx, name: str, version: str) -> Any:
    mv = A.get_model_version(ctx, name, version)["modelVersion"]
    uuid_ = (mv.get("checkpoint") or {}).get("uuid")
    if not uuid_:
        raise HTTPError(404, "the model version has no checkpoint")
    out = metrics_by_checkpoint(ctx, uuid_)
    out["metrics"][0].update(modelId=mv.get("model", {}).get("id"), modelVersion=mv.get("version"))
    return out


@route("GET", "/api/v1/trials/{tid}/workloads")
def trial_workloads(ctx: Ctx, tid: str) -> Any:
    """Training / validation reports and checkpoints of a trial as one time-ordered list of workloads."""
    A._trial_row(ctx, tid)
    name = ctx.q("metricType")
    want = {"METRIC_TYPE_TRAINING": {"training"}, "METRIC_TYPE_VALIDATION": {"validation"}}.get(name or "", {"training", "validation", "checkpoint"})
    items: List[Tuple[float, Dict[str, Any]]] = []
    for r in ctx.m.db.all("SELECT * FROM metrics WHERE trial_id=? AND grp IN ('training','validation')", (int(tid),)):
        if r["grp"] in want:
            items.append((r["report_time"], {r["grp"]: {"endTime": A._ts(r["report_time"]), "totalBatches": r["steps_completed"],
                                                        "metrics": {"avgMetrics": dbmod.U(r["metrics"], {})}}}))
    if "checkpoint" in want and ctx.q("includeBatchMetrics") != "only":
        for r in ctx.m.db.all("SELECT * FROM checkpoints WHERE trial_id=?", (int(tid),)):
            items.append((r["report_time"], {"checkpoint": {"uuid": r["uuid"], "endTime": A._ts(r["report_time"]), "totalBatches": r["steps_completed"],
                                                            "state": A._state("STATE", r["state"]), "resources": dbmod.U(r["resources"], {})}}))
    items.sort(key=lambda t: t[0], reverse=ctx.q("orderBy") == "ORDER_BY_DESC")
    rows, pg = A.page([w for _, w in items], ctx)
    return {"workloads": rows, "pagination": pg}


# ------------------------------------------------------------------------------------------------------------------ profiler
PROFILER_GROUPS = ("gpu", "cpu", "memory", "disk", "network", "profiler")


@route("GET", "/api/v1/trials/{tid}/profiler/available_series")
def profiler_series(ctx: Ctx, tid: str) -> Any:
    """Names of the system metrics the trial's profiler has reported, as ``(group, name, agent)`` labels."""
    labels, seen = [], set()
    for r in ctx.m.db.all("SELECT grp, metrics FROM metrics WHERE trial_id=? AND grp IN (%s)" % ",".join("?" * len(PROFILER_GROUPS)), (int(tid),) + PROFILER_GROUPS):
        for agent, vals in _by_agent(dbmod.U(r["metrics"], {})):
            for name in vals:
                key = (r["grp"], name, agent)
                if key not in seen:
                    seen.add(key)
                    labels.append({"trialId": int(tid), "name": name, "agentId": agent, "metricType": "PROFILER_METRIC_TYPE_SYSTEM", "group": r["grp"]})
    return [{"result": {"labels": labels}}]


def _by_agent(m: Dict[str, Any]) -> List[Tuple[str, Dict[str, Any]]]:
    """Profiler reports are ``{agent id: {metric: value}}`` or flat ``{metric: value}``."""
    if m and all(isinstance(v, dict) for v in m.values()):
        return [(str(a), v) for a, v in m.items()]
    return [("", m)]


@route("GET", "/api/v1/trials/{tid}/profiler/metrics")
def profiler_metrics(ctx: Ctx, tid: str) -> Any:
    name, agent = ctx.q("labels.name"), ctx.q("labels.agentId")
    if not name:
        raise HTTPError(400, "labels.name is required")
    times, batches, values = [], [], []
    for r in ctx.m.db.all("SELECT * FROM metrics WHERE trial_id=? AND grp IN (%s) ORDER BY report_time, id" % ",".join("?" * len(PROFILER_GROUPS)), (int(tid),) + PROFILER_GROUPS):
        for a, vals in _by_agent(dbmod.U(r["metrics"], {})):
            if name in vals and (not agent or a == agent) and isinstance(vals[name], (int, float)):
                times.append(A._ts(r["report_time"]))
                batches.append(r["steps_completed"])
                values.append(float(vals[name]))
    return [{"result": {"batch": {"values": values, "batches": batches, "timestamps": times,
                                  "labels": {"trialId": int(tid), "name": name, "agentId": agent or "", "metricType": "PROFILER_METRIC_TYPE_SYSTEM"}}}}]


@route("POST", "/api/v1/trials/profiler/metrics")
def post_profiler_batches(ctx: Ctx) -> Any:
    """Profiler samples in the batched form (``[{labels, values, batches, timestamps}]``): one metrics row per sample."""
    n = 0
    for batch in (ctx.body or {}).get("batches") or []:
        lab = batch.get("labels") or {}
        tid, name, agent = int(lab.get("trialId") or 0), str(lab.get("name") or ""), str(lab.get("agentId") or "")
        if not tid or not name:
            raise HTTPError(400, "every batch needs labels.trialId and labels.name")
        stamps = batch.get("timestamps") or []
        steps = batch.get("batches") or []
        for i, v in enumerate(batch.get("values") or []):
            ts = A._log_time(stamps[i]) if i < len(stamps) else None
            ctx.m.db.exec("INSERT INTO metrics(trial_id,trial_run_id,grp,steps_completed,metrics,batch_metrics,report_time) VALUES(?,?,?,?,?,?,?)",
                          (tid, 0, "profiler", int(steps[i]) if i < len(steps) else 0, dbmod.J({agent: {name: float(v)}}), None, ts or time.time()))
            n += 1
    return {"stored": n}


# --------------------------------------------------------------------------------------------------- allocations and tasks
@route("POST", "/api/v1/allocations/{aid}/acceleratorData")
def post_accelerator_data(ctx: Ctx, aid: str) -> Any:
    d = (ctx.body or {}).get("acceleratorData") or ctx.body or {}
    rows = M.get_attr(ctx.m, "allocation", aid, "acceleratorData") or []
    rows.append({"containerId": d.get("containerId"), "allocationId": aid, "nodeName": d.get("nodeName"), "taskId": d.get("taskId"),
                 "acceleratorType": d.get("acceleratorType"), "acceleratorUuids": list(d.get("acceleratorUuids") or []), "resourcePool": d.get("resourcePool")})
    M.set_attr(ctx.m, "allocation", aid, "acceleratorData", rows)
    if d.get("taskId"):
        idx = M.get_attr(ctx.m, "task", d["taskId"], "allocations") or []
        if aid not in idx:
            M.set_attr(ctx.m, "task", d["taskId"], "allocations", idx + [aid])
    return {}


@route("GET", "/api/v1/tasks/{task_id}/acceleratorData")
def task_accelerator_data(ctx: Ctx, task_id: str) -> Any:
    out = []
    for aid in M.get_attr(ctx.m, "task", task_id, "allocations") or []:
        out += M.get_attr(ctx.m, "allocation", aid, "acceleratorData") or []
    return {"acceleratorData": out}


@route("POST", "/api/v1/allocations/{aid}/notify_container_running")
def notify_container_running(ctx: Ctx, aid: str) -> Any:
    """Barrier of a multi-container allocation: every container reports in; the answer says whether all have.  One host per
    task here, so ``numPeers`` ranks of one node check in."""
    b = ctx.body or {}
    seen = set(M.get_attr(ctx.m, "allocation", aid, "containersRunning") or [])
    seen.add(str(b.get("rank", b.get("requestUuid", len(seen)))))
    M.set_attr(ctx.m, "allocation", aid, "containersRunning", sorted(seen))
    return {"data": {"allRunning": len(seen) >= int(b.get("numPeers") or 1), "running": len(seen)}}


@route("GET", "/api/v1/tasks/count")
def active_tasks_count(ctx: Ctx) -> Any:
    n = {"COMMAND": 0, "NOTEBOOK": 0, "SHELL": 0, "TENSORBOARD": 0}
    for t in ctx.m.commands.tasks.values():
        if t.state not in ("TERMINATED",):
            n[t.kind] = n.get(t.kind, 0) + 1
    return {"commands": n["COMMAND"], "notebooks": n["NOTEBOOK"], "shells": n["SHELL"], "tensorboards": n["TENSORBOARD"]}


@route("GET", "/api/v1/tasks/{task_id}")
def get_task(ctx: Ctx, task_id: str) -> Any:
    allocs = []
    if task_id.startswith("trial-"):
        for r in ctx.m.db.all("SELECT * FROM allocations WHERE task_id=? ORDER BY start_time", (task_id,)):
            allocs.append({"taskId": task_id, "allocationId": r["id"], "state": A._state("STATE", r["state"]), "isReady": r["state"] == "RUNNING",
                           "startTime": A._ts(r["start_time"]), "endTime": A._ts(r["end_time"]) if "end_time" in r.keys() else None})
        if not allocs and ctx.m.db.one("SELECT id FROM trials WHERE task_id=?", (task_id,)) is None:
            raise HTTPError(404, f"task {task_id} not found")
        return {"task": {"taskId": task_id, "taskType": "TASK_TYPE_TRIAL", "allocations": allocs}}
    t = M._any_task(ctx, task_id)
    return {"task": {"taskId": task_id, "taskType": f"TASK_TYPE_{t.kind}", "parentId": getattr(t, "parent_id", None),
                     "allocations": [{"taskId": task_id, "allocationId": task_id + ".1", "state": A._state("STATE", t.state), "isReady": t.state == "RUNNING"}]}}


def _log_fields(ctx: Ctx, task_id: str) -> Any:
    """The values the log filters of this task can take (one snapshot: a finished stream in the reference's terms)."""
    f: Dict[str, set] = {"agentIds": set(), "containerIds": set(), "rankIds": set(), "stdtypes": set(), "sources": set()}
    vals = ctx.m.db.log_field_values(task_id)
    for k, col in (("containerIds", "allocation_id"), ("rankIds", "rank_id"), ("stdtypes", "stdtype"), ("sources", "source")):
        f[k].update(v for v in vals.get(col, []) if v is not None)
    for r in ctx.m.db.all("SELECT DISTINCT id FROM allocations WHERE task_id=?", (task_id,)):
        aid = r["id"]
        live = ctx.m.allocations.get(aid) if hasattr(ctx.m, "allocations") else None
        if live is not None and getattr(live, "agent_id", None):
            f["agentIds"].add(live.agent_id)
    return [{"result": {k: sorted(v, key=str) for k, v in f.items()}}]


route("GET", "/api/v1/tasks/{task_id}/logs/fields")(_log_fields)
route("GET", "/api/v1/trials/{tid}/logs/fields")(lambda ctx, tid: _log_fields(ctx, f"trial-{tid}"))


@route("PUT", "/api/v1/notebooks/{task_id}/report_idle")
def notebook_idle(ctx: Ctx, task_id: str) -> Any:
    """The notebook's own idle watcher reports in: ``idle: false`` counts as activity for the idle timeout."""
    t = ctx.m.commands.get(task_id, "NOTEBOOK")
    if not (ctx.body or {}).get("idle", False):
        t.last_activity = time.time()
    return {}


@route("GET", "/api/v1/job-queues/stats")
def job_queue_stats(ctx: Ctx) -> Any:
    want = ctx.query.get("resourcePools") or []
    out = []
    with ctx.m.lock:
        for p in ctx.m.pools.values():
            if want and p.name not in want:
                continue
            rows = p.job_queue()
            jobs = {r["job_id"] for r in rows}
            sched = {r["job_id"] for r in rows if r["state"] == "SCHEDULED"}
            out.append({"resourcePool": p.name, "stats": {"queuedCount": len(jobs - sched), "scheduledCount": len(sched), "preemptibleCount": len(sched) if p.preemption else 0}})
    return {"results": out}


@route("GET", "/api/v1/agents/{agent_id}/slots/{slot_id}")
def get_slot(ctx: Ctx, agent_id: str, slot_id: str) -> Any:
    for s in A.get_slots(ctx, agent_id)["slots"]:
        if str(s.get("id")) == str(slot_id):
            return {"slot": s}
    raise HTTPError(404, f"slot {slot_id} of agent {agent_id} not found")


@route("GET", "/api/v1/master/telemetry")
def telemetry(ctx: Ctx) -> Any:
    t = getattr(ctx.m, "telemetry", None)
    out = {"enabled": bool(t and t.enabled), "segmentKey": (t.key if t and t.enabled else "")}
    if ctx.q("events") and t is not None:
        if not ctx.user["admin"]:
            raise HTTPError(403, "admin only")
        out["events"] = list(t.recent)
    return out


# ------------------------------------------------------------------------------------------------ unmanaged experiments / trials
def _by_external(ctx: Ctx, ext_exp: str) -> Optional[Any]:
    for r in ctx.m.db.all("SELECT id, state, config, original_config, owner_id, start_time, end_time, progress, archived, parent_id, notes, project_id, unmanaged, job_id FROM experiments WHERE unmanaged=1 ORDER BY id"):
        if (dbmod.U(r["config"], {}).get("data") or {}).get("external_experiment_id") == ext_exp:
            return r
    return None


@route("PUT", "/api/v1/experiments/by-external-id/{ext}")
def put_experiment(ctx: Ctx, ext: str) -> Any:
    """Create the unmanaged experiment known outside as ``ext``, or return it when it exists (idempotent)."""
    row = _by_external(ctx, ext)
    if row is not None:
        return {"experiment": A.exp_json(ctx.m, row), "config": dbmod.U(row["config"], {})}
    b = dict((ctx.body or {}).get("createExperimentRequest") or ctx.body or {})
    from determined_b200.master import expconf

    cfg = expconf.parse(b.get("config", "")) if isinstance(b.get("config"), str) else dict(b.get("config") or {})
    cfg.setdefault("data", {})["external_experiment_id"] = ext
    ctx.body = dict(b, config=cfg, unmanaged=True)
    return A.create_experiment(ctx)


def _trial_by_external(ctx: Ctx, eid: int, ext_trial: str) -> Optional[Any]:
    for t in ctx.m.db.all("SELECT * FROM trials WHERE experiment_id=? ORDER BY id", (eid,)):
        if dbmod.U(t["hparams"], {}).get("__external_trial_id") == ext_trial:
            return t
    return None


@route("GET", "/api/v1/trials/by-external-id/{ext_exp}/{ext_trial}")
def trial_by_external(ctx: Ctx, ext_exp: str, ext_trial: str) -> Any:
    e = _by_external(ctx, ext_exp)
    t = _trial_by_external(ctx, e["id"], ext_trial) if e is not None else None
    if t is None:
        raise HTTPError(404, f"no trial {ext_trial!r} in experiment {ext_exp!r}")
    return {"trial": A.trial_json(ctx.m, t)}


@route("POST", "/api/v1/trials")
def create_trial(ctx: Ctx) -> Any:
    b = ctx.body or {}
    if not b.get("unmanaged", True):
        raise HTTPError(400, "only unmanaged trials are created through the API")
    ctx.body = {"hparams": b.get("hparams") or {}}
    return A.create_unmanaged_trial(ctx, str(int(b.get("experimentId") or 0)))


@route("PUT", "/api/v1/trials")
def put_trial(ctx: Ctx) -> Any:
    b = ctx.body or {}
    req = b.get("createTrialRequest") or {}
    eid, ext = int(req.get("experimentId") or 0), str(b.get("externalTrialId") or "")
    if not ext:
        raise HTTPError(400, "externalTrialId is required")
    t = _trial_by_external(ctx, eid, ext)
    if t is not None:
        return {"trial": A.trial_json(ctx.m, t)}
    ctx.body = {"hparams": dict(req.get("hparams") or {}, __external_trial_id=ext)}
    return A.create_unmanaged_trial(ctx, str(eid))


@route("POST", "/api/v1/trials/{tid}/start")
def start_trial(ctx: Ctx, tid: str) -> Any:
    """An unmanaged trial (re)starts: a new run id, the latest checkpoint to resume from, the steps already done."""
    t = A._trial_row(ctx, tid)
    resume = bool((ctx.body or {}).get("resume"))
    run_id = int(M.get_attr(ctx.m, "trial", tid, "runId") or 0) + 1
    M.set_attr(ctx.m, "trial", tid, "runId", run_id)
    ck = ctx.m.db.one("SELECT uuid FROM checkpoints WHERE trial_id=? AND state='COMPLETED' ORDER BY steps_completed DESC LIMIT 1", (int(tid),)) if resume else None
    ctx.m.db.exec("UPDATE trials SET state='ACTIVE', last_activity=? WHERE id=?", (time.time(), int(tid)))
    return {"trialRunId": run_id, "latestCheckpoint": ck["uuid"] if ck else None, "stepsCompleted": int(t["total_batches"] or 0) if resume else 0}


# --------------------------------------------------------------------------------------------------------------------- users
@route("GET", "/api/v1/auth/user")
def current_user(ctx: Ctx) -> Any:
    return A.me(ctx)


@route("GET", "/api/v1/users/{username}/by-username")
def user_by_username(ctx: Ctx, username: str) -> Any:
    return M.get_user(ctx, username)


@route("PATCH", "/api/v1/users")
def patch_users(ctx: Ctx) -> Any:
    """Activate / deactivate many users at once."""
    if not ctx.user["admin"]:
        raise HTTPError(403, "admin only")
    b = ctx.body or {}
    results = []
    for uid in [int(x) for x in b.get("userIds") or []]:
        if ctx.m.db.one("SELECT id FROM users WHERE id=?", (uid,)) is None:
            results.append({"id": uid, "error": "not found"})
            continue
        ctx.m.db.exec("UPDATE users SET active=?, modified_at=? WHERE id=?", (int(bool(b.get("activate"))), time.time(), uid))
        results.append({"id": uid, "error": ""})
    return {"results": results}


@route("PATCH", "/api/v1/users/assignments")
def assign_multiple_groups(ctx: Ctx) -> Any:
    """Add and remove several users to / from several groups in one call."""
    if not ctx.user["admin"]:
        raise HTTPError(403, "admin only")
    b = ctx.body or {}
    uids = [int(u) for u in b.get("userIds") or []]
    for gid in [int(g) for g in list(b.get("addGroups") or []) + list(b.get("removeGroups") or [])]:
        if ctx.m.db.one("SELECT id FROM groups WHERE id=?", (gid,)) is None:
            raise HTTPError(404, f"group {gid} not found")
    for uid in uids:
        for gid in b.get("addGroups") or []:
            ctx.m.db.exec("INSERT OR IGNORE INTO group_members(group_id,user_id) VALUES(?,?)", (int(gid), uid))
        for gid in b.get("removeGroups") or []:
            ctx.m.db.exec("DELETE FROM group_members WHERE group_id=? AND user_id=?", (int(gid), uid))
    return {}


@route("GET", "/api/v1/users/setting")
def get_user_settings(ctx: Ctx) -> Any:
    s = M.get_attr(ctx.m, "user", ctx.user["id"], "settings") or {}
    return {"settings": [{"key": k.split("\0", 1)[1], "storagePath": k.split("\0", 1)[0], "value": v} for k, v in sorted(s.items())]}


@route("POST", "/api/v1/users/setting")
def post_user_settings(ctx: Ctx) -> Any:
    s = M.get_attr(ctx.m, "user", ctx.user["id"], "settings") or {}
    b = ctx.body or {}
    for it in b.get("settings") or ([b["setting"]] if b.get("setting") else []):
        k = f"{it.get('storagePath') or ''}\0{it.get('key') or ''}"
        if it.get("value") in (None, ""):
            s.pop(k, None)
        else:
            s[k] = it["value"]
    M.set_attr(ctx.m, "user", ctx.user["id"], "settings", s)
    return {}


@route("POST", "/api/v1/users/setting/reset")
def reset_user_settings(ctx: Ctx) -> Any:
    M.set_attr(ctx.m, "user", ctx.user["id"], "settings", {})
    return {}


@route("POST", "/api/v1/users/activity")
def post_user_activity(ctx: Ctx) -> Any:
    b = ctx.body or {}
    if b.get("entityType") in ("ENTITY_TYPE_PROJECT", "project") and b.get("entityId"):
        act = M.get_attr(ctx.m, "user", ctx.user["id"], "projectActivity") or {}
        act[str(int(b["entityId"]))] = time.time()
        M.set_attr(ctx.m, "user", ctx.user["id"], "projectActivity", act)
    return {}


@route("GET", "/api/v1/user/projects/activity")
def projects_by_activity(ctx: Ctx) -> Any:
    act = M.get_attr(ctx.m, "user", ctx.user["id"], "projectActivity") or {}
    out = []
    for pid, _ in sorted(act.items(), key=lambda kv: -kv[1])[:ctx.qi("limit", 0) or None]:
        try:
            out.append(A.get_project(ctx, pid)["project"])
        except HTTPError:
            continue
    return {"projects": out}


# ------------------------------------------------------------------------------------------------- workspaces and projects
def _pin(ctx: Ctx, wid: str, on: bool) -> Any:
    A.get_workspace(ctx, wid)
    pins = M.get_attr(ctx.m, "user", ctx.user["id"], "pinnedWorkspaces") or {}
    if on:
        pins[str(int(wid))] = time.time()
    else:
        pins.pop(str(int(wid)), None)
    M.set_attr(ctx.m, "user", ctx.user["id"], "pinnedWorkspaces", pins)
    return {}


route("POST", "/api/v1/workspaces/{wid}/pin")(lambda ctx, wid: _pin(ctx, wid, True))
route("POST", "/api/v1/workspaces/{wid}/unpin")(lambda ctx, wid: _pin(ctx, wid, False))


def _notes(ctx: Ctx, pid: str) -> List[Dict[str, str]]:
    p = A.get_project(ctx, pid)["project"]
    notes = p.get("notes")
    return list(notes) if isinstance(notes, list) else (M.get_attr(ctx.m, "project", int(pid), "notes") or [])


@route("POST", "/api/v1/projects/{pid}/notes")
def add_project_note(ctx: Ctx, pid: str) -> Any:
    b = ctx.body or {}
    note = b.get("note") or b
    notes = _notes(ctx, pid) + [{"name": str(note.get("name", "")), "contents": str(note.get("contents", ""))}]
    M.set_attr(ctx.m, "project", int(pid), "notes", notes)
    return {"notes": notes}


@route("PUT", "/api/v1/projects/{pid}/notes")
def put_project_notes(ctx: Ctx, pid: str) -> Any:
    A.get_project(ctx, pid)
    notes = [{"name": str(n.get("name", "")), "contents": str(n.get("contents", ""))} for n in (ctx.body or {}).get("notes") or []]
    M.set_attr(ctx.m, "project", int(pid), "notes", notes)
    return {"notes": notes}


@route("POST", "/api/v1/projects/{pid}/move")
def move_project(ctx: Ctx, pid: str) -> Any:
    dst = int((ctx.body or {}).get("destinationWorkspaceId") or 0)
    A.get_project(ctx, pid)
    if int(pid) == 1:
        raise HTTPError(400, "the Uncategorized project cannot be moved")
    if ctx.m.db.one("SELECT id FROM workspaces WHERE id=?", (dst,)) is None:
        raise HTTPError(404, f"workspace {dst} not found")
    ctx.m.db.exec("UPDATE projects SET workspace_id=? WHERE id=?", (dst, int(pid)))
    return {}


def _project_metrics(ctx: Ctx, pid: str) -> Dict[Tuple[str, str], List[float]]:
    """(group, metric name) -> every numeric value reported by the project's trials (latest report per trial and group)."""
    vals: Dict[Tuple[str, str], List[float]] = {}
    sql = ("SELECT m.grp g, m.metrics mm FROM metrics m JOIN trials t ON t.id=m.trial_id JOIN experiments e ON e.id=t.experiment_id "
           "WHERE e.project_id=? AND m.grp IN ('training','validation')")
    for r in ctx.m.db.all(sql, (int(pid),)):
        for k, v in dbmod.U(r["mm"], {}).items():
            if isinstance(v, (int, float)) and not isinstance(v, bool):
                vals.setdefault((r["g"], k), []).append(float(v))
    return vals


@route("GET", "/api/v1/projects/{pid}/experiments/metric-ranges")
def project_metric_ranges(ctx: Ctx, pid: str) -> Any:
    A.get_project(ctx, pid)
    return {"ranges": [{"metricsName": f"{g}.{k}", "min": min(v), "max": max(v)} for (g, k), v in sorted(_project_metrics(ctx, pid).items())]}


GENERAL_COLUMNS = [("id", "ID", "COLUMN_TYPE_NUMBER"), ("name", "Name", "COLUMN_TYPE_TEXT"), ("description", "Description", "COLUMN_TYPE_TEXT"),
                   ("tags", "Tags", "COLUMN_TYPE_TEXT"), ("forkedFrom", "Forked", "COLUMN_TYPE_NUMBER"), ("startTime", "Start time", "COLUMN_TYPE_DATE"),
                   ("duration", "Duration", "COLUMN_TYPE_NUMBER"), ("numTrials", "Trial count", "COLUMN_TYPE_NUMBER"), ("state", "State", "COLUMN_TYPE_TEXT"),
                   ("searcherType", "Searcher", "COLUMN_TYPE_TEXT"), ("resourcePool", "Resource pool", "COLUMN_TYPE_TEXT"), ("progress", "Progress", "COLUMN_TYPE_NUMBER"),
                   ("checkpointSize", "Checkpoint size", "COLUMN_TYPE_NUMBER"), ("checkpointCount", "Checkpoints", "COLUMN_TYPE_NUMBER"), ("user", "User", "COLUMN_TYPE_TEXT"),
                   ("searcherMetric", "Searcher metric", "COLUMN_TYPE_TEXT"), ("searcherMetricsVal", "Searcher metric value", "COLUMN_TYPE_NUMBER"),
                   ("externalExperimentId", "External experiment id", "COLUMN_TYPE_TEXT"), ("externalTrialId", "External trial id", "COLUMN_TYPE_TEXT")]


@route("GET", "/api/v1/projects/{pid}/columns")
def project_columns(ctx: Ctx, pid: str) -> Any:
    """What an experiment table of this project can show: the general columns, one per hyperparameter (dotted path) and one
    per reported metric."""
    A.get_project(ctx, pid)
    cols = [{"column": c, "location": "LOCATION_TYPE_EXPERIMENT", "type": t, "displayName": d} for c, d, t in GENERAL_COLUMNS]
    hps: Dict[str, str] = {}
    for r in ctx.m.db.all("SELECT t.hparams h FROM trials t JOIN experiments e ON e.id=t.experiment_id WHERE e.project_id=?", (int(pid),)):
        for k, v in A._flat_hparams(dbmod.U(r["h"], {})).items():
            kind = "COLUMN_TYPE_NUMBER" if isinstance(v, (int, float)) and not isinstance(v, bool) else "COLUMN_TYPE_TEXT"
            hps[k] = kind if hps.get(k, kind) == kind else "COLUMN_TYPE_UNSPECIFIED"
    cols += [{"column": f"hp.{k}", "location": "LOCATION_TYPE_HYPERPARAMETERS", "type": t} for k, t in sorted(hps.items())]
    for (g, k) in sorted(_project_metrics(ctx, pid)):
        cols.append({"column": f"{g}.{k}", "location": "LOCATION_TYPE_TRAINING" if g == "training" else "LOCATION_TYPE_VALIDATIONS", "type": "COLUMN_TYPE_NUMBER"})
    return {"columns": cols}


# --------------------------------------------------------------------------------------------------------------------- roles
@route("POST", "/api/v1/roles/search/by-ids")
def roles_by_ids(ctx: Ctx) -> Any:
    want = {int(x) for x in (ctx.body or {}).get("roleIds") or []}
    return {"roles": [M.describe_role(ctx, str(r["roleId"])) for r in A.ROLES if r["roleId"] in want]}


@route("POST", "/api/v1/roles/search/by-assignability")
def roles_by_assignability(ctx: Ctx) -> Any:
    """Roles that may be assigned at a scope: any role cluster-wide; inside a workspace only those without a permission
    that exists cluster-wide only."""
    wid = (ctx.body or {}).get("workspaceId")
    out = []
    for r in A.ROLES:
        if wid and set(M.ROLE_PERMISSIONS.get(r["name"], [])) & M._GLOBAL_ONLY:
            continue
        out.append(M._role_json(r["name"]))
    return {"roles": out, "pagination": {"total": len(out)}}


# -------------------------------------------------------------------------------------------------------------------- search
def _match_filter(e: Dict[str, Any], flt: Any) -> bool:
    """The table filter language of the experiment list: ``{"filterGroup": {"conjunction": "and"|"or", "children": [...]}}``
    with leaves ``{"kind": "field", "columnName": ..., "operator": ..., "value": ...}``."""
    if not flt:
        return True
    node = flt.get("filterGroup", flt)
    kind = node.get("kind", "group" if "children" in node else "field")
    if kind == "group":
        kids = [_match_filter(e, c) for c in node.get("children") or []]
        if not kids:
            return True
        return all(kids) if node.get("conjunction", "and") == "and" else any(kids)
    col, op, val = node.get("columnName"), node.get("operator", "="), node.get("value")
    have: Any = e.get(col)
    if col and col.startswith("hp."):
        have = A._flat_hparams((e.get("config") or {}).get("hyperparameters") or {}).get(col[3:])
        if isinstance(have, dict) and "val" in have:
            have = have["val"]
    if col == "tags":
        have = e.get("labels")
    if col == "user":
        have = e.get("username")
    if op in ("isEmpty", "notEmpty"):
        return (have in (None, "", [])) == (op == "isEmpty")
    if op in ("contains", "notContains"):
        hit = (val in have) if isinstance(have, list) else str(val).lower() in str(have or "").lower()
        return hit == (op == "contains")
    try:
        a, b = (float(have), float(val)) if isinstance(have, (int, float)) and not isinstance(have, bool) else (str(have), str(val))
    except (TypeError, ValueError):
        a, b = str(have), str(val)
    return {"=": a == b, "!=": a != b, ">": a > b, ">=": a >= b, "<": a < b, "<=": a <= b}.get(op, False)


def _searched(ctx: Ctx) -> List[Dict[str, Any]]:
    pid = ctx.qi("projectId", 0)
    flt = json.loads(ctx.q("filter")) if ctx.q("filter") else None
    out = []
    for r in ctx.m.db.all("SELECT id, state, config, original_config, owner_id, start_time, end_time, progress, archived, parent_id, notes, project_id, unmanaged, job_id FROM experiments ORDER BY id"):
        if pid and r["project_id"] != pid:
            continue
        e = A.exp_json(ctx.m, r)
        if _match_filter(e, flt):
            out.append(e)
    sort = ctx.q("sort")
    if sort:
        for part in reversed(sort.split(",")):
            col, _, direction = part.partition("=")
            out.sort(key=lambda e: (e.get(col) is None, e.get(col) if isinstance(e.get(col), (int, float)) else str(e.get(col))), reverse=direction == "desc")
    return out


@route("GET", "/api/v1/experiments-search")
def search_experiments(ctx: Ctx) -> Any:
    """Experiments of a project with their best trial, filtered and sorted the way the experiment table asks."""
    exps, pg = A.page(_searched(ctx), ctx)
    out = []
    for e in exps:
        smaller = bool(((e.get("config") or {}).get("searcher") or {}).get("smaller_is_better", True))
        ts = [A.trial_json(ctx.m, t) for t in ctx.m.db.all("SELECT * FROM trials WHERE experiment_id=?", (e["id"],))]
        scored = [t for t in ts if t.get("searcherMetricValue") is not None]
        best = (min if smaller else max)(scored, key=lambda t: t["searcherMetricValue"]) if scored else (ts[0] if ts else None)
        out.append({"experiment": e, "bestTrial": best})
    return {"experiments": out, "pagination": pg}


@route("GET", "/api/v1/runs")
def search_runs(ctx: Ctx) -> Any:
    """Trials as flat runs (one row per trial with its experiment's facts folded in)."""
    runs = []
    for e in _searched(ctx):
        for t in ctx.m.db.all("SELECT * FROM trials WHERE experiment_id=? ORDER BY id", (e["id"],)):
            tj = A.trial_json(ctx.m, t)
            runs.append({"id": tj["id"], "startTime": tj["startTime"], "endTime": tj["endTime"], "state": tj["state"], "labels": e["labels"],
                         "hyperparameters": tj["hparams"], "searcherMetricValue": tj.get("searcherMetricValue"), "projectId": e["projectId"],
                         "userId": e["userId"], "parentArchived": e["archived"], "restarts": tj.get("restarts", 0),
                         "experiment": {"id": e["id"], "name": e["name"], "description": e["description"], "searcherType": e["searcherType"],
                                        "searcherMetric": e["searcherMetric"], "resourcePool": e["resourcePool"], "unmanaged": e["unmanaged"]}})
    rows, pg = A.page(runs, ctx)
    return {"runs": rows, "pagination": pg}


@route("POST", "/api/v1/runs/move")
def move_runs(ctx: Ctx) -> Any:
    """Runs move with their experiment (a multi-trial experiment cannot be split across projects)."""
    b = ctx.body or {}
    dst = int(b.get("destinationProjectId") or 0)
    if ctx.m.db.one("SELECT id FROM projects WHERE id=?", (dst,)) is None:
        raise HTTPError(404, f"project {dst} not found")
    results = []
    for rid in [int(x) for x in b.get("runIds") or []]:
        t = ctx.m.db.one("SELECT experiment_id FROM trials WHERE id=?", (rid,))
        if t is None:
            results.append({"id": rid, "error": "not found"})
            continue
        n = ctx.m.db.one("SELECT COUNT(*) c FROM trials WHERE experiment_id=?", (t["experiment_id"],))["c"]
        others = [x for x in b.get("runIds") or []]
        mine = [r["id"] for r in ctx.m.db.all("SELECT id FROM trials WHERE experiment_id=?", (t["experiment_id"],))]
        if n > 1 and not set(mine) <= {int(x) for x in others} and b.get("skipMultitrial"):
            results.append({"id": rid, "error": "run belongs to a multi-trial experiment; move all of its runs together"})
            continue
        ctx.m.db.exec("UPDATE experiments SET project_id=? WHERE id=?", (dst, t["experiment_id"]))
        results.append({"id": rid, "error": ""})
    return {"results": results}
