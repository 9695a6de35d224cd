"""More of the REST surface: the routes behind the CLI commands that edit, archive, bind and account.

Parity (reference handlers): ``api_project.go`` (patch / archive / unarchive), ``api_workspace.go`` (patch incl. agent user,
checkpoint-storage override and default pools; archive; ``ListRPsBoundToWorkspace``), ``api_resourcepool.go`` workspace
bindings, ``api_rbac.go`` (role details, roles of a group, permission summary), ``api_resources.go`` allocation accounting
(``ResourceAllocationRaw`` / ``ResourceAllocationAggregated``), ``api_master.go`` (master logs, ``PatchMasterConfig`` for the
log level), ``api_experiment.go`` (``DeleteTensorboardFiles``, ``PutExperimentRetainLogs``), ``api_trials.go``
(``PutTrialRetainLogs``), ``api_tasks.go`` / ``api_generic_tasks.go`` (task config, generic tasks, pause / unpause),
``api_user.go`` (agent user), and the OAuth client registry the ``det oauth`` commands talk to.

Registered on import by ``master/api.py`` with the same ``route`` decorator and ``Ctx``.
"""
import collections
import datetime
import json
import logging
import os
import secrets
import shutil
import signal
import time
from typing import Any, Deque, Dict, List, Optional, Tuple

from determined_b200.master import core
from determined_b200.master import db as dbmod
from determined_b200.master.api import Ctx, HTTPError, ROLES, _exp_row, _trial_row, _ts, page, route, user_json

logger = logging.getLogger("determined.master.api")

def _ensure_schema(m: core.Master) -> None:
    """The tables used here (``attrs``: free-form attributes of workspaces / users / trials / tasks; ``rp_bindings``;
    ``oauth_clients``) are part of ``master/db.py``'s schema."""


def get_attr(m: core.Master, kind: str, id_: Any, key: str, default: Any = None) -> Any:
    _ensure_schema(m)
    r = m.db.one("SELECT value FROM attrs WHERE kind=? AND id=? AND key=?", (kind, str(id_), key))
    return default if r is None else json.loads(r["value"])


def set_attr(m: core.Master, kind: str, id_: Any, key: str, value: Any) -> None:
    _ensure_schema(m)
    if value is None:
        m.db.exec("DELETE FROM attrs WHERE kind=? AND id=? AND key=?", (kind, str(id_), key))
    else:
        m.db.exec("INSERT OR REPLACE INTO attrs(kind,id,key,value) VALUES(?,?,?,?)", (kind, str(id_), key, json.dumps(value)))


def _admin(ctx: Ctx) -> None:
    if not ctx.user.get("admin"):
        raise HTTPError(403, "only admin privileged users can do this")


# ---------------------------------------------------------------------------------------------------------------------
# projects and workspaces
# ---------------------------------------------------------------------------------------------------------------------
def _project(ctx: Ctx, pid: str) -> Any:
    r = ctx.m.db.one("SELECT * FROM projects WHERE id=?", (int(pid),))
    if r is None:
        raise HTTPError(404, "project not found")
    return r


def _project_json(ctx: Ctx, r: Any) -> Dict[str, Any]:
    return {"id": r["id"], "name": r["name"], "workspaceId": r["workspace_id"], "description": r["description"], "archived": bool(r["archived"]),
            "immutable": bool(r["immutable"]), "userId": r["user_id"], "notes": dbmod.U(r["notes"], []),
            "numExperiments": ctx.m.db.one("SELECT COUNT(*) c FROM experiments WHERE project_id=?", (r["id"],))["c"]}


@route("PATCH", "/api/v1/projects/{pid}")
def patch_project(ctx: Ctx, pid: str) -> Any:
    r = _project(ctx, pid)
    if r["immutable"]:
        raise HTTPError(400, "the built-in project cannot be changed")
    b = (ctx.body or {}).get("project", ctx.body or {})
    for key, col in (("name", "name"), ("description", "description")):
        if b.get(key) is not None:
            try:
                ctx.m.db.exec(f"UPDATE projects SET {col}=? WHERE id=?", (b[key], int(pid)))
            except Exception as e:  # noqa: BLE001  (unique name per workspace)
                raise HTTPError(409, f"cannot change the project: {e}")
    if b.get("notes") is not None:
        ctx.m.db.exec("UPDATE projects SET notes=? WHERE id=?", (dbmod.J(b["notes"]), int(pid)))
    return {"project": _project_json(ctx, _project(ctx, pid))}


def _set_project_archived(ctx: Ctx, pid: str, flag: bool) -> Any:
    r = _project(ctx, pid)
    if r["immutable"]:
        raise HTTPError(400, "the built-in project cannot be archived")
    ctx.m.db.exec("UPDATE projects SET archived=? WHERE id=?", (int(flag), int(pid)))
    return {}


route("POST", "/api/v1/projects/{pid}/archive")(lambda ctx, pid: _set_project_archived(ctx, pid, True))
route("POST", "/api/v1/projects/{pid}/unarchive")(lambda ctx, pid: _set_project_archived(ctx, pid, False))


def _workspace(ctx: Ctx, wid: str) -> Any:
    r = ctx.m.db.one("SELECT * FROM workspaces WHERE id=?", (int(wid),))
    if r is None:
        raise HTTPError(404, "workspace not found")
    return r


def workspace_json(ctx: Ctx, r: Any) -> Dict[str, Any]:
    m = ctx.m
    out = {"id": r["id"], "name": r["name"], "archived": bool(r["archived"]), "immutable": bool(r["immutable"]), "userId": r["user_id"],
           "numProjects": m.db.one("SELECT COUNT(*) c FROM projects WHERE workspace_id=?", (r["id"],))["c"]}
    for key in ("agentUserGroup", "checkpointStorageConfig", "defaultComputePool", "defaultAuxPool"):
        v = get_attr(m, "workspace", r["id"], key)
        if v is not None:
            out[key] = v
    return out


@route("PATCH", "/api/v1/workspaces/{wid}")
def patch_workspace(ctx: Ctx, wid: str) -> Any:
    r = _workspace(ctx, wid)
    if r["immutable"]:
        raise HTTPError(400, "the built-in workspace cannot be changed")
    b = (ctx.body or {}).get("workspace", ctx.body or {})
    if b.get("name") is not None:
        try:
            ctx.m.db.exec("UPDATE workspaces SET name=? WHERE id=?", (b["name"], int(wid)))
        except Exception as e:  # noqa: BLE001
            raise HTTPError(409, f"cannot rename the workspace: {e}")
    for key in ("defaultComputePool", "defaultAuxPool"):
        if key in b:
            if b[key] and b[key] not in ctx.m.pools:
                raise HTTPError(400, f"unknown resource pool {b[key]!r}")
            if b[key] and not _pool_usable(ctx.m, b[key], int(wid)):
                raise HTTPError(400, f"resource pool {b[key]!r} is not bound to this workspace")
            set_attr(ctx.m, "workspace", wid, key, b[key] or None)
    if "agentUserGroup" in b:
        _admin(ctx)
        set_attr(ctx.m, "workspace", wid, "agentUserGroup", b["agentUserGroup"] or None)
    if "checkpointStorageConfig" in b:
        cs = b["checkpointStorageConfig"]
        if cs and not isinstance(cs, dict):
            raise HTTPError(400, "checkpointStorageConfig must be an object")
        if cs and cs.get("type") not in ("shared_fs", "directory", "s3", "gcs", "azure"):
            raise HTTPError(400, f"unknown checkpoint storage type {cs.get('type')!r}")
        set_attr(ctx.m, "workspace", wid, "checkpointStorageConfig", cs or None)
    return {"workspace": workspace_json(ctx, _workspace(ctx, wid))}


def _set_workspace_archived(ctx: Ctx, wid: str, flag: bool) -> Any:
    r = _workspace(ctx, wid)
    if r["immutable"]:
        raise HTTPError(400, "the built-in workspace cannot be archived")
    ctx.m.db.exec("UPDATE workspaces SET archived=? WHERE id=?", (int(flag), int(wid)))
    ctx.m.db.exec("UPDATE projects SET archived=? WHERE workspace_id=?", (int(flag), int(wid)))
    return {}


route("POST", "/api/v1/workspaces/{wid}/archive")(lambda ctx, wid: _set_workspace_archived(ctx, wid, True))
route("POST", "/api/v1/workspaces/{wid}/unarchive")(lambda ctx, wid: _set_workspace_archived(ctx, wid, False))


# ---------------------------------------------------------------------------------------------------------------------
# resource pool <-> workspace bindings: a pool without bindings is open to every workspace; with bindings only to those
# ---------------------------------------------------------------------------------------------------------------------
def _bound(m: core.Master, pool: str) -> List[int]:
    _ensure_schema(m)
    return [r["workspace_id"] for r in m.db.all("SELECT workspace_id FROM rp_bindings WHERE pool=? ORDER BY workspace_id", (pool,))]


def _pool_usable(m: core.Master, pool: str, workspace_id: int) -> bool:
    b = _bound(m, pool)
    return not b or workspace_id in b


def pool_allowed_for_project(m: core.Master, pool: str, project_id: int) -> bool:
    r = m.db.one("SELECT workspace_id FROM projects WHERE id=?", (project_id,))
    return r is None or _pool_usable(m, pool, r["workspace_id"])


def default_pool_for_project(m: core.Master, project_id: int, slots: int) -> Optional[str]:
    """The workspace's default compute / aux pool, when it names one."""
    r = m.db.one("SELECT workspace_id FROM projects WHERE id=?", (project_id,))
    if r is None:
        return None
    return get_attr(m, "workspace", r["workspace_id"], "defaultComputePool" if slots > 0 else "defaultAuxPool")


def _workspace_ids(ctx: Ctx, b: Dict[str, Any]) -> List[int]:
    ids = [int(x) for x in b.get("workspaceIds") or []]
    for name in b.get("workspaceNames") or []:
        r = ctx.m.db.one("SELECT id FROM workspaces WHERE name=?", (name,))
        if r is None:
            raise HTTPError(404, f"workspace {name!r} not found")
        ids.append(r["id"])
    for i in ids:
        if ctx.m.db.one("SELECT id FROM workspaces WHERE id=?", (i,)) is None:
            raise HTTPError(404, f"workspace {i} not found")
    return ids


def _pool(ctx: Ctx, name: str) -> str:
    if name not in ctx.m.pools:
        raise HTTPError(404, f"resource pool {name!r} not found")
    return name


@route("GET", "/api/v1/resource-pools/{name}/workspace-bindings")
def list_bindings(ctx: Ctx, name: str) -> Any:
    ids = _bound(ctx.m, _pool(ctx, name))
    items, pg = page(ids, ctx)
    return {"workspaceIds": items, "pagination": pg}


@route("POST", "/api/v1/resource-pools/{name}/workspace-bindings")
def add_bindings(ctx: Ctx, name: str) -> Any:
    _admin(ctx)
    _pool(ctx, name)
    if name == "default" and False:
        pass
    for i in _workspace_ids(ctx, ctx.body or {}):
        ctx.m.db.exec("INSERT OR IGNORE INTO rp_bindings(pool, workspace_id) VALUES(?,?)", (name, i))
    return {}


@route("PUT", "/api/v1/resource-pools/{name}/workspace-bindings")
def replace_bindings(ctx: Ctx, name: str) -> Any:
    _admin(ctx)
    _pool(ctx, name)
    ids = _workspace_ids(ctx, ctx.body or {})
    _ensure_schema(ctx.m)
    ctx.m.db.exec("DELETE FROM rp_bindings WHERE pool=?", (name,))
    for i in ids:
        ctx.m.db.exec("INSERT OR IGNORE INTO rp_bindings(pool, workspace_id) VALUES(?,?)", (name, i))
    return {}


@route("DELETE", "/api/v1/resource-pools/{name}/workspace-bindings")
def remove_bindings(ctx: Ctx, name: str) -> Any:
    _admin(ctx)
    _pool(ctx, name)
    for i in _workspace_ids(ctx, ctx.body or {}):
        ctx.m.db.exec("DELETE FROM rp_bindings WHERE pool=? AND workspace_id=?", (name, i))
    return {}


@route("GET", "/api/v1/workspaces/{wid}/available-resource-pools")
def workspace_pools(ctx: Ctx, wid: str) -> Any:
    _workspace(ctx, wid)
    names = [p for p in ctx.m.pools if _pool_usable(ctx.m, p, int(wid))]
    items, pg = page(names, ctx)
    return {"resourcePools": [{"name": n} for n in items], "pagination": pg}


# ---------------------------------------------------------------------------------------------------------------------
# RBAC read side
# ---------------------------------------------------------------------------------------------------------------------
_VIEW = ["PERMISSION_TYPE_VIEW_EXPERIMENT_ARTIFACTS", "PERMISSION_TYPE_VIEW_EXPERIMENT_METADATA", "PERMISSION_TYPE_VIEW_PROJECT",
         "PERMISSION_TYPE_VIEW_WORKSPACE", "PERMISSION_TYPE_VIEW_MODEL_REGISTRY", "PERMISSION_TYPE_VIEW_NSC", "PERMISSION_TYPE_VIEW_TEMPLATES"]
_EDIT = ["PERMISSION_TYPE_CREATE_EXPERIMENT", "PERMISSION_TYPE_UPDATE_EXPERIMENT", "PERMISSION_TYPE_UPDATE_EXPERIMENT_METADATA",
         "PERMISSION_TYPE_DELETE_EXPERIMENT", "PERMISSION_TYPE_CREATE_NSC", "PERMISSION_TYPE_UPDATE_NSC", "PERMISSION_TYPE_CREATE_PROJECT",
         "PERMISSION_TYPE_UPDATE_PROJECT", "PERMISSION_TYPE_DELETE_PROJECT", "PERMISSION_TYPE_EDIT_MODEL_REGISTRY",
         "PERMISSION_TYPE_CREATE_MODEL_REGISTRY", "PERMISSION_TYPE_DELETE_MODEL_REGISTRY", "PERMISSION_TYPE_CREATE_TEMPLATES",
         "PERMISSION_TYPE_UPDATE_TEMPLATES", "PERMISSION_TYPE_DELETE_TEMPLATES"]
_WS_ADMIN = ["PERMISSION_TYPE_UPDATE_WORKSPACE", "PERMISSION_TYPE_DELETE_WORKSPACE", "PERMISSION_TYPE_SET_WORKSPACE_AGENT_USER_GROUP",
             "PERMISSION_TYPE_SET_WORKSPACE_CHECKPOINT_STORAGE_CONFIG", "PERMISSION_TYPE_SET_WORKSPACE_DEFAULT_RESOURCE_POOL",
             "PERMISSION_TYPE_ASSIGN_ROLES", "PERMISSION_TYPE_UPDATE_ROLES"]
_CLUSTER = ["PERMISSION_TYPE_ADMINISTRATE_USER", "PERMISSION_TYPE_UPDATE_GROUP", "PERMISSION_TYPE_CREATE_WORKSPACE", "PERMISSION_TYPE_UPDATE_AGENTS",
            "PERMISSION_TYPE_UPDATE_MASTER_CONFIG", "PERMISSION_TYPE_VIEW_MASTER_LOGS", "PERMISSION_TYPE_VIEW_CLUSTER_USAGE",
            "PERMISSION_TYPE_CONTROL_STRICT_JOB_QUEUE", "PERMISSION_TYPE_EDIT_WEBHOOKS", "PERMISSION_TYPE_MODIFY_RP_WORKSPACE_BINDINGS",
            "PERMISSION_TYPE_VIEW_SENSITIVE_AGENT_INFO", "PERMISSION_TYPE_VIEW_MASTER_CONFIG"]
ROLE_PERMISSIONS: Dict[str, List[str]] = {
    "ClusterAdmin": _VIEW + _EDIT + _WS_ADMIN + _CLUSTER,
    "WorkspaceAdmin": _VIEW + _EDIT + _WS_ADMIN,
    "WorkspaceCreator": ["PERMISSION_TYPE_CREATE_WORKSPACE"],
    "Viewer": _VIEW,
    "Editor": _VIEW + _EDIT,
    "ModelRegistryViewer": ["PERMISSION_TYPE_VIEW_MODEL_REGISTRY"],
}
_GLOBAL_ONLY = set(_CLUSTER) | {"PERMISSION_TYPE_CREATE_WORKSPACE"}


def _role_json(name: str) -> Dict[str, Any]:
    rid = next((r["roleId"] for r in ROLES if r["name"] == name), None)
    return {"roleId": rid, "name": name,
            "permissions": [{"id": p, "name": p.replace("PERMISSION_TYPE_", "").replace("_", " ").lower(), "scopeTypeMask": {"cluster": True, "workspace": p not in _GLOBAL_ONLY}}
                            for p in ROLE_PERMISSIONS.get(name, [])]}


@route("GET", "/api/v1/roles/{rid}")
def describe_role(ctx: Ctx, rid: str) -> Any:
    role = next((r for r in ROLES if str(r["roleId"]) == rid or r["name"] == rid), None)
    if role is None:
        raise HTTPError(404, f"role {rid!r} not found")
    _ensure_schema(ctx.m)
    assigned = ctx.m.db.all("SELECT * FROM role_assignments WHERE role=?", (role["name"],))
    return {"role": _role_json(role["name"]),
            "userRoleAssignments": [{"userId": a["user_id"], "roleAssignment": {"role": role, "scopeWorkspaceId": a["workspace_id"]}} for a in assigned if a["user_id"]],
            "groupRoleAssignments": [{"groupId": a["group_id"], "roleAssignment": {"role": role, "scopeWorkspaceId": a["workspace_id"]}} for a in assigned if a["group_id"]]}


@route("GET", "/api/v1/roles/search/by-group/{gid}")
def roles_of_group(ctx: Ctx, gid: str) -> Any:
    ids = {r["name"]: r["roleId"] for r in ROLES}
    rows = ctx.m.db.all("SELECT * FROM role_assignments WHERE group_id=?", (int(gid),))
    return {"roles": [{"role": {"roleId": ids.get(r["role"]), "name": r["role"]}, "scopeWorkspaceId": r["workspace_id"], "groupId": r["group_id"]} for r in rows]}


@route("GET", "/api/v1/permissions/summary")
def my_permissions(ctx: Ctx) -> Any:
    uid = ctx.user["id"]
    rows = list(ctx.m.db.all("SELECT role, workspace_id FROM role_assignments WHERE user_id=?", (uid,)))
    rows += list(ctx.m.db.all("SELECT role, workspace_id FROM role_assignments WHERE group_id IN (SELECT group_id FROM group_members WHERE user_id=?)", (uid,)))
    by_role: Dict[str, Dict[str, Any]] = {}
    if ctx.user.get("admin"):
        by_role["ClusterAdmin"] = {"scopeCluster": True, "scopeWorkspaceIds": []}
    for r in rows:
        e = by_role.setdefault(r["role"], {"scopeCluster": False, "scopeWorkspaceIds": []})
        if r["workspace_id"] is None:
            e["scopeCluster"] = True
        elif r["workspace_id"] not in e["scopeWorkspaceIds"]:
            e["scopeWorkspaceIds"].append(r["workspace_id"])
    ids = {r["name"]: r["roleId"] for r in ROLES}
    return {"roles": [_role_json(n) for n in by_role],
            "assignments": [{"roleId": ids.get(n), "scopeCluster": e["scopeCluster"], "scopeWorkspaceIds": sorted(e["scopeWorkspaceIds"])} for n, e in by_role.items()]}


# ---------------------------------------------------------------------------------------------------------------------
# resource accounting from the allocations table
# ---------------------------------------------------------------------------------------------------------------------
def _parse_time(s: Optional[str], default: float) -> float:
    if not s:
        return default
    try:
        return float(s)
    except ValueError:
        pass
    s = s.replace("Z", "+00:00")
    try:
        d = datetime.datetime.fromisoformat(s)
    except ValueError:
        raise HTTPError(400, f"cannot parse time {s!r}")
    if d.tzinfo is None:
        d = d.replace(tzinfo=datetime.timezone.utc)
    return d.timestamp()


def _allocation_rows(ctx: Ctx, t0: float, t1: float) -> List[Dict[str, Any]]:
    out = []
    now = time.time()
    for a in ctx.m.db.all("SELECT * FROM allocations WHERE start_time IS NOT NULL AND start_time < ? AND COALESCE(end_time, ?) > ? ORDER BY start_time", (t1, now, t0)):
        t = ctx.m.db.one("SELECT t.experiment_id, e.owner_id, e.config FROM trials t JOIN experiments e ON e.id=t.experiment_id WHERE t.id=?", (a["trial_id"],)) if a["trial_id"] else None
        owner = ctx.m.db.one("SELECT username FROM users WHERE id=?", (t["owner_id"],)) if t else None
        cfg = dbmod.U(t["config"], {}) if t else {}
        start, end = max(a["start_time"], t0), min(a["end_time"] or now, t1)
        out.append({"kind": "trial" if a["trial_id"] else "task", "allocationId": a["id"], "taskId": a["task_id"], "experimentId": t["experiment_id"] if t else None,
                    "username": owner["username"] if owner else None, "labels": cfg.get("labels") or [], "resourcePool": a["resource_pool"] or "default",
                    "slots": a["slots"] or 0, "startTime": _ts(a["start_time"]), "endTime": _ts(a["end_time"]),
                    "seconds": max(end - start, 0.0) * (a["slots"] or 0)})
    return out


@route("GET", "/api/v1/resources/allocation/raw")
def allocation_raw(ctx: Ctx) -> Any:
    """One entry per allocation that overlaps ``[timestampAfter, timestampBefore]``; ``seconds`` are slot-seconds inside it."""
    t0 = _parse_time(ctx.q("timestampAfter"), 0.0)
    t1 = _parse_time(ctx.q("timestampBefore"), time.time())
    return {"resourceEntries": _allocation_rows(ctx, t0, t1)}


@route("GET", "/api/v1/resources/allocation/aggregated")
def allocation_aggregated(ctx: Ctx) -> Any:
    """Slot-seconds per day (or month) between ``startDate`` and ``endDate`` (inclusive, ``YYYY-MM-DD`` / ``YYYY-MM``), in
    total and broken down by user, label, experiment and resource pool."""
    period = ctx.q("period", "RESOURCE_ALLOCATION_AGGREGATION_PERIOD_DAILY")
    monthly = period.endswith("MONTHLY")
    s, e = ctx.q("startDate"), ctx.q("endDate")
    if not s or not e:
        raise HTTPError(400, "startDate and endDate are required")

    def day(text: str, end: bool) -> datetime.datetime:
        parts = [int(x) for x in text.split("-")]
        if len(parts) == 2:
            parts.append(1)
            d = datetime.datetime(*parts, tzinfo=datetime.timezone.utc)
            if end:
                d = (d.replace(day=28) + datetime.timedelta(days=4)).replace(day=1) - datetime.timedelta(days=1)
            return d
        return datetime.datetime(*parts, tzinfo=datetime.timezone.utc)

    try:
        d0, d1 = day(s, False), day(e, True)
    except (ValueError, TypeError):
        raise HTTPError(400, "dates must look like YYYY-MM-DD (or YYYY-MM)")
    entries = []
    # one look at the table says where there is anything to add up: periods outside that span are written without a query
    span = ctx.m.db.one("SELECT MIN(start_time) lo, MAX(COALESCE(end_time, ?)) hi FROM allocations WHERE start_time IS NOT NULL", (time.time(),))
    lo, hi = (span["lo"], span["hi"]) if span is not None and span["lo"] is not None else (None, None)
    cur = d0
    while cur <= d1:
        if monthly:
            nxt = (cur.replace(day=28) + datetime.timedelta(days=4)).replace(day=1)
            label = cur.strftime("%Y-%m")
        else:
            nxt = cur + datetime.timedelta(days=1)
            label = cur.strftime("%Y-%m-%d")
        p0, p1 = cur.timestamp(), min(nxt, d1 + datetime.timedelta(days=1)).timestamp()
        rows = _allocation_rows(ctx, p0, p1) if lo is not None and p1 > lo and p0 < hi else []
        agg: Dict[str, Dict[str, float]] = {k: collections.defaultdict(float) for k in ("byUsername", "byExperimentLabel", "byResourcePool", "byAgentLabel")}
        total = 0.0
        for r in rows:
            total += r["seconds"]
            agg["byUsername"][r["username"] or ""] += r["seconds"]
            agg["byResourcePool"][r["resourcePool"]] += r["seconds"]
            for lb in r["labels"]:
                agg["byExperimentLabel"][lb] += r["seconds"]
        entries.append({"periodStart": label, "period": period, "seconds": total, **{k: dict(v) for k, v in agg.items()}})
        cur = nxt
    return {"resourceEntries": entries}


# ---------------------------------------------------------------------------------------------------------------------
# master logs and live log level
# ---------------------------------------------------------------------------------------------------------------------
class _RingHandler(logging.Handler):
    def __init__(self, capacity: int = 5000) -> None:
        super().__init__()
        self.records: Deque[Tuple[int, Dict[str, Any]]] = collections.deque(maxlen=capacity)
        self.next_id = 1

    def emit(self, record: logging.LogRecord) -> None:
        try:
            entry = {"id": self.next_id, "message": record.getMessage(), "timestamp": _ts(record.created),
                     "level": "LOG_LEVEL_" + {"WARN": "WARNING", "FATAL": "CRITICAL"}.get(record.levelname, record.levelname)}
            self.records.append((self.next_id, entry))
            self.next_id += 1
        except Exception:  # noqa: BLE001
            pass


MASTER_LOG = _RingHandler()
logging.getLogger("determined").addHandler(MASTER_LOG)


@route("GET", "/api/v1/master/logs")
def master_logs(ctx: Ctx) -> Any:
    """The last lines the master logged (``offset``: only entries with a larger id; negative: that many from the end;
    ``follow``: hold the request up to ``timeoutSeconds`` until something new is there)."""
    off, lim = ctx.qi("offset", 0), ctx.qi("limit", 0) or 1000
    follow = ctx.q("follow") in ("true", "1")
    deadline = time.time() + min(ctx.qi("timeoutSeconds", 20), 60)
    while True:
        recs = [e for _, e in list(MASTER_LOG.records)]
        out = recs[off:] if off < 0 else [e for e in recs if e["id"] > off]
        if out or not follow or time.time() > deadline:
            return {"logs": out[:lim]}
        time.sleep(0.25)


@route("PATCH", "/api/v1/master/config")
def patch_master_config(ctx: Ctx) -> Any:
    """Only the log section can change while the master runs (reference ``PatchMasterConfig`` with field mask ``log``)."""
    _admin(ctx)
    b = (ctx.body or {}).get("config", ctx.body or {})
    log = b.get("log") or {}
    level = log.get("level")
    if level is not None:
        name = str(level).replace("LOG_LEVEL_", "").lower()
        py = {"trace": "DEBUG", "debug": "DEBUG", "info": "INFO", "warn": "WARNING", "warning": "WARNING", "error": "ERROR", "fatal": "CRITICAL", "critical": "CRITICAL"}.get(name)
        if py is None:
            raise HTTPError(400, f"unknown log level {level!r}")
        logging.getLogger("determined").setLevel(getattr(logging, py))
        logging.getLogger().setLevel(getattr(logging, py))
        if ctx.m.config is not None:
            ctx.m.config.data["log"]["level"] = name
        logger.warning("log level set to %s by %s", name, ctx.user["username"])
    if log.get("color") is not None and ctx.m.config is not None:
        ctx.m.config.data["log"]["color"] = bool(log["color"])
    unknown = set(b) - {"log"}
    if unknown:
        raise HTTPError(400, f"only `log` can be changed at run time, not {sorted(unknown)}")
    return {}


# ---------------------------------------------------------------------------------------------------------------------
# experiments / trials: tensorboard files, log retention
# ---------------------------------------------------------------------------------------------------------------------
@route("DELETE", "/api/v1/experiments/{eid}/tensorboard-files")
def delete_tb_files(ctx: Ctx, eid: str) -> Any:
    row = _exp_row(ctx, eid)
    cfg = dbmod.U(row["config"], {})
    st = cfg.get("checkpoint_storage") or {}
    removed = 0
    if st.get("type") in ("shared_fs", "directory"):
        base = st.get("host_path") or st.get("container_path") or ""
        if st.get("storage_path"):
            base = os.path.join(base, st["storage_path"]) if not os.path.isabs(st["storage_path"]) else st["storage_path"]
        d = os.path.join(base, "tensorboard", f"experiment-{int(eid)}") if st.get("type") == "shared_fs" else os.path.join(base, "tensorboard", f"experiment-{int(eid)}")
        for cand in (d, os.path.join(base, ctx.m.cluster_id, "tensorboard", "experiment", str(int(eid)))):
            if os.path.isdir(cand):
                removed += sum(len(fs) for _, _, fs in os.walk(cand))
                shutil.rmtree(cand, ignore_errors=True)
    return {"removedFiles": removed}


def _retention(ctx: Ctx) -> Optional[int]:
    b = ctx.body or {}
    n = b.get("numDays")
    if n is None or int(n) < -1 or int(n) > 32767:
        raise HTTPError(400, "numDays must be between -1 (forever) and 32767")
    return int(n)


@route("PUT", "/api/v1/experiments/{eid}/retain_logs")
def exp_retain_logs(ctx: Ctx, eid: str) -> Any:
    _exp_row(ctx, eid)
    n = _retention(ctx)
    for t in ctx.m.db.all("SELECT id FROM trials WHERE experiment_id=?", (int(eid),)):
        set_attr(ctx.m, "trial", t["id"], "logRetentionDays", n)
    set_attr(ctx.m, "experiment", eid, "logRetentionDays", n)
    return {}


@route("PUT", "/api/v1/trials/{tid}/retain_logs")
def trial_retain_logs(ctx: Ctx, tid: str) -> Any:
    _trial_row(ctx, tid)
    set_attr(ctx.m, "trial", tid, "logRetentionDays", _retention(ctx))
    return {}


def sweep_trial_logs(m: core.Master, now: Optional[float] = None) -> int:
    """Per-trial retention set through the two routes above (the cluster-wide sweep is ``Master.sweep_old_logs``)."""
    _ensure_schema(m)
    now = time.time() if now is None else now
    n = 0
    for r in m.db.all("SELECT id, value FROM attrs WHERE kind='trial' AND key='logRetentionDays'"):
        days = json.loads(r["value"])
        if days is None or days < 0:
            continue
        n += m.db.delete_task_logs(f"trial-{int(r['id'])}", now - 86400.0 * days)
    return n


@route("POST", "/api/v1/cleanup_logs")
def cleanup_logs(ctx: Ctx) -> Any:
    _admin(ctx)
    return {"removedCount": ctx.m.sweep_old_logs() + sweep_trial_logs(ctx.m)}


# ---------------------------------------------------------------------------------------------------------------------
# tasks: config, generic tasks, pause / unpause
# ---------------------------------------------------------------------------------------------------------------------
def _any_task(ctx: Ctx, task_id: str) -> Any:
    t = ctx.m.commands.tasks.get(task_id)
    if t is None:
        raise HTTPError(404, f"task {task_id} not found")
    return t


@route("GET", "/api/v1/tasks/{task_id}/config")
def task_config(ctx: Ctx, task_id: str) -> Any:
    if task_id.startswith("trial-"):
        row = ctx.m.db.one("SELECT e.config FROM trials t JOIN experiments e ON e.id=t.experiment_id WHERE t.id=?", (int(task_id.split("-")[1]),))
        if row is None:
            raise HTTPError(404, f"task {task_id} not found")
        return {"config": dbmod.U(row["config"], {})}
    return {"config": _any_task(ctx, task_id).config}


@route("POST", "/api/v1/generic-tasks")
def create_generic_task(ctx: Ctx) -> Any:
    """A task that is just a command with a context directory (reference ``CreateGenericTask``); ``forkedFrom`` /
    ``parentId`` + ``inheritContext`` reuse the configuration or context of an earlier task."""
    import yaml as _yaml

    b = ctx.body or {}
    cfg = b.get("config") or {}
    if isinstance(cfg, str):
        cfg = _yaml.safe_load(cfg) or {}
    src = b.get("forkedFrom") or None
    if src:
        base = _any_task(ctx, src)
        merged = json.loads(json.dumps(base.config))
        merged.update(cfg)
        cfg = merged
    parent = b.get("parentId") or None
    body = dict(b)
    if parent and b.get("inheritContext") and not b.get("contextDirectory"):
        body["files"] = getattr(_any_task(ctx, parent), "files", None) or []
    elif b.get("contextDirectory"):
        body["files"] = b["contextDirectory"]
    if not cfg.get("entrypoint"):
        raise HTTPError(400, "a generic task needs an entrypoint")
    try:
        t = ctx.m.commands.launch("command", cfg, ctx.user["id"], ctx.user["username"], body)
    except ValueError as e:
        raise HTTPError(400, str(e))
    t.generic = True
    t.parent_id = parent
    if parent:
        set_attr(ctx.m, "task", t.id, "parentId", parent)
    return {"taskId": t.id, "warnings": []}


def _signal_task(ctx: Ctx, task_id: str, sig: int, state_word: str) -> Any:
    t = _any_task(ctx, task_id)
    if t.state != "RUNNING" or not getattr(t, "proc", None):
        raise HTTPError(409, f"task {task_id} is not running here (state {t.state})")
    try:
        os.killpg(os.getpgid(t.proc.pid), sig)
    except (ProcessLookupError, PermissionError) as e:
        raise HTTPError(409, f"cannot signal task {task_id}: {e}")
    t.paused = sig == signal.SIGSTOP
    t.last_activity = time.time()
    now = time.time()
    ctx.m.db.add_logs([(t.id, t.id + ".1", now, "INFO", None, "stdout", "master", f"task {state_word} by {ctx.user['username']}\n")])
    return {}


@route("POST", "/api/v1/tasks/{task_id}/kill")
def kill_generic_task(ctx: Ctx, task_id: str) -> Any:
    """Kill a task and every task forked from it (reference ``KillGenericTask``); ``killFromRoot`` starts at the first
    ancestor instead, taking the whole tree down."""
    tasks = ctx.m.commands.tasks
    _any_task(ctx, task_id)
    parent_of = {tid: (getattr(t, "parent_id", None) or get_attr(ctx.m, "task", tid, "parentId")) for tid, t in tasks.items()}
    root = task_id
    if (ctx.body or {}).get("killFromRoot"):
        seen = {root}
        while parent_of.get(root) in tasks and parent_of[root] not in seen:
            root = parent_of[root]
            seen.add(root)
    tree, frontier = [root], [root]
    while frontier:
        cur = frontier.pop()
        for tid, par in parent_of.items():
            if par == cur and tid not in tree:
                tree.append(tid)
                frontier.append(tid)
    killed = []
    for tid in tree:
        t = tasks[tid]
        if t.state not in ("TERMINATED",):
            t.kill()
            killed.append(tid)
    return {"killed": killed}


route("POST", "/api/v1/tasks/{task_id}/pause")(lambda ctx, task_id: _signal_task(ctx, task_id, signal.SIGSTOP, "paused"))
route("POST", "/api/v1/tasks/{task_id}/unpause")(lambda ctx, task_id: _signal_task(ctx, task_id, signal.SIGCONT, "resumed"))


# ---------------------------------------------------------------------------------------------------------------------
# users: the agent-side identity tasks of a user run as
# ---------------------------------------------------------------------------------------------------------------------
@route("PUT", "/api/v1/users/{uid}/agent-user-group")
def set_agent_user_group(ctx: Ctx, uid: str) -> Any:
    _admin(ctx)
    row = ctx.m.db.one("SELECT * FROM users WHERE id=?", (int(uid),))
    if row is None:
        raise HTTPError(404, "user not found")
    aug = (ctx.body or {}).get("agentUserGroup") or {}
    for k in ("agentUid", "agentGid"):
        if aug.get(k) is not None and (not isinstance(aug[k], int) or aug[k] < 0):
            raise HTTPError(400, f"{k} must be a non-negative integer")
    set_attr(ctx.m, "user", uid, "agentUserGroup", aug or None)
    return {"user": dict(user_json(row), agentUserGroup=aug or None)}


@route("GET", "/api/v1/users/{uid}")
def get_user(ctx: Ctx, uid: str) -> Any:
    row = ctx.m.db.one("SELECT * FROM users WHERE id=? OR username=?", (int(uid) if uid.isdigit() else -1, uid))
    if row is None:
        raise HTTPError(404, "user not found")
    out = user_json(row)
    aug = get_attr(ctx.m, "user", row["id"], "agentUserGroup")
    if aug:
        out["agentUserGroup"] = aug
    return {"user": out}


# ---------------------------------------------------------------------------------------------------------------------
# OAuth client registry (what `det oauth client ...` manages; tokens are issued by /oauth2/token for SCIM-style integrations)
# ---------------------------------------------------------------------------------------------------------------------
@route("GET", "/oauth2/clients")
def oauth_list(ctx: Ctx) -> Any:
    _admin(ctx)
    _ensure_schema(ctx.m)
    return [{"id": r["id"], "name": r["name"], "domain": r["domain"]} for r in ctx.m.db.all("SELECT * FROM oauth_clients ORDER BY created")]


@route("POST", "/oauth2/clients")
def oauth_add(ctx: Ctx) -> Any:
    _admin(ctx)
    _ensure_schema(ctx.m)
    b = ctx.body or {}
    if not b.get("name") or not str(b.get("domain") or "").startswith(("http://", "https://")):
        raise HTTPError(400, "an OAuth client needs a name and an http(s) redirect domain")
    cid, sec = secrets.token_hex(8), secrets.token_urlsafe(24)
    ctx.m.db.exec("INSERT INTO oauth_clients(id,name,domain,secret,created) VALUES(?,?,?,?,?)", (cid, b["name"], b["domain"], sec, time.time()))
    return {"id": cid, "name": b["name"], "domain": b["domain"], "secret": sec}       # the secret is shown once, here


@route("DELETE", "/oauth2/clients/{cid}")
def oauth_remove(ctx: Ctx, cid: str) -> Any:
    _admin(ctx)
    _ensure_schema(ctx.m)
    if ctx.m.db.one("SELECT id FROM oauth_clients WHERE id=?", (cid,)) is None:
        raise HTTPError(404, "client not found")
    ctx.m.db.exec("DELETE FROM oauth_clients WHERE id=?", (cid,))
    return {}


# ---------------------------------------------------------------------------------------------------------------------
# model versions: rename, annotate
# ---------------------------------------------------------------------------------------------------------------------
@route("PATCH", "/api/v1/models/{name}/versions/{version}")
def patch_model_version(ctx: Ctx, name: str, version: str) -> Any:
    from determined_b200.master.api import _model, _mv_json

    mdl = _model(ctx, name)
    v = ctx.m.db.one("SELECT * FROM model_versions WHERE model_id=? AND version=?", (mdl["id"], int(version)))
    if v is None:
        raise HTTPError(404, f"model {name!r} has no version {version}")
    b = (ctx.body or {}).get("modelVersion", ctx.body or {})
    for col, enc in (("name", None), ("comment", None), ("notes", None), ("metadata", dbmod.J), ("labels", dbmod.J)):
        if b.get(col) is not None:
            ctx.m.db.exec(f"UPDATE model_versions SET {col}=? WHERE id=?", (enc(b[col]) if enc else b[col], v["id"]))
    ctx.m.db.exec("UPDATE models SET last_updated_time=? WHERE id=?", (time.time(), mdl["id"]))
    return {"modelVersion": _mv_json(ctx.m, ctx.m.db.one("SELECT * FROM model_versions WHERE id=?", (v["id"],)))}
