"""REST API of the single-box master (grpc-gateway-style JSON over HTTP, camelCase fields,
uint64 as strings) — the wire contract in SURVEY §10.

Parity: reference ``proto/src/determined/api/v1/api.proto`` + ``master/internal/api_*.go``.
Implemented: the endpoints the harness, CLI and SDK call — auth (login/logout/me), master info,
experiments (create/get/list/activate/pause/cancel/kill/archive/delete/patch, trials, checkpoints,
best metric, model-def download), trials (get, searcher operation/completed_operation/progress/
early_exit/runner metadata, metrics report + query, logs), checkpoints (report/get/rm/patch),
allocations (preemption long-poll + ack + pending, rendezvous, all_gather, proxy address, daemon
mark), runs/start, task logs, agents/slots, resource pools + job queue, templates, model registry,
workspaces/projects, users, webhooks, trial-source-info, preview-search, custom-searcher events.
Built on ``http.server.ThreadingHTTPServer`` (stdlib only).
"""
import base64
import json
import os
import sys
import logging
import re
import secrets
import threading
import time
import urllib.parse
import uuid as uuid_mod
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Any, Callable, Dict, List, Optional, Tuple

from determined_b200 import __version__
from determined_b200.common import tracing
from determined_b200.common.api import authentication as auth
from determined_b200.master import core, db as dbmod, expconf
from determined_b200.master import searcher as S

logger = logging.getLogger("determined.master.api")
audit_logger = logging.getLogger("determined.master.audit")


class HTTPError(Exception):
    def __init__(self, status: int, message: str) -> None:
        super().__init__(message)
        self.status, self.message = status, message


class Redirect(Exception):
    """A handler answers with ``302 Location: <to>``."""

    def __init__(self, to: str) -> None:
        super().__init__(to)
        self.to = to


ROUTES: List[Tuple[str, "re.Pattern[str]", Callable, bool]] = []
# conformance tests set this to see every (method, route pattern, request body, handler result) the server answers with 200
RECORDER: Optional[Callable[[str, str, Any, Any], None]] = None


def route(method: str, pattern: str, auth_required: bool = True) -> Callable:
    rx = re.compile("^" + re.sub(r"\{(\w+)\}", r"(?P<\1>[^/]+)", pattern) + "/?$")

    def deco(fn: Callable) -> Callable:
        ROUTES.append((method, rx, fn, auth_required))
        return fn
    return deco


def _ts(t: Optional[float]) -> Optional[str]:
    if t is None:
        return None
    return time.strftime("%Y-%m-%dT%H:%M:%S", time.gmtime(t)) + (".%06dZ" % int((t % 1) * 1e6))


def _state(prefix: str, s: Optional[str]) -> str:
    return f"{prefix}_{s}" if s else f"{prefix}_UNSPECIFIED"


class Ctx:
    def __init__(self, master: core.Master, user: Optional[Dict[str, Any]], query: Dict[str, List[str]], body: Any,
                 raw: bytes) -> None:
        self.m, self.user, self.query, self.body, self.raw = master, user, query, body, raw
        self.headers: Any = {}

    def q(self, name: str, default: Any = None) -> Any:
        v = self.query.get(name)
        return v[0] if v else default

    def qi(self, name: str, default: int = 0) -> int:
        try:
            return int(self.q(name, default))
        except (TypeError, ValueError):
            return default


# ---------------------------------------------------------------------------------------------
# serializers
# ---------------------------------------------------------------------------------------------
def exp_json(m: core.Master, row: Any) -> Dict[str, Any]:
    cfg = dbmod.U(row["config"], {})
    u = m.db.one("SELECT username FROM users WHERE id=?", (row["owner_id"],))
    ntr = m.db.one("SELECT COUNT(*) c FROM trials WHERE experiment_id=?", (row["id"],))["c"]
    live = m.experiments.get(row["id"])
    state = live.state if live is not None else row["state"]
    return {"id": row["id"], "name": cfg.get("name"), "description": cfg.get("description"), "labels": cfg.get("labels") or [],
            "state": _state("STATE", state), "startTime": _ts(row["start_time"]), "endTime": _ts(row["end_time"]),
            "archived": bool(row["archived"]), "numTrials": ntr, "progress": row["progress"], "username": u["username"] if u else None,
            "userId": row["owner_id"], "searcherType": (cfg.get("searcher") or {}).get("name"),
            "searcherMetric": (cfg.get("searcher") or {}).get("metric"), "resourcePool": (cfg.get("resources") or {}).get("resource_pool") or "default",
            "projectId": row["project_id"], "parentId": row["parent_id"], "notes": row["notes"], "jobId": row["job_id"],
            "config": cfg, "originalConfig": row["original_config"], "unmanaged": bool(row["unmanaged"])}


def trial_json(m: core.Master, row: Any) -> Dict[str, Any]:
    live = m.trials.get(row["id"])
    state = live.state if live is not None else row["state"]
    lv = m.db.one("SELECT steps_completed, metrics, report_time FROM metrics WHERE trial_id=? AND grp='validation' ORDER BY id DESC LIMIT 1", (row["id"],))
    lt = m.db.one("SELECT steps_completed, metrics, report_time FROM metrics WHERE trial_id=? AND grp='training' ORDER BY id DESC LIMIT 1", (row["id"],))
    out = {"id": row["id"], "experimentId": row["experiment_id"], "state": _state("STATE", state), "hparams": dbmod.U(row["hparams"], {}),
           "startTime": _ts(row["start_time"]), "endTime": _ts(row["end_time"]), "restarts": row["restarts"],
           "totalBatchesProcessed": row["total_batches"], "latestCheckpoint": row["latest_checkpoint"],
           "warmStartCheckpointUuid": row["warm_start_checkpoint"], "taskId": row["task_id"], "runnerState": row["runner_state"],
           "searcherMetricValue": row["best_validation"], "bestValidation": None, "latestValidation": None, "latestTraining": None,
           "seed": row["seed"], "requestId": row["request_id"]}
    if lv is not None:
        out["latestValidation"] = {"totalBatches": lv["steps_completed"], "endTime": _ts(lv["report_time"]),
                                   "metrics": {"avgMetrics": dbmod.U(lv["metrics"], {})}}
    if lt is not None:
        out["latestTraining"] = {"totalBatches": lt["steps_completed"], "endTime": _ts(lt["report_time"]),
                                 "metrics": {"avgMetrics": dbmod.U(lt["metrics"], {})}}
    return out


def ckpt_json(m: core.Master, row: Any) -> Dict[str, Any]:
    t = m.db.one("SELECT experiment_id, hparams FROM trials WHERE id=?", (row["trial_id"],))
    exp_cfg = None
    if t is not None:
        e = m.db.one("SELECT config FROM experiments WHERE id=?", (t["experiment_id"],))
        exp_cfg = dbmod.U(e["config"], {}) if e is not None else None
    val = m.db.one("SELECT metrics FROM metrics WHERE trial_id=? AND grp='validation' AND steps_completed=? ORDER BY id DESC LIMIT 1",
                   (row["trial_id"], row["steps_completed"]))
    metric_name = ((exp_cfg or {}).get("searcher") or {}).get("metric")
    vm = dbmod.U(val["metrics"], {}) if val is not None else {}
    return {"uuid": row["uuid"], "taskId": row["task_id"], "allocationId": row["allocation_id"], "reportTime": _ts(row["report_time"]),
            "state": _state("STATE", row["state"]), "resources": {k: str(v) for k, v in dbmod.U(row["resources"], {}).items()},
            "metadata": dbmod.U(row["metadata"], {}), "storageId": row["storage_id"],
            "training": {"trialId": row["trial_id"], "experimentId": t["experiment_id"] if t else None, "experimentConfig": exp_cfg,
                         "hparams": dbmod.U(t["hparams"], {}) if t else None, "validationMetrics": {"avgMetrics": vm},
                         "searcherMetric": vm.get(metric_name) if metric_name else None}}


REMOTE_USER_MARK = "!remote:"           # password_hash of a user who signs in through an identity provider only: no hash looks like it


def user_json(row: Any) -> Dict[str, Any]:
    return {"id": row["id"], "username": row["username"], "admin": bool(row["admin"]), "active": bool(row["active"]),
            "displayName": row["display_name"], "modifiedAt": _ts(row["modified_at"]), "remote": str(row["password_hash"] or "").startswith(REMOTE_USER_MARK)}


def sort_rows(rows: List[Any], key: Callable[[Any], Any], order_by: Optional[str]) -> List[Any]:
    """Stable sort on ``key``; rows without a value go last whatever the direction."""
    have = [r for r in rows if key(r) is not None]
    none = [r for r in rows if key(r) is None]
    have.sort(key=key, reverse=order_by == "ORDER_BY_DESC")
    return have + none


def page(items: List[Any], ctx: Ctx) -> Tuple[List[Any], Dict[str, Any]]:
    off, lim = ctx.qi("offset", 0), ctx.qi("limit", 0)
    total = len(items)
    if off < 0:
        off = max(total + off, 0)
    sub = items[off: off + lim] if lim > 0 else items[off:]
    return sub, {"offset": off, "limit": lim, "startIndex": off, "endIndex": off + len(sub), "total": total}


# ---------------------------------------------------------------------------------------------
# auth / master
# ---------------------------------------------------------------------------------------------
@route("POST", "/api/v1/auth/login", auth_required=False)
def login(ctx: Ctx) -> Any:
    b = ctx.body or {}
    row = ctx.m.db.one("SELECT * FROM users WHERE username=?", (b.get("username"),))
    pw = b.get("password") or ""
    if not b.get("isHashed", False) and pw:
        pw = auth.salt_and_hash(pw)
    if row is not None and str(row["password_hash"] or "").startswith(REMOTE_USER_MARK):
        raise HTTPError(401, "this user signs in through the identity provider, not with a password")
    if row is None or not row["active"] or (row["password_hash"] or "") != (pw or auth.salt_and_hash("")) and (row["password_hash"] or "") != pw:
        raise HTTPError(401, "invalid credentials")
    tok = secrets.token_hex(24)
    ctx.m.db.exec("INSERT INTO sessions(token,user_id,expiry) VALUES(?,?,?)", (tok, row["id"], time.time() + 30 * 86400))
    return {"token": tok, "user": user_json(row)}


@route("POST", "/api/v1/auth/logout")
def logout(ctx: Ctx) -> Any:
    ctx.m.db.exec("DELETE FROM sessions WHERE token=?", (ctx.user["token"],))
    return {}


@route("GET", "/api/v1/me")
def me(ctx: Ctx) -> Any:
    return {"user": user_json(ctx.m.db.one("SELECT * FROM users WHERE id=?", (ctx.user["id"],)))}


@route("GET", "/api/v1/master", auth_required=False)
@route("GET", "/info", auth_required=False)
def master_info(ctx: Ctx) -> Any:
    from determined_b200.master import sso

    tel = getattr(ctx.m, "telemetry", None)
    return {"version": __version__, "masterId": ctx.m.cluster_id, "clusterId": ctx.m.cluster_id, "clusterName": ctx.m.cluster_name,
            "telemetryEnabled": bool(tel and tel.enabled), "ssoProviders": sso.providers(ctx.m), "externalLoginUri": "", "rbacEnabled": authz.enabled(ctx.m), "branding": "determined_b200",
            "strictJobQueueControl": bool(getattr(ctx.m, "strict_job_queue_control", False)), "userManagementEnabled": True, "product": "PRODUCT_COMMUNITY"}


@route("GET", "/api/v1/master/config")
def master_config(ctx: Ctx) -> Any:
    """The configuration the master runs with, secrets masked; the scheduler / pool sections always describe the live pools."""
    live = {"resource_manager": {"type": "agent", "scheduler": {"type": ctx.m.pool.scheduler, "fitting_policy": ctx.m.pool.policy,
                                                                 "preemption": {"enabled": ctx.m.pool.preemption}}},
            "resource_pools": [{"pool_name": pl.name, "description": ctx.m.pool_descriptions.get(pl.name, ""),
                                "scheduler": {"type": pl.scheduler, "fitting_policy": pl.policy, "preemption": {"enabled": pl.preemption}}}
                               for pl in ctx.m.pools.values()],
            "checkpoint_storage": ctx.m.cluster_defaults.get("checkpoint_storage")}
    if ctx.m.config is not None:
        out = ctx.m.config.printable()
        out["resource_manager"] = dict(out.get("resource_manager") or {}, **live["resource_manager"])
        out["resource_pools"] = live["resource_pools"]
        return {"config": out}
    return {"config": live}


# -- users ---------------------------------------------------------------------------------------
@route("GET", "/api/v1/users")
def list_users(ctx: Ctx) -> Any:
    rows = [user_json(r) for r in ctx.m.db.all("SELECT * FROM users ORDER BY id")]
    items, pg = page(rows, ctx)
    return {"users": items, "pagination": pg}


@route("POST", "/api/v1/users")
def create_user(ctx: Ctx) -> Any:
    if not ctx.user["admin"]:
        raise HTTPError(403, "admin only")
    b = ctx.body or {}
    u = b.get("user") or {}
    pw = b.get("password") or ""
    if not b.get("isHashed", False) and pw:
        pw = auth.salt_and_hash(pw)
    if u.get("remote"):
        pw = REMOTE_USER_MARK + secrets.token_hex(16)
    try:
        uid = ctx.m.db.exec("INSERT INTO users(username,password_hash,admin,active,display_name,modified_at) VALUES(?,?,?,?,?,?)",
                            (u.get("username"), pw or auth.salt_and_hash(""), int(bool(u.get("admin"))), int(u.get("active", True)), u.get("displayName"), time.time()))
    except Exception as e:
        raise HTTPError(409, f"cannot create user: {e}")
    return {"user": user_json(ctx.m.db.one("SELECT * FROM users WHERE id=?", (uid,)))}


@route("PATCH", "/api/v1/users/{uid}")
def patch_user(ctx: Ctx, uid: str) -> Any:
    """Reference api_user.go:429-466: ``admin`` / ``active`` / ``username`` may only be changed by an admin
    (CanSetUsersAdmin / CanSetUsersActive / CanSetUsersUsername); a non-admin may change its own display name and
    password and nothing else."""
    is_admin = bool(ctx.user["admin"])
    if not is_admin and int(uid) != ctx.user["id"]:
        raise HTTPError(403, "forbidden")
    b = ctx.body or {}
    u = b.get("user", b)
    if not isinstance(u, dict):
        raise HTTPError(400, "user patch must be an object")
    if not is_admin and any(u.get(k) is not None for k in ("admin", "active", "username", "remote")):
        raise HTTPError(403, "only admin privileged users can update users")
    if ctx.m.db.one("SELECT id FROM users WHERE id=?", (int(uid),)) is None:
        raise HTTPError(404, f"user {uid} not found")
    for col, key in (("active", "active"), ("admin", "admin"), ("display_name", "displayName"), ("username", "username")):
        if key in u and u[key] is not None:
            ctx.m.db.exec(f"UPDATE users SET {col}=?, modified_at=? WHERE id=?", (int(u[key]) if isinstance(u[key], bool) else u[key], time.time(), int(uid)))
    pw = u.get("password")
    if pw is None and isinstance(b, dict) and b is not u:
        pw = b.get("password")
    if pw is not None:
        hashed = bool(u.get("isHashed") or (isinstance(b, dict) and b.get("isHashed")))
        ctx.m.db.exec("UPDATE users SET password_hash=?, modified_at=? WHERE id=?",
                      (pw if hashed else auth.salt_and_hash(str(pw)), time.time(), int(uid)))
    return {"user": user_json(ctx.m.db.one("SELECT * FROM users WHERE id=?", (int(uid),)))}


@route("POST", "/api/v1/users/{uid}/password")
def set_password(ctx: Ctx, uid: str) -> Any:
    if not ctx.user["admin"] and int(uid) != ctx.user["id"]:
        raise HTTPError(403, "forbidden")
    pw = ctx.body if isinstance(ctx.body, str) else (ctx.body or {}).get("password", "")
    ctx.m.db.exec("UPDATE users SET password_hash=?, modified_at=? WHERE id=?", (pw, time.time(), int(uid)))
    return {"user": user_json(ctx.m.db.one("SELECT * FROM users WHERE id=?", (int(uid),)))}


# ---------------------------------------------------------------------------------------------
# groups and role assignments (reference master/internal/usergroup, master/internal/rbac; the OSS edition ships the
# same API with a fixed role catalogue).  Roles gate nothing beyond admin-only mutations here: single-tenant box.
# ---------------------------------------------------------------------------------------------
ROLES = [{"roleId": 1, "name": "ClusterAdmin"}, {"roleId": 2, "name": "WorkspaceAdmin"}, {"roleId": 4, "name": "WorkspaceCreator"},
         {"roleId": 5, "name": "Viewer"}, {"roleId": 6, "name": "Editor"}, {"roleId": 7, "name": "ModelRegistryViewer"}]


def _group_json(ctx: Ctx, row: Any) -> Dict[str, Any]:
    members = ctx.m.db.all("SELECT u.* FROM users u JOIN group_members g ON g.user_id=u.id WHERE g.group_id=?", (row["id"],))
    return {"groupId": row["id"], "name": row["name"], "users": [user_json(u) for u in members]}


@route("POST", "/api/v1/groups")
def create_group(ctx: Ctx) -> Any:
    if not ctx.user["admin"]:
        raise HTTPError(403, "admin only")
    b = ctx.body or {}
    try:
        gid = ctx.m.db.exec("INSERT INTO groups(name) VALUES(?)", (b.get("name"),))
    except Exception as e:
        raise HTTPError(409, f"cannot create group: {e}")
    for uid in b.get("addUsers") or []:
        ctx.m.db.exec("INSERT OR IGNORE INTO group_members(group_id,user_id) VALUES(?,?)", (gid, int(uid)))
    return {"group": _group_json(ctx, ctx.m.db.one("SELECT * FROM groups WHERE id=?", (gid,)))}


@route("POST", "/api/v1/groups/search")
@route("GET", "/api/v1/groups")
def list_groups(ctx: Ctx) -> Any:
    rows = ctx.m.db.all("SELECT * FROM groups ORDER BY id", ())
    uid = (ctx.body or {}).get("userId") if isinstance(ctx.body, dict) else None
    out = [_group_json(ctx, r) for r in rows]
    if uid:
        out = [g for g in out if any(u["id"] == int(uid) for u in g["users"])]
    return {"groups": [{"group": g, "numMembers": len(g["users"])} for g in out]}


@route("GET", "/api/v1/groups/{gid}")
def get_group(ctx: Ctx, gid: str) -> Any:
    row = ctx.m.db.one("SELECT * FROM groups WHERE id=?", (int(gid),))
    if row is None:
        raise HTTPError(404, f"group {gid} not found")
    return {"group": _group_json(ctx, row)}


@route("PUT", "/api/v1/groups/{gid}")
def update_group(ctx: Ctx, gid: str) -> Any:
    if not ctx.user["admin"]:
        raise HTTPError(403, "admin only")
    b = ctx.body or {}
    if ctx.m.db.one("SELECT id FROM groups WHERE id=?", (int(gid),)) is None:
        raise HTTPError(404, f"group {gid} not found")
    if b.get("name"):
        ctx.m.db.exec("UPDATE groups SET name=? WHERE id=?", (b["name"], int(gid)))
    for uid in b.get("addUsers") or []:
        ctx.m.db.exec("INSERT OR IGNORE INTO group_members(group_id,user_id) VALUES(?,?)", (int(gid), int(uid)))
    for uid in b.get("removeUsers") or []:
        ctx.m.db.exec("DELETE FROM group_members WHERE group_id=? AND user_id=?", (int(gid), int(uid)))
    return {"group": _group_json(ctx, ctx.m.db.one("SELECT * FROM groups WHERE id=?", (int(gid),)))}


@route("DELETE", "/api/v1/groups/{gid}")
def delete_group(ctx: Ctx, gid: str) -> Any:
    if not ctx.user["admin"]:
        raise HTTPError(403, "admin only")
    ctx.m.db.exec("DELETE FROM group_members WHERE group_id=?", (int(gid),))
    ctx.m.db.exec("DELETE FROM role_assignments WHERE group_id=?", (int(gid),))
    ctx.m.db.exec("DELETE FROM groups WHERE id=?", (int(gid),))
    return {}


@route("POST", "/api/v1/roles/search")
@route("GET", "/api/v1/roles")
def list_roles(ctx: Ctx) -> Any:
    return {"roles": ROLES}


@route("POST", "/api/v1/roles/add-assignments")
def assign_roles(ctx: Ctx) -> Any:
    if not authz.can_assign_roles(ctx.m, ctx.user, ctx.body):
        raise HTTPError(403, "admin only")
    b = ctx.body or {}
    names = {r["roleId"]: r["name"] for r in ROLES}
    for a in b.get("userRoleAssignments") or []:
        ra = a.get("roleAssignment") or {}
        ctx.m.db.exec("INSERT INTO role_assignments(role,user_id,group_id,workspace_id) VALUES(?,?,?,?)",
                      (names.get(int((ra.get("role") or {}).get("roleId", 0)), "Viewer"), int(a["userId"]), None, ra.get("scopeWorkspaceId")))
    for a in b.get("groupRoleAssignments") or []:
        ra = a.get("roleAssignment") or {}
        ctx.m.db.exec("INSERT INTO role_assignments(role,user_id,group_id,workspace_id) VALUES(?,?,?,?)",
                      (names.get(int((ra.get("role") or {}).get("roleId", 0)), "Viewer"), None, int(a["groupId"]), ra.get("scopeWorkspaceId")))
    return {}


@route("POST", "/api/v1/roles/remove-assignments")
def remove_roles(ctx: Ctx) -> Any:
    if not authz.can_assign_roles(ctx.m, ctx.user, ctx.body):
        raise HTTPError(403, "admin only")
    b = ctx.body or {}
    names = {r["roleId"]: r["name"] for r in ROLES}
    for a in b.get("userRoleAssignments") or []:
        ra = a.get("roleAssignment") or {}
        ctx.m.db.exec("DELETE FROM role_assignments WHERE role=? AND user_id=? AND workspace_id IS ?",
                      (names.get(int((ra.get("role") or {}).get("roleId", 0))), int(a["userId"]), ra.get("scopeWorkspaceId")))
    for a in b.get("groupRoleAssignments") or []:
        ra = a.get("roleAssignment") or {}
        ctx.m.db.exec("DELETE FROM role_assignments WHERE role=? AND group_id=? AND workspace_id IS ?",
                      (names.get(int((ra.get("role") or {}).get("roleId", 0))), int(a["groupId"]), ra.get("scopeWorkspaceId")))
    return {}


@route("GET", "/api/v1/roles/search/by-user/{uid}")
def roles_of_user(ctx: Ctx, uid: str) -> Any:
    direct = ctx.m.db.all("SELECT * FROM role_assignments WHERE user_id=?", (int(uid),))
    via = ctx.m.db.all("SELECT r.* FROM role_assignments r JOIN group_members g ON g.group_id=r.group_id WHERE g.user_id=?", (int(uid),))
    ids = {r["name"]: r["roleId"] for r in ROLES}
    def fmt(r: Any) -> Dict[str, Any]:
        return {"role": {"roleId": ids.get(r["role"]), "name": r["role"]}, "scopeWorkspaceId": r["workspace_id"], "groupId": r["group_id"]}
    return {"roles": [fmt(r) for r in list(direct) + list(via)]}


# ---------------------------------------------------------------------------------------------
# experiments
# ---------------------------------------------------------------------------------------------
@route("POST", "/api/v1/experiments")
def create_experiment(ctx: Ctx) -> Any:
    b = ctx.body or {}
    files = b.get("modelDefinition") or []
    blob = None
    if files:
        import io
        import tarfile

        buf = io.BytesIO()
        with tarfile.open(fileobj=buf, mode="w:gz") as tf:
            for f in files:
                data = base64.b64decode(f.get("content") or "")
                ti = tarfile.TarInfo(f["path"])
                if f.get("type") == 53 or str(f.get("type")) == "53" or f["path"].endswith("/"):
                    ti.type = tarfile.DIRTYPE
                    ti.mode = f.get("mode", 0o755) & 0o7777 or 0o755
                    tf.addfile(ti)
                    continue
                ti.size = len(data)
                ti.mode = (f.get("mode", 0o644) & 0o7777) or 0o644
                ti.mtime = int(f.get("mtime", time.time()))
                tf.addfile(ti, io.BytesIO(data))
        blob = buf.getvalue()
    from determined_b200.master import api_more

    try:
        early = expconf.parse(b.get("config", "")) if b.get("config") else {}
    except Exception:  # noqa: BLE001  (reported properly by create_experiment below)
        early = {}
    pid = int(b.get("projectId") or 0)
    if not pid and isinstance(early, dict) and (early.get("workspace") or early.get("project")):
        # the configuration itself may say where the experiment belongs: `workspace: name` + `project: name`
        if not (early.get("workspace") and early.get("project")):
            raise HTTPError(400, "workspace and project must both be given in the experiment configuration, or neither")
        prow = ctx.m.db.one("SELECT p.id FROM projects p JOIN workspaces w ON w.id=p.workspace_id WHERE w.name=? AND p.name=?", (early["workspace"], early["project"]))
        if prow is None:
            raise HTTPError(404, f"project {early['project']!r} not found in workspace {early['workspace']!r}")
        pid = int(prow["id"])
        b = dict(b, projectId=pid)
    pid = pid or 1
    res = (early.get("resources") or {}) if isinstance(early, dict) else {}
    if isinstance(early, dict) and not res.get("resource_pool"):
        dflt = api_more.default_pool_for_project(ctx.m, pid, int(res.get("slots_per_trial", 1) or 0))
        if dflt:
            early.setdefault("resources", {})["resource_pool"] = dflt
            b = dict(b, config=early)
            res = early["resources"]
    pool_name = res.get("resource_pool") or "default"
    if pool_name in ctx.m.pools and not api_more.pool_allowed_for_project(ctx.m, pool_name, pid):
        raise HTTPError(400, f"resource pool {pool_name!r} is not bound to the workspace of project {pid}")
    try:
        exp, cfg = ctx.m.create_experiment(b.get("config", ""), blob, owner=ctx.user["username"], activate=bool(b.get("activate", True)),
                                           parent_id=b.get("parentId"), template=b.get("template"), validate_only=bool(b.get("validateOnly")),
                                           project_id=int(b.get("projectId") or 1), unmanaged=bool(b.get("unmanaged")))
    except expconf.ConfigError as e:
        raise HTTPError(400, "invalid experiment configuration: " + "; ".join(e.errors))
    except core.NotFound as e:
        raise HTTPError(404, str(e))
    except ValueError as e:
        raise HTTPError(400, f"invalid experiment: {e}")
    if exp is None:
        return {"experiment": None, "config": cfg, "warnings": []}
    row = ctx.m.db.one("SELECT id, state, config, original_config, owner_id, start_time, end_time, progress, archived, parent_id, notes, project_id, unmanaged, job_id FROM experiments WHERE id=?", (exp.id,))
    return {"experiment": exp_json(ctx.m, row), "config": cfg, "warnings": []}


@route("GET", "/api/v1/experiments")
def list_experiments(ctx: Ctx) -> Any:
    where, args = ["1=1"], []
    if ctx.q("archived") is not None:
        where.append("archived=?")
        args.append(1 if ctx.q("archived") in ("true", "1", True) else 0)
    if ctx.q("projectId"):
        where.append("project_id=?")
        args.append(ctx.qi("projectId"))
    states = ctx.query.get("states") or []
    rows = ctx.m.db.all(f"SELECT id, state, config, original_config, owner_id, start_time, end_time, progress, archived, parent_id, notes, project_id, unmanaged, job_id FROM experiments WHERE {' AND '.join(where)} ORDER BY id " + ("DESC" if ctx.q("orderBy") == "ORDER_BY_DESC" else "ASC"), args)
    out = [exp_json(ctx.m, r) for r in rows]
    if states:
        out = [e for e in out if e["state"] in states]
    users = ctx.query.get("users") or []
    if users:
        out = [e for e in out if e["username"] in users]
    if ctx.q("name"):
        out = [e for e in out if ctx.q("name") in (e["name"] or "")]
    items, pg = page(out, ctx)
    return {"experiments": items, "pagination": pg}


def _exp_row(ctx: Ctx, eid: str) -> Any:
    row = ctx.m.db.one("SELECT id, state, config, original_config, owner_id, start_time, end_time, progress, archived, parent_id, notes, project_id, unmanaged, job_id FROM experiments WHERE id=?", (int(eid),))
    if row is None:
        raise HTTPError(404, f"experiment {eid} not found")
    return row


@route("GET", "/api/v1/experiments/{eid}")
def get_experiment(ctx: Ctx, eid: str) -> Any:
    row = _exp_row(ctx, eid)
    return {"experiment": exp_json(ctx.m, row), "jobSummary": None}


def _exp_action(name: str) -> Callable:
    def handler(ctx: Ctx, eid: str) -> Any:
        _exp_row(ctx, eid)
        with ctx.m.lock:
            exp = ctx.m.experiments.get(int(eid))
            if exp is None:
                raise HTTPError(409, f"experiment {eid} is in a terminal state")
            getattr(exp, name)()
        ctx.m.kick()
        return {}
    return handler


route("POST", "/api/v1/experiments/{eid}/activate")(_exp_action("activate"))
route("POST", "/api/v1/experiments/{eid}/pause")(_exp_action("pause"))
route("POST", "/api/v1/experiments/{eid}/cancel")(_exp_action("cancel"))
route("POST", "/api/v1/experiments/{eid}/kill")(_exp_action("kill"))


@route("POST", "/api/v1/experiments/continue")
def continue_experiment(ctx: Ctx) -> Any:
    """Re-open a finished single-trial experiment, optionally with config overrides (``overrideConfig``: YAML text or object)."""
    b = ctx.body or {}
    try:
        exp = ctx.m.continue_experiment(int(b.get("id")), b.get("overrideConfig") or None)
    except core.Conflict as e:
        raise HTTPError(409, str(e))
    row = ctx.m.db.one("SELECT id, state, config, original_config, owner_id, start_time, end_time, progress, archived, parent_id, notes, project_id, unmanaged, job_id FROM experiments WHERE id=?", (exp.id,))
    return {"experiment": exp_json(ctx.m, row), "warnings": []}


@route("POST", "/api/v1/experiments/{eid}/move")
def move_experiment(ctx: Ctx, eid: str) -> Any:
    """Move an experiment to another project (``destinationProjectId``)."""
    ctx.m.move_experiment(int(eid), int((ctx.body or {}).get("destinationProjectId")))
    return {}


@route("POST", "/api/v1/experiments/{eid}/archive")
def archive_experiment(ctx: Ctx, eid: str) -> Any:
    row = _exp_row(ctx, eid)
    if row["state"] not in core.TERMINAL:
        raise HTTPError(409, "only terminal experiments can be archived")
    ctx.m.db.exec("UPDATE experiments SET archived=1 WHERE id=?", (int(eid),))
    return {}


@route("POST", "/api/v1/experiments/{eid}/unarchive")
def unarchive_experiment(ctx: Ctx, eid: str) -> Any:
    _exp_row(ctx, eid)
    ctx.m.db.exec("UPDATE experiments SET archived=0 WHERE id=?", (int(eid),))
    return {}


@route("DELETE", "/api/v1/experiments/{eid}")
def delete_experiment(ctx: Ctx, eid: str) -> Any:
    row = _exp_row(ctx, eid)
    if row["state"] not in core.TERMINAL:
        raise HTTPError(409, "only terminal experiments can be deleted")
    e = int(eid)
    from determined_b200.common import storage

    cfg = dbmod.U(row["config"], {})
    try:
        mgr = storage.build(cfg["checkpoint_storage"], None)
        for c in ctx.m.db.all("SELECT c.uuid FROM checkpoints c JOIN trials t ON t.id=c.trial_id WHERE t.experiment_id=?", (e,)):
            try:
                mgr.delete(c["uuid"], ["**/*"])
            except Exception:
                pass
    except Exception:
        pass
    tids = [r["id"] for r in ctx.m.db.all("SELECT id FROM trials WHERE experiment_id=?", (e,))]
    for t in tids:
        ctx.m.db.exec("DELETE FROM metrics WHERE trial_id=?", (t,))
        ctx.m.db.exec("DELETE FROM checkpoints WHERE trial_id=?", (t,))
        ctx.m.db.delete_task_logs(f"trial-{t}")
    ctx.m.db.exec("DELETE FROM trials WHERE experiment_id=?", (e,))
    ctx.m.db.exec("DELETE FROM experiments WHERE id=?", (e,))
    ctx.m.experiments.pop(e, None)
    return {}


@route("PATCH", "/api/v1/experiments/{eid}")
def patch_experiment(ctx: Ctx, eid: str) -> Any:
    row = _exp_row(ctx, eid)
    b = (ctx.body or {}).get("experiment", ctx.body or {})
    cfg = dbmod.U(row["config"], {})
    changed = False
    for k in ("name", "description", "labels"):
        if k in b and b[k] is not None:
            cfg[k] = b[k]
            changed = True
    if "notes" in b:
        ctx.m.db.exec("UPDATE experiments SET notes=? WHERE id=?", (b["notes"], int(eid)))
    res = b.get("resources") or {}
    with ctx.m.lock:
        live = ctx.m.experiments.get(int(eid))
        def on_all_pools(name: str) -> Callable[[str, Any], None]:
            def apply(job_id: str, value: Any) -> None:
                for pl in ctx.m.pools.values():
                    getattr(pl, name)(job_id, value)
            return apply

        for key, setter in (("priority", on_all_pools("set_priority")), ("weight", on_all_pools("set_weight")), ("maxSlots", on_all_pools("set_max_slots"))):
            if res.get(key) is not None:
                ck = {"maxSlots": "max_slots"}.get(key, key)
                cfg.setdefault("resources", {})[ck] = res[key]
                changed = True
                if live is not None:
                    live.config.setdefault("resources", {})[ck] = res[key]
                    setter(live.job_id, res[key])
        if b.get("checkpointStorage"):
            for k, ck in (("saveExperimentBest", "save_experiment_best"), ("saveTrialBest", "save_trial_best"), ("saveTrialLatest", "save_trial_latest")):
                if k in b["checkpointStorage"]:
                    cfg.setdefault("checkpoint_storage", {})[ck] = b["checkpointStorage"][k]
                    changed = True
        if changed:
            ctx.m.db.exec("UPDATE experiments SET config=? WHERE id=?", (dbmod.J(cfg), int(eid)))
            if live is not None:
                for k in ("name", "description", "labels"):
                    if k in cfg:
                        live.config[k] = cfg[k]
    ctx.m.kick()
    return {"experiment": exp_json(ctx.m, _exp_row(ctx, eid))}


@route("GET", "/api/v1/experiments/{eid}/trials")
def exp_trials(ctx: Ctx, eid: str) -> Any:
    _exp_row(ctx, eid)
    rows = ctx.m.db.all("SELECT * FROM trials WHERE experiment_id=? ORDER BY id", (int(eid),))
    trials = [trial_json(ctx.m, r) for r in rows]
    states = set(ctx.query.get("states", []))
    if states:
        trials = [t for t in trials if t["state"] in states]
    smaller = bool((dbmod.U(_exp_row(ctx, eid)["config"], {}).get("searcher") or {}).get("smaller_is_better", True))

    def duration(t: Dict[str, Any]) -> Optional[float]:
        a, b = _log_time(t.get("startTime")), _log_time(t.get("endTime"))
        return (b if b is not None else time.time()) - a if a is not None else None

    keys: Dict[str, Callable[[Dict[str, Any]], Any]] = {
        "SORT_BY_ID": lambda t: t["id"], "SORT_BY_START_TIME": lambda t: t.get("startTime"), "SORT_BY_END_TIME": lambda t: t.get("endTime"),
        "SORT_BY_STATE": lambda t: t["state"], "SORT_BY_BATCHES_PROCESSED": lambda t: t.get("totalBatchesProcessed"), "SORT_BY_DURATION": duration,
        "SORT_BY_RESTARTS": lambda t: t.get("restarts"), "SORT_BY_CHECKPOINT_SIZE": lambda t: t.get("totalCheckpointSize"),
        "SORT_BY_BEST_VALIDATION_METRIC": lambda t: (t.get("searcherMetricValue") if smaller else (-t["searcherMetricValue"] if t.get("searcherMetricValue") is not None else None)),
        "SORT_BY_LATEST_VALIDATION_METRIC": lambda t: ((t.get("latestValidation") or {}).get("metrics", {}).get("avgMetrics", {}) or {}).get(
            (dbmod.U(_exp_row(ctx, eid)["config"], {}).get("searcher") or {}).get("metric")),
        "SORT_BY_LOG_RETENTION_DAYS": lambda t: t.get("logRetentionDays")}
    trials = sort_rows(trials, keys.get(ctx.q("sortBy") or "", keys["SORT_BY_ID"]), ctx.q("orderBy"))
    items, pg = page(trials, ctx)
    return {"trials": items, "pagination": pg}


@route("GET", "/api/v1/experiments/{eid}/checkpoints")
def exp_checkpoints(ctx: Ctx, eid: str) -> Any:
    _exp_row(ctx, eid)
    rows = ctx.m.db.all("SELECT c.* FROM checkpoints c JOIN trials t ON t.id=c.trial_id WHERE t.experiment_id=? ORDER BY c.report_time", (int(eid),))
    out = [ckpt_json(ctx.m, r) for r in rows]
    states = ctx.query.get("states") or []
    if states:
        out = [c for c in out if c["state"] in states]
    if ctx.q("sortByMetric") or ctx.q("sortBy") == "SORT_BY_SEARCHER_METRIC":
        sib = bool((dbmod.U(_exp_row(ctx, eid)["config"], {}).get("searcher") or {}).get("smaller_is_better", True))
        key = ctx.q("sortByMetric")
        def metric_of(c: Dict[str, Any]) -> float:
            v = (c["training"]["validationMetrics"]["avgMetrics"] or {}).get(key) if key else c["training"]["searcherMetric"]
            return float("inf") if v is None else (float(v) if sib else -float(v))
        out.sort(key=metric_of)
    if ctx.q("orderBy") == "ORDER_BY_DESC" and not (ctx.q("sortByMetric") or ctx.q("sortBy") == "SORT_BY_SEARCHER_METRIC"):
        out.reverse()
    items, pg = page(out, ctx)
    return {"checkpoints": items, "pagination": pg}


@route("GET", "/api/v1/experiments/{eid}/searcher/best_searcher_validation_metric")
def best_metric(ctx: Ctx, eid: str) -> Any:
    row = _exp_row(ctx, eid)
    cfg = dbmod.U(row["config"], {})
    metric = (cfg.get("searcher") or {}).get("metric")
    sib = bool((cfg.get("searcher") or {}).get("smaller_is_better", True))
    best = None
    for r in ctx.m.db.all("SELECT m.metrics FROM metrics m JOIN trials t ON t.id=m.trial_id WHERE t.experiment_id=? AND m.grp='validation'", (int(eid),)):
        v = dbmod.U(r["metrics"], {}).get(metric)
        if isinstance(v, (int, float)) and (best is None or (v < best) == sib):
            best = float(v)
    if best is None:
        raise HTTPError(404, "no validations for experiment")
    return {"metric": best}


@route("GET", "/api/v1/experiments/{eid}/model_def")
def model_def(ctx: Ctx, eid: str) -> Any:
    _exp_row(ctx, eid)
    row = ctx.m.db.one("SELECT model_definition FROM experiments WHERE id=?", (int(eid),))
    return {"b64Tgz": base64.b64encode(row["model_definition"] or b"").decode()}


@route("GET", "/api/v1/tasks/{task_id}/context_directory")
def context_directory(ctx: Ctx, task_id: str) -> Any:
    t = ctx.m.db.one("SELECT experiment_id FROM trials WHERE task_id=?", (task_id,))
    if t is None:
        raise HTTPError(404, "task not found")
    row = ctx.m.db.one("SELECT model_definition FROM experiments WHERE id=?", (t["experiment_id"],))
    return {"b64Tgz": base64.b64encode(row["model_definition"] or b"").decode()}


@route("POST", "/api/v1/preview-hp-search")
def preview_search(ctx: Ctx) -> Any:
    b = ctx.body or {}
    cfg = expconf.parse(b.get("config", "{}"))
    expconf.validate_sanity(cfg)
    cfg = expconf.with_defaults(cfg)
    res = S.simulate(cfg["searcher"], cfg.get("hyperparameters") or {}, seed=int(b.get("seed", 0)))
    unit = S.new_search_method(cfg["searcher"]).unit() or "batches"
    trials = [{"occurrences": n, "operations": [{"validateAfter": {"length": str(length)}}]} for length, n in sorted(res["results"].items())]
    return {"simulation": {"config": cfg, "seed": int(b.get("seed", 0)), "trials": trials, "unit": unit}}


# -- hyperparameter visualisation (reference api_experiment.go: MetricNames, TrialsSnapshot, TrialsSample; api_trials.go: Summarize / Compare)
def _flat_hparams(hp: Any, prefix: str = "") -> Dict[str, Any]:
    out: Dict[str, Any] = {}
    for k, v in (hp or {}).items():
        if isinstance(v, dict):
            out.update(_flat_hparams(v, f"{prefix}{k}."))
        else:
            out[f"{prefix}{k}"] = v
    return out


@route("GET", "/api/v1/experiments/{eid}/metric-names")
@route("GET", "/api/v1/experiments/metrics-stream/metric-names")      # registered first (decorators apply bottom-up): the literal path wins
def metric_names(ctx: Ctx, eid: Optional[str] = None) -> Any:
    """Names of every training / validation metric the experiment(s) reported, plus the searcher metric."""
    ids = [int(eid)] if eid else [int(x) for x in ctx.query.get("ids", [])]
    out = []
    for i in ids:
        row = _exp_row(ctx, str(i))
        names: Dict[str, set] = {"training": set(), "validation": set()}
        for r in ctx.m.db.all("SELECT DISTINCT grp, metrics FROM metrics WHERE trial_id IN (SELECT id FROM trials WHERE experiment_id=?)", (i,)):
            if r["grp"] in names:
                names[r["grp"]].update(k for k, v in dbmod.U(r["metrics"], {}).items() if isinstance(v, (int, float)))
        cfg = dbmod.U(row["config"], {})
        out.append({"experimentId": i, "searcherMetric": (cfg.get("searcher") or {}).get("metric"), "trainingMetrics": sorted(names["training"]),
                    "validationMetrics": sorted(names["validation"]),
                    "metricNames": [{"group": g, "name": n} for g in ("training", "validation") for n in sorted(names[g])]})
    return out[0] if eid else {"results": out}


def _metric_points(ctx: Ctx, eid: str, name: str, group: str) -> Dict[int, List[Tuple[int, float]]]:
    pts: Dict[int, List[Tuple[int, float]]] = {}
    for r in ctx.m.db.all("SELECT trial_id, steps_completed, metrics FROM metrics WHERE grp=? AND trial_id IN (SELECT id FROM trials WHERE experiment_id=?) "
                          "ORDER BY steps_completed, id", (group, int(eid))):
        v = dbmod.U(r["metrics"], {}).get(name)
        if isinstance(v, (int, float)):
            pts.setdefault(r["trial_id"], []).append((r["steps_completed"], float(v)))
    return pts


@route("GET", "/api/v1/experiments/{eid}/metrics-stream/trials-snapshot")
@route("GET", "/api/v1/experiments/{eid}/trials-snapshot")
def trials_snapshot(ctx: Ctx, eid: str) -> Any:
    """One point per trial: its hyperparameters and the value of ``metricName`` at ``batchesProcessed`` (default: each trial's
    latest report) — the data behind the parallel-coordinates / scatter / heat-map views."""
    _exp_row(ctx, eid)
    name, group = ctx.q("metricName"), ctx.q("group") or {"METRIC_TYPE_TRAINING": "training"}.get(ctx.q("metricType", ""), "validation")
    if not name:
        raise HTTPError(400, "metricName is required")
    at = ctx.q("batchesProcessed")
    trials = []
    hp_of = {r["id"]: _flat_hparams(dbmod.U(r["hparams"], {})) for r in ctx.m.db.all("SELECT id, hparams FROM trials WHERE experiment_id=?", (int(eid),))}
    for tid, series in sorted(_metric_points(ctx, eid, name, group).items()):
        pick = [p for p in series if at is None or p[0] <= int(at)]
        if pick:
            trials.append({"trialId": tid, "hparams": hp_of.get(tid, {}), "metric": pick[-1][1], "batchesProcessed": pick[-1][0]})
    return {"trials": trials}


@route("GET", "/api/v1/experiments/{eid}/metrics-stream/trials-sample")
@route("GET", "/api/v1/experiments/{eid}/trials-sample")
def trials_sample(ctx: Ctx, eid: str) -> Any:
    """Learning curves of up to ``maxTrials`` trials (the best by final value first), each thinned to ``maxDatapoints`` points."""
    row = _exp_row(ctx, eid)
    name, group = ctx.q("metricName"), ctx.q("group") or {"METRIC_TYPE_TRAINING": "training"}.get(ctx.q("metricType", ""), "validation")
    if not name:
        raise HTTPError(400, "metricName is required")
    smaller = bool((dbmod.U(row["config"], {}).get("searcher") or {}).get("smaller_is_better", True))
    max_trials, max_pts = ctx.qi("maxTrials", 25), max(ctx.qi("maxDatapoints", 1000), 2)
    series = _metric_points(ctx, eid, name, group)
    order = sorted(series, key=lambda t: series[t][-1][1], reverse=not smaller)[:max_trials]
    hp_of = {r["id"]: _flat_hparams(dbmod.U(r["hparams"], {})) for r in ctx.m.db.all("SELECT id, hparams FROM trials WHERE experiment_id=?", (int(eid),))}
    trials = []
    for tid in order:
        pts = series[tid]
        if len(pts) > max_pts:
            step = (len(pts) - 1) / (max_pts - 1)
            pts = [pts[round(i * step)] for i in range(max_pts)]
        trials.append({"trialId": tid, "hparams": hp_of.get(tid, {}), "data": [{"batches": b, "value": v} for b, v in pts]})
    return {"trials": trials, "promotedTrials": [], "demotedTrials": []}


@route("GET", "/api/v1/trials/{tid}/summarize")
def summarize_trial(ctx: Ctx, tid: str) -> Any:
    """A trial with its metric series thinned to ``maxDatapoints`` per metric (reference ``SummarizeTrial``)."""
    row = _trial_row(ctx, tid)
    want, max_pts = set(ctx.query.get("metricNames", [])), max(ctx.qi("maxDatapoints", 1000), 2)
    series: Dict[Tuple[str, str], List[Dict[str, Any]]] = {}
    for r in ctx.m.db.all("SELECT grp, steps_completed, metrics, report_time FROM metrics WHERE trial_id=? ORDER BY steps_completed, id", (int(tid),)):
        for k, v in dbmod.U(r["metrics"], {}).items():
            if isinstance(v, (int, float)) and (not want or k in want):
                series.setdefault((r["grp"], k), []).append({"batches": r["steps_completed"], "value": float(v), "time": _ts(r["report_time"])})
    out = []
    for (g, k), pts in sorted(series.items()):
        if len(pts) > max_pts:
            step = (len(pts) - 1) / (max_pts - 1)
            pts = [pts[round(i * step)] for i in range(max_pts)]
        out.append({"name": k, "type": "METRIC_TYPE_TRAINING" if g == "training" else "METRIC_TYPE_VALIDATION", "group": g, "data": pts})
    return {"trial": trial_json(ctx.m, row), "metrics": out}


@route("GET", "/api/v1/trials/compare")
def compare_trials(ctx: Ctx) -> Any:
    """``summarize`` for several trials at once (``trialIds=1&trialIds=2``)."""
    out = []
    for t in ctx.query.get("trialIds", []):
        s_ = summarize_trial(ctx, t)
        out.append({"trial": s_["trial"], "metrics": s_["metrics"]})
    return {"trials": out}


# -- custom searcher events ------------------------------------------------------------------------
@route("GET", "/api/v1/experiments/{eid}/searcher_events")
def searcher_events(ctx: Ctx, eid: str) -> Any:
    """Long poll for a client-side search method.  Pending events, oldest first; an experiment that is finished (or
    paused with nothing pending when the poll runs out) answers with one ``experimentInactive`` event, which is how the
    runner learns that it can stop (or should tell its user to activate the experiment)."""
    deadline = time.time() + min(ctx.qi("timeoutSeconds", ctx.qi("timeout_seconds", 60)), 120)
    inactive_id = 2 ** 31 - 1          # not part of the numbered queue: never acknowledged, never replayed
    while True:
        with ctx.m.lock:
            exp = ctx.m.experiments.get(int(eid))
            if exp is None:
                row = _exp_row(ctx, eid)
                return {"searcherEvents": [{"id": inactive_id, "experimentInactive": {"experimentState": _state("STATE", row["state"])}}]}
            meth = exp.searcher.method
            if not isinstance(meth, S.CustomSearch):
                raise HTTPError(400, "experiment does not use a custom searcher")
            if meth.events:
                return {"searcherEvents": list(meth.events)}
            if exp.state in core.TERMINAL:
                return {"searcherEvents": [{"id": inactive_id, "experimentInactive": {"experimentState": _state("STATE", exp.state)}}]}
            paused = exp.state == core.PAUSED
        if time.time() > deadline:
            if paused:
                return {"searcherEvents": [{"id": inactive_id, "experimentInactive": {"experimentState": "STATE_PAUSED"}}]}
            return {"searcherEvents": []}
        time.sleep(0.2)


@route("POST", "/api/v1/experiments/{eid}/searcher_operations")
def searcher_operations(ctx: Ctx, eid: str) -> Any:
    b = ctx.body or {}
    with ctx.m.lock:
        exp = ctx.m.experiments.get(int(eid))
        if exp is None:
            raise HTTPError(404, "experiment not active")
        meth = exp.searcher.method
        if not isinstance(meth, S.CustomSearch):
            raise HTTPError(400, "experiment does not use a custom searcher")
        ops: List[Any] = []
        for o in b.get("searcherOperations") or []:
            if "createTrial" in o:
                c = o["createTrial"]
                hp = c.get("hyperparams")
                hp = json.loads(hp) if isinstance(hp, str) else (hp or {})
                ops.append(S.Create(c["requestId"], exp.searcher.rand.int64(0, (1 << 31) - 1), hp, c.get("checkpoint")))
            elif "trialOperation" in o and "validateAfter" in o["trialOperation"]:
                v = o["trialOperation"]["validateAfter"]
                ops.append(S.ValidateAfter(v["requestId"], int(v["length"])))
            elif "closeTrial" in o:
                ops.append(S.Close(o["closeTrial"]["requestId"]))
            elif "shutDown" in o:
                sd = o["shutDown"] or {}
                ops.append(S.Shutdown(bool(sd.get("cancel")), bool(sd.get("failure"))))
            elif "setSearcherProgress" in o:
                meth.progress_value = float(o["setSearcherProgress"].get("progress", 0))
        trig = b.get("triggeredByEvent") or {}
        if trig.get("id") is not None:
            meth.pop_events_through(int(trig["id"]))
        exp.process_operations(exp.searcher.record_external_ops(ops))
    return {}


# ---------------------------------------------------------------------------------------------
# trials
# ---------------------------------------------------------------------------------------------
def _trial_row(ctx: Ctx, tid: str) -> Any:
    row = ctx.m.db.one("SELECT * FROM trials WHERE id=?", (int(tid),))
    if row is None:
        raise HTTPError(404, f"trial {tid} not found")
    return row


@route("GET", "/api/v1/trials/{tid}")
def get_trial(ctx: Ctx, tid: str) -> Any:
    return {"trial": trial_json(ctx.m, _trial_row(ctx, tid))}


@route("PATCH", "/api/v1/trials/{tid}")
def patch_trial(ctx: Ctx, tid: str) -> Any:
    """Unmanaged-trial heartbeat / state (reference ``core/_heartbeat.py:67-69,129-131``)."""
    _trial_row(ctx, tid)
    st = (ctx.body or {}).get("state")
    if st:
        s = st.replace("STATE_", "")
        ctx.m.db.exec("UPDATE trials SET state=?, last_activity=?, end_time=? WHERE id=?",
                      (s, time.time(), time.time() if s in core.TERMINAL else None, int(tid)))
        if s in core.TERMINAL:
            t = ctx.m.db.one("SELECT experiment_id FROM trials WHERE id=?", (int(tid),))
            ctx.m.db.exec("UPDATE experiments SET state=?, end_time=? WHERE id=? AND unmanaged=1", (s, time.time(), t["experiment_id"]))
    else:
        ctx.m.db.exec("UPDATE trials SET last_activity=? WHERE id=?", (time.time(), int(tid)))
    return {"trial": trial_json(ctx.m, _trial_row(ctx, tid))}


@route("POST", "/api/v1/experiments/{eid}/trials")
def create_unmanaged_trial(ctx: Ctx, eid: str) -> Any:
    """``core_v2`` unmanaged mode: the client creates its own trial row."""
    row = _exp_row(ctx, eid)
    b = ctx.body or {}
    tid = ctx.m.db.exec("INSERT INTO trials(experiment_id,request_id,state,hparams,seed,start_time,searcher_state) VALUES(?,?,?,?,?,?,?)",
                        (int(eid), secrets.token_hex(8), "ACTIVE", dbmod.J(b.get("hparams") or {}), 0, time.time(), "{}"))
    ctx.m.db.exec("UPDATE trials SET task_id=? WHERE id=?", (f"trial-{tid}", tid))
    return {"trial": trial_json(ctx.m, _trial_row(ctx, str(tid)))}


@route("GET", "/api/v1/trials/{tid}/searcher/operation")
def trial_searcher_op(ctx: Ctx, tid: str) -> Any:
    with ctx.m.lock:
        try:
            t = ctx.m.get_trial(int(tid))
        except core.NotFound as e:
            raise HTTPError(404, str(e))
        op = t.current_op()
        if op is None or t.experiment.state not in (core.ACTIVE, core.PAUSED):
            return {"op": None, "completed": True}
        return {"op": {"validateAfter": {"length": str(op)}}, "completed": False}


@route("POST", "/api/v1/trials/{tid}/searcher/completed_operation")
def trial_completed_op(ctx: Ctx, tid: str) -> Any:
    b = ctx.body or {}
    length = int((b.get("op") or {}).get("length") or (b.get("completedOperation", {}).get("op", {}) or {}).get("length", 0))
    metric = b.get("searcherMetric", (b.get("completedOperation") or {}).get("searcherMetric"))
    with ctx.m.lock:
        t = ctx.m.get_trial(int(tid))
        try:
            t.experiment.op_completed(t, length, metric)
        except core.Conflict as e:
            raise HTTPError(409, str(e))
    return {}


@route("POST", "/api/v1/trials/{tid}/progress")
def trial_progress(ctx: Ctx, tid: str) -> Any:
    v = ctx.body
    if isinstance(v, dict):
        v = v.get("progress", 0)
    with ctx.m.lock:
        t = ctx.m.get_trial(int(tid))
        t.experiment.trial_progress(t, float(v or 0))
    return {}


@route("POST", "/api/v1/trials/{tid}/early_exit")
def trial_early_exit(ctx: Ctx, tid: str) -> Any:
    b = ctx.body or {}
    reason = b.get("reason") or (b.get("earlyExit") or {}).get("reason") or ""
    reason = reason.replace("EXITED_REASON_", "")
    if reason not in (S.ExitedReason.INVALID_HP, S.ExitedReason.INIT_INVALID_HP, "USER_REQUESTED_STOP"):
        raise HTTPError(400, f"unknown exited reason {reason!r}")
    if reason == "USER_REQUESTED_STOP":
        reason = S.ExitedReason.USER_CANCELED
    with ctx.m.lock:
        t = ctx.m.get_trial(int(tid))
        try:
            t.experiment.trial_early_exit(t, reason)
        except core.Conflict as e:
            raise HTTPError(400, str(e))
    return {}


@route("POST", "/api/v1/trials/{tid}/runner/metadata")
def trial_runner_metadata(ctx: Ctx, tid: str) -> Any:
    st = (ctx.body or {}).get("state", "")
    with ctx.m.lock:
        t = ctx.m.trials.get(int(tid))
        if t is not None:
            t.runner_state = st
    ctx.m.db.exec("UPDATE trials SET runner_state=? WHERE id=?", (st, int(tid)))
    return {}


@route("POST", "/api/v1/trials/{tid}/metrics")
def report_metrics(ctx: Ctx, tid: str) -> Any:
    b = ctx.body or {}
    grp = b.get("group", "training")
    mm = b.get("metrics") or {}
    inner = mm.get("metrics") or {}
    steps = int(mm.get("stepsCompleted") or 0)
    ctx.m.db.exec("INSERT INTO metrics(trial_id,trial_run_id,grp,steps_completed,metrics,batch_metrics,report_time) VALUES(?,?,?,?,?,?,?)",
                  (int(tid), int(mm.get("trialRunId") or 0), grp, steps, dbmod.J(inner.get("avgMetrics") or {}),
                   dbmod.J(inner.get("batchMetrics")) if inner.get("batchMetrics") else None, time.time()))
    if grp in ("training", "validation"):
        with ctx.m.lock:
            t = ctx.m.trials.get(int(tid))
            if t is not None:
                if grp == "training":
                    t.total_batches = max(t.total_batches, steps)
                else:
                    t.latest_validation_batches = steps
                t.persist()
            else:
                ctx.m.db.exec("UPDATE trials SET total_batches=MAX(total_batches, ?) WHERE id=?", (steps, int(tid)))
    return {}


@route("GET", "/api/v1/trials/{tid}/metrics")
@route("GET", "/api/v1/trials/metrics/trial_metrics")
def get_metrics(ctx: Ctx, tid: Optional[str] = None) -> Any:
    ids = [int(tid)] if tid else [int(x) for x in ctx.query.get("trialIds", [])]
    grp = ctx.q("group")
    out = []
    for i in ids:
        sql, args = "SELECT * FROM metrics WHERE trial_id=?", [i]
        if grp:
            sql += " AND grp=?"
            args.append(grp)
        for r in ctx.m.db.all(sql + " ORDER BY steps_completed, id", args):
            out.append({"trialId": r["trial_id"], "group": r["grp"], "totalBatches": r["steps_completed"], "trialRunId": r["trial_run_id"],
                        "endTime": _ts(r["report_time"]), "metrics": {"avgMetrics": dbmod.U(r["metrics"], {}), "batchMetrics": dbmod.U(r["batch_metrics"])}})
    return {"metrics": out}


@route("GET", "/api/v1/trials/{tid}/checkpoints")
def trial_checkpoints(ctx: Ctx, tid: str) -> Any:
    _trial_row(ctx, tid)
    rows = ctx.m.db.all("SELECT * FROM checkpoints WHERE trial_id=? ORDER BY report_time", (int(tid),))
    out = [ckpt_json(ctx.m, r) for r in rows]
    states = ctx.query.get("states") or []
    if states:
        out = [c for c in out if c["state"] in states]
    if ctx.q("orderBy") == "ORDER_BY_DESC":
        out.reverse()
    items, pg = page(out, ctx)
    return {"checkpoints": items, "pagination": pg}


@route("POST", "/api/v1/trials/{tid}/kill")
def kill_trial(ctx: Ctx, tid: str) -> Any:
    with ctx.m.lock:
        t = ctx.m.get_trial(int(tid))
        if t.allocation is not None:
            t.allocation.kill()
        t.ops = []
        t.close_requested = True
        t.early_exit_reason = t.early_exit_reason or S.ExitedReason.USER_CANCELED
        t.persist()
    return {}


def _log_time(s: Optional[str]) -> Optional[float]:
    if not s:
        return None
    try:
        return float(s)
    except ValueError:
        import datetime

        try:
            d = datetime.datetime.fromisoformat(s.replace("Z", "+00:00"))
        except ValueError:
            raise HTTPError(400, f"cannot parse time {s!r}")
        return (d if d.tzinfo else d.replace(tzinfo=datetime.timezone.utc)).timestamp()


def _logs(ctx: Ctx, task_id: str) -> Any:
    follow = ctx.q("follow") in ("true", "1")
    after = ctx.qi("offset", 0)
    limit = ctx.qi("limit", 0) or 5000
    deadline = time.time() + (min(ctx.qi("timeout_seconds", 20), 60) if follow else 0)
    # filters (reference TrialLogs / TaskLogs request fields): levels, rankIds, sources, stdtypes, allocationIds (containerIds
    # is accepted as a synonym: one process group per allocation here), timestampAfter / timestampBefore, searchText
    levels = {x.replace("LOG_LEVEL_", "") for x in ctx.query.get("levels", [])}
    ranks = {int(x) for x in ctx.query.get("rankIds", []) if str(x).lstrip("-").isdigit()}
    sources, stdtypes = set(ctx.query.get("sources", [])), set(ctx.query.get("stdtypes", []))
    allocs = set(ctx.query.get("allocationIds", [])) | set(ctx.query.get("containerIds", []))
    t_after = _log_time(ctx.q("timestampAfter"))
    t_before = _log_time(ctx.q("timestampBefore"))
    needle = ctx.q("searchText")
    filtered = bool(levels or ranks or sources or stdtypes or allocs or t_after or t_before or needle)

    def keep(r: Any) -> bool:
        return ((not levels or (r["level"] or "INFO") in levels) and (not ranks or r["rank_id"] in ranks) and (not sources or r["source"] in sources)
                and (not stdtypes or r["stdtype"] in stdtypes) and (not allocs or r["allocation_id"] in allocs)
                and (t_after is None or r["ts"] > t_after) and (t_before is None or r["ts"] < t_before) and (not needle or needle in (r["log"] or "")))

    while True:
        rows = ctx.m.db.get_logs(task_id, after, limit if not filtered else 1_000_000)
        if filtered:
            rows = [r for r in rows if keep(r)][:limit]
        if rows or not follow or time.time() > deadline:
            break
        running = ctx.m.db.one("SELECT COUNT(*) c FROM allocations WHERE task_id=? AND state='RUNNING'", (task_id,))["c"]
        if not running:
            break
        time.sleep(0.25)
    return {"logs": [{"id": r["id"], "taskId": r["task_id"], "allocationId": r["allocation_id"], "timestamp": _ts(r["ts"]), "level": "LOG_LEVEL_" + (r["level"] or "INFO"),
                      "rankId": r["rank_id"], "stdtype": r["stdtype"], "source": r["source"], "log": r["log"], "message": r["log"]} for r in rows]}


@route("GET", "/api/v1/trials/{tid}/logs")
def trial_logs(ctx: Ctx, tid: str) -> Any:
    _trial_row(ctx, tid)
    return _logs(ctx, f"trial-{int(tid)}")


@route("GET", "/api/v1/tasks/{task_id}/logs")
def task_logs(ctx: Ctx, task_id: str) -> Any:
    return _logs(ctx, task_id)


@route("POST", "/task-logs")
@route("POST", "/api/v1/task/logs")
def post_task_logs(ctx: Ctx) -> Any:
    rows = []
    items = ctx.body if isinstance(ctx.body, list) else (ctx.body or {}).get("logs", [])
    for it in items:
        rows.append((it.get("task_id") or it.get("taskId"), it.get("allocation_id"), time.time(), it.get("level", "INFO"),
                     it.get("rank_id", it.get("rank")), it.get("stdtype", "stdout"), it.get("source", "task"), it.get("log", "")))
    if rows:
        ctx.m.db.add_logs(rows)
    return {}


# ---------------------------------------------------------------------------------------------
# checkpoints
# ---------------------------------------------------------------------------------------------
@route("POST", "/api/v1/checkpoints")
def report_checkpoint(ctx: Ctx) -> Any:
    b = ctx.body or {}
    ck = b.get("checkpoint", b)
    uuid_ = ck.get("uuid")
    if not uuid_:
        raise HTTPError(400, "checkpoint uuid required")
    # the id becomes a directory name under checkpoint storage (GC and DELETE rmtree it): only real UUIDs, as in the
    # reference's uuid-typed column (master/internal/api_checkpoint.go, checkpoints_v2.uuid)
    try:
        uuid_ = str(uuid_mod.UUID(str(uuid_)))
    except (ValueError, AttributeError, TypeError):
        raise HTTPError(400, f"checkpoint uuid {uuid_!r} is not a UUID")
    md = ck.get("metadata") or {}
    res = {k: int(v) for k, v in (ck.get("resources") or {}).items()}
    t = ctx.m.db.one("SELECT id FROM trials WHERE task_id=?", (ck.get("taskId"),))
    tid = t["id"] if t is not None else None
    steps = int(md.get("steps_completed", 0) or 0)
    state = (ck.get("state") or "STATE_COMPLETED").replace("STATE_", "")
    ctx.m.db.exec("INSERT OR REPLACE INTO checkpoints(uuid,trial_id,task_id,allocation_id,state,resources,metadata,report_time,steps_completed,storage_id,size) VALUES(?,?,?,?,?,?,?,?,?,?,?)",
                  (uuid_, tid, ck.get("taskId"), ck.get("allocationId"), state, dbmod.J(res), dbmod.J(md), time.time(), steps, ck.get("storageId"), sum(res.values())))
    if tid is not None and state == "COMPLETED":
        with ctx.m.lock:
            tr = ctx.m.trials.get(tid)
            if tr is not None:
                tr.latest_checkpoint = uuid_
                tr.total_batches = max(tr.total_batches, steps)
                tr.persist()
            else:
                ctx.m.db.exec("UPDATE trials SET latest_checkpoint=? WHERE id=?", (uuid_, tid))
    return {}


def _ckpt_row(ctx: Ctx, uuid_: str) -> Any:
    row = ctx.m.db.one("SELECT * FROM checkpoints WHERE uuid=?", (uuid_,))
    if row is None:
        raise HTTPError(404, f"checkpoint {uuid_} not found")
    return row


@route("GET", "/api/v1/checkpoints/{uuid_}")
def get_checkpoint(ctx: Ctx, uuid_: str) -> Any:
    return {"checkpoint": ckpt_json(ctx.m, _ckpt_row(ctx, uuid_))}


@route("POST", "/api/v1/checkpoints/{uuid_}/metadata")
def post_checkpoint_metadata(ctx: Ctx, uuid_: str) -> Any:
    row = _ckpt_row(ctx, uuid_)
    md = dbmod.U(row["metadata"], {})
    md.update(((ctx.body or {}).get("checkpoint") or {}).get("metadata") or (ctx.body or {}).get("metadata") or {})
    ctx.m.db.exec("UPDATE checkpoints SET metadata=? WHERE uuid=?", (dbmod.J(md), uuid_))
    return {"checkpoint": ckpt_json(ctx.m, _ckpt_row(ctx, uuid_))}


@route("PATCH", "/api/v1/checkpoints")
def patch_checkpoints(ctx: Ctx) -> Any:
    for c in (ctx.body or {}).get("checkpoints", []):
        row = ctx.m.db.one("SELECT * FROM checkpoints WHERE uuid=?", (c.get("uuid"),))
        if row is None:
            continue
        if "resources" in c and c["resources"] is not None:
            res = (c["resources"] or {}).get("resources", c["resources"])
            res = {k: int(v) for k, v in (res or {}).items()}
            ctx.m.db.exec("UPDATE checkpoints SET resources=?, size=?, state=? WHERE uuid=?",
                          (dbmod.J(res), sum(res.values()), "COMPLETED" if res else "DELETED", c["uuid"]))
    return {}


def _delete_ckpts(ctx: Ctx, uuids: List[str], globs: List[str]) -> None:
    from determined_b200.common import storage

    for u in uuids:
        row = ctx.m.db.one("SELECT c.*, t.experiment_id FROM checkpoints c LEFT JOIN trials t ON t.id=c.trial_id WHERE c.uuid=?", (u,))
        if row is None:
            raise HTTPError(404, f"checkpoint {u} not found")
        if ctx.m.db.one("SELECT id FROM model_versions WHERE checkpoint_uuid=?", (u,)) is not None:
            raise HTTPError(400, f"checkpoint {u} is registered in the model registry and cannot be deleted")
        e = ctx.m.db.one("SELECT config FROM experiments WHERE id=?", (row["experiment_id"],))
        cfg = dbmod.U(e["config"], {}) if e is not None else {"checkpoint_storage": ctx.m.cluster_defaults["checkpoint_storage"]}
        mgr = storage.build(cfg["checkpoint_storage"], None)
        left = mgr.delete(u, globs)
        full = "**/*" in globs or not left
        ctx.m.db.exec("UPDATE checkpoints SET resources=?, size=?, state=? WHERE uuid=?",
                      (dbmod.J(left), sum(left.values()), "DELETED" if full else ("PARTIALLY_DELETED" if len(left) < len(dbmod.U(row["resources"], {})) else row["state"]), u))


@route("DELETE", "/api/v1/checkpoints")
def delete_checkpoints(ctx: Ctx) -> Any:
    _delete_ckpts(ctx, (ctx.body or {}).get("checkpointUuids", []), ["**/*"])
    return {}


@route("POST", "/api/v1/checkpoints/rm")
def rm_checkpoints(ctx: Ctx) -> Any:
    b = ctx.body or {}
    _delete_ckpts(ctx, b.get("checkpointUuids", []), b.get("checkpointGlobs") or ["**/*"])
    return {}


@route("POST", "/api/v1/runs/start")
def runs_start(ctx: Ctx) -> Any:
    return {"storageId": None}


@route("POST", "/api/v1/trial-source-info")
def trial_source_info(ctx: Ctx) -> Any:
    i = (ctx.body or {}).get("trialSourceInfo") or {}
    ctx.m.db.exec("INSERT INTO trial_source_infos(trial_id,checkpoint_uuid,type,model_id,model_version) VALUES(?,?,?,?,?)",
                  (i.get("trialId"), i.get("checkpointUuid"), i.get("trialSourceInfoType"), i.get("modelId"), i.get("modelVersion")))
    return {"trialId": i.get("trialId"), "checkpointUuid": i.get("checkpointUuid")}


# ---------------------------------------------------------------------------------------------
# allocations
# ---------------------------------------------------------------------------------------------
def _alloc(ctx: Ctx, aid: str) -> core.Allocation:
    try:
        return ctx.m.get_allocation(aid)
    except core.NotFound as e:
        raise HTTPError(404, str(e))


@route("GET", "/api/v1/allocations/{aid}/signals/preemption")
def preemption_signal(ctx: Ctx, aid: str) -> Any:
    a = _alloc(ctx, aid)
    timeout = min(float(ctx.q("timeout_seconds", ctx.q("timeoutSeconds", 60)) or 0), 120.0)
    return {"preempt": a.wait_preemption(timeout)}


@route("POST", "/api/v1/allocations/{aid}/signals/ack_preemption")
def ack_preemption(ctx: Ctx, aid: str) -> Any:
    a = ctx.m.allocations.get(aid)
    if a is not None:
        a.preempt_acked = True
    return {}


@route("POST", "/api/v1/allocations/{aid}/signals/pending_preemption")
def pending_preemption(ctx: Ctx, aid: str) -> Any:
    _alloc(ctx, aid).preempt()
    return {}


@route("GET", "/api/v1/allocations/{aid}/resources/{rid}/rendezvous")
def rendezvous(ctx: Ctx, aid: str, rid: str) -> Any:
    a = _alloc(ctx, aid)
    return {"rendezvousInfo": {"addresses": ["127.0.0.1"], "rank": 0, "slots": [len(a.slot_ids)]}}


@route("POST", "/api/v1/allocations/{aid}/all_gather")
def all_gather(ctx: Ctx, aid: str) -> Any:
    a = _alloc(ctx, aid)
    b = ctx.body or {}
    try:
        data = a.all_gather(b.get("requestUuid", ""), int(b.get("rank", len(a.allgathers.get(b.get("requestUuid", ""), {})))),
                            int(b.get("numPeers", 1)), b.get("data"))
    except TimeoutError as e:
        raise HTTPError(408, str(e))
    return {"data": data}


@route("POST", "/api/v1/allocations/{aid}/proxy_address")
@route("POST", "/api/v1/allocations/{aid}/resources/{rid}/daemon")
@route("POST", "/api/v1/allocations/{aid}/ready")
@route("POST", "/api/v1/allocations/{aid}/waiting")
def alloc_noop(ctx: Ctx, aid: str, rid: Optional[str] = None) -> Any:
    _alloc(ctx, aid)
    return {}


@route("GET", "/api/v1/allocations/{aid}")
def get_allocation(ctx: Ctx, aid: str) -> Any:
    r = ctx.m.db.one("SELECT * FROM allocations WHERE id=?", (aid,))
    if r is None:
        raise HTTPError(404, "allocation not found")
    return {"allocation": {"allocationId": r["id"], "taskId": r["task_id"], "state": "STATE_" + r["state"], "slots": r["slots"],
                           "startTime": _ts(r["start_time"]), "endTime": _ts(r["end_time"]), "exitCode": r["exit_code"]}}


# ---------------------------------------------------------------------------------------------
# cluster: agents / slots / pools / job queue
# ---------------------------------------------------------------------------------------------
def _agent_json(m: core.Master) -> Dict[str, Any]:
    ag = m.pool.agents[m.agent_id]
    slots = {}
    for i in range(ag.num_slots):
        g = m.gpus[i] if i < len(m.gpus) else None
        owner = ag.used.get(i)
        slots[str(i)] = {"id": str(i), "enabled": (ag.slot_enabled or [True] * ag.num_slots)[i],
                         "device": {"id": i, "brand": g.name if g else "cpu", "uuid": g.uuid if g else f"cpu-{i}", "type": "DEVICE_TYPE_CUDA" if g else "DEVICE_TYPE_CPU"},
                         "container": {"id": owner, "state": "STATE_RUNNING"} if owner else None, "draining": False}
    return {"id": m.agent_id, "registeredTime": _ts(m.start_time), "slots": slots, "resourcePools": [m.pool.name], "enabled": ag.enabled,
            "draining": False, "addresses": ["127.0.0.1"], "version": __version__, "label": ag.label}


def _any_agent_json(m: core.Master, agent_id: str) -> Dict[str, Any]:
    if agent_id == m.agent_id:
        return _agent_json(m)
    ra = m.agents.get(agent_id)
    if ra is None:
        raise HTTPError(404, "agent not found")
    owner = m.pool_of_agent(agent_id)
    j = ra.to_json(owner.agents.get(agent_id) if owner is not None else None)
    j["registeredTime"] = _ts(ra.registered)
    return j


@route("GET", "/api/v1/agents")
def agents(ctx: Ctx) -> Any:
    m = ctx.m
    out = [_agent_json(m)] if m.pool.agents[m.agent_id].num_slots or not m.agents.agents else []
    out += [_any_agent_json(m, a) for a in sorted(m.agents.agents)]
    return {"agents": out, "pagination": {"total": len(out)}}


@route("GET", "/api/v1/agents/{agent_id}")
def get_agent(ctx: Ctx, agent_id: str) -> Any:
    return {"agent": _any_agent_json(ctx.m, agent_id)}


@route("GET", "/api/v1/agents/{agent_id}/slots")
def get_slots(ctx: Ctx, agent_id: str) -> Any:
    return {"slots": list(_any_agent_json(ctx.m, agent_id)["slots"].values())}


def _agent_auth(ctx: Ctx) -> None:
    """Agents are part of the cluster, not users: they present the cluster's agent token if one is configured
    (``DET_AGENT_TOKEN`` / ``--agent-token``), like the reference's internal agent endpoint they need no login."""
    import os

    want = getattr(ctx.m, "agent_token", None) or os.environ.get("DET_AGENT_TOKEN")
    if want and (ctx.body or {}).get("token") != want:
        raise HTTPError(403, "bad agent token")


@route("POST", "/api/v1/agents/register", auth_required=False)
def agent_register(ctx: Ctx) -> Any:
    """Agent protocol (reference ``aproto.AgentStarted`` over the agent websocket): announce slots and the tasks still running."""
    _agent_auth(ctx)
    try:
        return ctx.m.agents.register(ctx.body or {})
    except ValueError as e:
        raise HTTPError(400, str(e))


@route("POST", "/api/v1/agents/{agent_id}/poll", auth_required=False)
def agent_poll(ctx: Ctx, agent_id: str) -> Any:
    """Events up (task logs, exits), actions down (start / kill); held open up to ``timeoutSeconds`` when idle."""
    _agent_auth(ctx)
    b = ctx.body or {}
    return ctx.m.agents.poll(agent_id, b, float(b.get("timeoutSeconds", 10)))


def _set_enabled(ctx: Ctx, agent_id: str, slot: Optional[str], enabled: bool) -> Any:
    owner = ctx.m.pool_of_agent(agent_id)
    if owner is None:
        raise HTTPError(404, "agent not found")
    with ctx.m.lock:
        ag = owner.agents[agent_id]
        if slot is None:
            ag.enabled = enabled
        else:
            if ag.slot_enabled is None:
                ag.slot_enabled = [True] * ag.num_slots
            ag.slot_enabled[int(slot)] = enabled
    ctx.m.kick()
    return {"agent": _any_agent_json(ctx.m, agent_id)}


route("POST", "/api/v1/agents/{agent_id}/enable")(lambda ctx, agent_id: _set_enabled(ctx, agent_id, None, True))
route("POST", "/api/v1/agents/{agent_id}/disable")(lambda ctx, agent_id: _set_enabled(ctx, agent_id, None, False))
route("POST", "/api/v1/agents/{agent_id}/slots/{slot}/enable")(lambda ctx, agent_id, slot: _set_enabled(ctx, agent_id, slot, True))
route("POST", "/api/v1/agents/{agent_id}/slots/{slot}/disable")(lambda ctx, agent_id, slot: _set_enabled(ctx, agent_id, slot, False))


@route("GET", "/api/v1/resource-pools")
def resource_pools(ctx: Ctx) -> Any:
    out = []
    for p in ctx.m.pools.values():
        live = [a for a in p.agents.values() if a.enabled and a.num_slots]
        out.append({"name": p.name, "description": ctx.m.pool_descriptions.get(p.name, ""), "type": "RESOURCE_POOL_TYPE_STATIC", "numAgents": len(live),
                    "slotsAvailable": sum(a.num_slots for a in live), "slotsUsed": sum(a.num_used() for a in p.agents.values()),
                    "slotType": "DEVICE_TYPE_CUDA" if ctx.m.num_gpus else "DEVICE_TYPE_CPU", "defaultComputePool": p.name == "default",
                    "defaultAuxPool": p.name == "default", "preemptible": p.preemption,
                    "schedulerType": "SCHEDULER_TYPE_" + (ctx.m.backends[p.name].NAME if p.name in getattr(ctx.m, "backends", {}) else p.scheduler).upper(),
                    "schedulerFittingPolicy": "FITTING_POLICY_" + p.policy.upper(), "slotsPerAgent": max((a.num_slots for a in p.agents.values()), default=0),
                    "location": "on-prem", "auxContainerCapacity": 100 * max(len(p.agents), 1),
                    "auxContainersRunning": sum(len(a.zero_slot_tasks) for a in p.agents.values())})
    return {"resourcePools": out, "pagination": {"total": len(out)}}


@route("GET", "/api/v1/job-queues-v2")
@route("GET", "/api/v1/job-queues")
def job_queue(ctx: Ctx) -> Any:
    """Jobs of one pool (``resourcePool=``) or of all pools."""
    want = ctx.q("resourcePool")
    jobs: Dict[str, Dict[str, Any]] = {}
    with ctx.m.lock:
        per_pool = [(p.name, p.job_queue()) for p in ctx.m.pools.values() if not want or p.name == want]
    for pname, rows in per_pool:
        for r in rows:
            j = jobs.setdefault(r["job_id"], {"jobId": r["job_id"], "type": "TYPE_EXPERIMENT", "entityId": r["job_id"].replace("exp-", ""), "priority": r["priority"],
                                              "weight": r["weight"], "allocatedSlots": 0, "requestedSlots": 0, "resourcePool": pname, "name": r["name"],
                                              "summary": {"state": "STATE_QUEUED", "jobsAhead": 0}})
            j["requestedSlots"] += r["slots"]
            if r["state"] == "SCHEDULED":
                j["allocatedSlots"] += r["slots"]
                j["summary"]["state"] = "STATE_SCHEDULED"
    return {"jobs": list(jobs.values()), "pagination": {"total": len(jobs)}}


@route("POST", "/api/v1/job-queues")
def update_job_queue(ctx: Ctx) -> Any:
    # security.authz.strict_job_queue_control: only administrators (or, under rbac, holders of control_strict_job_queue) reorder jobs
    # or change their priority / weight; otherwise owners may tune their own jobs like any other field of their experiments
    if getattr(ctx.m, "strict_job_queue_control", False) and not (
            ctx.user["admin"] or authz.allowed(ctx.m, ctx.user, "PERMISSION_TYPE_CONTROL_STRICT_JOB_QUEUE", None) and authz.enabled(ctx.m)):
        raise HTTPError(403, "the job queue is under strict control: only administrators change priorities, weights and positions")
    with ctx.m.lock:
        for u in (ctx.body or {}).get("updates", []):
            jid = u.get("jobId")
            for p in ctx.m.pools.values():
                if u.get("priority") is not None:
                    p.set_priority(jid, int(u["priority"]))
                if u.get("weight") is not None:
                    p.set_weight(jid, float(u["weight"]))
            if u.get("resourcePool"):
                # queued requests of the job change queues now; what is running stays where it is and asks the new pool next time
                dst = ctx.m.pools.get(str(u["resourcePool"]))
                if dst is None:
                    raise HTTPError(404, f"resource pool {u['resourcePool']} not found")
                for p in ctx.m.pools.values():
                    if p is dst:
                        continue
                    for aid, req in list(p.pending.items()):
                        if req.job_id == jid:
                            del p.pending[aid]
                            dst.allocate(req)
                for exp in ctx.m.experiments.values():
                    if exp.job_id == jid:
                        exp.config.setdefault("resources", {})["resource_pool"] = dst.name
                        ctx.m.db.exec("UPDATE experiments SET config=? WHERE id=?", (dbmod.J(exp.config), exp.id))
            ahead, behind = u.get("aheadOf") or (u.get("queuePosition") or {}).get("aheadOf"), u.get("behindOf") or (u.get("queuePosition") or {}).get("behindOf")
            if ahead or behind:
                moved = False
                for p in ctx.m.pools.values():
                    try:
                        p.move_job(jid, ahead_of=ahead, behind_of=behind)
                        moved = True
                    except KeyError:
                        continue
                    except ValueError as e:
                        raise HTTPError(400, str(e))
                if not moved:
                    raise HTTPError(404, f"job {jid} and its anchor share no resource pool queue")
            for exp in ctx.m.experiments.values():
                if exp.job_id == jid:
                    if u.get("priority") is not None:
                        exp.config.setdefault("resources", {})["priority"] = int(u["priority"])
                    if u.get("weight") is not None:
                        exp.config.setdefault("resources", {})["weight"] = float(u["weight"])
    ctx.m.kick()
    return {}


# ---------------------------------------------------------------------------------------------
# templates, model registry, workspaces / projects, webhooks
# ---------------------------------------------------------------------------------------------
@route("GET", "/api/v1/templates")
def list_templates(ctx: Ctx) -> Any:
    return {"templates": [{"name": r["name"], "config": dbmod.U(r["config"], {}), "workspaceId": r["workspace_id"]} for r in ctx.m.db.all("SELECT * FROM templates ORDER BY name")]}


@route("GET", "/api/v1/templates/{name}")
def get_template(ctx: Ctx, name: str) -> Any:
    r = ctx.m.db.one("SELECT * FROM templates WHERE name=?", (name,))
    if r is None:
        raise HTTPError(404, "template not found")
    return {"template": {"name": r["name"], "config": dbmod.U(r["config"], {}), "workspaceId": r["workspace_id"]}}


@route("PUT", "/api/v1/templates/{name}")
@route("POST", "/api/v1/templates/{name}")
@route("PATCH", "/api/v1/templates/{name}")
def put_template(ctx: Ctx, name: str) -> Any:
    b = ctx.body or {}
    cfg = b.get("config", b.get("template", {}).get("config", b))
    if isinstance(cfg, str):
        cfg = expconf.parse(cfg)
    ctx.m.db.exec("INSERT INTO templates(name,config) VALUES(?,?) ON CONFLICT(name) DO UPDATE SET config=excluded.config", (name, dbmod.J(cfg)))
    wid = b.get("workspaceId") or (b.get("template") or {}).get("workspaceId")
    if wid:
        if ctx.m.db.one("SELECT id FROM workspaces WHERE id=?", (int(wid),)) is None:
            raise HTTPError(404, f"workspace {wid} not found")
        ctx.m.db.exec("UPDATE templates SET workspace_id=? WHERE name=?", (int(wid), name))
    r = ctx.m.db.one("SELECT * FROM templates WHERE name=?", (name,))
    return {"template": {"name": name, "config": cfg, "workspaceId": r["workspace_id"]}}


@route("DELETE", "/api/v1/templates/{name}")
def delete_template(ctx: Ctx, name: str) -> Any:
    ctx.m.db.exec("DELETE FROM templates WHERE name=?", (name,))
    return {}


def _model_json(m: core.Master, r: Any) -> Dict[str, Any]:
    n = m.db.one("SELECT COUNT(*) c FROM model_versions WHERE model_id=?", (r["id"],))["c"]
    u = m.db.one("SELECT username FROM users WHERE id=?", (r["user_id"],))
    return {"id": r["id"], "name": r["name"], "description": r["description"], "metadata": dbmod.U(r["metadata"], {}), "labels": dbmod.U(r["labels"], []),
            "creationTime": _ts(r["creation_time"]), "lastUpdatedTime": _ts(r["last_updated_time"]), "numVersions": n, "username": u["username"] if u else None,
            "userId": r["user_id"], "archived": bool(r["archived"]), "notes": r["notes"], "workspaceId": r["workspace_id"]}


def _mv_json(m: core.Master, r: Any) -> Dict[str, Any]:
    model = m.db.one("SELECT * FROM models WHERE id=?", (r["model_id"],))
    ck = m.db.one("SELECT * FROM checkpoints WHERE uuid=?", (r["checkpoint_uuid"],))
    return {"id": r["id"], "version": r["version"], "model": _model_json(m, model), "checkpoint": ckpt_json(m, ck) if ck is not None else {"uuid": r["checkpoint_uuid"]},
            "name": r["name"], "comment": r["comment"], "metadata": dbmod.U(r["metadata"], {}), "labels": dbmod.U(r["labels"], []),
            "creationTime": _ts(r["creation_time"]), "notes": r["notes"], "userId": r["user_id"]}


def _model(ctx: Ctx, name: str) -> Any:
    r = ctx.m.db.one("SELECT * FROM models WHERE name=? OR CAST(id AS TEXT)=?", (name, name))
    if r is None:
        raise HTTPError(404, f"model {name!r} not found")
    return r


@route("GET", "/api/v1/models")
def list_models(ctx: Ctx) -> Any:
    rows = [_model_json(ctx.m, r) for r in ctx.m.db.all("SELECT * FROM models ORDER BY id")]
    if ctx.q("name"):
        rows = [r for r in rows if ctx.q("name") in r["name"]]
    if ctx.q("description"):
        rows = [r for r in rows if ctx.q("description") in (r["description"] or "")]
    if ctx.query.get("labels"):
        rows = [r for r in rows if set(ctx.query["labels"]) <= set(r["labels"])]
    if ctx.q("archived") is not None:
        rows = [r for r in rows if r["archived"] == (ctx.q("archived") in ("true", "1"))]
    if ctx.query.get("workspaceIds"):
        rows = [r for r in rows if r["workspaceId"] in {int(x) for x in ctx.query["workspaceIds"]}]
    ws_names = {w["id"]: w["name"] for w in ctx.m.db.all("SELECT id, name FROM workspaces")}
    keys = {"SORT_BY_NAME": lambda r: r["name"], "SORT_BY_DESCRIPTION": lambda r: r["description"], "SORT_BY_CREATION_TIME": lambda r: r["creationTime"],
            "SORT_BY_LAST_UPDATED_TIME": lambda r: r["lastUpdatedTime"], "SORT_BY_NUM_VERSIONS": lambda r: r["numVersions"],
            "SORT_BY_WORKSPACE": lambda r: ws_names.get(r["workspaceId"], "")}
    rows = sort_rows(rows, keys.get(ctx.q("sortBy") or "", lambda r: r["id"]), ctx.q("orderBy"))
    items, pg = page(rows, ctx)
    return {"models": items, "pagination": pg}


@route("POST", "/api/v1/models")
def create_model(ctx: Ctx) -> Any:
    b = ctx.body or {}
    try:
        mid = ctx.m.db.exec("INSERT INTO models(name,description,metadata,labels,user_id,creation_time,last_updated_time,notes) VALUES(?,?,?,?,?,?,?,?)",
                            (b.get("name"), b.get("description", ""), dbmod.J(b.get("metadata") or {}), dbmod.J(b.get("labels") or []), ctx.user["id"], time.time(), time.time(), b.get("notes", "")))
    except Exception as e:
        raise HTTPError(409, f"cannot create model: {e}")
    wid = b.get("workspaceId")
    if not wid and b.get("workspaceName"):
        w = ctx.m.db.one("SELECT id FROM workspaces WHERE name=?", (b["workspaceName"],))
        if w is None:
            raise HTTPError(404, f"workspace {b['workspaceName']!r} not found")
        wid = w["id"]
    if wid:
        if ctx.m.db.one("SELECT id FROM workspaces WHERE id=?", (int(wid),)) is None:
            raise HTTPError(404, f"workspace {wid} not found")
        ctx.m.db.exec("UPDATE models SET workspace_id=? WHERE id=?", (int(wid), mid))
    return {"model": _model_json(ctx.m, ctx.m.db.one("SELECT * FROM models WHERE id=?", (mid,)))}


@route("GET", "/api/v1/models/{name}")
def get_model(ctx: Ctx, name: str) -> Any:
    return {"model": _model_json(ctx.m, _model(ctx, urllib.parse.unquote(name)))}


@route("PATCH", "/api/v1/models/{name}")
def patch_model(ctx: Ctx, name: str) -> Any:
    r = _model(ctx, urllib.parse.unquote(name))
    b = (ctx.body or {}).get("model", ctx.body or {})
    for col, key, enc in (("description", "description", None), ("metadata", "metadata", dbmod.J), ("labels", "labels", dbmod.J), ("notes", "notes", None), ("name", "name", None)):
        if key in b and b[key] is not None:
            ctx.m.db.exec(f"UPDATE models SET {col}=?, last_updated_time=? WHERE id=?", (enc(b[key]) if enc else b[key], time.time(), r["id"]))
    wid = b.get("workspaceId")
    if wid is None and b.get("workspaceName"):
        w = ctx.m.db.one("SELECT id FROM workspaces WHERE name=?", (b["workspaceName"],))
        if w is None:
            raise HTTPError(404, f"workspace {b['workspaceName']!r} not found")
        wid = w["id"]
    if wid is not None:             # move the model (reference MoveModel)
        if ctx.m.db.one("SELECT id FROM workspaces WHERE id=?", (int(wid),)) is None:
            raise HTTPError(404, f"workspace {wid} not found")
        ctx.m.db.exec("UPDATE models SET workspace_id=?, last_updated_time=? WHERE id=?", (int(wid), time.time(), r["id"]))
    return {"model": _model_json(ctx.m, ctx.m.db.one("SELECT * FROM models WHERE id=?", (r["id"],)))}


@route("POST", "/api/v1/models/{name}/archive")
def archive_model(ctx: Ctx, name: str) -> Any:
    ctx.m.db.exec("UPDATE models SET archived=1 WHERE id=?", (_model(ctx, urllib.parse.unquote(name))["id"],))
    return {}


@route("POST", "/api/v1/models/{name}/unarchive")
def unarchive_model(ctx: Ctx, name: str) -> Any:
    ctx.m.db.exec("UPDATE models SET archived=0 WHERE id=?", (_model(ctx, urllib.parse.unquote(name))["id"],))
    return {}


@route("DELETE", "/api/v1/models/{name}")
def delete_model(ctx: Ctx, name: str) -> Any:
    r = _model(ctx, urllib.parse.unquote(name))
    ctx.m.db.exec("DELETE FROM model_versions WHERE model_id=?", (r["id"],))
    ctx.m.db.exec("DELETE FROM models WHERE id=?", (r["id"],))
    return {}


@route("GET", "/api/v1/models/{name}/versions")
def list_model_versions(ctx: Ctx, name: str) -> Any:
    r = _model(ctx, urllib.parse.unquote(name))
    rows = ctx.m.db.all("SELECT * FROM model_versions WHERE model_id=? ORDER BY version", (r["id"],))
    out = [_mv_json(ctx.m, v) for v in rows]
    if ctx.q("orderBy") == "ORDER_BY_DESC":
        out.reverse()
    items, pg = page(out, ctx)
    return {"model": _model_json(ctx.m, r), "modelVersions": items, "pagination": pg}


@route("POST", "/api/v1/models/{name}/versions")
def create_model_version(ctx: Ctx, name: str) -> Any:
    r = _model(ctx, urllib.parse.unquote(name))
    b = ctx.body or {}
    if ctx.m.db.one("SELECT uuid FROM checkpoints WHERE uuid=?", (b.get("checkpointUuid"),)) is None:
        raise HTTPError(404, f"checkpoint {b.get('checkpointUuid')} not found")
    last = ctx.m.db.one("SELECT MAX(version) v FROM model_versions WHERE model_id=?", (r["id"],))["v"] or 0
    vid = ctx.m.db.exec("INSERT INTO model_versions(model_id,version,checkpoint_uuid,name,comment,metadata,labels,user_id,creation_time,notes) VALUES(?,?,?,?,?,?,?,?,?,?)",
                        (r["id"], last + 1, b["checkpointUuid"], b.get("name", ""), b.get("comment", ""), dbmod.J(b.get("metadata") or {}), dbmod.J(b.get("labels") or []),
                         ctx.user["id"], time.time(), b.get("notes", "")))
    ctx.m.db.exec("UPDATE models SET last_updated_time=? WHERE id=?", (time.time(), r["id"]))
    return {"modelVersion": _mv_json(ctx.m, ctx.m.db.one("SELECT * FROM model_versions WHERE id=?", (vid,)))}


@route("GET", "/api/v1/models/{name}/versions/{version}")
def get_model_version(ctx: Ctx, name: str, version: str) -> Any:
    r = _model(ctx, urllib.parse.unquote(name))
    v = ctx.m.db.one("SELECT * FROM model_versions WHERE model_id=? AND version=?", (r["id"], int(version)))
    if v is None:
        raise HTTPError(404, "model version not found")
    return {"modelVersion": _mv_json(ctx.m, v)}


@route("DELETE", "/api/v1/models/{name}/versions/{version}")
def delete_model_version(ctx: Ctx, name: str, version: str) -> Any:
    r = _model(ctx, urllib.parse.unquote(name))
    ctx.m.db.exec("DELETE FROM model_versions WHERE model_id=? AND version=?", (r["id"], int(version)))
    return {}


@route("GET", "/api/v1/workspaces")
def list_workspaces(ctx: Ctx) -> Any:
    rows = [{"id": r["id"], "name": r["name"], "archived": bool(r["archived"]), "immutable": bool(r["immutable"]), "userId": r["user_id"],
             "numProjects": ctx.m.db.one("SELECT COUNT(*) c FROM projects WHERE workspace_id=?", (r["id"],))["c"]} for r in ctx.m.db.all("SELECT * FROM workspaces ORDER BY id")]
    if ctx.q("name"):
        rows = [r for r in rows if r["name"] == ctx.q("name")]
    items, pg = page(rows, ctx)
    return {"workspaces": items, "pagination": pg}


@route("POST", "/api/v1/workspaces")
def create_workspace(ctx: Ctx) -> Any:
    try:
        wid = ctx.m.db.exec("INSERT INTO workspaces(name,user_id) VALUES(?,?)", ((ctx.body or {}).get("name"), ctx.user["id"]))
    except Exception as e:
        raise HTTPError(409, f"cannot create workspace: {e}")
    return {"workspace": {"id": wid, "name": (ctx.body or {}).get("name"), "archived": False, "immutable": False, "userId": ctx.user["id"], "numProjects": 0}}


@route("GET", "/api/v1/workspaces/{wid}")
def get_workspace(ctx: Ctx, wid: str) -> Any:
    r = ctx.m.db.one("SELECT * FROM workspaces WHERE id=?", (int(wid),))
    if r is None:
        raise HTTPError(404, "workspace not found")
    from determined_b200.master import api_more

    return {"workspace": api_more.workspace_json(ctx, r)}


@route("DELETE", "/api/v1/workspaces/{wid}")
def delete_workspace(ctx: Ctx, wid: str) -> Any:
    if int(wid) == 1:
        raise HTTPError(400, "cannot delete the default workspace")
    if ctx.m.db.one("SELECT id FROM experiments WHERE project_id IN (SELECT id FROM projects WHERE workspace_id=?)", (int(wid),)) is not None:
        raise HTTPError(409, "workspace still has experiments")
    ctx.m.db.exec("DELETE FROM projects WHERE workspace_id=?", (int(wid),))
    ctx.m.db.exec("DELETE FROM workspaces WHERE id=?", (int(wid),))
    return {"completed": True}


@route("GET", "/api/v1/workspaces/{wid}/projects")
def list_projects(ctx: Ctx, wid: str) -> Any:
    rows = [{"id": r["id"], "name": r["name"], "workspaceId": r["workspace_id"], "description": r["description"], "archived": bool(r["archived"]),
             "immutable": bool(r["immutable"]), "userId": r["user_id"], "notes": dbmod.U(r["notes"], []),
             "numExperiments": ctx.m.db.one("SELECT COUNT(*) c FROM experiments WHERE project_id=?", (r["id"],))["c"]}
            for r in ctx.m.db.all("SELECT * FROM projects WHERE workspace_id=? ORDER BY id", (int(wid),))]
    items, pg = page(rows, ctx)
    return {"projects": items, "pagination": pg}


@route("POST", "/api/v1/workspaces/{wid}/projects")
def create_project(ctx: Ctx, wid: str) -> Any:
    b = ctx.body or {}
    try:
        pid = ctx.m.db.exec("INSERT INTO projects(name,workspace_id,description,user_id) VALUES(?,?,?,?)", (b.get("name"), int(wid), b.get("description", ""), ctx.user["id"]))
    except Exception as e:
        raise HTTPError(409, f"cannot create project: {e}")
    return {"project": {"id": pid, "name": b.get("name"), "workspaceId": int(wid), "description": b.get("description", ""), "archived": False, "immutable": False,
                        "userId": ctx.user["id"], "notes": [], "numExperiments": 0}}


@route("GET", "/api/v1/projects/{pid}")
def get_project(ctx: Ctx, pid: str) -> Any:
    r = ctx.m.db.one("SELECT * FROM projects WHERE id=?", (int(pid),))
    if r is None:
        raise HTTPError(404, "project not found")
    return {"project": {"id": r["id"], "name": r["name"], "workspaceId": r["workspace_id"], "description": r["description"], "archived": bool(r["archived"]),
                        "immutable": bool(r["immutable"]), "userId": r["user_id"], "notes": dbmod.U(r["notes"], [])}}


@route("DELETE", "/api/v1/projects/{pid}")
def delete_project(ctx: Ctx, pid: str) -> Any:
    if int(pid) == 1:
        raise HTTPError(400, "cannot delete the default project")
    if ctx.m.db.one("SELECT id FROM experiments WHERE project_id=?", (int(pid),)) is not None:
        raise HTTPError(409, "project still has experiments")
    ctx.m.db.exec("DELETE FROM projects WHERE id=?", (int(pid),))
    return {"completed": True}


@route("POST", "/api/v1/experiments/{eid}/move")
def move_experiment(ctx: Ctx, eid: str) -> Any:
    _exp_row(ctx, eid)
    ctx.m.db.exec("UPDATE experiments SET project_id=? WHERE id=?", (int((ctx.body or {}).get("destinationProjectId", 1)), int(eid)))
    return {}


@route("GET", "/api/v1/webhooks")
def list_webhooks(ctx: Ctx) -> Any:
    return {"webhooks": [{"id": r["id"], "url": r["url"], "webhookType": r["webhook_type"], "triggers": dbmod.U(r["triggers"], [])} for r in ctx.m.db.all("SELECT * FROM webhooks")]}


@route("POST", "/api/v1/webhooks")
def create_webhook(ctx: Ctx) -> Any:
    b = ctx.body or {}
    if not str(b.get("url") or "").startswith(("http://", "https://")):
        raise HTTPError(400, "webhook url must be http(s)")
    wid = ctx.m.db.exec("INSERT INTO webhooks(url,webhook_type,triggers) VALUES(?,?,?)", (b.get("url"), b.get("webhookType", "WEBHOOK_TYPE_DEFAULT"), dbmod.J(b.get("triggers") or [])))
    ctx.m.webhooks_changed()
    return {"webhook": {"id": wid, "url": b.get("url"), "webhookType": b.get("webhookType", "WEBHOOK_TYPE_DEFAULT"), "triggers": b.get("triggers") or []}}


@route("DELETE", "/api/v1/webhooks/{wid}")
def delete_webhook(ctx: Ctx, wid: str) -> Any:
    ctx.m.db.exec("DELETE FROM webhooks WHERE id=?", (int(wid),))
    ctx.m.webhooks_changed()
    return {}


@route("POST", "/api/v1/webhooks/{wid}/test")
def test_webhook(ctx: Ctx, wid: str) -> Any:
    """Deliver a sample event to one webhook (reference ``TestWebhook``)."""
    r = ctx.m.db.one("SELECT * FROM webhooks WHERE id=?", (int(wid),))
    if r is None:
        raise HTTPError(404, "webhook not found")
    ctx.m._post_webhook(dict(r), "EXPERIMENT_STATE_CHANGE", {"experiment_id": 0, "state": "COMPLETED", "test": True})
    return {"completed": True}


@route("GET", "/api/v1/master/webhook_signing_key")
def webhook_signing_key(ctx: Ctx) -> Any:
    """The key receivers use to verify ``X-Determined-AuthSignature`` (admins only)."""
    if not ctx.user["admin"]:
        raise HTTPError(403, "admin only")
    key = ctx.m.db.get_meta("webhook_signing_key")
    if not key:
        key = secrets.token_hex(16)
        ctx.m.db.set_meta("webhook_signing_key", key)
    return {"key": key}


@route("GET", "/api/v1/tasks")
def list_tasks(ctx: Ctx) -> Any:
    """Every running allocation: trials and command / shell / notebook / tensorboard tasks."""
    m = ctx.m
    out = {a.id: {"taskId": a.trial.task_id, "allocationId": a.id, "name": f"trial {a.trial.id}", "slotsNeeded": a.slots, "state": a.state,
                  "resources": [{"agentDevices": {getattr(a, "agent_id", m.agent_id): {"devices": a.slot_ids}}}]} for a in m.allocations.values()}
    for t in m.commands.tasks.values():
        if t.state in ("QUEUED", "RUNNING"):
            out[t.id + ".1"] = {"taskId": t.id, "allocationId": t.id + ".1", "name": t.description, "slotsNeeded": t.slots, "state": t.state,
                                "resources": [{"agentDevices": {getattr(t, "agent_id", m.agent_id): {"devices": t.slot_ids}}}]}
    return {"allocationIdToSummary": out}


# ---------------------------------------------------------------------------------------------
# commands / shells / notebooks / tensorboards (reference api_command.go, api_shell.go, api_notebook.go, api_tensorboard.go)
# ---------------------------------------------------------------------------------------------
def _ntsc_routes(kind: str, plural: str) -> None:
    key = kind

    @route("POST", f"/api/v1/{plural}")
    def launch(ctx: Ctx) -> Any:
        b = ctx.body or {}
        cfg = b.get("config") or {}
        if isinstance(cfg, str):
            import yaml as _yaml
            cfg = _yaml.safe_load(cfg) or {}
        if b.get("templateName"):
            row = ctx.m.db.one("SELECT config FROM templates WHERE name=?", (b["templateName"],))
            if row is None:
                raise HTTPError(404, f"template {b['templateName']} not found")
            from determined_b200.master import expconf as _expconf
            cfg = _expconf.merge(cfg, dbmod.U(row["config"], {}))
        try:
            t = ctx.m.commands.launch(kind, cfg, ctx.user["id"], ctx.user["username"], b)
        except ValueError as e:
            raise HTTPError(400, str(e))
        return {key: t.to_json(), "config": cfg}

    @route("GET", f"/api/v1/{plural}")
    def list_(ctx: Ctx) -> Any:
        return {plural: [t.to_json() for t in ctx.m.commands.list(kind)]}

    @route("GET", f"/api/v1/{plural}/{{task_id}}")
    def get(ctx: Ctx, task_id: str) -> Any:
        try:
            return {key: ctx.m.commands.get(task_id, kind).to_json()}
        except KeyError:
            raise HTTPError(404, f"{kind} {task_id} not found")

    @route("POST", f"/api/v1/{plural}/{{task_id}}/kill")
    def kill(ctx: Ctx, task_id: str) -> Any:
        try:
            t = ctx.m.commands.get(task_id, kind)
        except KeyError:
            raise HTTPError(404, f"{kind} {task_id} not found")
        t.kill()
        return {key: t.to_json()}

    @route("POST", f"/api/v1/{plural}/{{task_id}}/set_priority")
    def prio(ctx: Ctx, task_id: str) -> Any:
        try:
            t = ctx.m.commands.get(task_id, kind)
        except KeyError:
            raise HTTPError(404, f"{kind} {task_id} not found")
        t.config.setdefault("resources", {})["priority"] = int((ctx.body or {}).get("priority", 42))
        return {key: t.to_json()}


for _k, _p in (("command", "commands"), ("shell", "shells"), ("notebook", "notebooks"), ("tensorboard", "tensorboards")):
    _ntsc_routes(_k, _p)


# ---------------------------------------------------------------------------------------------
# streaming updates (reference master/internal/stream/*: websocket publisher for projects / models / model versions;
# here the same startup / sync / upsert / delete messages over a long-poll POST)
# ---------------------------------------------------------------------------------------------
class _StreamState:
    """seq numbers per (kind, id), bumped when the row's content changes (detected lazily at poll time)."""
    lock = threading.Lock()
    seq = 0
    rows: Dict[Tuple[str, int], Tuple[int, str]] = {}


def _stream_scan(ctx: Ctx) -> Dict[str, Dict[int, Dict[str, Any]]]:
    db = ctx.m.db
    cur: Dict[str, Dict[int, Dict[str, Any]]] = {"projects": {}, "models": {}, "model_versions": {}}
    for r in db.all("SELECT * FROM projects", ()):
        cur["projects"][r["id"]] = {"id": r["id"], "name": r["name"], "description": r["description"] or "", "archived": bool(r["archived"]) if "archived" in r.keys() else False,
                                    "workspace_id": r["workspace_id"], "user_id": r["user_id"] if "user_id" in r.keys() else None, "state": "UNSPECIFIED"}
    for r in db.all("SELECT * FROM models", ()):
        cur["models"][r["id"]] = {"id": r["id"], "name": r["name"], "description": r["description"] or "", "archived": bool(r["archived"]) if "archived" in r.keys() else False,
                                  "labels": dbmod.U(r["labels"], []) if "labels" in r.keys() else [], "workspace_id": r["workspace_id"] if "workspace_id" in r.keys() else 1,
                                  "user_id": r["user_id"] if "user_id" in r.keys() else None}
    for r in db.all("SELECT * FROM model_versions", ()):
        cur["model_versions"][r["id"]] = {"id": r["id"], "name": r["name"] if "name" in r.keys() else "", "version": r["version"], "checkpoint_uuid": r["checkpoint_uuid"],
                                          "model_id": r["model_id"], "comment": r["comment"] if "comment" in r.keys() else "", "labels": [],
                                          "user_id": r["user_id"] if "user_id" in r.keys() else None}
    with _StreamState.lock:
        for kind, rows in cur.items():
            for i, row in rows.items():
                h = json.dumps(row, sort_keys=True, default=str)
                old = _StreamState.rows.get((kind, i))
                if old is None or old[1] != h:
                    _StreamState.seq += 1
                    _StreamState.rows[(kind, i)] = (_StreamState.seq, h)
                row["seq"] = _StreamState.rows[(kind, i)][0]
        for key in [k for k in _StreamState.rows if k[1] not in cur[k[0]]]:
            del _StreamState.rows[key]
    return cur


def _stream_match(kind: str, row: Dict[str, Any], spec: Dict[str, Any]) -> bool:
    def hit(field: str, key: str) -> bool:
        ids = spec.get(key)
        return bool(ids) and row.get(field) in ids
    if kind == "projects":
        return hit("workspace_id", "workspace_ids") or hit("id", "project_ids")
    if kind == "models":
        return hit("workspace_id", "workspace_ids") or hit("id", "model_ids") or hit("user_id", "user_ids")
    return hit("model_id", "model_ids") or hit("id", "model_version_ids") or hit("user_id", "user_ids")


_WIRE_NAME = {"projects": "project", "models": "model", "model_versions": "modelversion"}


class _StreamSession:
    """One subscriber of the streaming-updates channel: what it subscribed to and which rows it has been told about.

    ``startup(msg)`` answers a startup message: ``sync_msg(complete=False)``, deletions of rows the client knows
    (``known``: range-encoded ids per kind) that it may no longer see, every visible row that is new to it or changed
    after ``since``, ``sync_msg(complete=True)``.  ``tick()`` is the online phase: rows that changed since the last
    message, and deletions of rows that disappeared or fell out of the subscription."""

    def __init__(self, ctx: Ctx) -> None:
        self.ctx = ctx
        self.sub: Dict[str, Dict[str, Any]] = {}
        self.seq = 0
        self.sent: Dict[str, set] = {}

    def _visible(self, cur: Dict[str, Dict[int, Dict[str, Any]]]) -> Dict[str, Dict[int, Dict[str, Any]]]:
        return {kind: {i: r for i, r in cur[kind].items() if _stream_match(kind, r, spec)} for kind, spec in self.sub.items() if kind in cur}

    def startup(self, b: Dict[str, Any]) -> List[Dict[str, Any]]:
        from determined_b200.common.streams import _util as su

        self.sub = {k: v for k, v in (b.get("subscribe") or {}).items() if isinstance(v, dict)}
        cur = _stream_scan(self.ctx)
        msgs: List[Dict[str, Any]] = [{"sync_msg": {"sync_id": b.get("sync_id"), "complete": False}}]
        self.sent = {}
        for kind, rows in self._visible(cur).items():
            since = int(self.sub[kind].get("since") or 0)
            known = set(su.range_encoded_keys((b.get("known") or {}).get(kind, "")))
            gone = known - set(rows)
            if gone:
                msgs.append({kind + "_deleted": su.encode_keys(gone)})
            for i, r in sorted(rows.items()):
                if r["seq"] > since or i not in known:
                    msgs.append({_WIRE_NAME[kind]: r})
            self.sent[kind] = set(rows)
        msgs.append({"sync_msg": {"sync_id": b.get("sync_id"), "complete": True}})
        self.seq = _StreamState.seq
        return msgs

    def tick(self) -> List[Dict[str, Any]]:
        from determined_b200.common.streams import _util as su

        if not self.sub:
            return []
        cur = _stream_scan(self.ctx)
        msgs: List[Dict[str, Any]] = []
        for kind, rows in self._visible(cur).items():
            gone = self.sent.get(kind, set()) - set(rows)
            if gone:
                msgs.append({kind + "_deleted": su.encode_keys(gone)})
            for i, r in sorted(rows.items()):
                if r["seq"] > self.seq:
                    msgs.append({_WIRE_NAME[kind]: r})
            self.sent[kind] = set(rows)
        self.seq = _StreamState.seq
        return msgs


@route("POST", "/stream")
def stream(ctx: Ctx) -> Any:
    """The streaming-updates protocol for clients that cannot hold a websocket: a startup message gets its whole offline
    answer at once; ``poll_since`` asks for rows that changed after a sequence number (no online deletions in this form:
    the websocket at ``GET /stream`` has them)."""
    b = ctx.body or {}
    sess = _StreamSession(ctx)
    if "poll_since" not in b:
        return {"messages": sess.startup(b), "seq": _StreamState.seq}
    sess.sub = {k: v for k, v in (b.get("subscribe") or {}).items() if isinstance(v, dict)}
    sess.seq = int(b["poll_since"])
    sess.sent = {}
    return {"messages": [m for m in sess.tick() if not any(k.endswith("_deleted") for k in m)], "seq": _StreamState.seq}


def _ws_stream(handler: Any, user: Dict[str, Any]) -> None:
    """``GET /stream`` with ``Upgrade: websocket`` (reference ``master/internal/stream/publisher_set.go``): JSON text
    frames both ways; the client may send a new startup message at any time to change its subscription."""
    import queue

    from determined_b200.common import websocket

    ws = websocket.server_handshake(handler)
    sess = _StreamSession(Ctx(handler.master, user, {}, None, b""))
    inbox: "queue.Queue[Any]" = queue.Queue()

    def reader() -> None:
        try:
            while True:
                inbox.put(ws.recv_json())
        except Exception as e:  # noqa: BLE001  (closed, lost, malformed: all end the session)
            inbox.put(e)

    threading.Thread(target=reader, daemon=True, name="ws-stream-reader").start()
    try:
        while not handler.master._stop.is_set():
            try:
                msg = inbox.get(timeout=0.25)
            except queue.Empty:
                msg = None
            if isinstance(msg, Exception):
                break
            out = sess.startup(msg) if isinstance(msg, dict) else sess.tick()
            for m in out:
                ws.send_json(m)
    except (OSError, websocket.WebSocketError):
        pass
    finally:
        ws.close()


@route("GET", "/metrics", auth_required=False)
@route("GET", "/prom/det-state-metrics", auth_required=False)
def prometheus(ctx: Ctx) -> Any:
    """Prometheus text exposition (reference ``master/internal/prom``)."""
    m = ctx.m
    lines = ["# TYPE det_slots_total gauge", "# TYPE det_slots_used gauge", "# TYPE det_agent_connected gauge"]
    for aid, ag in sorted((a, g) for pl in m.pools.values() for a, g in pl.agents.items()):
        if aid == m.agent_id and not ag.num_slots:
            continue
        ra = m.agents.get(aid)
        lines += [f'det_slots_total{{agent="{aid}"}} {ag.num_slots}', f'det_slots_used{{agent="{aid}"}} {ag.num_used()}',
                  f'det_agent_connected{{agent="{aid}"}} {int(ra.connected) if ra is not None else 1}']
    lines += ["# TYPE det_allocations_running gauge", f"det_allocations_running {len(m.allocations)}",
              "# TYPE det_pending_requests gauge", f"det_pending_requests {sum(len(pl.pending) for pl in m.pools.values())}", "# TYPE det_experiments gauge"]
    by_state: Dict[str, int] = {}
    for e in m.experiments.values():
        by_state[e.state] = by_state.get(e.state, 0) + 1
    by_state.setdefault(core.ACTIVE, 0)
    lines += [f'det_experiments{{state="{k}"}} {v}' for k, v in sorted(by_state.items())]
    lines += ["# TYPE det_experiments_active gauge", f"det_experiments_active {by_state[core.ACTIVE]}",
              "# TYPE det_tasks_running gauge", f"det_tasks_running {sum(1 for t in m.commands.tasks.values() if t.state == 'RUNNING')}",
              "# TYPE det_master_uptime_seconds counter", f"det_master_uptime_seconds {time.time() - m.start_time:.0f}"]
    return ("text/plain; version=0.0.4", "\n".join(lines) + "\n")


# ---------------------------------------------------------------------------------------------
# server
# ---------------------------------------------------------------------------------------------
@route("GET", "/", auth_required=False)
@route("GET", "/det", auth_required=False)
@route("GET", "/index.html", auth_required=False)
def webui(ctx: Ctx) -> Any:
    """The dashboard: one static page that talks to this API (reference ``webui/react``)."""
    import pathlib

    return "text/html; charset=utf-8", (pathlib.Path(__file__).parent / "static" / "index.html").read_bytes()


@route("GET", "/api/v1/openapi.json", auth_required=False)
@route("GET", "/api/v1/swagger.json", auth_required=False)
def openapi_doc(ctx: Ctx) -> Any:
    """Machine-readable description of every route (feeds the binding generator)."""
    from determined_b200.master import openapi

    return openapi.build(ROUTES)


class Handler(BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"
    # headers and body leave in separate writes: with Nagle's algorithm on, the second waits for the client's delayed ACK of the
    # first (~40 ms per request on a kept-alive connection)
    disable_nagle_algorithm = True
    master: core.Master = None  # type: ignore

    def log_message(self, fmt: str, *args: Any) -> None:
        logger.debug("%s %s", self.address_string(), fmt % args)

    def _user(self) -> Optional[Dict[str, Any]]:
        h = self.headers.get("Authorization", "")
        tok = h[7:] if h.startswith("Bearer ") else None
        if not tok:
            return None
        seen = getattr(self, "_user_seen", None)
        if seen is not None and seen[0] == tok:          # asked more than once per request (dispatch, rbac, audit)
            return seen[1]
        out = self._lookup_user(tok)
        self._user_seen = (tok, out)
        return out

    def _lookup_user(self, tok: str) -> Optional[Dict[str, Any]]:
        r = self.master.db.one("SELECT s.token, s.expiry, s.task_id session_task, u.* FROM sessions s JOIN users u ON u.id=s.user_id WHERE s.token=?", (tok,))
        if r is None or (r["expiry"] or 0) < time.time() or not r["active"]:
            return None
        return {"id": r["id"], "username": r["username"], "admin": bool(r["admin"]), "token": tok, "task_id": r["session_task"]}

    def _send(self, status: int, payload: Any, ctype: str = "application/json") -> None:
        if isinstance(payload, (dict, list)) or payload is None:
            from determined_b200 import util

            data = util.json_encode(payload if payload is not None else {}).encode()
        elif isinstance(payload, str):
            data = payload.encode()
        else:
            data = payload
        self.send_response(status)
        self.send_header("Content-Type", ctype)
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        try:
            self.wfile.write(data)
        except (BrokenPipeError, ConnectionResetError):
            pass

    def _dispatch(self, method: str) -> None:
        u = urllib.parse.urlparse(self.path)
        path = u.path
        n = int(self.headers.get("Content-Length") or 0)
        raw = self.rfile.read(n) if n else b""
        body: Any = None
        if raw:
            try:
                body = json.loads(raw)
            except ValueError:
                body = raw.decode(errors="replace")
        if path.startswith("/proxy/"):
            self._proxy(method, path, u.query, raw)
            return
        if method == "GET" and path.startswith("/tunnel/"):
            self._tunnel(path)
            return
        if method == "GET" and path.rstrip("/") == "/stream":
            from determined_b200.common import websocket

            if websocket.is_upgrade(self.headers):
                user = self._user()
                if user is None:
                    self._send(401, {"error": {"code": 16, "message": "unauthenticated"}, "message": "unauthenticated"})
                    return
                _ws_stream(self, user)
                return
        # the most literal pattern that matches wins (/users/setting before /users/{uid}), whatever the registration order
        hits = [(rx.groups, i, rx, fn, need_auth, mt) for i, (meth, rx, fn, need_auth) in enumerate(ROUTES)
                if meth == method for mt in (rx.match(path),) if mt is not None]
        for _, _, rx, fn, need_auth, mt in sorted(hits, key=lambda h: h[:2])[:1]:
            user = self._user()
            if need_auth and user is None:
                self._send(401, {"error": {"code": 16, "message": "unauthenticated"}, "message": "unauthenticated"})
                return
            ctx = Ctx(self.master, user, urllib.parse.parse_qs(u.query), body, raw)
            ctx.headers = self.headers
            if need_auth and user is not None and authz.enabled(self.master):
                denied = authz.check(self.master, user, method, path, body)
                if denied:
                    self._send(403, {"error": {"code": 7, "message": denied}, "message": denied})
                    return
            try:
                with tracing.span(f"{method} {rx.pattern[1:-3]}", traceparent=self.headers.get("traceparent"), **{"http.method": method, "http.target": path}):
                    out = fn(ctx, **{k: urllib.parse.unquote(v) for k, v in mt.groupdict().items()})
                if RECORDER is not None:
                    RECORDER(method, rx.pattern[1:-3], body, out)
                if need_auth and user is not None:
                    out = authz.filter_listing(self.master, user, method, path, out)
                if isinstance(out, tuple) and isinstance(out[0], int):          # (status, JSON payload or None)
                    if out[1] is None:
                        self.send_response(out[0])
                        self.send_header("Content-Length", "0")
                        self.end_headers()
                    else:
                        self._send(out[0], out[1])
                elif isinstance(out, tuple):                                     # (content type, payload)
                    self._send(200, out[1], out[0])
                else:
                    self._send(200, out)
            except Redirect as r:
                self.send_response(302)
                self.send_header("Location", r.to)
                self.send_header("Content-Length", "0")
                self.end_headers()
            except HTTPError as e:
                self._send(e.status, getattr(e, "body", None) or {"error": {"code": e.status, "message": e.message}, "message": e.message})
            except core.NotFound as e:
                self._send(404, {"error": {"code": 5, "message": str(e)}, "message": str(e)})
            except expconf.ConfigError as e:
                self._send(400, {"error": {"code": 3, "message": str(e)}, "message": str(e)})
            except Exception as e:  # noqa
                logger.exception("handler for %s %s failed", method, path)
                self._send(500, {"error": {"code": 13, "message": f"{type(e).__name__}: {e}"}, "message": f"{type(e).__name__}: {e}"})
            return
        self._send(404, {"error": {"code": 5, "message": f"no route for {method} {path}"}, "message": f"no route for {method} {path}"})

    def _proxy(self, method: str, path: str, query: str, raw: bytes) -> None:
        """``/proxy/<task id>/<rest>``: reverse proxy to the HTTP service of a running task (tensorboard, notebook, a command
        with ``proxy_ports``) at the port the master's registry gave it.  Reference ``master/internal/proxy``: there the
        task registers its address; here the master chose host and port itself, so there is nothing to trust.  Plain HTTP
        request/response only (no websocket upgrade)."""
        import http.client

        parts = path.split("/", 3)
        task_id = parts[2] if len(parts) > 2 else ""
        rest = "/" + (parts[3] if len(parts) > 3 else "")
        qs = urllib.parse.parse_qs(query)
        user = self._user()
        if user is None and qs.get("token"):
            r = self.master.db.one("SELECT user_id FROM sessions WHERE token=? AND expiry>?", (qs["token"][0], time.time()))
            user = {"id": r["user_id"]} if r is not None else None
        if user is None:
            self._send(401, {"message": "unauthenticated"})
            return
        t = self.master.commands.tasks.get(task_id)
        if t is None or t.state != "RUNNING" or not t.service_port:
            self._send(404, {"message": f"task {task_id} has no running service"})
            return
        t.last_activity = time.time()
        fwd_q = urllib.parse.urlencode([(k, v) for k, vs in qs.items() if k != "token" for v in vs])
        from determined_b200.common import websocket as _ws

        if _ws.is_upgrade(self.headers):
            self._proxy_tunnel(t, method, rest + ("?" + fwd_q if fwd_q else ""), task_id)
            return
        hdrs = {k: v for k, v in self.headers.items() if k.lower() not in ("host", "authorization", "connection", "content-length")}
        hdrs["X-Forwarded-Prefix"] = f"/proxy/{task_id}"
        try:
            conn = http.client.HTTPConnection(t.service_host, t.service_port, timeout=30)
            conn.request(method, rest + ("?" + fwd_q if fwd_q else ""), body=raw or None, headers=hdrs)
            resp = conn.getresponse()
            data = resp.read()
        except OSError as e:
            self._send(502, {"message": f"service of task {task_id} is not answering: {e}"})
            return
        self.send_response(resp.status)
        for k, v in resp.getheaders():
            if k.lower() not in ("transfer-encoding", "connection", "content-length"):
                self.send_header(k, v)
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        try:
            self.wfile.write(data)
        except (BrokenPipeError, ConnectionResetError):
            pass
        conn.close()

    def _tunnel(self, path: str) -> None:
        """``GET /tunnel/<task id>[/<port>]`` with ``Upgrade: websocket``: the binary frames of the websocket are the bytes of one
        TCP connection to the task's service port, or to a port it declared in ``proxy_ports`` (reference ``master/internal/proxy``
        TCP proxying behind ``python -m determined.cli.tunnel``)."""
        import socket as _socket
        import threading as _threading

        from determined_b200.common import websocket

        parts = path.strip("/").split("/")
        task_id = parts[1] if len(parts) > 1 else ""
        if self._user() is None:
            self._send(401, {"message": "unauthenticated"})
            return
        if not websocket.is_upgrade(self.headers):
            self._send(400, {"message": "the tunnel is a websocket: send an Upgrade request"})
            return
        t = self.master.commands.tasks.get(task_id)
        if t is None or t.state != "RUNNING":
            self._send(404, {"message": f"task {task_id} is not running"})
            return
        allowed = {int(t.service_port)} if t.service_port else set()
        for pp in (t.config.get("proxy_ports") or []) if isinstance(t.config.get("proxy_ports"), list) else []:
            if isinstance(pp, dict) and pp.get("proxy_port"):
                allowed.add(int(pp["proxy_port"]))
        port = int(parts[2]) if len(parts) > 2 and parts[2].isdigit() else (int(t.service_port) if t.service_port else 0)
        if not port or port not in allowed:
            self._send(403, {"message": f"task {task_id} exposes no port {port or ''}: declare it in proxy_ports".replace("  ", " ")})
            return
        try:
            back = _socket.create_connection((t.service_host, port), timeout=15)
        except OSError as e:
            self._send(502, {"message": f"port {port} of task {task_id} is not answering: {e}"})
            return
        back.setsockopt(_socket.IPPROTO_TCP, _socket.TCP_NODELAY, 1)
        ws = websocket.server_handshake(self)
        wlock = _threading.Lock()
        done = _threading.Event()

        def down() -> None:                      # task -> client
            try:
                while not done.is_set():
                    data = back.recv(65536)
                    if not data:
                        break
                    with wlock:
                        ws.send_binary(data)
                    t.last_activity = time.time()
            except (OSError, websocket.WebSocketError):
                pass
            finally:
                done.set()

        th = _threading.Thread(target=down, daemon=True, name=f"tunnel-{task_id[:8]}")
        th.start()
        try:
            while not done.is_set() and not self.master._stop.is_set():
                try:
                    _, data = ws.recv(timeout=1.0)
                except (_socket.timeout, TimeoutError):
                    continue
                if data:
                    back.sendall(data)
                    t.last_activity = time.time()
        except (websocket.WebSocketError, OSError):
            pass
        finally:
            done.set()
            for c in (back,):
                try:
                    c.shutdown(_socket.SHUT_RDWR)
                except OSError:
                    pass
                c.close()
            try:
                with wlock:
                    ws.close()
            except Exception:  # noqa: BLE001
                pass

    def _proxy_tunnel(self, t: Any, method: str, target: str, task_id: str) -> None:
        """An upgraded connection (a websocket of a notebook kernel, of tensorboard's live reload): replay the request to
        the task's service and then copy bytes both ways until either side closes.  The master does not look inside."""
        import select
        import socket as _socket

        try:
            back = _socket.create_connection((t.service_host, t.service_port), timeout=30)
        except OSError as e:
            self._send(502, {"message": f"service of task {task_id} is not answering: {e}"})
            return
        lines = [f"{method} {target} HTTP/1.1"]
        for k, v in self.headers.items():
            if k.lower() in ("authorization",):
                continue
            lines.append(f"{k}: {f'{t.service_host}:{t.service_port}' if k.lower() == 'host' else v}")
        lines.append(f"X-Forwarded-Prefix: /proxy/{task_id}")
        back.sendall(("\r\n".join(lines) + "\r\n\r\n").encode())
        self.close_connection = True
        front = self.connection
        # bytes the client sent right behind its request sit in rfile's buffer, where select() cannot see them
        front.settimeout(0)
        try:
            pending = self.rfile.peek(1) if hasattr(self.rfile, "peek") else b""
        except (BlockingIOError, OSError):
            pending = b""
        if pending:
            back.sendall(self.rfile.read(len(pending)))
        front.settimeout(None)
        back.settimeout(None)
        try:
            while not self.master._stop.is_set():
                ready, _, _ = select.select([front, back], [], [], 1.0)
                for src in ready:
                    data = src.recv(65536)
                    if not data:
                        return
                    (back if src is front else front).sendall(data)
                    t.last_activity = time.time()
        except OSError:
            pass
        finally:
            back.close()

    def send_response(self, code: int, message: Optional[str] = None) -> None:
        self._status = code
        super().send_response(code, message)

    def _audited(self, method: str) -> None:
        """Every API request leaves an audit line (reference ``master/internal/audit.go``): who, from where, what, with which
        outcome — at INFO for writes and failures, at DEBUG for reads; static pages and proxied task traffic are left out."""
        self._status = 0
        self._user_seen = None
        try:
            self._dispatch(method)
        finally:
            path = self.path.split("?", 1)[0]
            if path.startswith(("/api/", "/oauth2/", "/scim/", "/oidc/", "/stream", "/tunnel/")) and not path.startswith("/api/v1/agents/"):
                try:
                    u = self._user()
                except Exception:  # noqa: BLE001
                    u = None
                status = getattr(self, "_status", 0)
                audit_logger.log(logging.INFO if method != "GET" or status >= 400 else logging.DEBUG,
                                 "audit: user=%s remote_ip=%s method=%s path=%s status=%s unauthorized=%s", (u or {}).get("username"),
                                 self.client_address[0], method, path, status, status in (401, 403))

    def do_GET(self) -> None:
        self._audited("GET")

    def do_POST(self) -> None:
        self._audited("POST")

    def do_PATCH(self) -> None:
        self._audited("PATCH")

    def do_PUT(self) -> None:
        self._audited("PUT")

    def do_DELETE(self) -> None:
        self._audited("DELETE")


class _QuietServer(ThreadingHTTPServer):
    def handle_error(self, request: Any, client_address: Any) -> None:
        """A client that goes away, or one that rejects our certificate during the handshake, is its own business: one
        debug line instead of a traceback on stderr."""
        import ssl

        e = sys.exc_info()[1]
        if isinstance(e, (ssl.SSLError, ConnectionError, TimeoutError)):
            logger.debug("connection from %s dropped: %s", client_address, e)
            return
        super().handle_error(request, client_address)


class APIServer:
    def __init__(self, master: core.Master, host: str = "127.0.0.1", port: int = 8080,
                 tls_cert: Optional[str] = None, tls_key: Optional[str] = None) -> None:
        """``tls_cert`` / ``tls_key`` (PEM files; master config ``security.tls.cert`` / ``.key``) make the server speak
        HTTPS; tasks then get ``DET_MASTER_CERT_FILE`` so that the harness trusts exactly this certificate."""
        handler = type("BoundHandler", (Handler,), {"master": master})
        self.httpd = _QuietServer((host, port), handler)
        self.httpd.daemon_threads = True
        if tls_cert:
            import ssl

            ctx = ssl.SSLContext(ssl.PROTOCOL_TLS_SERVER)
            ctx.minimum_version = ssl.TLSVersion.TLSv1_2
            ctx.load_cert_chain(tls_cert, tls_key or None)
            # handshake in the worker thread, not in accept(): a client that connects and stays silent must not block others
            self.httpd.socket = ctx.wrap_socket(self.httpd.socket, server_side=True, do_handshake_on_connect=False)
            master.tls_cert_file = os.path.abspath(tls_cert)
        master.host, master.port = host, self.httpd.server_address[1]
        self.thread: Optional[threading.Thread] = None

    @property
    def port(self) -> int:
        return self.httpd.server_address[1]

    def start(self) -> "APIServer":
        self.thread = threading.Thread(target=self.httpd.serve_forever, daemon=True, name="api-server")
        self.thread.start()
        return self

    def stop(self) -> None:
        self.httpd.shutdown()
        self.httpd.server_close()


from determined_b200.master import authz  # noqa: E402
from determined_b200.master import api_more  # noqa: E402,F401  (registers its routes)
from determined_b200.master import api_rest  # noqa: E402,F401  (the rest of the REST surface)
from determined_b200.master import sso  # noqa: E402,F401  (/oidc/sso, /oidc/callback)
from determined_b200.master import scim  # noqa: E402,F401  (/scim/v2/*, /oauth2/token)
