"""Role-based access control (``security.authz.type: rbac``).

Parity: the reference's enterprise authorisation (``master/internal/rbac``, ``*_authz_rbac.go``): a user may do something in a
workspace when a role assigned to them — directly or through a group, for that workspace or cluster-wide — carries the
permission.  With ``type: basic`` (the default) every signed-in user may use every workspace and only ``admin`` is special.

One table (`RULES`) maps request shapes to ``(permission, where)``; the dispatcher asks :func:`check` before it calls a
handler and passes listings through :func:`filter_listing` afterwards.  Administrators, holders of ``ClusterAdmin`` and task
sessions (the harness inside an allocation, acting for the owner who was allowed to start it) are not restricted.
"""
import re
from typing import Any, Callable, Dict, List, Optional, Set, Tuple

GLOBAL = None       # the cluster-wide scope


def enabled(m: Any) -> bool:
    return getattr(m, "authz_type", "basic") == "rbac"


def permissions_of(m: Any, uid: int) -> Dict[Optional[int], Set[str]]:
    """workspace id (``None``: everywhere) -> permissions the user holds there."""
    from determined_b200.master.api_more import ROLE_PERMISSIONS

    rows = list(m.db.all("SELECT role, workspace_id FROM role_assignments WHERE user_id=?", (uid,)))
    rows += list(m.db.all("SELECT r.role, r.workspace_id FROM role_assignments r JOIN group_members g ON g.group_id=r.group_id WHERE g.user_id=?", (uid,)))
    out: Dict[Optional[int], Set[str]] = {}
    for r in rows:
        out.setdefault(r["workspace_id"], set()).update(ROLE_PERMISSIONS.get(r["role"], []))
    return out


def unrestricted(m: Any, user: Dict[str, Any]) -> bool:
    if user.get("admin") or user.get("task_id"):
        return True
    return "PERMISSION_TYPE_ADMINISTRATE_USER" in permissions_of(m, user["id"]).get(GLOBAL, set())


def allowed(m: Any, user: Dict[str, Any], perm: str, workspace_id: Optional[int]) -> bool:
    if unrestricted(m, user):
        return True
    p = permissions_of(m, user["id"])
    return perm in p.get(GLOBAL, set()) or (workspace_id is not None and perm in p.get(int(workspace_id), set()))


def viewable_workspaces(m: Any, user: Dict[str, Any], perm: str = "PERMISSION_TYPE_VIEW_WORKSPACE") -> Optional[Set[int]]:
    """Workspace ids the user may see, or ``None`` for all of them."""
    if unrestricted(m, user):
        return None
    p = permissions_of(m, user["id"])
    if perm in p.get(GLOBAL, set()):
        return None
    return {int(w) for w, perms in p.items() if w is not None and perm in perms}


# ------------------------------------------------------------------------------------------------------------------ scopes
def _ws_of_project(m: Any, pid: Any) -> Optional[int]:
    r = m.db.one("SELECT workspace_id FROM projects WHERE id=?", (int(pid),))
    return int(r["workspace_id"]) if r is not None else None


def _ws_of_exp(m: Any, eid: Any) -> Optional[int]:
    r = m.db.one("SELECT project_id FROM experiments WHERE id=?", (int(eid),))
    return _ws_of_project(m, r["project_id"]) if r is not None else None


def _ws_of_trial(m: Any, tid: Any) -> Optional[int]:
    r = m.db.one("SELECT experiment_id FROM trials WHERE id=?", (int(tid),))
    return _ws_of_exp(m, r["experiment_id"]) if r is not None else None


def _ws_of_ckpt(m: Any, uuid_: str) -> Optional[int]:
    r = m.db.one("SELECT trial_id FROM checkpoints WHERE uuid=?", (uuid_,))
    return _ws_of_trial(m, r["trial_id"]) if r is not None and r["trial_id"] else None


def _ws_of_model(m: Any, name: str) -> Optional[int]:
    import urllib.parse

    r = m.db.one("SELECT workspace_id FROM models WHERE name=?", (urllib.parse.unquote(name),))
    return int(r["workspace_id"] or 1) if r is not None else None


Scope = Callable[[Any, "re.Match[str]", Any], Optional[int]]
_first: Callable[[Callable[[Any, Any], Optional[int]]], Scope] = lambda f: (lambda m, mt, body: f(m, mt.group(1)))      # noqa: E731
_body_project: Scope = lambda m, mt, body: _ws_of_project(m, (body or {}).get("projectId") or 1) if isinstance(body, dict) else 1      # noqa: E731
_body_workspace: Scope = lambda m, mt, body: int((body or {}).get("workspaceId") or 1) if isinstance(body, dict) else 1      # noqa: E731
_global: Scope = lambda m, mt, body: GLOBAL      # noqa: E731
_self: Scope = lambda m, mt, body: int(mt.group(1))      # noqa: E731
WRITE = ("POST", "PUT", "PATCH", "DELETE")

# (methods, path pattern, permission, scope); the first match decides
RULES: List[Tuple[Tuple[str, ...], "re.Pattern[str]", str, Scope]] = [(ms, re.compile(rx), perm, sc) for ms, rx, perm, sc in [
    (("POST",), r"^/api/v1/experiments/?$", "PERMISSION_TYPE_CREATE_EXPERIMENT", _body_project),
    (("GET",), r"^/api/v1/experiments/(\d+)(?:/.*)?$", "PERMISSION_TYPE_VIEW_EXPERIMENT_METADATA", _first(_ws_of_exp)),
    (("DELETE",), r"^/api/v1/experiments/(\d+)/?$", "PERMISSION_TYPE_DELETE_EXPERIMENT", _first(_ws_of_exp)),
    (WRITE, r"^/api/v1/experiments/(\d+)(?:/.*)?$", "PERMISSION_TYPE_UPDATE_EXPERIMENT", _first(_ws_of_exp)),
    (("GET",), r"^/api/v1/trials/(\d+)(?:/.*)?$", "PERMISSION_TYPE_VIEW_EXPERIMENT_METADATA", _first(_ws_of_trial)),
    (WRITE, r"^/api/v1/trials/(\d+)(?:/.*)?$", "PERMISSION_TYPE_UPDATE_EXPERIMENT", _first(_ws_of_trial)),
    (("GET",), r"^/api/v1/checkpoints/([0-9a-fA-F-]{36})(?:/.*)?$", "PERMISSION_TYPE_VIEW_EXPERIMENT_ARTIFACTS", _first(_ws_of_ckpt)),
    (WRITE, r"^/api/v1/checkpoints/([0-9a-fA-F-]{36})(?:/.*)?$", "PERMISSION_TYPE_UPDATE_EXPERIMENT", _first(_ws_of_ckpt)),
    (("GET",), r"^/api/v1/projects/(\d+)(?:/.*)?$", "PERMISSION_TYPE_VIEW_PROJECT", _first(_ws_of_project)),
    (("DELETE",), r"^/api/v1/projects/(\d+)/?$", "PERMISSION_TYPE_DELETE_PROJECT", _first(_ws_of_project)),
    (WRITE, r"^/api/v1/projects/(\d+)(?:/.*)?$", "PERMISSION_TYPE_UPDATE_PROJECT", _first(_ws_of_project)),
    (("POST",), r"^/api/v1/workspaces/(\d+)/projects/?$", "PERMISSION_TYPE_CREATE_PROJECT", _self),
    (("POST",), r"^/api/v1/workspaces/?$", "PERMISSION_TYPE_CREATE_WORKSPACE", _global),
    (("GET",), r"^/api/v1/workspaces/(\d+)(?:/.*)?$", "PERMISSION_TYPE_VIEW_WORKSPACE", _self),
    (("POST",), r"^/api/v1/workspaces/(\d+)/(?:pin|unpin)$", "PERMISSION_TYPE_VIEW_WORKSPACE", _self),
    (("DELETE",), r"^/api/v1/workspaces/(\d+)/?$", "PERMISSION_TYPE_DELETE_WORKSPACE", _self),
    (WRITE, r"^/api/v1/workspaces/(\d+)(?:/.*)?$", "PERMISSION_TYPE_UPDATE_WORKSPACE", _self),
    (("POST",), r"^/api/v1/models/?$", "PERMISSION_TYPE_CREATE_MODEL_REGISTRY", _body_workspace),
    (("GET",), r"^/api/v1/models/([^/]+)(?:/.*)?$", "PERMISSION_TYPE_VIEW_MODEL_REGISTRY", _first(_ws_of_model)),
    (("DELETE",), r"^/api/v1/models/([^/]+)(?:/.*)?$", "PERMISSION_TYPE_DELETE_MODEL_REGISTRY", _first(_ws_of_model)),
    (WRITE, r"^/api/v1/models/([^/]+)(?:/.*)?$", "PERMISSION_TYPE_EDIT_MODEL_REGISTRY", _first(_ws_of_model)),
    (("POST",), r"^/api/v1/(?:commands|shells|notebooks|tensorboards|generic-tasks)/?$", "PERMISSION_TYPE_CREATE_NSC", _body_workspace),
    (WRITE, r"^/api/v1/webhooks(?:/.*)?$", "PERMISSION_TYPE_EDIT_WEBHOOKS", _global),
    (("GET",), r"^/api/v1/master/logs$", "PERMISSION_TYPE_VIEW_MASTER_LOGS", _global),
    (("GET",), r"^/api/v1/resources/allocation/(?:raw|aggregated)$", "PERMISSION_TYPE_VIEW_CLUSTER_USAGE", _global),
]]


def check(m: Any, user: Dict[str, Any], method: str, path: str, body: Any) -> Optional[str]:
    """``None`` when the request may go ahead, else the sentence for the 403."""
    if not enabled(m) or unrestricted(m, user):
        return None
    bulk = re.match(r"^/api/v1/experiments/(activate|pause|cancel|kill|archive|unarchive|move|delete|retain_logs)/?$", path)
    if bulk and isinstance(body, dict):
        perm = "PERMISSION_TYPE_DELETE_EXPERIMENT" if bulk.group(1) == "delete" else "PERMISSION_TYPE_UPDATE_EXPERIMENT"
        for eid in body.get("experimentIds") or []:
            ws = _ws_of_exp(m, eid)
            if ws is not None and not allowed(m, user, perm, ws):
                return f"access denied: {user['username']} lacks {perm.replace('PERMISSION_TYPE_', '').lower()} in workspace {ws} (experiment {eid})"
        if bulk.group(1) == "move" and body.get("destinationProjectId"):
            ws = _ws_of_project(m, body["destinationProjectId"])
            if ws is not None and not allowed(m, user, "PERMISSION_TYPE_CREATE_EXPERIMENT", ws):
                return f"access denied: {user['username']} lacks create_experiment in workspace {ws}"
        if not body.get("experimentIds") and body.get("filters") is not None and viewable_workspaces(m, user, perm) is not None:
            return "access denied: bulk actions by filter need the permission cluster-wide; name the experiments instead"
        return None
    for methods, rx, perm, scope in RULES:
        if method not in methods:
            continue
        mt = rx.match(path)
        if mt is None:
            continue
        try:
            ws = scope(m, mt, body)
        except (TypeError, ValueError):
            return None                          # malformed ids: the handler answers 400 / 404
        if ws is None and scope is not _global:
            return None                          # the object does not exist: the handler answers 404
        if not allowed(m, user, perm, ws):
            where = "cluster-wide" if ws is None else f"in workspace {ws}"
            return f"access denied: {user['username']} lacks {perm.replace('PERMISSION_TYPE_', '').lower()} {where}"
        return None
    return None


def filter_listing(m: Any, user: Dict[str, Any], method: str, path: str, out: Any) -> Any:
    """Listings only show what the user may view."""
    if method != "GET" or not isinstance(out, dict) or not enabled(m) or unrestricted(m, user):
        return out
    p = path.rstrip("/")
    if p == "/api/v1/workspaces" and "workspaces" in out:
        ok = viewable_workspaces(m, user)
        if ok is not None:
            out = dict(out, workspaces=[w for w in out["workspaces"] if int(w["id"]) in ok])
    elif p in ("/api/v1/experiments", "/api/v1/experiments-search") and "experiments" in out:
        ok = viewable_workspaces(m, user, "PERMISSION_TYPE_VIEW_EXPERIMENT_METADATA")
        if ok is not None:
            ws_of = {r["id"]: r["workspace_id"] for r in m.db.all("SELECT id, workspace_id FROM projects")}
            pick = (lambda e: e.get("experiment", e))
            out = dict(out, experiments=[e for e in out["experiments"] if ws_of.get(pick(e).get("projectId")) in ok])
    elif p == "/api/v1/models" and "models" in out:
        ok = viewable_workspaces(m, user, "PERMISSION_TYPE_VIEW_MODEL_REGISTRY")
        if ok is not None:
            out = dict(out, models=[x for x in out["models"] if int(x.get("workspaceId") or 1) in ok])
    elif p == "/api/v1/runs" and "runs" in out:
        ok = viewable_workspaces(m, user, "PERMISSION_TYPE_VIEW_EXPERIMENT_METADATA")
        if ok is not None:
            ws_of = {r["id"]: r["workspace_id"] for r in m.db.all("SELECT id, workspace_id FROM projects")}
            out = dict(out, runs=[r for r in out["runs"] if ws_of.get(r.get("projectId")) in ok])
    if isinstance(out.get("pagination"), dict):
        for key in ("workspaces", "experiments", "models", "runs"):
            if key in out and out["pagination"].get("total", 0) > len(out[key]) and not out["pagination"].get("limit"):
                out = dict(out, pagination=dict(out["pagination"], total=len(out[key])))
    return out


def can_assign_roles(m: Any, user: Dict[str, Any], body: Any) -> bool:
    """Administrators always; with rbac also whoever holds ``assign_roles`` in every workspace the request names (a
    WorkspaceAdmin hands out roles inside the workspace, never cluster-wide ones)."""
    if user.get("admin"):
        return True
    if not enabled(m):
        return False
    if unrestricted(m, user):
        return True
    scopes = [(a.get("roleAssignment") or {}).get("scopeWorkspaceId") for key in ("userRoleAssignments", "groupRoleAssignments") for a in (body or {}).get(key) or []]
    return bool(scopes) and all(ws is not None and allowed(m, user, "PERMISSION_TYPE_ASSIGN_ROLES", int(ws)) for ws in scopes)
