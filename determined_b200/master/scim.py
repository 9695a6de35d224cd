"""SCIM 2.0 user provisioning (RFC 7643 / 7644): an identity provider creates, updates and deactivates users here.

Parity: the reference's enterprise ``master/internal/plugin/scim`` + ``internal/user/postgres_scim_users.go``: ``/scim/v2/Users``
(list with ``filter=userName eq "..."``, ``startIndex`` / ``count``; create; get; replace; patch ``active`` / ``userName`` /
``displayName`` / ``name`` / ``emails``; delete = deactivate), ``/scim/v2/Groups`` (empty list: groups are managed here),
``/scim/v2/ServiceProviderConfig``.  Provisioned users are *remote* users: they sign in through the identity provider
(``master/sso.py``), never with a password.  ``master.yaml``::

    scim:
      enabled: true
      auth: {type: basic, username: scim, password: ...}       # or {type: oauth}: bearer tokens from POST /oauth2/token

``POST /oauth2/token`` implements the client-credentials grant for the clients registered with ``det oauth client add``.
"""
import base64
import hmac
import secrets
import time
from typing import Any, Dict, List, Optional

from determined_b200.master import api as A
from determined_b200.master import api_more as M

route, Ctx = A.route, A.Ctx
USER_SCHEMA = "urn:ietf:params:scim:schemas:core:2.0:User"
LIST_SCHEMA = "urn:ietf:params:scim:api:messages:2.0:ListResponse"
ERROR_SCHEMA = "urn:ietf:params:scim:api:messages:2.0:Error"
PATCH_SCHEMA = "urn:ietf:params:scim:api:messages:2.0:PatchOp"
TOKEN_TTL_S = 3600


class ScimError(A.HTTPError):
    def __init__(self, status: int, detail: str, scim_type: Optional[str] = None) -> None:
        super().__init__(status, detail)
        self.body = {"schemas": [ERROR_SCHEMA], "status": str(status), "detail": detail}
        if scim_type:
            self.body["scimType"] = scim_type


def _cfg(m: Any) -> Dict[str, Any]:
    c = getattr(m, "scim", None) or {}
    if not c.get("enabled"):
        raise ScimError(404, "SCIM is not enabled on this master")
    return c


# ---------------------------------------------------------------------------------------------------------------- OAuth token
@route("POST", "/oauth2/token", auth_required=False)
def oauth_token(ctx: Ctx) -> Any:
    """Client-credentials grant: ``Authorization: Basic client_id:secret`` (or the two as form fields) -> a bearer token."""
    import urllib.parse

    form = {k: v[0] for k, v in urllib.parse.parse_qs((ctx.raw or b"").decode(errors="replace")).items()} if not isinstance(ctx.body, dict) else dict(ctx.body)
    cid, sec = form.get("client_id"), form.get("client_secret")
    h = ctx.headers.get("Authorization", "") if getattr(ctx, "headers", None) else ""
    if h.startswith("Basic "):
        try:
            cid, _, sec = base64.b64decode(h[6:]).decode().partition(":")
        except Exception:  # noqa: BLE001
            pass
    if form.get("grant_type") != "client_credentials":
        raise A.HTTPError(400, "unsupported_grant_type")
    row = ctx.m.db.one("SELECT * FROM oauth_clients WHERE id=?", (cid or "",))
    if row is None or not hmac.compare_digest(str(row["secret"]), str(sec or "")):
        raise A.HTTPError(401, "invalid_client")
    tok = secrets.token_urlsafe(32)
    M.set_attr(ctx.m, "oauth_token", tok, "grant", {"client": cid, "expiry": time.time() + TOKEN_TTL_S})
    return {"access_token": tok, "token_type": "bearer", "expires_in": TOKEN_TTL_S}


def _authorize(ctx: Ctx) -> None:
    c = _cfg(ctx.m)
    auth = c.get("auth") or {}
    h = ctx.headers.get("Authorization", "") if getattr(ctx, "headers", None) else ""
    if auth.get("type", "basic") == "basic":
        if h.startswith("Basic "):
            try:
                user, _, pw = base64.b64decode(h[6:]).decode().partition(":")
            except Exception:  # noqa: BLE001
                user, pw = "", ""
            if hmac.compare_digest(user, str(auth.get("username") or "")) and hmac.compare_digest(pw, str(auth.get("password") or "")) and auth.get("password"):
                return
    elif h.startswith("Bearer "):
        g = M.get_attr(ctx.m, "oauth_token", h[7:], "grant")
        if g and g.get("expiry", 0) > time.time():
            return
    raise ScimError(401, "the identity provider's credentials were not accepted")


# ----------------------------------------------------------------------------------------------------------------------- users
def _extra(ctx: Ctx, uid: int) -> Dict[str, Any]:
    return M.get_attr(ctx.m, "user", uid, "scim") or {}


def _user(ctx: Ctx, row: Any) -> Dict[str, Any]:
    x = _extra(ctx, row["id"])
    base = (getattr(ctx.m, "url", "") or "").rstrip("/")
    out = {"schemas": [USER_SCHEMA], "id": x.get("id") or str(row["id"]), "userName": row["username"], "displayName": row["display_name"] or "",
           "active": bool(row["active"]), "meta": {"resourceType": "User", "location": f"{base}/scim/v2/Users/{x.get('id') or row['id']}"}}
    for k in ("externalId", "name", "emails"):
        if x.get(k) is not None:
            out[k] = x[k]
    return out


def _find(ctx: Ctx, scim_id: str) -> Any:
    for r in ctx.m.db.all("SELECT * FROM users ORDER BY id"):
        if (_extra(ctx, r["id"]).get("id") or str(r["id"])) == scim_id:
            return r
    raise ScimError(404, f"user {scim_id} not found")


def _apply(ctx: Ctx, row: Any, doc: Dict[str, Any]) -> Any:
    """Write the attributes of a SCIM user document that are present onto the user row and its SCIM side record."""
    uid = row["id"]
    if doc.get("userName") is not None and doc["userName"] != row["username"]:
        if ctx.m.db.one("SELECT id FROM users WHERE username=? AND id<>?", (doc["userName"], uid)) is not None:
            raise ScimError(409, f"userName {doc['userName']!r} is taken", "uniqueness")
        ctx.m.db.exec("UPDATE users SET username=? WHERE id=?", (doc["userName"], uid))
    display = doc.get("displayName")
    if display is None and isinstance(doc.get("name"), dict) and (doc["name"].get("givenName") or doc["name"].get("familyName")):
        display = " ".join(x for x in (doc["name"].get("givenName"), doc["name"].get("familyName")) if x)
    if display is not None:
        ctx.m.db.exec("UPDATE users SET display_name=? WHERE id=?", (display, uid))
    if doc.get("active") is not None:
        ctx.m.db.exec("UPDATE users SET active=? WHERE id=?", (int(bool(doc["active"])), uid))
        if not doc["active"]:
            ctx.m.db.exec("DELETE FROM sessions WHERE user_id=?", (uid,))          # a deprovisioned user is signed out everywhere
    x = _extra(ctx, uid)
    for k in ("externalId", "name", "emails"):
        if k in doc:
            x[k] = doc[k]
    M.set_attr(ctx.m, "user", uid, "scim", x)
    ctx.m.db.exec("UPDATE users SET modified_at=? WHERE id=?", (time.time(), uid))
    return ctx.m.db.one("SELECT * FROM users WHERE id=?", (uid,))


@route("GET", "/scim/v2/Users", auth_required=False)
def list_users(ctx: Ctx) -> Any:
    _authorize(ctx)
    rows = list(ctx.m.db.all("SELECT * FROM users ORDER BY id"))
    flt = ctx.q("filter")
    if flt:
        parts = flt.split(None, 2)
        if len(parts) != 3 or parts[1].lower() != "eq" or parts[0] not in ("userName", "externalId"):
            raise ScimError(400, 'only filters of the form: userName eq "x" / externalId eq "x" are supported', "invalidFilter")
        want = parts[2].strip().strip('"')
        rows = [r for r in rows if (r["username"] if parts[0] == "userName" else _extra(ctx, r["id"]).get("externalId")) == want]
    start, count = max(ctx.qi("startIndex", 1), 1), ctx.qi("count", 100)
    page = rows[start - 1:start - 1 + max(count, 0)]
    return {"schemas": [LIST_SCHEMA], "totalResults": len(rows), "startIndex": start, "itemsPerPage": len(page), "Resources": [_user(ctx, r) for r in page]}


@route("POST", "/scim/v2/Users", auth_required=False)
def create_user(ctx: Ctx) -> Any:
    _authorize(ctx)
    doc = ctx.body or {}
    if not doc.get("userName"):
        raise ScimError(400, "userName is required", "invalidValue")
    if ctx.m.db.one("SELECT id FROM users WHERE username=?", (doc["userName"],)) is not None:
        raise ScimError(409, f"userName {doc['userName']!r} is taken", "uniqueness")
    uid = ctx.m.db.exec("INSERT INTO users(username,password_hash,admin,active,display_name,modified_at) VALUES(?,?,?,?,?,?)",
                        (doc["userName"], A.REMOTE_USER_MARK + secrets.token_hex(16), 0, 1, None, time.time()))
    M.set_attr(ctx.m, "user", uid, "scim", {"id": secrets.token_hex(16)})
    row = _apply(ctx, ctx.m.db.one("SELECT * FROM users WHERE id=?", (uid,)), dict(doc, active=doc.get("active", True)))
    return (201, _user(ctx, row))


@route("GET", "/scim/v2/Users/{sid}", auth_required=False)
def get_user(ctx: Ctx, sid: str) -> Any:
    _authorize(ctx)
    return _user(ctx, _find(ctx, sid))


@route("PUT", "/scim/v2/Users/{sid}", auth_required=False)
def put_user(ctx: Ctx, sid: str) -> Any:
    _authorize(ctx)
    return _user(ctx, _apply(ctx, _find(ctx, sid), ctx.body or {}))


@route("PATCH", "/scim/v2/Users/{sid}", auth_required=False)
def patch_user(ctx: Ctx, sid: str) -> Any:
    _authorize(ctx)
    row = _find(ctx, sid)
    doc: Dict[str, Any] = {}
    ops: List[Dict[str, Any]] = (ctx.body or {}).get("Operations") or []
    for op in ops:
        kind = str(op.get("op", "")).lower()
        if kind not in ("replace", "add"):
            raise ScimError(400, f"operation {op.get('op')!r} is not supported", "invalidSyntax")
        if op.get("path"):
            doc[str(op["path"])] = op.get("value")
        elif isinstance(op.get("value"), dict):
            doc.update(op["value"])
    if isinstance(doc.get("active"), str):
        doc["active"] = doc["active"].lower() == "true"            # some providers send booleans as strings
    return _user(ctx, _apply(ctx, row, doc))


@route("DELETE", "/scim/v2/Users/{sid}", auth_required=False)
def delete_user(ctx: Ctx, sid: str) -> Any:
    """Users own experiments, so they are never removed: deleting deactivates."""
    _authorize(ctx)
    _apply(ctx, _find(ctx, sid), {"active": False})
    return (204, None)


@route("GET", "/scim/v2/Groups", auth_required=False)
def list_groups(ctx: Ctx) -> Any:
    _authorize(ctx)
    return {"schemas": [LIST_SCHEMA], "totalResults": 0, "startIndex": 1, "itemsPerPage": 0, "Resources": []}


@route("GET", "/scim/v2/ServiceProviderConfig", auth_required=False)
def service_provider_config(ctx: Ctx) -> Any:
    _cfg(ctx.m)
    return {"schemas": ["urn:ietf:params:scim:schemas:core:2.0:ServiceProviderConfig"], "patch": {"supported": True}, "bulk": {"supported": False},
            "filter": {"supported": True, "maxResults": 1000}, "changePassword": {"supported": False}, "sort": {"supported": False}, "etag": {"supported": False},
            "authenticationSchemes": [{"type": "httpbasic", "name": "HTTP Basic"}, {"type": "oauthbearertoken", "name": "OAuth Bearer Token"}]}
