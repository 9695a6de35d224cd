"""Single sign-on through an OpenID Connect provider (authorization-code flow).

Parity: the reference's enterprise ``master/internal/plugin/oidc``: ``GET /oidc/sso`` sends the browser to the provider,
``GET /oidc/callback`` trades the code for tokens, reads the configured claim, maps it to a user and hands the browser (or the
CLI's local listener, through ``relayState``) a session token.  ``master.yaml``::

    oidc:
      enabled: true
      provider: okta                       # the name shown by `det auth list-providers`
      client_id: ...
      client_secret: ...
      idp_sso_url: https://idp/authorize   # authorization endpoint
      token_url: https://idp/token
      userinfo_url: https://idp/userinfo
      idp_recipient_url: https://det.example.com      # this master as the provider reaches it (redirect_uri = <this>/oidc/callback)
      authentication_claim: email          # the userinfo field that holds the user name
      auto_provision_users: false          # create unknown users as remote users on first sign-in
      always_redirect: false

The identity comes from the provider's userinfo endpoint, fetched by the master itself over TLS with the access token it just
received for its own client id — so no token signature has to be checked locally.  ``state`` is an HMAC-signed, time-limited
blob that carries the relay target, so the callback cannot be replayed with a forged destination.
"""
import base64
import hashlib
import hmac
import json
import secrets
import time
import urllib.parse
from typing import Any, Dict, Optional

import requests

from determined_b200.master import api as A

route, Ctx, HTTPError = A.route, A.Ctx, A.HTTPError
STATE_TTL_S = 600


def cfg_of(m: Any) -> Optional[Dict[str, Any]]:
    c = getattr(m, "oidc", None)
    return c if c and c.get("enabled") else None


def providers(m: Any) -> Any:
    c = cfg_of(m)
    return [{"name": c.get("provider") or "oidc", "ssoUrl": "/oidc/sso", "type": "OIDC", "alwaysRedirect": bool(c.get("always_redirect"))}] if c else []


def _key(m: Any) -> bytes:
    k = m.db.get_meta("oidc_state_key")
    if not k:
        k = secrets.token_hex(32)
        m.db.set_meta("oidc_state_key", k)
    return k.encode()


def sign_state(m: Any, relay: str, now: Optional[float] = None) -> str:
    body = base64.urlsafe_b64encode(json.dumps({"r": relay, "t": int(now if now is not None else time.time()), "n": secrets.token_hex(8)}).encode()).decode()
    return body + "." + hmac.new(_key(m), body.encode(), hashlib.sha256).hexdigest()


def read_state(m: Any, state: str) -> str:
    body, _, mac = state.rpartition(".")
    if not body or not hmac.compare_digest(mac, hmac.new(_key(m), body.encode(), hashlib.sha256).hexdigest()):
        raise HTTPError(400, "the sign-in state is not ours")
    d = json.loads(base64.urlsafe_b64decode(body.encode()))
    if time.time() - int(d["t"]) > STATE_TTL_S:
        raise HTTPError(400, "the sign-in took too long; start again")
    return str(d["r"])


def _relay_ok(relay: str) -> bool:
    """Where a token may be delivered: a path on this master, or a listener on this machine's loopback (the CLI)."""
    if relay.startswith("/") and not relay.startswith("//"):
        return True
    u = urllib.parse.urlparse(relay)
    return u.scheme == "http" and u.hostname in ("127.0.0.1", "localhost", "::1")


@route("GET", "/oidc/sso", auth_required=False)
def sso_start(ctx: Ctx) -> Any:
    c = cfg_of(ctx.m)
    if c is None:
        raise HTTPError(404, "single sign-on is not configured")
    relay = ctx.q("relayState") or "/det/login"
    if not _relay_ok(relay):
        raise HTTPError(400, "relayState must be a path on this master or a loopback address")
    q = {"response_type": "code", "client_id": c["client_id"], "redirect_uri": c["idp_recipient_url"].rstrip("/") + "/oidc/callback",
         "scope": c.get("scope") or "openid profile email", "state": sign_state(ctx.m, relay)}
    sep = "&" if "?" in c["idp_sso_url"] else "?"
    raise A.Redirect(c["idp_sso_url"] + sep + urllib.parse.urlencode(q))


@route("GET", "/oidc/callback", auth_required=False)
def sso_callback(ctx: Ctx) -> Any:
    c = cfg_of(ctx.m)
    if c is None:
        raise HTTPError(404, "single sign-on is not configured")
    if ctx.q("error"):
        raise HTTPError(401, f"the identity provider refused: {ctx.q('error')} {ctx.q('error_description') or ''}".strip())
    relay = read_state(ctx.m, ctx.q("state") or "")
    code = ctx.q("code")
    if not code:
        raise HTTPError(400, "no authorization code")
    try:
        tr = requests.post(c["token_url"], data={"grant_type": "authorization_code", "code": code,
                                                 "redirect_uri": c["idp_recipient_url"].rstrip("/") + "/oidc/callback"},
                           auth=(c["client_id"], c.get("client_secret") or ""), timeout=15)
        tr.raise_for_status()
        access = tr.json()["access_token"]
        ur = requests.get(c["userinfo_url"], headers={"Authorization": f"Bearer {access}"}, timeout=15)
        ur.raise_for_status()
        info = ur.json()
    except (requests.RequestException, KeyError, ValueError) as e:
        raise HTTPError(502, f"the identity provider did not complete the sign-in: {e}")
    claim = c.get("authentication_claim") or "email"
    name = info.get(claim)
    if not name:
        raise HTTPError(401, f"the identity provider sent no {claim!r} claim")
    row = ctx.m.db.one("SELECT * FROM users WHERE username=?", (str(name),))
    if row is None:
        if not c.get("auto_provision_users"):
            raise HTTPError(401, f"user {name!r} does not exist here; ask an administrator to create it (det user create --remote)")
        uid = ctx.m.db.exec("INSERT INTO users(username,password_hash,admin,active,display_name,modified_at) VALUES(?,?,?,?,?,?)",
                            (str(name), A.REMOTE_USER_MARK + secrets.token_hex(16), 0, 1, info.get("name"), time.time()))
        row = ctx.m.db.one("SELECT * FROM users WHERE id=?", (uid,))
    if not row["active"]:
        raise HTTPError(401, f"user {name!r} is deactivated")
    tok = secrets.token_hex(24)
    ctx.m.db.exec("INSERT INTO sessions(token,user_id,expiry) VALUES(?,?,?)", (tok, row["id"], time.time() + 30 * 86400))
    sep = "&" if "?" in relay else "?"
    raise A.Redirect(relay + sep + urllib.parse.urlencode({"jwt": tok, "user": row["username"]}))
