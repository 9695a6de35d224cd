"""Login + on-disk token cache.

Parity: reference ``harness/determined/common/api/authentication.py`` (``login``,
``TokenStore`` at ``~/.config/determined/auth.json`` keyed by master URL, ``salt_and_hash``,
``DET_USER``/``DET_PASS``/``DET_USER_TOKEN`` env overrides, session-token fast path inside tasks).
"""
import contextlib
import getpass
import hashlib
import json
import os
import pathlib
from typing import Any, Dict, List, Optional, Tuple

from determined_b200.common.api import _session, errors

PASSWORD_SALT = "GubPEmmotfiK9TMD6Zdw"
DEFAULT_USER = "determined"


def salt_and_hash(password: str) -> str:
    if not password:
        return password
    return hashlib.sha512((PASSWORD_SALT + password).encode()).hexdigest()


def _config_dir() -> pathlib.Path:
    base = os.environ.get("DET_DEBUG_CONFIG_PATH") or os.environ.get("DET_CONFIG_DIR") or os.path.join(
        os.environ.get("XDG_CONFIG_HOME", os.path.expanduser("~/.config")), "determined_b200")
    return pathlib.Path(base)


class TokenStore:
    """``{"version":2, "masters": {url: {"active_user": u, "tokens": {u: token}}}}``"""

    def __init__(self, master_url: str, path: Optional[pathlib.Path] = None) -> None:
        self.master_url = _session.canonicalize_master_url(master_url)
        self.path = path or (_config_dir() / "auth.json")
        self._d = self._load()

    def _load(self) -> Dict[str, Any]:
        try:
            d = json.loads(self.path.read_text())
        except FileNotFoundError:
            return {"version": 2, "masters": {}}
        except json.JSONDecodeError:
            raise errors.CorruptTokenCacheException()
        if not isinstance(d, dict):
            raise errors.CorruptTokenCacheException()
        version = d.get("version", 0)
        if version == 0 and "masters" not in d:
            # the oldest layout knew one master only: {"active_user": ..., "tokens": {...}} belongs to the master asked for
            entry = {"tokens": dict(d.get("tokens") or {})}
            if d.get("active_user"):
                entry["active_user"] = d["active_user"]
            d = {"version": 2, "masters": {self.master_url: entry}}
        elif version == 1:
            d = {"version": 2, "masters": _merge_v1_masters(d.get("masters") or {})}
        if "masters" not in d or not isinstance(d["masters"], dict):
            raise errors.CorruptTokenCacheException()
        return d

    def _save(self) -> None:
        self.path.parent.mkdir(parents=True, exist_ok=True)
        tmp = self.path.with_suffix(".tmp")
        tmp.write_text(json.dumps(self._d, indent=1))
        os.chmod(tmp, 0o600)
        os.replace(tmp, self.path)

    def _m(self) -> Dict[str, Any]:
        return self._d["masters"].setdefault(self.master_url, {"tokens": {}})

    def get_active_user(self) -> Optional[str]:
        return self._m().get("active_user")

    def get_token(self, user: str) -> Optional[str]:
        return self._m()["tokens"].get(user)

    def get_all_users(self) -> List[str]:
        return sorted(self._m()["tokens"])

    def set_token(self, user: str, token: str) -> None:
        self._m()["tokens"][user] = token
        self._save()

    def set_active(self, user: str) -> None:
        self._m()["active_user"] = user
        self._save()

    def drop_user(self, user: str) -> None:
        self._m()["tokens"].pop(user, None)
        if self._m().get("active_user") == user:
            self._m().pop("active_user")
        self._save()

    def clear_active(self) -> None:
        self._m().pop("active_user", None)
        self._save()


def _merge_v1_masters(masters: Dict[str, Any]) -> Dict[str, Any]:
    """Version 1 keyed masters by whatever the user had typed (``host``, ``http://host:80#x``, ``https://host/?k=v``); version
    2 keys by the canonical URL.  Entries that meet under one key are merged: tokens are united (either one wins on a clash),
    one of the active users survives.  URLs nobody can parse are dropped rather than failing the whole file."""
    import urllib.parse

    out: Dict[str, Any] = {}
    for raw, entry in masters.items():
        try:
            bare = "://" not in raw
            u = urllib.parse.urlparse(("http://" + raw) if bare else raw)
            host = u.hostname or ""
            if ":" in host:
                host = f"[{host}]"
            port = f":{u.port}" if u.port else ""
            key = _session.canonicalize_master_url((f"{host}{port}" if bare else f"{u.scheme}://{host}{port}") + u.path)
        except ValueError:
            continue
        if not isinstance(entry, dict):
            continue
        dst = out.setdefault(key, {"tokens": {}})
        for user, tok in (entry.get("tokens") or {}).items():
            dst["tokens"].setdefault(user, tok)
        if entry.get("active_user") and "active_user" not in dst:
            dst["active_user"] = entry["active_user"]
    return out


def login(master: str, username: str, password: str, cert: Any = None) -> _session.Session:
    s = _session.UnauthSession(master, cert)
    r = s.post("/api/v1/auth/login", json={"username": username, "password": salt_and_hash(password),
                                            "isHashed": True})
    token = r.json()["token"]
    return _session.Session(master, username, token, cert)


def _token_valid(master: str, token: str) -> bool:
    try:
        _session.Session(master, None, token, None, max_retries=1).get("/api/v1/me")
        return True
    except (errors.UnauthenticatedException, errors.ForbiddenException):
        return False


def login_with_cache(master: str, requested_user: Optional[str] = None, password: Optional[str] = None,
                     cert: Any = None) -> _session.Session:
    """Order: task session token (``DET_SESSION_TOKEN``) → ``DET_USER_TOKEN`` → cached token →
    ``DET_USER``/``DET_PASS`` → default user with blank password → interactive prompt."""
    env = os.environ
    if env.get("DET_SESSION_TOKEN") and env.get("DET_MASTER") and requested_user is None:
        return _session.Session(master, env.get("DET_USER", DEFAULT_USER), env["DET_SESSION_TOKEN"], cert)
    if env.get("DET_USER_TOKEN") and requested_user is None:
        return _session.Session(master, env.get("DET_USER", DEFAULT_USER), env["DET_USER_TOKEN"], cert)
    store = None
    with contextlib.suppress(errors.CorruptTokenCacheException, OSError):
        store = TokenStore(master)
    user = requested_user or env.get("DET_USER") or (store.get_active_user() if store else None) or DEFAULT_USER
    if store is not None and password is None:
        tok = store.get_token(user)
        if tok and _token_valid(master, tok):
            return _session.Session(master, user, tok, cert)
    if password is None:
        password = env.get("DET_PASS")
    if password is None:
        # default accounts ship with a blank password: try that before bothering the user
        try:
            sess = login(master, user, "", cert)
        except errors.UnauthenticatedException:
            import sys

            if not sys.stdin or not sys.stdin.isatty():
                raise
            password = getpass.getpass(f"Password for user '{user}': ")
            sess = login(master, user, password, cert)
    else:
        sess = login(master, user, password, cert)
    if store is not None:
        with contextlib.suppress(OSError):
            store.set_token(user, sess.token or "")
            if requested_user is None and store.get_active_user() is None:
                store.set_active(user)     # `-u other` is a one-off; `det user login` switches
    return sess


def logout(master: str, user: Optional[str] = None, cert: Any = None) -> None:
    store = TokenStore(master)
    user = user or store.get_active_user() or DEFAULT_USER
    tok = store.get_token(user)
    if tok:
        with contextlib.suppress(errors.APIException, errors.MasterNotFoundException):
            _session.Session(master, user, tok, cert, max_retries=0).post("/api/v1/auth/logout")
    store.drop_user(user)


def get_det_username_from_env() -> Optional[str]:
    return os.environ.get("DET_USER")


def get_det_user_token_from_env() -> Optional[str]:
    return os.environ.get("DET_USER_TOKEN")


def get_det_password_from_env() -> Optional[str]:
    return os.environ.get("DET_PASS")


def default_load_user_password(requested_user: Optional[str], password: Optional[str], token_store: "TokenStore") -> Tuple[Optional[str], Optional[str], bool]:
    """(user, password, was the user the fall-back default?) from the arguments, the environment, the active user of the
    token store, or the built-in ``determined`` user with its blank password."""
    if requested_user:
        return requested_user, password, False
    if get_det_username_from_env() and get_det_password_from_env() is not None:
        return get_det_username_from_env(), get_det_password_from_env(), False
    if token_store.get_active_user():
        return token_store.get_active_user(), password, False
    return "determined", password if password is not None else "", True


def logout_all(master: str, cert: Any = None) -> None:
    """Drop every session this machine holds on ``master``."""
    for user in TokenStore(master).get_all_users():
        logout(master, user, cert) if "cert" in logout.__code__.co_varnames else logout(master, user)
