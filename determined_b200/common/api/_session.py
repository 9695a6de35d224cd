"""HTTP session to the master with bearer-token auth and bounded retries.

Parity: reference ``harness/determined/common/api/_session.py:90-284`` (``Session`` with
``get/post/put/patch/delete``, retry policy on connection errors and 502/503/504).
"""
import json as _json
import logging
import time
from typing import Any, Dict, Optional, Tuple, Union

from determined_b200.common.api import errors

logger = logging.getLogger("determined.api")

RETRY_STATUSES = (502, 503, 504)
DEFAULT_MAX_RETRIES = 5


def canonicalize_master_url(url: str) -> str:
    """``host`` -> ``http://host:8080``; explicit schemes get their own default port (80 / 443); a path is kept without its
    trailing slash; credentials, queries and fragments are refused, as is an empty host."""
    import urllib.parse

    bare = "://" not in url
    u = urllib.parse.urlparse(("http://" + url) if bare else url)
    if u.username or u.password or u.query or u.fragment or u.params:
        raise ValueError(f"master url must not contain username, password, query, or fragment: {url!r}")
    if not u.hostname:
        raise ValueError(f"master url must contain a nonempty hostname: {url!r}")
    if u.scheme not in ("http", "https"):
        raise ValueError(f"master url must be http or https: {url!r}")
    port = u.port or (8080 if bare else 443 if u.scheme == "https" else 80)
    host = f"[{u.hostname}]" if ":" in u.hostname else u.hostname
    return f"{u.scheme}://{host}:{port}{u.path.rstrip('/')}"


class BaseSession:
    def __init__(self, master: str, token: Optional[str] = None, cert: Any = None,
                 max_retries: int = DEFAULT_MAX_RETRIES, timeout: Optional[float] = None) -> None:
        self.master = canonicalize_master_url(master)
        self.token = token
        self.cert = cert
        self._max_retries = max_retries
        self._timeout = timeout
        self._http = None
        self._verify_kw: Dict[str, Any] = {}

    def _client(self) -> Any:
        if self._http is None:
            import requests

            self._http = requests.Session()
            cert = self._effective_cert()
            self._verify_kw = {}
            if cert is not None:
                if cert.bundle is not None:
                    # per request, not on the session: REQUESTS_CA_BUNDLE / CURL_CA_BUNDLE in the environment would
                    # otherwise replace a session-level setting
                    self._verify_kw = {"verify": cert.bundle}
                if cert.name:
                    from determined_b200.common.api import certs

                    self._http.mount("https://", certs.host_name_adapter(cert.name))
        return self._http

    def _effective_cert(self) -> Any:
        """The ``cert=`` given, else the process-wide default (``certs.cli_cert``), else — for an https master — whatever
        ``certs.default_load`` finds (environment, accepted-certificate store)."""
        if self.cert is not None:
            return self.cert
        from determined_b200.common.api import certs

        if certs.cli_cert is not None:
            return certs.cli_cert
        if self.master.startswith("https://"):
            return certs.default_load(self.master)
        return None

    def _headers(self, extra: Optional[Dict[str, str]]) -> Dict[str, str]:
        h = {"Content-Type": "application/json"}
        if self.token:
            h["Authorization"] = f"Bearer {self.token}"
        from determined_b200.common import tracing

        tp = tracing.current_traceparent()
        if tp:
            h["traceparent"] = tp       # the master's span for this request becomes a child of the caller's
        if extra:
            h.update(extra)
        return h

    def _do_request(self, method: str, path: str, params: Optional[Dict[str, Any]] = None,
                    json: Any = None, data: Any = None, headers: Optional[Dict[str, str]] = None,
                    timeout: Optional[Union[Tuple, float]] = None, stream: bool = False) -> Any:
        import requests

        url = self.master + (path if path.startswith("/") else "/" + path)
        if json is not None and data is None:
            from determined_b200 import util

            data = util.json_encode(json)
        attempt, delay = 0, 0.2
        while True:
            try:
                r = self._client().request(method, url, params=params, data=data, headers=self._headers(headers),
                                           timeout=timeout or self._timeout, stream=stream, **self._verify_kw)
            except requests.exceptions.SSLError:
                raise                    # a certificate problem does not go away by retrying; the CLI offers to trust it
            except (requests.ConnectionError, requests.Timeout) as e:
                if attempt >= self._max_retries:
                    raise errors.MasterNotFoundException(f"cannot reach master at {self.master}: {e}") from e
                attempt += 1
                time.sleep(delay)
                delay = min(delay * 2, 5.0)
                continue
            if r.status_code in RETRY_STATUSES and attempt < self._max_retries:
                attempt += 1
                time.sleep(delay)
                delay = min(delay * 2, 5.0)
                continue
            break
        if r.status_code >= 400:
            try:
                body = r.json()
                msg = body.get("message") or body.get("error", {}).get("message") or r.text
            except Exception:
                body, msg = None, r.text
            if r.status_code == 404:
                raise errors.NotFoundException(msg, body)
            if r.status_code == 401:
                raise errors.UnauthenticatedException(msg, body)
            if r.status_code == 403:
                raise errors.ForbiddenException(msg, body)
            raise errors.APIException(r.status_code, msg, body)
        return r

    def get(self, path: str, **kw: Any) -> Any:
        return self._do_request("GET", path, **kw)

    def delete(self, path: str, **kw: Any) -> Any:
        return self._do_request("DELETE", path, **kw)

    def post(self, path: str, **kw: Any) -> Any:
        return self._do_request("POST", path, **kw)

    def patch(self, path: str, **kw: Any) -> Any:
        return self._do_request("PATCH", path, **kw)

    def put(self, path: str, **kw: Any) -> Any:
        return self._do_request("PUT", path, **kw)

    def close(self) -> None:
        if self._http is not None:
            self._http.close()
            self._http = None


class UnauthSession(BaseSession):
    def __init__(self, master: str, cert: Any = None, max_retries: int = DEFAULT_MAX_RETRIES) -> None:
        super().__init__(master, None, cert, max_retries)


class Session(BaseSession):
    def __init__(self, master: str, username: Optional[str] = None, token: Optional[str] = None,
                 cert: Any = None, max_retries: int = DEFAULT_MAX_RETRIES) -> None:
        super().__init__(master, token, cert, max_retries)
        self.username = username

    def with_retry(self, max_retries: int) -> "Session":
        return Session(self.master, self.username, self.token, self.cert, max_retries)
