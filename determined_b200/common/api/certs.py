"""Which certificate a client trusts when the master serves HTTPS.

Parity: reference ``harness/determined/common/api/certs.py`` (``Cert(cert_pem=, noverify=, name=)``, ``default_load`` from
``DET_MASTER_CERT_FILE`` / ``DET_MASTER_CERT_NAME`` or the per-master store next to the token cache, ``CertStore``) and the
CLI's trust-on-first-use flow (``harness/determined/cli/cli.py``: on an ``SSLError`` the server's certificate is fetched,
its fingerprint shown, and it is stored once the user agrees).

A :class:`Cert` is what ``Session(..., cert=...)`` takes:

* ``Cert()`` — the system's trust roots (what ``requests`` does by default);
* ``Cert(cert_pem=...)`` — trust exactly this PEM bundle (a self-signed master certificate, or a private CA);
* ``Cert(noverify=True)`` — no verification at all;
* ``name=`` — the host name the certificate must carry when it differs from the one in the URL (a master reached through
  an IP address or a tunnel).  Implemented with a transport adapter that sets the TLS server name and the name to check.
"""
import contextlib
import hashlib
import json
import os
import pathlib
import ssl
import tempfile
import urllib.parse
from typing import Any, Dict, Optional, Tuple, Union


class Cert:
    def __init__(self, cert_pem: Optional[str] = None, noverify: bool = False, name: Optional[str] = None) -> None:
        self._pem = cert_pem
        self._noverify = noverify
        self._name = name
        self._bundle_path: Optional[str] = None

    @property
    def bundle(self) -> Union[None, str, bool]:
        """``requests``' ``verify=``: False (no verification), a CA-bundle path, or None (system roots)."""
        if self._noverify:
            return False
        if self._pem is None:
            return None
        if self._bundle_path is None or not os.path.exists(self._bundle_path):
            # requests wants a file; it lives as long as the process (one small file per distinct Cert)
            fd, path = tempfile.mkstemp(prefix="det-master-cert-", suffix=".pem")
            with os.fdopen(fd, "w") as f:
                f.write(self._pem)
            self._bundle_path = path
        return self._bundle_path

    @property
    def name(self) -> Optional[str]:
        return self._name

    @property
    def pem(self) -> Optional[str]:
        return self._pem

    def ssl_context(self) -> ssl.SSLContext:
        """For code that opens sockets itself (the websocket client, ``urllib``)."""
        if self._noverify:
            ctx = ssl.create_default_context()
            ctx.check_hostname = False
            ctx.verify_mode = ssl.CERT_NONE
            return ctx
        ctx = ssl.create_default_context()
        if self._pem is not None:
            ctx.load_verify_locations(cadata=self._pem)
        return ctx

    def __repr__(self) -> str:
        kind = "noverify" if self._noverify else (f"pem:{fingerprint(self._pem)[:16]}" if self._pem else "system")
        return f"Cert({kind}{', name=' + self._name if self._name else ''})"


def fingerprint(pem: str) -> str:
    """SHA-256 over the DER form of the first certificate of a PEM bundle, colon-separated hex."""
    der = ssl.PEM_cert_to_DER_cert(pem[pem.index("-----BEGIN CERTIFICATE-----"):pem.index("-----END CERTIFICATE-----") + 25])
    h = hashlib.sha256(der).hexdigest().upper()
    return ":".join(h[i:i + 2] for i in range(0, len(h), 2))


def _config_dir() -> pathlib.Path:
    from determined_b200.common.api import authentication

    return authentication._config_dir()


class CertStore:
    """Certificates the user accepted, keyed by master URL: ``{url: {"cert_pem": ..., "name": ...}}`` in ``certs.json`` next
    to the token cache (reference keeps them in the same directory as ``auth.json``)."""

    def __init__(self, path: Optional[pathlib.Path] = None) -> None:
        self.path = pathlib.Path(path) if path is not None else _config_dir() / "certs.json"

    def _load(self) -> Dict[str, Any]:
        try:
            d = json.loads(self.path.read_text())
            return d if isinstance(d, dict) else {}
        except (FileNotFoundError, json.JSONDecodeError):
            return {}

    def get_cert(self, url: str) -> Optional[Tuple[str, Optional[str]]]:
        e = self._load().get(_key(url))
        return (e["cert_pem"], e.get("name")) if e and e.get("cert_pem") else None

    def set_cert(self, url: str, cert_pem: str, name: Optional[str] = None) -> None:
        d = self._load()
        d[_key(url)] = {"cert_pem": cert_pem, "name": name}
        self.path.parent.mkdir(parents=True, exist_ok=True)
        tmp = self.path.with_suffix(".tmp")
        tmp.write_text(json.dumps(d, indent=1))
        os.chmod(tmp, 0o600)
        os.replace(tmp, self.path)

    def delete_cert(self, url: str) -> None:
        d = self._load()
        if d.pop(_key(url), None) is not None:
            self.path.write_text(json.dumps(d, indent=1))


def _key(url: str) -> str:
    from determined_b200.common.api import _session

    return _session.canonicalize_master_url(url)


def default_load(master_url: str, explicit_path: Optional[str] = None, explicit_cert_name: Optional[str] = None,
                 explicit_noverify: bool = False) -> Cert:
    """The certificate to use for ``master_url``: arguments first, then ``DET_MASTER_CERT_FILE`` (a path, or the literal
    ``noverify``) / ``DET_MASTER_CERT_NAME`` (what tasks get from the master), then the store, then the system roots."""
    path = explicit_path or os.environ.get("DET_MASTER_CERT_FILE")
    name = explicit_cert_name or os.environ.get("DET_MASTER_CERT_NAME") or None
    noverify = explicit_noverify
    if path is not None and path.lower() == "noverify":
        noverify, path = True, None
    if noverify:
        return Cert(noverify=True, name=name)
    if path:
        return Cert(cert_pem=pathlib.Path(path).read_text(), name=name)
    with contextlib.suppress(OSError, ValueError):
        old_file = _config_dir() / "master.crt"          # the layout before the per-master store: one certificate, one master
        if old_file.exists():
            CertStore().set_cert(master_url, old_file.read_text())
            old_file.unlink()
    with contextlib.suppress(OSError, ValueError):
        got = CertStore().get_cert(master_url)
        if got is not None:
            return Cert(cert_pem=got[0], name=name or got[1])
    return Cert(name=name)


def fetch_server_cert(master_url: str, timeout: float = 10.0) -> str:
    """The PEM certificate the server at ``master_url`` presents (unverified: for showing its fingerprint to the user)."""
    u = urllib.parse.urlparse(_key(master_url))
    assert u.hostname and u.port
    return ssl.get_server_certificate((u.hostname, u.port), timeout=timeout)


def trust_on_first_use(master_url: str, confirm: Any = None, store: Optional[CertStore] = None) -> Cert:
    """Fetch the master's certificate, ask ``confirm(fingerprint) -> bool`` (default: a terminal prompt), store it.
    Raises ``PermissionError`` when the user declines."""
    pem = fetch_server_cert(master_url)
    fp = fingerprint(pem)
    if confirm is None:
        def confirm(f: str) -> bool:  # noqa: E306
            ans = input(f"The master at {master_url} presents a certificate this machine does not trust.\n"
                        f"SHA256 fingerprint: {f}\nTrust it from now on? [y/N] ")
            return ans.strip().lower() in ("y", "yes")
    if not confirm(fp):
        raise PermissionError(f"certificate of {master_url} ({fp}) was not accepted")
    (store or CertStore()).set_cert(master_url, pem)
    return Cert(cert_pem=pem)


# the process-wide default: what ``Session(master)`` uses when no ``cert=`` is given (reference ``certs.cli_cert``)
cli_cert: Optional[Cert] = None


def host_name_adapter(name: str) -> Any:
    """A ``requests`` transport adapter that connects to the URL's host but checks the certificate against ``name``."""
    import requests.adapters

    class _Adapter(requests.adapters.HTTPAdapter):
        def init_poolmanager(self, *a: Any, **kw: Any) -> Any:
            kw["assert_hostname"] = name
            kw["server_hostname"] = name
            return super().init_poolmanager(*a, **kw)

    return _Adapter()
