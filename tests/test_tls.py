"""HTTPS between clients, tasks, agents and the master (reference ``harness/tests/common/test_tls.py`` checks ``certs.Cert``
against a test server; here the server is the master itself and the chain goes all the way into a trial process)."""
import base64
import datetime
import ipaddress
import pathlib
import time

import pytest
import requests
import yaml

from determined_b200.common import api
from determined_b200.common.api import certs
from determined_b200.master import api as mapi
from determined_b200.master import core

FIX = pathlib.Path(__file__).parent / "fixtures" / "onevar_exp"


def make_cert(tmp: pathlib.Path, common_name: str = "127.0.0.1", stem: str = "master"):
    """A self-signed certificate for 127.0.0.1 (IP SAN) and ``common_name`` (DNS SAN)."""
    from cryptography import x509
    from cryptography.hazmat.primitives import hashes, serialization
    from cryptography.hazmat.primitives.asymmetric import ec
    from cryptography.x509.oid import NameOID

    key = ec.generate_private_key(ec.SECP256R1())
    name = x509.Name([x509.NameAttribute(NameOID.COMMON_NAME, common_name)])
    now = datetime.datetime.now(datetime.timezone.utc)
    san = [x509.IPAddress(ipaddress.ip_address("127.0.0.1"))] if common_name == "127.0.0.1" else [x509.DNSName(common_name)]
    cert = (x509.CertificateBuilder().subject_name(name).issuer_name(name).public_key(key.public_key())
            .serial_number(x509.random_serial_number()).not_valid_before(now - datetime.timedelta(minutes=5))
            .not_valid_after(now + datetime.timedelta(days=2))
            .add_extension(x509.SubjectAlternativeName(san), critical=False)
            .add_extension(x509.BasicConstraints(ca=True, path_length=None), critical=True)
            .sign(key, hashes.SHA256()))
    cpath, kpath = tmp / f"{stem}-cert.pem", tmp / f"{stem}-key.pem"
    cpath.write_bytes(cert.public_bytes(serialization.Encoding.PEM))
    kpath.write_bytes(key.private_bytes(serialization.Encoding.PEM, serialization.PrivateFormat.PKCS8, serialization.NoEncryption()))
    return cpath, kpath


@pytest.fixture()
def tls_cluster(tmp_path, monkeypatch):
    monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "config"))
    monkeypatch.delenv("DET_MASTER_CERT_FILE", raising=False)
    monkeypatch.delenv("DET_MASTER_CERT_NAME", raising=False)
    monkeypatch.setattr(certs, "cli_cert", None)
    cpath, kpath = make_cert(tmp_path)
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}}, scheduler_type="priority")
    srv = mapi.APIServer(m, "127.0.0.1", 0, tls_cert=str(cpath), tls_key=str(kpath)).start()
    m.start()
    try:
        yield m, cpath, tmp_path
    finally:
        m.stop()
        srv.stop()


def test_cert_kinds_against_an_https_master(tls_cluster):
    m, cpath, tmp = tls_cluster
    assert m.url.startswith("https://") and m.master_env()["DET_MASTER_CERT_FILE"] == str(cpath)
    pem = cpath.read_text()
    other_pem = make_cert(tmp, stem="other")[0].read_text()
    for kwargs, ok in [({"noverify": True}, True), ({"cert_pem": pem}, True), ({}, False), ({"noverify": False}, False), ({"cert_pem": other_pem}, False)]:
        sess = api.UnauthSession(m.url, cert=certs.Cert(**kwargs), max_retries=0)
        if ok:
            assert sess.get("/api/v1/master").json()["clusterId"] == m.cluster_id
        else:
            with pytest.raises(requests.exceptions.SSLError):
                sess.get("/api/v1/master")
    # plain http against the TLS port is not a certificate problem, and not an answer either
    with pytest.raises((api.errors.MasterNotFoundException, requests.exceptions.RequestException)):
        api.UnauthSession(m.url.replace("https://", "http://"), max_retries=0).get("/api/v1/master")
    assert len(certs.fingerprint(pem).split(":")) == 32 and certs.fingerprint(pem) != certs.fingerprint(other_pem)
    assert "pem:" in repr(certs.Cert(cert_pem=pem)) and repr(certs.Cert(noverify=True)) == "Cert(noverify)"


def test_default_load_order_and_the_store_of_accepted_certificates(tls_cluster, monkeypatch):
    m, cpath, tmp = tls_cluster
    assert certs.default_load(m.url).bundle is None                     # nothing configured: system roots
    # trust on first use: the fingerprint shown is the server's, a refusal stores nothing
    shown = []
    with pytest.raises(PermissionError):
        certs.trust_on_first_use(m.url, confirm=lambda fp: shown.append(fp) or False)
    assert shown == [certs.fingerprint(cpath.read_text())] and certs.CertStore().get_cert(m.url) is None
    cert = certs.trust_on_first_use(m.url, confirm=lambda fp: True)
    assert certs.fingerprint(cert.pem) == shown[0] and certs.CertStore().get_cert(m.url)[0] == cert.pem
    # from now on a Session without cert= finds it by itself
    sess = api.authentication.login(m.url, "determined", "")
    assert sess.get("/api/v1/me").json()["user"]["username"] == "determined"
    assert certs.fingerprint(certs.default_load(m.url).pem) == shown[0]
    # the environment (what tasks get) wins over the store; "noverify" is a value of the file variable
    monkeypatch.setenv("DET_MASTER_CERT_FILE", "noverify")
    assert certs.default_load(m.url).bundle is False
    monkeypatch.setenv("DET_MASTER_CERT_FILE", str(cpath))
    monkeypatch.setenv("DET_MASTER_CERT_NAME", "det-master.internal")
    c = certs.default_load(m.url)
    assert c.name == "det-master.internal" and open(c.bundle).read() == cpath.read_text()
    # arguments win over the environment
    assert certs.default_load(m.url, explicit_noverify=True).bundle is False
    certs.CertStore().delete_cert(m.url)
    assert certs.CertStore().get_cert(m.url) is None


def test_cert_name_checks_the_certificate_against_another_host_name(tmp_path, monkeypatch):
    """A master reached by IP whose certificate carries a DNS name: ``name=`` is what the certificate is checked against."""
    monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "config"))
    monkeypatch.setattr(certs, "cli_cert", None)
    cpath, kpath = make_cert(tmp_path, common_name="det-master.internal")
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"), launch=False)
    srv = mapi.APIServer(m, "127.0.0.1", 0, tls_cert=str(cpath), tls_key=str(kpath)).start()
    try:
        pem = cpath.read_text()
        with pytest.raises(requests.exceptions.SSLError):          # right CA, wrong name
            api.UnauthSession(m.url, cert=certs.Cert(cert_pem=pem), max_retries=0).get("/api/v1/master")
        ok = api.UnauthSession(m.url, cert=certs.Cert(cert_pem=pem, name="det-master.internal"), max_retries=0)
        assert ok.get("/api/v1/master").json()["clusterId"] == m.cluster_id
        with pytest.raises(requests.exceptions.SSLError):
            api.UnauthSession(m.url, cert=certs.Cert(cert_pem=pem, name="someone-else.internal"), max_retries=0).get("/api/v1/master")
    finally:
        srv.stop()


def test_cli_takes_the_certificate_from_its_flags(tls_cluster, capsys):
    from determined_b200.cli import cli

    m, cpath, tmp = tls_cluster
    assert cli.main(["-m", m.url, "--cert-file", str(cpath), "master", "info"]) == 0
    assert m.cluster_id in capsys.readouterr().out
    certs.cli_cert = None
    assert cli.main(["-m", m.url, "--noverify", "user", "whoami"]) == 0
    certs.cli_cert = None
    assert cli.main(["-m", m.url, "master", "info"]) == 1               # not a terminal: no prompt, a clear failure
    assert "not trusted" in capsys.readouterr().err


def test_a_trial_runs_against_an_https_master(tls_cluster):
    """The whole chain: REST over TLS, the trial process trusts the master through DET_MASTER_CERT_FILE."""
    m, cpath, tmp = tls_cluster
    sess = api.authentication.login(m.url, "determined", "", cert=certs.Cert(cert_pem=cpath.read_text()))
    cfg = {"name": "tls", "entrypoint": "model_def:OneVarTrial", "searcher": {"name": "single", "metric": "val_loss", "max_length": {"batches": 8}},
           "hyperparameters": {"learning_rate": 0.1, "batch_size": 8, "global_batch_size": 8}, "scheduling_unit": 8, "resources": {"slots_per_trial": 1},
           "max_restarts": 0}
    files = [{"path": p.name, "type": 48, "content": base64.b64encode(p.read_bytes()).decode(), "mode": 0o644, "mtime": int(time.time())}
             for p in FIX.iterdir() if p.is_file()]
    eid = sess.post("/api/v1/experiments", json={"config": yaml.safe_dump(cfg), "modelDefinition": files, "activate": True}).json()["experiment"]["id"]
    deadline = time.time() + 180
    while True:
        st = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"]
        if st in ("STATE_COMPLETED", "STATE_ERROR"):
            break
        assert time.time() < deadline
        time.sleep(0.5)
    trial = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]
    logs = "\n".join(l["log"] for l in sess.get(f"/api/v1/trials/{trial['id']}/logs").json()["logs"][-30:])
    assert st == "STATE_COMPLETED" and trial["totalBatchesProcessed"] == 8, logs
