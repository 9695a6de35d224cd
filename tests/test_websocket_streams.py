"""Websockets: the frame codec on a socket pair, and the streaming-updates channel of a live master over ``GET /stream``
(reference: ``master/internal/stream`` + ``harness/determined/common/streams``; its client tests drive a fake websocket)."""
import socket
import threading
import time

import pytest

from determined_b200.common import api, streams, websocket
from determined_b200.master import api as mapi
from determined_b200.master import core


def _pair():
    a, b = socket.socketpair()
    return websocket.WebSocket(a, is_client=True), websocket.WebSocket(b, is_client=False)


def test_frames_round_trip_in_both_directions():
    c, s = _pair()
    assert websocket.accept_key("dGhlIHNhbXBsZSBub25jZQ==") == "s3pPLMBiTxaQ9kYGzzhZRbK+xOo="      # the RFC's own example
    for payload in ("", "hello", "x" * 125, "y" * 126, "z" * 70000, "ünïcödé"):
        c.send_text(payload)
        assert s.recv(2) == (websocket.OP_TEXT, payload.encode())
        s.send_text(payload)
        assert c.recv(2) == (websocket.OP_TEXT, payload.encode())
    c.send_binary(bytes(range(256)) * 20)
    assert s.recv(2) == (websocket.OP_BINARY, bytes(range(256)) * 20)
    c.send_json({"a": [1, 2, {"b": None}]})
    assert s.recv_json(2) == {"a": [1, 2, {"b": None}]}
    # a ping is answered behind the caller's back; the next data message still arrives
    s.ping(b"are you there")
    s.send_text("after ping")
    assert c.recv(2)[1] == b"after ping"
    c.send_text("again")
    assert s.recv(2)[1] == b"again"          # (the pong the client sent in between is dropped)
    # fragmented message
    s._send_frame = websocket.WebSocket._send_frame.__get__(s)
    s.sock.sendall(bytes([0x01, 3]) + b"abc" + bytes([0x00, 3]) + b"def" + bytes([0x80, 2]) + b"gh")
    assert c.recv(2) == (websocket.OP_TEXT, b"abcdefgh")
    # close handshake
    c.close(1000, "bye")
    with pytest.raises(websocket.ConnectionClosed) as e:
        s.recv(2)
    assert e.value.code == 1000 and e.value.reason == "bye"


def test_server_rejects_unmasked_client_frames_and_lost_connections_raise():
    c, s = _pair()
    c.sock.sendall(bytes([0x81, 2]) + b"hi")                  # a client frame without a mask
    with pytest.raises(websocket.WebSocketError, match="masked"):
        s.recv(2)
    c2, s2 = _pair()
    c2.sock.shutdown(socket.SHUT_RDWR)
    with pytest.raises(websocket.ConnectionClosed):
        s2.recv(2)


@pytest.fixture()
def cluster(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"), launch=False)
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    sess = api.authentication.login(m.url, "determined", "")
    yield m, sess, srv, tmp_path
    m.stop()
    srv.stop()


def _drain_until(st, pred, timeout=10.0):
    got = []
    deadline = time.time() + timeout
    while time.time() < deadline:
        ev = st.next_nowait(timeout=0.5)
        if ev is not None:
            got.append(ev)
            if pred(ev):
                return got
    raise AssertionError(f"not seen within {timeout}s; got {got}")


def test_stream_over_a_websocket_offline_then_online_with_deletions(cluster):
    m, sess, srv, tmp = cluster
    w = sess.post("/api/v1/workspaces", json={"name": "ws-live"}).json()["workspace"]
    p1 = sess.post(f"/api/v1/workspaces/{w['id']}/projects", json={"name": "p1"}).json()["project"]
    st = streams.Stream.connect(sess)
    st.subscribe(sync_id="a", projects=streams.ProjectSpec(workspace_id=w["id"]))
    got = _drain_until(st, lambda e: e == streams.Sync("a", True))
    assert got[0] == streams.Sync("a", False) and [e.id for e in got[1:-1]] == [p1["id"]]
    # online: a new project is pushed without asking
    p2 = sess.post(f"/api/v1/workspaces/{w['id']}/projects", json={"name": "p2"}).json()["project"]
    ev = _drain_until(st, lambda e: isinstance(e, streams.wire.ProjectMsg))[-1]
    assert ev.id == p2["id"] and ev.name == "p2"
    # online: a deletion is pushed too (the long-poll form cannot do this)
    sess.delete(f"/api/v1/projects/{p1['id']}")
    ev = _drain_until(st, lambda e: isinstance(e, streams.wire.ProjectsDeleted))[-1]
    assert ev.keys == str(p1["id"])
    # changing the subscription on the same connection
    st.subscribe(sync_id="b", models=streams.ModelSpec(workspace_id=1))
    sess.post("/api/v1/models", json={"name": "m-live"})
    got = _drain_until(st, lambda e: isinstance(e, streams.wire.ModelMsg))
    assert got[-1].name == "m-live"
    if streams.Sync("b", True) not in got:        # the model can be created before the master has read the new subscription: then it is part of the sync
        got += _drain_until(st, lambda e: e == streams.Sync("b", True))
    assert streams.Sync("b", True) in got
    st.close()
    # no token, no stream
    with pytest.raises(websocket.WebSocketError, match="401"):
        websocket.connect(m.url + "/stream")


def test_stream_resumes_after_the_connection_is_lost(cluster):
    m, sess, srv, tmp = cluster
    w = sess.post("/api/v1/workspaces", json={"name": "ws-re"}).json()["workspace"]
    p1 = sess.post(f"/api/v1/workspaces/{w['id']}/projects", json={"name": "p1"}).json()["project"]
    st = streams.Stream.connect(sess)
    st.subscribe(sync_id="a", projects=streams.ProjectSpec(workspace_id=w["id"]))
    _drain_until(st, lambda e: e == streams.Sync("a", True))
    # cut the wire under the client, change things while it is away
    st._ws._ws.sock.shutdown(socket.SHUT_RDWR)
    p2 = sess.post(f"/api/v1/workspaces/{w['id']}/projects", json={"name": "p2"}).json()["project"]
    sess.delete(f"/api/v1/projects/{p1['id']}")
    got = _drain_until(st, lambda e: e == streams.Sync("a", True), timeout=15)
    assert st._ws.reconnects == 1
    kinds = {type(e).__name__: e for e in got}
    # the resumed subscription says what the client already knows: p1 comes back as deleted, p2 as new, nothing else
    assert kinds["ProjectsDeleted"].keys == str(p1["id"]) and kinds["ProjectMsg"].id == p2["id"]
    assert sum(isinstance(e, streams.wire.ProjectMsg) for e in got) == 1
    st.close()


def test_stream_over_tls(tmp_path, monkeypatch):
    from determined_b200.common.api import certs
    from tests.test_tls import make_cert

    monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "config"))
    monkeypatch.setattr(certs, "cli_cert", None)
    cpath, kpath = make_cert(tmp_path)
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"), launch=False)
    srv = mapi.APIServer(m, "127.0.0.1", 0, tls_cert=str(cpath), tls_key=str(kpath)).start()
    m.start()
    try:
        sess = api.authentication.login(m.url, "determined", "", cert=certs.Cert(cert_pem=cpath.read_text()))
        st = streams.Stream.connect(sess)
        st.subscribe(sync_id=1, projects=streams.ProjectSpec(workspace_id=1))
        got = _drain_until(st, lambda e: e == streams.Sync(1, True))
        assert got[0] == streams.Sync(1, False)
        st.close()
    finally:
        m.stop()
        srv.stop()


def test_proxy_tunnels_a_websocket_to_a_task_service(tmp_path):
    """A command that serves a websocket echo on its service port is reachable through ``/proxy/<task>/``."""
    import pathlib
    import sys

    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        sess = api.authentication.login(m.url, "determined", "")
        pathlib.Path(m.work_dir).mkdir(parents=True, exist_ok=True)
        serve = (
            "import http.server, os\n"
            "from determined_b200.common import websocket\n"
            "class H(http.server.BaseHTTPRequestHandler):\n"
            "    def log_message(self, *a): pass\n"
            "    def do_GET(self):\n"
            "        if not websocket.is_upgrade(self.headers):\n"
            "            self.send_response(200); self.send_header('Content-Length', '2'); self.end_headers(); self.wfile.write(b'ok'); return\n"
            "        ws = websocket.server_handshake(self)\n"
            "        ws.send_text('hello from ' + self.path + ' via ' + self.headers.get('X-Forwarded-Prefix', '?'))\n"
            "        try:\n"
            "            while True:\n"
            "                op, data = ws.recv()\n"
            "                ws.send_text('echo:' + data.decode())\n"
            "        except websocket.WebSocketError:\n"
            "            pass\n"
            "http.server.ThreadingHTTPServer(('127.0.0.1', int(os.environ['DET_SERVICE_PORT'])), H).serve_forever()\n")
        cmd = sess.post("/api/v1/commands", json={"config": {"entrypoint": [sys.executable, "-c", serve], "proxy_ports": True}}).json()["command"]
        deadline = time.time() + 60
        while True:
            c = sess.get(f"/api/v1/commands/{cmd['id']}").json()["command"]
            if c["state"] == "STATE_RUNNING" and c.get("serviceAddress"):
                try:
                    if sess.get(c["serviceAddress"]).text == "ok":
                        break
                except api.errors.APIException:
                    pass
            assert time.time() < deadline, c
            time.sleep(0.2)
        ws = websocket.connect(m.url + c["serviceAddress"] + "kernel/1?x=2", headers={"Authorization": f"Bearer {sess.token}"})
        assert ws.recv(10)[1].decode() == f"hello from /kernel/1?x=2 via /proxy/{cmd['id']}"
        for text in ("one", "two" * 5000):
            ws.send_text(text)
            assert ws.recv(10)[1].decode() == "echo:" + text
        ws.close()
        with pytest.raises(websocket.WebSocketError, match="401"):
            websocket.connect(m.url + c["serviceAddress"] + "kernel/1")
        sess.post(f"/api/v1/commands/{cmd['id']}/kill")
    finally:
        m.stop()
        srv.stop()
