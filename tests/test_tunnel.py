"""TCP through the master: a command task runs a raw TCP echo server on its service port; the tunnel module forwards a local
port to it over a websocket, and the stdio mode (ssh ProxyCommand) carries one connection."""
import os
import pathlib
import socket
import subprocess
import sys
import time

import pytest

from determined_b200.cli import proxy
from determined_b200.common import api, websocket
from determined_b200.master import api as mapi
from determined_b200.master import core

ROOT = pathlib.Path(__file__).resolve().parents[1]
ECHO = ("import os, socket\ns = socket.socket(); s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)\n"
        "s.bind(('127.0.0.1', int(os.environ['DET_SERVICE_PORT']))); s.listen(4); print('listening', flush=True)\n"
        "while True:\n    c, _ = s.accept()\n    while True:\n        d = c.recv(65536)\n        if not d: break\n        c.sendall(d.upper())\n    c.close()\n")


@pytest.fixture()
def echo_task(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    sess = api.authentication.login(m.url, "determined", "")
    cmd = sess.post("/api/v1/commands", json={"config": {"entrypoint": [sys.executable, "-c", ECHO], "resources": {"slots": 0}, "proxy_ports": True}}).json()["command"]
    for _ in range(200):
        if "listening" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cmd['id']}/logs").json()["logs"]):
            break
        time.sleep(0.1)
    yield m, sess, cmd["id"]
    m.stop()
    srv.stop()


def _roundtrip(port, payload):
    with socket.create_connection(("127.0.0.1", port), timeout=10) as c:
        c.sendall(payload)
        got = b""
        while len(got) < len(payload):
            d = c.recv(65536)
            if not d:
                break
            got += d
    return got


def test_listener_forwards_connections_and_refusals_are_clear(echo_task):
    m, sess, task = echo_task
    with proxy.http_tunnel_listener(sess, [proxy.ListenerConfig(task, 0)]) as ports:
        assert _roundtrip(ports[0], b"hello through the master") == b"HELLO THROUGH THE MASTER"
        big = bytes(range(97, 123)) * 20000                     # 520 kB: many frames each way
        assert _roundtrip(ports[0], big) == big.upper()
        assert _roundtrip(ports[0], b"again") == b"AGAIN", "a second connection gets its own websocket"
    with pytest.raises(websocket.WebSocketError, match="403"):
        proxy._open(sess, task, 1)                              # a port the task did not declare
    with pytest.raises(websocket.WebSocketError, match="404"):
        proxy._open(sess, "no-such-task", None)
    with pytest.raises(websocket.WebSocketError, match="401"):
        proxy._open(api.UnauthSession(m.url), task, None)


def test_stdio_mode_carries_one_connection(echo_task, tmp_path):
    m, sess, task = echo_task
    env = dict(os.environ, PYTHONPATH=str(ROOT), DET_CONFIG_DIR=str(tmp_path / "cfg"), DET_USER="determined", DET_PASS="")
    p = subprocess.Popen([sys.executable, "-m", "determined_b200.cli.tunnel", m.url, task, "--auth"], stdin=subprocess.PIPE, stdout=subprocess.PIPE, env=env)
    try:
        p.stdin.write(b"proxy command\n")
        p.stdin.flush()
        assert p.stdout.read(14) == b"PROXY COMMAND\n"
        p.stdin.close()
        assert p.wait(timeout=20) == 0
    finally:
        if p.poll() is None:
            p.kill()
