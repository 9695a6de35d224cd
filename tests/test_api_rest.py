"""The routes of ``master/api_rest.py``, the telemetry recorder and OIDC single sign-on, against an in-process master.
One short experiment runs so that metric / file / history routes have something to answer with."""
import base64
import http.server
import json
import threading
import time
import urllib.parse

import pytest
import requests

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import config as mconfig
from determined_b200.master import core

from tests.test_master_e2e import BASE, create, dump_logs, wait_state


@pytest.fixture()
def cluster(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=2, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}})
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    yield m, api.authentication.login(m.url, "admin", ""), tmp_path
    m.stop()
    srv.stop()


def test_reference_rpcs_all_have_a_route():
    """Every rpc of the reference's api.proto that carries an HTTP rule has a route with the same method and path shape."""
    import re

    ref = open("/root/reference/proto/src/determined/api/v1/api.proto").read()
    mine = {(meth, re.sub(r"\(\?P<\w+>\[\^/\]\+\)", "{}", rx.pattern[1:-3])) for meth, rx, _, _ in mapi.ROUTES}
    missing = []
    for m in re.finditer(r"rpc (\w+)\((?:stream )?\w+\)\s*returns\s*\((?:stream )?\w+\)\s*\{(.*?)\n  \}", ref, re.S):
        h = re.search(r"(get|post|put|patch|delete):\s*\"([^\"]+)\"", m.group(2))
        if h and (h.group(1).upper(), re.sub(r"\{[^}]+\}", "{}", h.group(2))) not in mine:
            missing.append(m.group(1))
    assert missing == []


def test_bulk_actions_labels_files_history_and_search(cluster):
    m, sess, tmp = cluster
    a = create(sess, dict(BASE, name="alpha", labels=["x"], hyperparameters=dict(BASE["hyperparameters"], learning_rate=0.1)))
    b = create(sess, dict(BASE, name="beta", labels=["x", "y"]), activate=False)
    assert wait_state(sess, a, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, a)

    assert sess.get("/api/v1/experiment/labels").json()["labels"] == ["x", "y"]
    assert sess.put(f"/api/v1/experiments/{b}/labels/z").json()["labels"] == ["x", "y", "z"]
    assert sess.delete(f"/api/v1/experiments/{b}/labels/x").json()["labels"] == ["y", "z"]

    r = sess.post("/api/v1/experiments/archive", json={"experimentIds": [a, b, 999]}).json()["results"]
    assert [x["error"] == "" for x in r] == [True, False, False], "only terminal experiments archive; unknown ids report an error"
    assert sess.post("/api/v1/experiments/unarchive", json={"filters": {"name": "alph"}}).json()["results"] == [{"id": a, "error": ""}]
    r = sess.post("/api/v1/experiments/kill", json={"experimentIds": [b]}).json()["results"]
    assert r == [{"id": b, "error": ""}] and wait_state(sess, b, {"STATE_CANCELED"}) == "STATE_CANCELED"

    tree = sess.get(f"/api/v1/experiments/{a}/file_tree").json()["files"]
    assert any(f["name"] == "model_def.py" and not f["isDir"] and f["contentLength"] > 0 for f in tree)
    src = base64.b64decode(sess.post(f"/api/v1/experiments/{a}/file", json={"path": "model_def.py"}).json()["file"]).decode()
    assert "OneVarTrial" in src
    with pytest.raises(api.errors.APIException):
        sess.post(f"/api/v1/experiments/{a}/file", json={"path": "nope.py"})

    hist = sess.get(f"/api/v1/experiments/{a}/validation-history").json()["validationHistory"]
    vals = [h["searcherMetric"] for h in hist]
    assert len(vals) >= 2 and vals == sorted(vals, reverse=True), "each entry improves on the one before"
    tid = sess.get(f"/api/v1/experiments/{a}/trials").json()["trials"][0]["id"]
    assert sess.get(f"/api/v1/experiments/{a}/metrics-stream/batches", params={"metricName": "val_loss"}).json()[0]["result"]["batches"] == [8, 16, 24]
    tm = sess.get("/api/v1/trials/metrics/training_metrics", params={"trialIds": [tid]}).json()[0]["result"]["metrics"]
    assert tm and all(x["group"] == "training" for x in tm)
    wl = sess.get(f"/api/v1/trials/{tid}/workloads").json()["workloads"]
    assert {next(iter(w)) for w in wl} == {"training", "validation", "checkpoint"}
    ck = sess.get(f"/api/v1/trials/{tid}/checkpoints").json()["checkpoints"][0]["uuid"]
    cm = sess.get(f"/api/v1/checkpoints/{ck}/metrics", params={"group": "validation"}).json()["metrics"][0]
    assert cm["checkpointUuid"] == ck and cm["metricReports"]
    assert sess.get(f"/api/v1/trials/{tid}/logs/fields").json()[0]["result"]["stdtypes"]
    assert sess.get(f"/api/v1/tasks/trial-{tid}").json()["task"]["allocations"]
    assert sess.get("/api/v1/tasks/count").json() == {"commands": 0, "notebooks": 0, "shells": 0, "tensorboards": 0}
    assert sess.get("/api/v1/trials/time-series", params={"trialIds": [tid]}).json()["trials"][0]["trial"]["id"] == tid

    # profiler samples in the batched form, then read back as series
    sess.post("/api/v1/trials/profiler/metrics", json={"batches": [{"labels": {"trialId": tid, "name": "gpu_util", "agentId": "box"}, "values": [10, 30],
                                                                   "batches": [1, 2], "timestamps": ["2026-01-01T00:00:00Z", "2026-01-01T00:00:01Z"]}]})
    labels = sess.get(f"/api/v1/trials/{tid}/profiler/available_series").json()[0]["result"]["labels"]
    assert {"trialId": tid, "name": "gpu_util", "agentId": "box", "metricType": "PROFILER_METRIC_TYPE_SYSTEM", "group": "profiler"} in labels
    pm = sess.get(f"/api/v1/trials/{tid}/profiler/metrics", params={"labels.name": "gpu_util"}).json()[0]["result"]["batch"]
    assert pm["values"] == [10.0, 30.0] and pm["batches"] == [1, 2]

    # the experiment table: filter language, sort, best trial; runs
    flt = {"filterGroup": {"conjunction": "and", "children": [{"kind": "field", "columnName": "tags", "operator": "contains", "value": "x"},
                                                              {"kind": "field", "columnName": "hp.learning_rate", "operator": "<=", "value": 0.1}]}}
    found = sess.get("/api/v1/experiments-search", params={"filter": json.dumps(flt), "sort": "id=desc"}).json()["experiments"]
    assert [e["experiment"]["id"] for e in found] == [a] and found[0]["bestTrial"]["id"] == tid
    runs = sess.get("/api/v1/runs", params={"projectId": 1}).json()["runs"]
    assert runs[0]["id"] == tid and runs[0]["experiment"]["name"] == "alpha" and {r["experiment"]["name"] for r in runs} == {"alpha", "beta"}
    cols = {c["column"]: c for c in sess.get("/api/v1/projects/1/columns").json()["columns"]}
    assert cols["hp.learning_rate"]["type"] == "COLUMN_TYPE_NUMBER" and "validation.val_loss" in cols and "name" in cols
    rng = {r["metricsName"]: r for r in sess.get("/api/v1/projects/1/experiments/metric-ranges").json()["ranges"]}
    assert rng["validation.val_loss"]["min"] <= rng["validation.val_loss"]["max"]

    # moves
    wid = sess.post("/api/v1/workspaces", json={"name": "w2"}).json()["workspace"]["id"]
    pid = sess.post(f"/api/v1/workspaces/{wid}/projects", json={"name": "p2"}).json()["project"]["id"]
    assert sess.post("/api/v1/runs/move", json={"runIds": [tid], "destinationProjectId": pid}).json()["results"] == [{"id": tid, "error": ""}]
    assert sess.get(f"/api/v1/experiments/{a}").json()["experiment"]["projectId"] == pid
    assert sess.post("/api/v1/experiments/move", json={"experimentIds": [a], "destinationProjectId": 1}).json()["results"][0]["error"] == ""
    sess.post(f"/api/v1/projects/{pid}/move", json={"destinationWorkspaceId": 1})
    assert sess.get(f"/api/v1/projects/{pid}").json()["project"]["workspaceId"] == 1
    assert sess.delete("/api/v1/experiments/delete", json={"experimentIds": [a, b]}).json()["results"] == [{"id": a, "error": ""}, {"id": b, "error": ""}]
    assert sess.get("/api/v1/experiments").json()["experiments"] == []


def test_users_settings_pins_notes_roles_and_unmanaged_ids(cluster):
    m, sess, tmp = cluster
    assert sess.get("/api/v1/auth/user").json()["user"]["username"] == "admin"
    assert sess.get("/api/v1/users/determined/by-username").json()["user"]["username"] == "determined"
    sess.post("/api/v1/users/setting", json={"settings": [{"key": "theme", "storagePath": "ui", "value": "dark"}, {"key": "cols", "storagePath": "table", "value": "[1]"}]})
    got = sess.get("/api/v1/users/setting").json()["settings"]
    assert {"key": "theme", "storagePath": "ui", "value": "dark"} in got and len(got) == 2
    sess.post("/api/v1/users/setting/reset")
    assert sess.get("/api/v1/users/setting").json()["settings"] == []
    u1 = sess.post("/api/v1/users", json={"user": {"username": "u1"}}).json()["user"]["id"]
    u2 = sess.post("/api/v1/users", json={"user": {"username": "u2"}}).json()["user"]["id"]
    assert [r["error"] for r in sess.patch("/api/v1/users", json={"userIds": [u1, u2, 12345], "activate": False}).json()["results"]] == ["", "", "not found"]
    assert not sess.get(f"/api/v1/users/{u1}").json()["user"]["active"]
    g = sess.post("/api/v1/groups", json={"name": "g"}).json()["group"]["groupId"]
    sess.patch("/api/v1/users/assignments", json={"userIds": [u1, u2], "addGroups": [g]})
    assert {u["id"] for u in sess.get(f"/api/v1/groups/{g}").json()["group"]["users"]} == {u1, u2}
    sess.patch("/api/v1/users/assignments", json={"userIds": [u1], "removeGroups": [g]})
    assert {u["id"] for u in sess.get(f"/api/v1/groups/{g}").json()["group"]["users"]} == {u2}

    wid = sess.post("/api/v1/workspaces", json={"name": "pinme"}).json()["workspace"]["id"]
    sess.post(f"/api/v1/workspaces/{wid}/pin")
    sess.post(f"/api/v1/workspaces/{wid}/unpin")
    pid = sess.post(f"/api/v1/workspaces/{wid}/projects", json={"name": "proj"}).json()["project"]["id"]
    assert sess.post(f"/api/v1/projects/{pid}/notes", json={"name": "n1", "contents": "hello"}).json()["notes"] == [{"name": "n1", "contents": "hello"}]
    assert sess.put(f"/api/v1/projects/{pid}/notes", json={"notes": [{"name": "only", "contents": "c"}]}).json()["notes"] == [{"name": "only", "contents": "c"}]
    sess.post("/api/v1/users/activity", json={"activityType": "ACTIVITY_TYPE_GET", "entityType": "ENTITY_TYPE_PROJECT", "entityId": pid})
    assert [p["id"] for p in sess.get("/api/v1/user/projects/activity").json()["projects"]] == [pid]

    roles = sess.get("/api/v1/roles").json()["roles"]
    byid = sess.post("/api/v1/roles/search/by-ids", json={"roleIds": [roles[0]["roleId"]]}).json()["roles"]
    assert byid[0]["role"]["name"] == roles[0]["name"] and byid[0]["role"]["permissions"]
    everywhere = {r["name"] for r in sess.post("/api/v1/roles/search/by-assignability", json={}).json()["roles"]}
    scoped = {r["name"] for r in sess.post("/api/v1/roles/search/by-assignability", json={"workspaceId": wid}).json()["roles"]}
    assert scoped < everywhere and "ClusterAdmin" not in scoped

    # unmanaged experiments / trials by external id are idempotent
    cfg = {"name": "ext", "entrypoint": "x", "searcher": {"name": "single", "metric": "loss", "max_length": 1}, "hyperparameters": {}}
    e1 = sess.put("/api/v1/experiments/by-external-id/run-42", json={"config": json.dumps(cfg)}).json()["experiment"]
    e2 = sess.put("/api/v1/experiments/by-external-id/run-42", json={"config": json.dumps(cfg)}).json()["experiment"]
    assert e1["id"] == e2["id"] and e1["unmanaged"]
    t1 = sess.put("/api/v1/trials", json={"createTrialRequest": {"experimentId": e1["id"], "hparams": {"lr": 1}}, "externalTrialId": "t-a"}).json()["trial"]
    t2 = sess.put("/api/v1/trials", json={"createTrialRequest": {"experimentId": e1["id"], "hparams": {"lr": 1}}, "externalTrialId": "t-a"}).json()["trial"]
    assert t1["id"] == t2["id"] and sess.get("/api/v1/trials/by-external-id/run-42/t-a").json()["trial"]["id"] == t1["id"]
    sess.post(f"/api/v1/trials/{t1['id']}/training_metrics", json={"trainingMetrics": {"stepsCompleted": 5, "metrics": {"avgMetrics": {"loss": 1.5}}}})
    sess.post(f"/api/v1/trials/{t1['id']}/validation_metrics", json={"validationMetrics": {"stepsCompleted": 5, "metrics": {"avgMetrics": {"loss": 1.25}}}})
    groups = [x["group"] for x in sess.get(f"/api/v1/trials/{t1['id']}/metrics").json()["metrics"]]
    assert groups == ["training", "validation"]
    s1 = sess.post(f"/api/v1/trials/{t1['id']}/start", json={"resume": True}).json()
    s2 = sess.post(f"/api/v1/trials/{t1['id']}/start", json={"resume": True}).json()
    assert (s1["trialRunId"], s2["trialRunId"]) == (1, 2) and s2["stepsCompleted"] == 5
    assert sess.get("/api/v1/job-queues/stats").json()["results"][0]["stats"] == {"queuedCount": 0, "scheduledCount": 0, "preemptibleCount": 0}
    sess.post("/api/v1/allocations/a1/acceleratorData", json={"acceleratorData": {"taskId": "tk", "nodeName": "box", "acceleratorType": "cuda", "acceleratorUuids": ["GPU-1"]}})
    assert sess.get("/api/v1/tasks/tk/acceleratorData").json()["acceleratorData"][0]["acceleratorUuids"] == ["GPU-1"]
    assert sess.post("/api/v1/allocations/a1/notify_container_running", json={"rank": 0, "numPeers": 2}).json()["data"]["allRunning"] is False
    assert sess.post("/api/v1/allocations/a1/notify_container_running", json={"rank": 1, "numPeers": 2}).json()["data"]["allRunning"] is True


class _Collector(http.server.BaseHTTPRequestHandler):
    got = []

    def do_POST(self):  # noqa: N802
        body = self.rfile.read(int(self.headers.get("Content-Length") or 0))
        type(self).got.append((self.headers.get("Authorization"), json.loads(body)))
        self.send_response(200)
        self.send_header("Content-Length", "0")
        self.end_headers()

    def log_message(self, *a):
        pass


def test_telemetry_is_off_by_default_and_posts_events_when_configured(cluster, tmp_path):
    m, sess, tmp = cluster
    assert sess.get("/api/v1/master/telemetry").json() == {"enabled": False, "segmentKey": ""}
    assert sess.get("/api/v1/master").json()["telemetryEnabled"] is False
    _Collector.got = []
    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), _Collector)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        cfg = mconfig.load(None, {}, {"telemetry.enabled": True, "telemetry.segment_url": f"http://127.0.0.1:{srv.server_address[1]}/v1/track", "telemetry.segment_master_key": "k3y"})
        m.apply_config(cfg)
        eid = create(sess, dict(BASE), activate=False)
        sess.post(f"/api/v1/experiments/{eid}/kill")
        wait_state(sess, eid, {"STATE_CANCELED"})
        m.telemetry.flush()
        for _ in range(50):
            if len(_Collector.got) >= 3:
                break
            time.sleep(0.1)
        events = [b["event"] for _, b in _Collector.got]
        assert events[:2] == ["master_started", "experiment_created"] and "experiment_state_changed" in events
        auth, first = _Collector.got[0]
        assert auth == "Basic " + base64.b64encode(b"k3y:").decode() and first["userId"] == m.cluster_id and first["properties"]["num_pools"] == 1
        assert sess.get("/api/v1/master/telemetry").json() == {"enabled": True, "segmentKey": "k3y"}
        assert [e["event"] for e in sess.get("/api/v1/master/telemetry", params={"events": 1}).json()["events"]][:2] == ["master_started", "experiment_created"]
        assert mconfig.load(None, {}, {"telemetry.segment_master_key": "k3y"}).printable()["telemetry"]["segment_master_key"] != "k3y"
    finally:
        srv.shutdown()
        srv.server_close()


class _IdP(http.server.BaseHTTPRequestHandler):
    """A minimal OpenID provider: /authorize approves at once, /token trades the code, /userinfo names the user."""
    user = {"email": "alice@example.com", "name": "Alice"}
    codes, tokens, seen = {}, {}, []

    def _json(self, obj, status=200):
        data = json.dumps(obj).encode()
        self.send_response(status)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        self.wfile.write(data)

    def do_GET(self):  # noqa: N802
        u = urllib.parse.urlparse(self.path)
        q = urllib.parse.parse_qs(u.query)
        if u.path == "/authorize":
            type(self).seen.append(q)
            code = f"code-{len(self.codes)}"
            self.codes[code] = q["redirect_uri"][0]
            self.send_response(302)
            self.send_header("Location", q["redirect_uri"][0] + "?" + urllib.parse.urlencode({"code": code, "state": q["state"][0]}))
            self.send_header("Content-Length", "0")
            self.end_headers()
        elif u.path == "/userinfo":
            tok = (self.headers.get("Authorization") or "").replace("Bearer ", "")
            self._json(self.user if tok in self.tokens else {"error": "invalid_token"}, 200 if tok in self.tokens else 401)
        else:
            self._json({}, 404)

    def do_POST(self):  # noqa: N802
        form = urllib.parse.parse_qs(self.rfile.read(int(self.headers.get("Content-Length") or 0)).decode())
        basic = base64.b64decode((self.headers.get("Authorization") or "Basic Og==").split()[1]).decode()
        code = form.get("code", [""])[0]
        if basic != "det-client:s3cret" or self.codes.pop(code, None) != form.get("redirect_uri", [""])[0]:
            return self._json({"error": "invalid_grant"}, 400)
        tok = f"at-{len(self.tokens)}"
        self.tokens[tok] = True
        self._json({"access_token": tok, "token_type": "Bearer"})

    def log_message(self, *a):
        pass


def test_oidc_single_sign_on_browser_flow_and_cli(cluster, tmp_path, monkeypatch):
    m, sess, tmp = cluster
    assert sess.get("/api/v1/master").json()["ssoProviders"] == []
    assert requests.get(m.url + "/oidc/sso", allow_redirects=False).status_code == 404
    idp = http.server.ThreadingHTTPServer(("127.0.0.1", 0), _IdP)
    threading.Thread(target=idp.serve_forever, daemon=True).start()
    base = f"http://127.0.0.1:{idp.server_address[1]}"
    try:
        with pytest.raises(ValueError):
            m.apply_config(mconfig.load(None, {}, {"oidc.enabled": True}))
        flags = {"oidc.enabled": True, "oidc.provider": "Okta", "oidc.client_id": "det-client", "oidc.client_secret": "s3cret", "oidc.idp_sso_url": base + "/authorize",
                 "oidc.token_url": base + "/token", "oidc.userinfo_url": base + "/userinfo", "oidc.idp_recipient_url": m.url}
        m.apply_config(mconfig.load(None, {}, flags))
        assert sess.get("/api/v1/master").json()["ssoProviders"] == [{"name": "Okta", "ssoUrl": "/oidc/sso", "type": "OIDC", "alwaysRedirect": False}]

        # an unknown user is refused unless users are provisioned on first sign-in
        r = requests.get(m.url + "/oidc/sso", params={"relayState": "/det/login"}, allow_redirects=True)
        assert r.status_code == 401 and "does not exist" in r.json()["message"]
        assert _IdP.seen[-1]["client_id"] == ["det-client"] and _IdP.seen[-1]["redirect_uri"] == [m.url + "/oidc/callback"]
        # a remote user created by an admin signs in; the browser ends on the relay path with a session token
        sess.post("/api/v1/users", json={"user": {"username": "alice@example.com", "remote": True}})
        r = requests.get(m.url + "/oidc/sso", params={"relayState": "/det/login"}, allow_redirects=False)
        r = requests.get(r.headers["Location"], allow_redirects=False)            # the provider approves
        r = requests.get(r.headers["Location"], allow_redirects=False)            # back at the master
        assert r.status_code == 302 and r.headers["Location"].startswith("/det/login?jwt=")
        tok = urllib.parse.parse_qs(urllib.parse.urlparse(r.headers["Location"]).query)["jwt"][0]
        assert api.Session(m.url, "alice@example.com", tok).get("/api/v1/me").json()["user"]["remote"] is True
        # forged state, foreign relay targets and replayed codes are refused
        assert requests.get(m.url + "/oidc/callback", params={"code": "x", "state": "abc.def"}).status_code == 400
        assert requests.get(m.url + "/oidc/sso", params={"relayState": "https://evil.example/"}, allow_redirects=False).status_code == 400
        # auto-provisioning creates the user as a remote one
        m.apply_config(mconfig.load(None, {}, dict(flags, **{"oidc.auto_provision_users": True})))
        _IdP.user = {"email": "bob@example.com", "name": "Bob"}
        r = requests.get(m.url + "/oidc/sso", allow_redirects=False)
        r = requests.get(r.headers["Location"], allow_redirects=False)
        r = requests.get(r.headers["Location"], allow_redirects=False)
        assert r.status_code == 302 and sess.get("/api/v1/users/bob@example.com/by-username").json()["user"]["remote"] is True

        # the CLI: a loopback listener is the relay target; the "browser" here is requests following the redirects
        from determined_b200.cli import cli

        monkeypatch.setenv("DET_MASTER", m.url)
        monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "config"))
        from determined_b200.cli import more

        orig = more.sso_login
        monkeypatch.setattr(more, "sso_login", lambda args: orig(args, open_url=lambda u: threading.Thread(target=requests.get, args=(u,), daemon=True).start()))
        assert cli.main(["auth", "login", "--provider", "okta"]) == 0
        store = api.authentication.TokenStore(m.url)
        assert store.get_active_user() == "bob@example.com" and store.get_token("bob@example.com")
        with pytest.raises(SystemExit):
            cli.main(["auth", "login", "--provider", "nope"])
    finally:
        idp.shutdown()
        idp.server_close()


def test_experiment_config_names_its_workspace_and_project(cluster):
    m, sess, tmp = cluster
    wid = sess.post("/api/v1/workspaces", json={"name": "by-name"}).json()["workspace"]["id"]
    pid = sess.post(f"/api/v1/workspaces/{wid}/projects", json={"name": "proj"}).json()["project"]["id"]
    eid = create(sess, dict(BASE, workspace="by-name", project="proj"), activate=False)
    assert sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["projectId"] == pid
    import yaml

    for cfg, status in ((dict(BASE, workspace="by-name"), 400), (dict(BASE, workspace="by-name", project="nope"), 404)):
        with pytest.raises(api.errors.APIException) as e:
            sess.post("/api/v1/experiments", json={"config": yaml.safe_dump(cfg), "modelDefinition": [], "activate": False})
        assert e.value.status_code == status
    # an explicit projectId wins over the names in the configuration
    r = sess.post("/api/v1/experiments", json={"config": yaml.safe_dump(dict(BASE, workspace="by-name", project="proj")), "modelDefinition": [], "activate": False, "projectId": 1})
    assert r.json()["experiment"]["projectId"] == 1
