"""Master end-to-end on CPU: REST create -> scheduler -> launched process -> harness -> searcher
ops -> checkpoints/metrics -> completion; restarts; pause/activate (preemption); ASHA; GC.
(test strategy: SURVEY §4 "E2E"; devcluster replaced by an in-process master on fake CPU slots.)"""
import base64
import json
import os
import pathlib
import time

import pytest
import yaml

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import core

FIX = pathlib.Path(__file__).parent / "fixtures" / "onevar_exp"


@pytest.fixture()
def cluster(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=4, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}},
                    scheduler_type="priority")
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    sess = api.authentication.login(m.url, "determined", "")
    yield m, sess, tmp_path
    m.stop()
    srv.stop()


def files():
    out = []
    for p in FIX.iterdir():
        if p.is_file():
            out.append({"path": p.name, "type": 48, "content": base64.b64encode(p.read_bytes()).decode(), "mode": 0o644, "mtime": int(time.time())})
    return out


def create(sess, cfg, activate=True):
    r = sess.post("/api/v1/experiments", json={"config": yaml.safe_dump(cfg), "modelDefinition": files(), "activate": activate}).json()
    return r["experiment"]["id"]


def wait_state(sess, eid, states, timeout=180):
    t0 = time.time()
    while time.time() - t0 < timeout:
        st = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"]
        if st in states:
            return st
        time.sleep(0.5)
    raise AssertionError(f"experiment {eid} stuck in {st}; logs:\n" + dump_logs(sess, eid))


def dump_logs(sess, eid):
    out = []
    for t in sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]:
        for l in sess.get(f"/api/v1/trials/{t['id']}/logs").json()["logs"][-30:]:
            out.append(f"[t{t['id']}] {l['log'].rstrip()}")
    return "\n".join(out)


BASE = {"name": "onevar", "entrypoint": "model_def:OneVarTrial",
        "searcher": {"name": "single", "metric": "val_loss", "max_length": {"batches": 24}},
        "hyperparameters": {"learning_rate": 0.1, "batch_size": 8, "global_batch_size": 8},
        "scheduling_unit": 8, "min_validation_period": {"batches": 8}, "max_restarts": 2,
        "resources": {"slots_per_trial": 1}}


def test_single_trial_runs_to_completion(cluster):
    m, sess, tmp = cluster
    eid = create(sess, BASE)
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    trials = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
    assert len(trials) == 1 and trials[0]["state"] == "STATE_COMPLETED"
    tid = trials[0]["id"]
    assert trials[0]["totalBatchesProcessed"] == 24
    mets = sess.get(f"/api/v1/trials/{tid}/metrics").json()["metrics"]
    train = [x for x in mets if x["group"] == "training"]
    val = [x for x in mets if x["group"] == "validation"]
    assert [x["totalBatches"] for x in train] == [8, 16, 24]
    assert [x["totalBatches"] for x in val] == [8, 16, 24]
    assert val[-1]["metrics"]["avgMetrics"]["val_loss"] < val[0]["metrics"]["avgMetrics"]["val_loss"]
    cks = sess.get(f"/api/v1/trials/{tid}/checkpoints").json()["checkpoints"]
    assert cks and cks[-1]["metadata"]["steps_completed"] == 24
    ck = cks[-1]
    d = tmp / "ckpt" / ck["uuid"]
    assert {"state_dict.pth", "metadata.json", "load_data.json", "trial_state.pkl", "code"} <= {p.name for p in d.iterdir()}
    assert set(ck["resources"]) >= {"state_dict.pth", "metadata.json"}
    best = sess.get(f"/api/v1/experiments/{eid}/searcher/best_searcher_validation_metric").json()["metric"]
    assert abs(best - val[-1]["metrics"]["avgMetrics"]["val_loss"]) < 1e-9
    logs = sess.get(f"/api/v1/trials/{tid}/logs").json()["logs"]
    assert any("trained" in l["log"] for l in logs)


def test_crash_restarts_from_checkpoint(cluster):
    m, sess, tmp = cluster
    cfg = json.loads(json.dumps(BASE))
    cfg["hyperparameters"]["crash_on_batch"] = 12
    cfg["min_checkpoint_period"] = {"batches": 8}
    eid = create(sess, cfg)
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    t = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]
    assert t["restarts"] == 1 and t["totalBatchesProcessed"] == 24
    logs = "".join(l["log"] for l in sess.get(f"/api/v1/trials/{t['id']}/logs").json()["logs"])
    assert "Restoring trial from checkpoint" in logs
    for f in pathlib.Path("/tmp").glob(".detb200_crashed_*"):
        f.unlink()


def test_pause_preempts_and_activate_resumes(cluster):
    m, sess, tmp = cluster
    cfg = json.loads(json.dumps(BASE))
    cfg["searcher"]["max_length"] = {"batches": 4000}
    cfg["scheduling_unit"] = 50
    cfg["min_validation_period"] = {"batches": 0}
    eid = create(sess, cfg)
    t0 = time.time()
    while time.time() - t0 < 90:
        tr = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
        if tr and tr[0]["totalBatchesProcessed"] >= 50:
            break
        time.sleep(0.3)
    sess.post(f"/api/v1/experiments/{eid}/pause")
    t0 = time.time()
    while time.time() - t0 < 60 and m.allocations:
        time.sleep(0.3)
    assert not m.allocations, "allocation did not exit after preemption"
    tr = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]
    assert tr["latestCheckpoint"], "preempted trial must leave a checkpoint"
    done_before = tr["totalBatchesProcessed"]
    assert 0 < done_before < 4000
    assert sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"] == "STATE_PAUSED"
    sess.post(f"/api/v1/experiments/{eid}/activate")
    sess.post(f"/api/v1/experiments/{eid}/kill")
    wait_state(sess, eid, {"STATE_CANCELED"}, 60)


def test_adaptive_asha_packs_trials_and_finishes(cluster):
    m, sess, tmp = cluster
    cfg = json.loads(json.dumps(BASE))
    cfg["searcher"] = {"name": "adaptive_asha", "metric": "val_loss", "max_length": {"batches": 32}, "max_trials": 6,
                       "max_concurrent_trials": 4, "divisor": 2, "max_rungs": 3}
    cfg["hyperparameters"]["learning_rate"] = {"type": "log", "minval": -3, "maxval": -0.5, "base": 10}
    cfg["min_validation_period"] = {"batches": 0}
    cfg["scheduling_unit"] = 16
    eid = create(sess, cfg)
    peak = 0
    t0 = time.time()
    while time.time() - t0 < 400:
        peak = max(peak, len(m.allocations))
        st = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"]
        if st in ("STATE_COMPLETED", "STATE_ERROR", "STATE_CANCELED"):
            break
        time.sleep(0.3)
    assert st == "STATE_COMPLETED", dump_logs(sess, eid)
    trials = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
    assert len(trials) == 6 and all(t["state"] == "STATE_COMPLETED" for t in trials)
    assert peak >= 2, "trials should run concurrently on the 4 slots"
    lengths = sorted(t["totalBatchesProcessed"] for t in trials)
    assert lengths[0] < lengths[-1], f"ASHA should stop some trials early: {lengths}"
    # checkpoint GC (save_trial_latest=1, save_trial_best=1, save_experiment_best=0)
    time.sleep(1.5)
    for t in trials:
        cks = sess.get(f"/api/v1/trials/{t['id']}/checkpoints").json()["checkpoints"]
        alive = [c for c in cks if c["state"] == "STATE_COMPLETED"]
        assert 1 <= len(alive) <= 2


def test_adaptive_asha_survives_chaos_with_max_restarts(cluster):
    """The reference's ``adaptive_chaos`` experiment (``e2e_tests/tests/fixtures/no_op/adaptive_chaos.yaml``: adaptive ASHA,
    ``max_restarts`` high, every batch may kill the trial process) scaled to the test box: scheduler packing, promotions,
    crash -> restart from the latest checkpoint and searcher bookkeeping all at once.  Every trial must end closed."""
    m, sess, tmp = cluster
    cfg = json.loads(json.dumps(BASE))
    cfg["searcher"] = {"name": "adaptive_asha", "metric": "val_loss", "max_length": {"batches": 48}, "max_trials": 8,
                       "max_concurrent_trials": 4, "divisor": 2, "max_rungs": 3}
    cfg["hyperparameters"]["learning_rate"] = {"type": "log", "minval": -3, "maxval": -0.5, "base": 10}
    cfg["hyperparameters"]["chaos_probability"] = 0.02
    cfg["min_validation_period"] = {"batches": 0}
    cfg["min_checkpoint_period"] = {"batches": 8}
    cfg["scheduling_unit"] = 8
    cfg["max_restarts"] = 1000
    eid = create(sess, cfg)
    st = wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR", "STATE_CANCELED"}, timeout=900)
    assert st == "STATE_COMPLETED", dump_logs(sess, eid)
    trials = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
    assert len(trials) == 8 and all(t["state"] == "STATE_COMPLETED" for t in trials), [(t["id"], t["state"]) for t in trials]
    restarts = sum(t["restarts"] for t in trials)
    assert restarts >= 1, "with p=0.02 per batch over a few hundred batches at least one trial must have been restarted"
    # Restarted or not, every trial ends exactly on a rung boundary of its bracket: its last validation and its last
    # checkpoint are both at its final length.  Rung lengths are cumulative in ASHA (reference
    # master/pkg/searcher/asha.go:56-78: unitsNeeded += max_length / divisor^(rungs-i-1)): 12, 36, 84 in the 3-rung bracket,
    # 24, 72 in the 2-rung one, 48 in the 1-rung one.
    boundaries = {12, 36, 84, 24, 72, 48}
    for t in trials:
        total = t["totalBatchesProcessed"]
        vals = [r["steps_completed"] for r in m.db.all("SELECT steps_completed FROM metrics WHERE trial_id=? AND grp='validation' ORDER BY id", (t["id"],))]
        cks = [c["metadata"].get("steps_completed") for c in sess.get(f"/api/v1/trials/{t['id']}/checkpoints").json()["checkpoints"]]
        assert total in boundaries and vals and vals[-1] == total and max(cks) == total, (t["id"], total, vals, cks)
        assert set(vals) <= boundaries, (t["id"], vals)
    assert max(t["totalBatchesProcessed"] for t in trials) in (84, 72, 48)
    assert not m.allocations


def test_invalid_hp_trial_is_closed_gracefully(cluster):
    m, sess, tmp = cluster
    cfg = json.loads(json.dumps(BASE))
    cfg["hyperparameters"]["invalid"] = True
    eid = create(sess, cfg)
    st = wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR", "STATE_CANCELED"})
    t = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]
    assert t["restarts"] == 0, "InvalidHP must not be retried"


def test_validation_errors_and_auth(cluster):
    m, sess, tmp = cluster
    with pytest.raises(api.errors.APIException) as e:
        sess.post("/api/v1/experiments", json={"config": "searcher: {name: bogus}", "modelDefinition": []})
    assert e.value.status_code == 400
    with pytest.raises(api.errors.UnauthenticatedException):
        api.Session(m.url, None, "bad-token", None, max_retries=0).get("/api/v1/experiments")
    r = sess.post("/api/v1/experiments", json={"config": yaml.safe_dump(BASE), "validateOnly": True}).json()
    assert r["experiment"] is None and r["config"]["max_restarts"] == 2


def test_non_admin_cannot_escalate_and_checkpoint_ids_must_be_uuids(cluster):
    """Reference api_user.go:429-466 (admin / active / username need an admin) and the uuid-typed checkpoint id."""
    m, _, tmp = cluster
    sess = api.authentication.login(m.url, "admin", "")
    uid = sess.post("/api/v1/users", json={"user": {"username": "mallory", "active": True}, "password": "pw"}).json()["user"]["id"]
    mal = api.authentication.login(m.url, "mallory", "pw")
    for patch in ({"user": {"admin": True}}, {"admin": True}, {"user": {"active": False}}, {"user": {"username": "root"}}):
        with pytest.raises(api.errors.APIException) as e:
            mal.patch(f"/api/v1/users/{uid}", json=patch)
        assert e.value.status_code == 403, patch
    assert not mal.get("/api/v1/me").json()["user"]["admin"]
    with pytest.raises(api.errors.APIException) as e:
        mal.post("/api/v1/users", json={"user": {"username": "x"}})
    assert e.value.status_code == 403
    # its own display name and password are its to change; the password is stored hashed
    assert mal.patch(f"/api/v1/users/{uid}", json={"user": {"displayName": "M", "password": "pw2"}}).json()["user"]["displayName"] == "M"
    api.authentication.login(m.url, "mallory", "pw2")
    # an admin setting a password through the dict body gets it hashed as well
    sess.patch(f"/api/v1/users/{uid}", json={"user": {"password": "pw3"}})
    api.authentication.login(m.url, "mallory", "pw3")
    # a reported checkpoint id names a directory that GC will rmtree: only UUIDs are taken
    outside = tmp / "precious"
    outside.mkdir()
    (outside / "f").write_text("x")
    for bad in ("../precious", "/etc", "not-a-uuid", "../../x"):
        with pytest.raises(api.errors.APIException) as e:
            sess.post("/api/v1/checkpoints", json={"checkpoint": {"uuid": bad, "resources": {}, "metadata": {"steps_completed": 1}}})
        assert e.value.status_code == 400, bad
    from determined_b200.common.storage import shared
    sm = shared.SharedFSStorageManager(str(tmp / "ckpt"))
    (tmp / "ckpt").mkdir(exist_ok=True)
    for bad in ("../precious", "..", ".", str(outside)):
        with pytest.raises(ValueError):
            sm.delete(bad, ["**/*"])
    assert (outside / "f").exists()


def test_master_restart_restores_experiments(tmp_path):
    dbp = str(tmp_path / "m.db")
    m = core.Master(db_path=dbp, num_slots=2, launch=False, work_dir=str(tmp_path / "w"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}})
    cfg = json.loads(json.dumps(BASE))
    cfg["searcher"] = {"name": "random", "metric": "val_loss", "max_length": 10, "max_trials": 5, "max_concurrent_trials": 2}
    exp, _ = m.create_experiment(yaml.safe_dump(cfg))
    m.schedule_once()
    assert len(m.allocations) == 2
    t = next(iter(exp.trials.values()))
    exp.op_completed(t, 10, 0.5)             # the trial finished its op; Close follows
    snap_trials = {tr.request_id: (tr.state, list(tr.ops)) for tr in exp.trials.values()}
    n_trials = len(exp.trials)
    m.db.close()
    m2 = core.Master(db_path=dbp, num_slots=2, launch=False, work_dir=str(tmp_path / "w2"),
                     cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}})
    e2 = m2.experiments[exp.id]
    # every original trial survived with its pending ops; the one that had finished its op and was
    # only waiting for its (now dead) allocation is closed on restore, which lets the random
    # searcher create its replacement
    for rid, (state, ops) in snap_trials.items():
        assert rid in e2.trials and list(e2.trials[rid].ops) == ops
    assert e2.trials[t.request_id].state == "COMPLETED"
    assert len(e2.trials) == n_trials + 1
    assert e2.searcher.trials_requested == 3
    m2.schedule_once()
    assert len(m2.allocations) == 2          # re-queued after restart


def test_continue_extends_a_finished_experiment_and_move_changes_project(cluster):
    m, sess, tmp = cluster
    eid = create(sess, BASE)
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    tid = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]["id"]
    # same length again: refused
    with pytest.raises(api.errors.APIException):
        sess.post("/api/v1/experiments/continue", json={"id": eid})
    out = sess.post("/api/v1/experiments/continue", json={"id": eid, "overrideConfig": yaml.safe_dump({"searcher": {"max_length": {"batches": 40}}})}).json()
    assert out["experiment"]["state"] in ("STATE_ACTIVE", "STATE_RUNNING", "STATE_QUEUED")
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    trials = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
    assert len(trials) == 1 and trials[0]["id"] == tid and trials[0]["totalBatchesProcessed"] == 40 and trials[0]["state"] == "STATE_COMPLETED"
    logs = "".join(l["log"] for l in sess.get(f"/api/v1/trials/{tid}/logs").json()["logs"])
    assert "Restoring trial from checkpoint" in logs, "the continued run must resume, not start over"
    val = [x["totalBatches"] for x in sess.get(f"/api/v1/trials/{tid}/metrics").json()["metrics"] if x["group"] == "validation"]
    assert val[-1] == 40 and 24 in val
    ws = sess.post("/api/v1/workspaces", json={"name": "w2"}).json()["workspace"]
    pj = sess.post(f"/api/v1/workspaces/{ws['id']}/projects", json={"name": "p2"}).json()["project"]
    sess.post(f"/api/v1/experiments/{eid}/move", json={"destinationProjectId": pj["id"]})
    assert sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["projectId"] == pj["id"]
    assert [e["id"] for e in sess.get("/api/v1/experiments", params={"projectId": pj["id"]}).json()["experiments"]] == [eid]


def test_two_slot_trial_runs_data_parallel_over_gloo(cluster):
    """slots_per_trial=2 on CPU slots: launcher -> two ranks -> gloo gradient averaging + object collectives -> one trial result."""
    m, sess, tmp = cluster
    cfg = json.loads(json.dumps(BASE))
    cfg["resources"] = {"slots_per_trial": 2}
    cfg["hyperparameters"]["global_batch_size"] = 8
    eid = create(sess, cfg)
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    t = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]
    assert t["totalBatchesProcessed"] == 24
    ranks = {l.get("rankId") for l in sess.get(f"/api/v1/trials/{t['id']}/logs").json()["logs"]}
    assert ranks & {0, 1, None}
    alloc = m.db.one("SELECT slots FROM allocations WHERE trial_id=?", (t["id"],))
    assert alloc["slots"] == 2


def test_cli_on_cluster_test_mode(cluster, capsys, monkeypatch):
    """`det experiment create --test` on a cluster: a one-batch copy of the experiment runs on real slots and reports success."""
    from determined_b200.cli import cli

    m, sess, tmp = cluster
    monkeypatch.setenv("DET_MASTER", m.url)
    cfg_path = tmp / "cfg.yaml"
    cfg_path.write_text(yaml.safe_dump(dict(BASE, searcher={"name": "adaptive_asha", "metric": "val_loss", "max_trials": 4, "max_length": {"batches": 64}})))
    rc = cli.main(["experiment", "create", str(cfg_path), str(FIX), "--test"])
    out = capsys.readouterr().out
    assert rc == 0 and "Model training test succeeded." in out, out[-2000:]
    exps = sess.get("/api/v1/experiments").json()["experiments"]
    assert len(exps) == 1 and exps[0]["description"].startswith("[test mode]")
    t = sess.get(f"/api/v1/experiments/{exps[0]['id']}/trials").json()["trials"]
    assert len(t) == 1 and t[0]["totalBatchesProcessed"] == 1


def test_webhooks_state_conditions_log_triggers_slack_and_signatures(cluster):
    import hashlib
    import hmac
    import http.server
    import threading

    m, sess, tmp = cluster
    got = []

    class H(http.server.BaseHTTPRequestHandler):
        def do_POST(self):
            n = int(self.headers.get("Content-Length") or 0)
            got.append((self.path, dict(self.headers), self.rfile.read(n)))
            self.send_response(200)
            self.send_header("Content-Length", "0")
            self.end_headers()

        def log_message(self, *a):
            pass

    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), H)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    base = f"http://127.0.0.1:{srv.server_address[1]}"
    try:
        mk = lambda path, typ, trig: sess.post("/api/v1/webhooks", json={"url": base + path, "webhookType": typ, "triggers": trig}).json()["webhook"]  # noqa: E731
        w1 = mk("/done", "WEBHOOK_TYPE_DEFAULT", [{"triggerType": "TRIGGER_TYPE_EXPERIMENT_STATE_CHANGE", "condition": {"state": "COMPLETED"}}])
        mk("/slack", "WEBHOOK_TYPE_SLACK", [{"triggerType": "TRIGGER_TYPE_EXPERIMENT_STATE_CHANGE", "condition": {"state": ["COMPLETED", "ERROR"]}}])
        mk("/log", "WEBHOOK_TYPE_DEFAULT", [{"triggerType": "TRIGGER_TYPE_TASK_LOG", "condition": {"regex": "Restoring trial|validat"}}])
        with pytest.raises(api.errors.APIException):
            sess.post("/api/v1/webhooks", json={"url": "file:///etc/passwd"})
        eid = create(sess, BASE)
        assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
        t0 = time.time()
        while time.time() - t0 < 10 and not ({p for p, _, _ in got} >= {"/done", "/slack"}):
            time.sleep(0.1)
        done = [(h, b) for p, h, b in got if p == "/done"]
        assert len(done) == 1, "only the COMPLETED transition matches the condition"
        body = json.loads(done[0][1])
        assert body["event_type"] == "EXPERIMENT_STATE_CHANGE" and body["data"]["state"] == "COMPLETED" and body["data"]["experiment_id"] == eid
        admin = api.authentication.login(m.url, "admin", "")
        key = admin.get("/api/v1/master/webhook_signing_key").json()["key"].encode()
        h = {k.lower(): v for k, v in done[0][0].items()}
        want = hmac.new(key, h["x-determined-authtimestamp"].encode() + b"," + done[0][1], hashlib.sha256).hexdigest()
        assert h["x-determined-authsignature"] == want
        logs = [json.loads(b) for p, _, b in got if p == "/log"]
        assert logs and logs[0]["event_type"] == "TASK_LOG" and logs[0]["data"]["task_id"].startswith("trial-"), dump_logs(sess, eid)
        slack = json.loads([b for p, _, b in got if p == "/slack"][0])
        assert "text" in slack and "blocks" in slack and "COMPLETED" in slack["text"]
        got.clear()
        sess.post(f"/api/v1/webhooks/{w1['id']}/test")
        t0 = time.time()
        while time.time() - t0 < 5 and not got:
            time.sleep(0.05)
        assert got and json.loads(got[0][2])["data"]["test"] is True
    finally:
        srv.shutdown()
