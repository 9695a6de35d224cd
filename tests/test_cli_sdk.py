"""CLI + SDK against an in-process master (no trial processes are launched)."""
import io
import json
import os
import pathlib
import sys
import contextlib

import pytest
import yaml

from determined_b200.cli import cli
from determined_b200.common import api
from determined_b200.experimental import client as sdk
from determined_b200.master import api as mapi
from determined_b200.master import core

FIX = pathlib.Path(__file__).parent / "fixtures" / "onevar_exp"
CFG = {"name": "cli-exp", "entrypoint": "model_def:OneVarTrial", "searcher": {"name": "random", "metric": "val_loss", "max_length": 20, "max_trials": 3, "max_concurrent_trials": 2},
       "hyperparameters": {"learning_rate": {"type": "double", "minval": 0.01, "maxval": 0.1}, "global_batch_size": 8}}


@pytest.fixture()
def cluster(tmp_path, monkeypatch):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=8, launch=False, work_dir=str(tmp_path / "w"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}})
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    monkeypatch.setenv("DET_MASTER", m.url)
    monkeypatch.delenv("DET_SESSION_TOKEN", raising=False)
    yield m, tmp_path
    m.stop()
    srv.stop()


def run(*argv):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        rc = cli.main(list(argv))
    return rc, buf.getvalue()


def test_cli_experiment_lifecycle(cluster):
    m, tmp = cluster
    cfgf = tmp / "cfg.yaml"
    cfgf.write_text(yaml.safe_dump(CFG))
    rc, out = run("experiment", "create", str(cfgf), str(FIX), "--config", "searcher.max_trials=4", "--config", "description=from-cli")
    assert rc == 0 and "Created experiment 1" in out
    rc, out = run("e", "list", "--json")
    rows = json.loads(out)
    assert rows[0]["ID"] == 1 and rows[0]["State"] == "ACTIVE" and rows[0]["Name"] == "cli-exp"
    rc, out = run("e", "describe", "1", "--json")
    d = json.loads(out)[0]
    assert d["config"]["searcher"]["max_trials"] == 4 and d["description"] == "from-cli" and len(d["trials"]) == 2
    m.schedule_once()
    rc, out = run("job", "list", "--json")
    assert json.loads(out)[0]["Slots (acquired/needed)"] == "2/2"
    rc, out = run("slot", "list", "--json")
    assert sum(1 for s in json.loads(out) if s["Allocation ID"]) == 2
    assert run("e", "pause", "1")[0] == 0
    assert json.loads(run("e", "list", "--json")[1])[0]["State"] == "PAUSED"
    assert run("e", "activate", "1")[0] == 0
    assert run("e", "set", "priority", "1", "7")[0] == 0 and m.experiments[1].config["resources"]["priority"] == 7
    assert run("e", "label", "add", "1", "blue")[0] == 0
    assert "blue" in json.loads(run("e", "describe", "1", "--json")[1])[0]["labels"]
    assert run("e", "kill", "1")[0] == 0
    # launch=False: allocations never exit by themselves
    for a in list(m.allocations.values()):
        a.state = "TERMINATED"
        a.exit_code = 137
        m.on_allocation_exit(a)
    assert json.loads(run("e", "list", "--json")[1])[0]["State"] == "CANCELED"
    assert run("e", "archive", "1")[0] == 0
    assert json.loads(run("e", "list", "--json")[1]) == []
    assert run("e", "delete", "1", "--yes")[0] == 0
    assert run("e", "describe", "1")[0] == 1


def test_cli_misc_commands(cluster):
    m, tmp = cluster
    assert run("master", "info")[0] == 0
    rc, out = run("agent", "list", "--json")
    assert json.loads(out)[0]["Slots"] == 8
    rc, out = run("resource-pool", "list", "--json")
    assert json.loads(out)[0]["Scheduler"] == "FAIR_SHARE"
    t = tmp / "t.yaml"
    t.write_text("max_restarts: 0\nresources: {slots_per_trial: 2}\n")
    assert run("template", "set", "small", str(t))[0] == 0
    assert "max_restarts: 0" in run("template", "describe", "small")[1]
    cfgf = tmp / "cfg.yaml"
    cfgf.write_text(yaml.safe_dump(CFG))
    assert run("e", "create", str(cfgf), str(FIX), "--template", "small", "--paused")[0] == 0
    assert m.experiments[1].config["max_restarts"] == 0 and m.experiments[1].slots_per_trial() == 2
    rc, out = run("preview-search", str(cfgf))
    assert rc == 0 and "train 20" in out
    assert run("user", "create", "alice")[0] == 1           # only admins may create users
    assert run("-u", "admin", "user", "create", "alice")[0] == 0
    assert any(u["Username"] == "alice" for u in json.loads(run("user", "list", "--json")[1]))
    assert "determined" in run("user", "whoami")[1]
    assert run("workspace", "create", "ws1")[0] == 0 and run("workspace", "create-project", "ws1", "p1")[0] == 0
    assert json.loads(run("workspace", "list-projects", "ws1", "--json")[1])[0]["Name"] == "p1"
    assert run("slot", "disable", m.agent_id, "3")[0] == 0
    assert m.pool.agents[m.agent_id].slot_enabled[3] is False


def test_sdk_checkpoints_and_models(cluster):
    m, tmp = cluster
    d = sdk.Determined(m.url)
    exp = d.create_experiment(CFG, str(FIX))
    assert exp.state == sdk.ExperimentState.ACTIVE and len(exp.list_trials()) == 2
    t = exp.list_trials()[0]
    sess = api.authentication.login(m.url, "determined", "")
    # fake what a harness would report for trial t
    ck_store = tmp / "ck"
    for i, (steps, loss) in enumerate([(10, 0.5), (20, 0.2)]):
        uuid = f"00000000-0000-0000-0000-00000000000{i}"
        (ck_store / uuid).mkdir(parents=True)
        (ck_store / uuid / "state_dict.pth").write_bytes(b"w" * 10)
        sess.post(f"/api/v1/trials/{t.id}/metrics", json={"group": "validation", "metrics": {"trialId": t.id, "stepsCompleted": steps, "metrics": {"avgMetrics": {"val_loss": loss}}}})
        sess.post("/api/v1/checkpoints", json={"uuid": uuid, "taskId": f"trial-{t.id}", "allocationId": "a", "state": "STATE_COMPLETED",
                                                "resources": {"state_dict.pth": "10"}, "metadata": {"steps_completed": steps}})
    best = exp.top_checkpoint()
    assert best.uuid.endswith("1") and best.training["searcherMetric"] == 0.2
    assert [c.uuid[-1] for c in exp.top_n_checkpoints(2)] == ["1", "0"]
    assert t.select_checkpoint(latest=True).uuid.endswith("1")
    assert t.top_checkpoint(sort_by="val_loss", smaller_is_better=False).uuid.endswith("0")
    path = best.download(str(tmp / "dl"))
    assert (pathlib.Path(path) / "state_dict.pth").exists() and json.loads((pathlib.Path(path) / "metadata.json").read_text())["steps_completed"] == 20
    best.add_metadata({"note": "good"})
    assert d.get_checkpoint(best.uuid).metadata["note"] == "good"
    model = d.create_model("resnet", "desc", labels=["cv"])
    v = model.register_version(best.uuid)
    assert v.model_version == 1 and d.get_model("resnet").get_version().checkpoint.uuid == best.uuid
    with pytest.raises(api.errors.APIException):
        best.delete()                       # registered checkpoints are protected
    other = d.get_checkpoint("00000000-0000-0000-0000-000000000000")
    other.delete()
    assert not (ck_store / other.uuid).exists()
    assert list(t.iter_metrics("validation"))[1]["metrics"] == {"val_loss": 0.2}
    rc, out = run("checkpoint", "describe", best.uuid, "--json")
    assert json.loads(out)["metadata"]["note"] == "good"
    rc, out = run("e", "list-checkpoints", str(exp.id), "--json")
    assert len(json.loads(out)) == 2


def test_local_test_mode_via_cli(tmp_path, monkeypatch):
    cfgf = tmp_path / "cfg.yaml"
    c = dict(CFG)
    c["hyperparameters"] = {"learning_rate": 0.05, "global_batch_size": 8}
    cfgf.write_text(yaml.safe_dump(c))
    cwd = os.getcwd()
    try:
        rc, out = run("e", "create", str(cfgf), str(FIX), "--local", "--test")
    finally:
        os.chdir(cwd)
        sys.path[:] = [p for p in sys.path if p != str(FIX)]
        sys.modules.pop("model_def", None)
    assert rc == 0 and "test succeeded" in out


def test_groups_and_role_assignments(cluster):
    m, tmp = cluster
    assert run("-u", "admin", "user", "create", "alice")[0] == 0
    rc, out = run("-u", "admin", "user-group", "create", "research", "--add-user", "alice")
    assert rc == 0 and "Created group research" in out
    rc, out = run("user-group", "describe", "research", "--json")
    assert [r["Username"] for r in json.loads(out)] == ["alice"]
    assert run("-u", "admin", "rbac", "assign-role", "Editor", "-g", "research", "-w", "1")[0] == 0
    assert run("-u", "admin", "rbac", "assign-role", "Viewer", "-u", "alice")[0] == 0
    rc, out = run("rbac", "list-users-roles", "alice", "--json")
    roles = sorted(r["Role"] for r in json.loads(out))
    assert roles == ["Editor", "Viewer"]
    assert run("-u", "admin", "rbac", "unassign-role", "Viewer", "-u", "alice")[0] == 0
    rc, out = run("rbac", "list-users-roles", "alice", "--json")
    assert [r["Role"] for r in json.loads(out)] == ["Editor"]
    assert run("-u", "admin", "user-group", "remove-user", "research", "alice")[0] == 0
    rc, out = run("user-group", "list", "--json")
    assert json.loads(out)[0]["# Members"] == 0
    assert run("-u", "admin", "user-group", "delete", "research")[0] == 0


def test_generated_bindings_are_fresh_and_work(cluster):
    """The committed bindings match the route table, and calling through them hits the master."""
    from determined_b200.common.api import bindings, generate_bindings

    assert generate_bindings.main(["--check"]) == 0
    from determined_b200.common import api

    sess = api.authentication.login(cluster[0].url, "determined", "")
    doc = bindings.get_openapi_json(sess)
    assert "/api/v1/experiments/{eid}" in doc["paths"] or any(p.startswith("/api/v1/experiments/{") for p in doc["paths"])
    assert "agents" in bindings.get_agents(sess)
    w = bindings.post_workspaces(sess, body={"name": "bindings-ws"})["workspace"]
    assert bindings.get_workspaces_by_wid(sess, wid=w["id"])["workspace"]["name"] == "bindings-ws"


def test_webui_is_served_and_its_api_calls_exist(cluster):
    """The dashboard page is served at / and every API path it fetches is a real route."""
    import re

    import requests

    from determined_b200.master import api as mapi_

    m = cluster[0]
    r = requests.get(m.url + "/", timeout=5)
    assert r.status_code == 200 and "text/html" in r.headers["Content-Type"] and "determined_b200" in r.text
    paths = set(re.findall(r"[\"`](/api/v1/[A-Za-z0-9_\-/${}.]*)", r.text))
    assert len(paths) >= 12
    for p in paths:
        concrete = p.replace("${kind}", "commands").replace("${b.dataset.k}", "commands").replace("${b.dataset.a}", "pause")
        concrete = re.sub(r"\$\{[^}]*\}", "1", concrete).split("?")[0]
        assert any(rx.match(concrete) for _, rx, _, _ in mapi_.ROUTES), p


def test_prometheus_endpoint(cluster):
    import requests

    m = cluster[0]
    txt = requests.get(m.url + "/prom/det-state-metrics", timeout=5).text
    assert f'det_slots_total{{agent="{m.agent_id}"}} 8' in txt and "det_experiments_active 0" in txt and "det_master_uptime_seconds" in txt


def test_cli_project_task_dev(cluster):
    def run(*argv):
        rc, out = globals()["run"](*argv)
        assert rc == 0, out
        return out

    out = run("workspace", "create", "ws1")
    assert "Created workspace ws1" in out
    assert "Created project p1" in run("project", "create", "ws1", "p1", "--description", "first")
    assert "p1" in run("project", "list", "ws1") and "first" in run("project", "describe", "ws1", "p1")
    assert "ID" in run("project", "list-experiments", "ws1", "p1")
    assert "Deleted project p1" in run("project", "delete", "ws1", "p1")
    assert "p1" not in run("project", "list", "ws1")
    assert "Task ID" in run("task", "list")
    got = json.loads(run("dev", "curl", "/api/v1/master"))
    assert "clusterId" in got or "version" in got
    assert "get_agents" in run("dev", "bindings", "list")
    ws = json.loads(run("dev", "bindings", "call", "post_workspaces", 'body={"name": "via-bindings"}'))
    assert ws["workspace"]["name"] == "via-bindings"
    one = json.loads(run("dev", "bindings", "call", "get_workspaces_by_wid", f"wid={ws['workspace']['id']}"))
    assert one["workspace"]["name"] == "via-bindings"


def test_sdk_users_workspaces_projects_pools(cluster):
    from determined_b200.experimental import client

    d = client.Determined(cluster[0].url, "admin", "")
    u = d.create_user("alice", password="pw1")
    assert u.username == "alice" and not u.admin and u.active
    u.change_display_name("Alice A.")
    u.deactivate()
    assert d.get_user_by_name("alice").active is False and d.get_user_by_id(u.user_id).display_name == "Alice A."
    assert [x.username for x in d.list_users(active=False)] == ["alice"]
    u.activate()
    u.rename("alice2")
    assert d.get_user_by_name("alice2").user_id == u.user_id and d.get_session_username() == "admin"
    client.Determined(cluster[0].url, "alice2", "pw1")                 # the password set at creation works
    ws = d.create_workspace("research")
    p = ws.create_project("vision", "cv models")
    assert [x.name for x in ws.list_projects()] == ["vision"] and ws.get_project("vision").description == "cv models"
    assert "research" in [w.name for w in d.list_workspaces()] and p.list_experiments() == []
    p.delete()
    d.delete_workspace("research")
    assert "research" not in [w.name for w in d.list_workspaces()]
    pools = d.list_resource_pools()
    assert pools[0].name == "default" and pools[0].slots_available == 8
    m = d.create_model("m1", labels=["a", "b"])
    assert d.get_model_labels() == ["a", "b"] and d.get_model_by_id(getattr(m, "model_id", getattr(m, "id", None))).name == "m1"
    assert list(d.iter_trials_metrics([], "training")) == []


def test_cli_agent_user_model_additions(cluster):
    def ok(*argv):
        rc, out = run(*argv)
        assert rc == 0, out
        return out

    m = cluster[0]
    assert "Disabled agent" in ok("agent", "disable", m.agent_id)
    assert not m.pool.agents[m.agent_id].enabled
    assert "Enabled agent" in ok("agent", "enable", m.agent_id)
    ok("-u", "admin", "user", "create", "bob")
    assert "Renamed user bob to robert" in ok("-u", "admin", "user", "rename", "bob", "robert")
    assert "robert" in ok("user", "list")
    ok("model", "create", "mm", "--description", "d")
    assert "Version" in ok("model", "list-versions", "mm")


def test_hp_visualisation_endpoints(cluster):
    """metric-names / trials-snapshot / trials-sample / summarize / compare over metrics reported for two trials."""
    m, tmp = cluster
    d = sdk.Determined(m.url)
    exp = d.create_experiment(CFG, str(FIX))
    sess = api.authentication.login(m.url, "determined", "")
    trials = exp.list_trials()
    for k, t in enumerate(trials):
        for steps in (10, 20, 30):
            sess.post(f"/api/v1/trials/{t.id}/metrics", json={"group": "validation", "metrics": {"stepsCompleted": steps, "trialId": t.id, "trialRunId": 0,
                                                                                                 "metrics": {"avgMetrics": {"val_loss": 1.0 / steps + k, "acc": 0.5}}}})
            sess.post(f"/api/v1/trials/{t.id}/metrics", json={"group": "training", "metrics": {"stepsCompleted": steps, "trialId": t.id, "trialRunId": 0,
                                                                                               "metrics": {"avgMetrics": {"loss": 2.0 / steps}}}})
    names = sess.get(f"/api/v1/experiments/{exp.id}/metric-names").json()
    assert names["validationMetrics"] == ["acc", "val_loss"] and names["trainingMetrics"] == ["loss"] and names["searcherMetric"] == "val_loss"
    multi = sess.get("/api/v1/experiments/metrics-stream/metric-names", params={"ids": [exp.id]}).json()["results"]
    assert multi[0]["experimentId"] == exp.id
    snap = sess.get(f"/api/v1/experiments/{exp.id}/trials-snapshot", params={"metricName": "val_loss", "batchesProcessed": 20}).json()["trials"]
    assert [round(x["metric"], 4) for x in snap] == [0.05, 1.05] and all(x["batchesProcessed"] == 20 and "hparams" in x for x in snap)
    sample = sess.get(f"/api/v1/experiments/{exp.id}/trials-sample", params={"metricName": "val_loss", "maxTrials": 1, "maxDatapoints": 2}).json()["trials"]
    assert len(sample) == 1 and sample[0]["trialId"] == trials[0].id and [p["batches"] for p in sample[0]["data"]] == [10, 30]
    summ = sess.get(f"/api/v1/trials/{trials[0].id}/summarize", params={"metricNames": ["loss"]}).json()
    assert [x["name"] for x in summ["metrics"]] == ["loss"] and len(summ["metrics"][0]["data"]) == 3 and summ["trial"]["id"] == trials[0].id
    cmp_ = sess.get("/api/v1/trials/compare", params={"trialIds": [t.id for t in trials]}).json()["trials"]
    assert len(cmp_) == 2 and {x["name"] for x in cmp_[1]["metrics"]} == {"acc", "loss", "val_loss"}


def test_sdk_list_trials_sorting(cluster):
    m, tmp = cluster
    d = sdk.Determined(m.url)
    exp = d.create_experiment(CFG, str(FIX))
    ids = [t.id for t in exp.list_trials()]
    assert ids == sorted(ids) and len(ids) == 2
    assert [t.id for t in exp.list_trials(sort_by=sdk.TrialSortBy.ID, order_by=sdk.OrderBy.DESC)] == sorted(ids, reverse=True)
    assert sdk.OrderBy.ASC is sdk.OrderBy.ASCENDING and sdk.TrialOrderBy is sdk.OrderBy
    assert len(exp.list_trials(sort_by=sdk.TrialSortBy.BEST_VALIDATION_METRIC)) == 2        # no validations yet: still listed
