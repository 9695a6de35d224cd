"""Options added by ``cli/flags.py`` (list views, confirmations, per-command extras) against an in-process master; one
short experiment runs so that trial / checkpoint options have something to look at."""
import json

import pytest

from determined_b200.cli import flags
from determined_b200.common import api

from tests.test_cli_more import CFG, FIX, admin, cluster, run  # noqa: F401
from tests.test_master_e2e import BASE, dump_logs, wait_state


def test_view_options_sort_and_cut_any_printed_list(cluster):
    rows = [[3, "b", None], [1, "c", 2.0], [2, "a", 1.0]]
    assert [r[0] for r in flags.apply_view(rows, ["ID", "Name", "# Things"], {"sort_by": "id"})] == [1, 2, 3]
    assert [r[0] for r in flags.apply_view(rows, ["ID", "Name", "# Things"], {"sort_by": "name", "order_by": "desc"})] == [1, 3, 2]
    assert [r[0] for r in flags.apply_view(rows, ["ID", "Name", "# Things"], {"sort_by": "things"})] == [2, 1, 3], "None sorts last"
    assert [r[0] for r in flags.apply_view(rows, ["ID"], {"offset": 1, "limit": 1, "pages_explicit": True})] == [1]
    with pytest.raises(SystemExit):
        flags.apply_view(rows, ["ID", "Name"], {"sort_by": "owner"})
    for n in ("zeta", "alpha", "mid"):
        assert run("workspace", "create", n)[0] == 0
    rc, out = run("workspace", "list", "--sort-by", "name", "--order-by", "desc", "--json")
    names = [w["Name"] for w in json.loads(out)]
    assert rc == 0 and names == sorted(names, reverse=True) and "zeta" == names[0]
    rc, out = run("workspace", "list", "--sort-by", "name", "--limit", "2", "--offset", "1", "--json")
    assert [w["Name"] for w in json.loads(out)] == sorted(names)[1:3]
    for p in ("p2", "p1"):
        assert run("workspace", "create-project", "alpha", p)[0] == 0
    rc, out = run("project", "list", "alpha", "--sort-by", "name", "--json")
    assert [p["Name"] for p in json.loads(out)] == ["p1", "p2"]
    rc, out = run("workspace", "list-projects", "alpha", "--sort-by", "name", "--order-by", "desc", "--limit", "1", "--json")
    assert [p["Name"] for p in json.loads(out)] == ["p2"]


def test_workspace_create_options_users_groups_and_roles(cluster):
    m, tmp = cluster
    rc, out = admin("workspace", "create", "lab", "--agent-uid", "1000", "--agent-user", "worker", "--default-compute-pool", "default", "--default-aux-pool", "aux",
                    "--checkpoint-storage-config", json.dumps({"type": "shared_fs", "host_path": str(tmp / "labck")}))
    assert rc == 0, out
    sess = api.authentication.login(m.url, "admin", "")
    w = [w for w in sess.get("/api/v1/workspaces").json()["workspaces"] if w["name"] == "lab"][0]
    w = sess.get(f"/api/v1/workspaces/{w['id']}").json()["workspace"]
    assert w["agentUserGroup"]["agentUid"] == 1000 and w["defaultAuxPool"] == "aux" and w["checkpointStorageConfig"]["type"] == "shared_fs"
    assert admin("workspace", "create", "bad", "--agent-uid", "5")[0] != 0, "uid without user name"

    assert admin("user", "create", "alice")[0] == 0 and admin("user", "create", "sso-bob", "--remote")[0] == 0
    with pytest.raises(api.errors.APIException):
        api.authentication.login(m.url, "sso-bob", "")
    assert admin("user", "deactivate", "alice")[0] == 0
    rc, out = admin("user", "list", "--json")
    assert "alice" not in [u["Username"] for u in json.loads(out)] and {u["Username"]: u["Remote"] for u in json.loads(out)}["sso-bob"] is True
    rc, out = admin("user", "list", "--all", "--json")
    assert "alice" in [u["Username"] for u in json.loads(out)]
    assert admin("user", "activate", "alice")[0] == 0

    import os
    os.environ["DET_NEW_PASS"] = "s3cret"
    try:
        assert admin("user", "change-password", "alice")[0] == 0
    finally:
        del os.environ["DET_NEW_PASS"]
    api.authentication.login(m.url, "alice", "s3cret")

    assert admin("user-group", "create", "g1", "--add-user", "alice")[0] == 0 and admin("user-group", "create", "g2")[0] == 0
    rc, out = admin("user-group", "list", "--groups-user-belongs-to", "alice", "--json")
    assert [g["Name"] for g in json.loads(out)] == ["g1"]
    rc, out = admin("user-group", "list", "--limit", "1", "--pages", "1", "--json")
    assert len(json.loads(out)) == 1
    assert admin("user-group", "delete", "g2", "--yes")[0] == 0

    rc, out = admin("rbac", "list-roles", "--json")
    every = [r["Role"] for r in json.loads(out)]
    rc, out = admin("rbac", "list-roles", "--exclude-global-roles", "--json")
    scoped = [r["Role"] for r in json.loads(out)]
    assert rc == 0 and set(scoped) < set(every) and "ClusterAdmin" not in scoped and "Editor" in scoped
    assert admin("rbac", "assign-role", "Editor", "--username-to-assign", "alice", "--workspace-name", "lab")[0] == 0
    rc, out = admin("rbac", "list-users-roles", "alice", "--json")
    assert any(r["Role"] == "Editor" and r["Workspace"] == w["id"] for r in json.loads(out))
    assert admin("rbac", "unassign-role", "Editor", "--username-to-assign", "alice", "--workspace-name", "lab")[0] == 0


def test_trial_checkpoint_and_experiment_options(cluster, tmp_path):
    m, tmp = cluster
    cfgp = tmp_path / "c.yaml"
    cfg = dict(BASE, min_checkpoint_period={"batches": 8}, checkpoint_policy="all",
               checkpoint_storage={"type": "shared_fs", "host_path": str(tmp / "ck"), "save_trial_latest": 5})
    cfgp.write_text(json.dumps(cfg))
    rc, out = run("experiment", "create", str(cfgp), str(FIX))
    assert rc == 0, out
    sess = api.authentication.login(m.url, "determined", "")
    eid = sess.get("/api/v1/experiments").json()["experiments"][0]["id"]
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}, timeout=120) == "STATE_COMPLETED", dump_logs(sess, eid)
    tid = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]["id"]

    rc, out = run("trial", "describe", str(tid), "--metrics-summary", "--json")
    summ = json.loads(out)
    assert rc == 0 and summ["validation.val_loss"]["count"] >= 2 and summ["validation.val_loss"]["min"] <= summ["validation.val_loss"]["max"]
    rc, out = run("trial", "describe", str(tid), "--metrics", "--limit", "2", "--pages", "1")
    assert rc == 0 and "Metrics" in out

    rc, out = run("experiment", "list-checkpoints", str(eid), "--best", "1", "--json")
    best = json.loads(out)
    cks = [c for c in sess.get(f"/api/v1/experiments/{eid}/checkpoints").json()["checkpoints"] if c["state"] == "STATE_COMPLETED"]
    assert len(best) == 1 and len(cks) >= 2
    assert best[0]["Validation Metric"] == min(c["training"]["searcherMetric"] for c in cks if c["training"]["searcherMetric"] is not None)

    rc, out = run("trial", "download", str(tid), "--sort-by", "val_loss", "--smaller-is-better", "false", "-o", str(tmp_path / "dl"), "--quiet")
    assert rc == 0 and out.strip() == str(tmp_path / "dl"), out
    worst = max((c for c in cks if c["training"]["searcherMetric"] is not None), key=lambda c: c["training"]["searcherMetric"])
    assert json.loads((tmp_path / "dl" / "metadata.json").read_text())["steps_completed"] == worst["metadata"]["steps_completed"]

    rc, out = run("experiment", "describe", str(eid), "--metrics", "--outdir", str(tmp_path / "desc"))
    assert rc == 0 and {p.name for p in (tmp_path / "desc").iterdir()} == {"experiments.csv", "trials.csv", "metrics.csv"}
    assert len((tmp_path / "desc" / "metrics.csv").read_text().splitlines()) > 3

    # project views: the default is the caller's own unarchived experiments, -a is everything
    assert run("experiment", "archive", str(eid))[0] == 0
    rc, out = run("project", "list-experiments", "Uncategorized", "Uncategorized", "--json")
    assert rc == 0 and json.loads(out) == []
    rc, out = run("project", "list-experiments", "Uncategorized", "Uncategorized", "-a", "--json")
    assert [e["ID"] for e in json.loads(out)] == [eid]

    # the model registry
    ck = best[0]["UUID"]
    assert run("workspace", "create", "reg")[0] == 0
    assert run("model", "create", "m-b")[0] == 0 and run("model", "create", "m-a")[0] == 0
    assert run("model", "move", "m-a", "-w", "reg")[0] == 0
    assert run("model", "register-version", "m-a", ck)[0] == 0
    rc, out = run("model", "list", "--sort-by", "name", "--json")
    assert [x["Name"] for x in json.loads(out)] == ["m-a", "m-b"]
    rc, out = run("model", "list", "--workspace-names", "reg", "--json")
    assert [x["Name"] for x in json.loads(out)] == ["m-a"]
    rc, out = run("model", "describe", "m-a", "--version", "1", "--json")
    assert rc == 0 and json.loads(out)["checkpoint"]["uuid"] == ck


def test_task_tree_kill_template_details_and_job_pool_move(cluster, tmp_path):
    m, tmp = cluster
    sess = api.authentication.login(m.url, "determined", "")
    cfgp = tmp_path / "t.yaml"
    cfgp.write_text(json.dumps({"entrypoint": ["/bin/sh", "-c", "sleep 300"], "resources": {"slots": 0}}))
    rc, out = run("task", "create", str(cfgp))
    root = out.split()[2]
    rc, out = run("task", "fork", root)
    child = out.split()[2]
    rc, out = run("task", "kill", child, "--root")
    assert rc == 0 and root in out and child in out, out
    for tid in (root, child):
        for _ in range(100):
            if m.commands.tasks[tid].state == "TERMINATED":
                break
            import time
            time.sleep(0.1)
        assert m.commands.tasks[tid].state == "TERMINATED"

    tp = tmp_path / "tpl.yaml"
    tp.write_text("resources:\n  slots_per_trial: 2\n")
    assert run("template", "create", "two-slots", str(tp))[0] == 0
    rc, out = run("template", "list", "--details")
    assert rc == 0 and "two-slots" in out and "slots_per_trial: 2" in out

    # a queued job changes pools: the first experiment holds both slots, the second waits behind it
    ep = tmp_path / "big.yaml"
    ep.write_text(json.dumps(dict(BASE, resources={"slots_per_trial": 2}, searcher={"name": "single", "metric": "val_loss", "max_length": {"batches": 1000000}})))
    assert run("experiment", "create", str(ep), str(FIX))[0] == 0
    first = sess.get("/api/v1/experiments").json()["experiments"][-1]["id"]
    import time
    for _ in range(200):
        if m.pools["default"].running:
            break
        time.sleep(0.1)
    assert run("experiment", "create", str(ep), str(FIX))[0] == 0
    eid = max(e["id"] for e in sess.get("/api/v1/experiments").json()["experiments"])
    for _ in range(100):
        if any(r.job_id == f"exp-{eid}" for r in m.pools["default"].pending.values()):
            break
        time.sleep(0.1)
    assert run("job", "update", f"exp-{eid}", "--resource-pool", "aux")[0] == 0
    assert not any(r.job_id == f"exp-{eid}" for r in m.pools["default"].pending.values())
    assert any(r.job_id == f"exp-{eid}" for r in m.pools["aux"].pending.values())
    assert sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["config"]["resources"]["resource_pool"] == "aux"
    run("experiment", "kill", str(first))
    run("experiment", "kill", str(eid))

    rc, out = run("dev", "bindings", "list")
    assert rc == 0 and "get_GetMaster" in out
    rc2, out2 = run("dev", "bindings", "list", "--show-unusable")
    assert rc2 == 0 and len(out2.splitlines()) >= len(out.splitlines())
    rc, out = run("dev", "curl", "-x", "/api/v1/me")
    assert rc == 0 and out.splitlines()[0].startswith("curl http://127.0.0.1:") and "Authorization: Bearer" in out and '"username"' in out
    assert run("workspace", "create", "tplws")[0] == 0
    assert run("template", "create", "scoped", str(tp), "--workspace-name", "tplws")[0] == 0
    wid = [w for w in sess.get("/api/v1/workspaces").json()["workspaces"] if w["name"] == "tplws"][0]["id"]
    assert sess.get("/api/v1/templates/scoped").json()["template"]["workspaceId"] == wid
    assert run("model", "create", "scoped-model", "--workspace-name", "tplws")[0] == 0
    assert sess.get("/api/v1/models/scoped-model").json()["model"]["workspaceId"] == wid
