"""The wider ``det`` tree (``cli/more.py``) and the routes behind it (``master/api_more.py``), against an in-process master.
One real command task runs for the generic-task / pause / context checks; no trials are launched."""
import contextlib
import io
import json
import pathlib
import sys
import tarfile
import time

import pytest
import yaml

from determined_b200.cli import cli
from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import core

FIX = pathlib.Path(__file__).parent / "fixtures" / "onevar_exp"
CFG = {"name": "cli-exp", "entrypoint": "model_def:OneVarTrial", "searcher": {"name": "single", "metric": "val_loss", "max_length": 20},
       "hyperparameters": {"learning_rate": 0.1, "global_batch_size": 8}, "labels": ["nightly"]}


@pytest.fixture()
def cluster(tmp_path, monkeypatch):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=2, launch=True, work_dir=str(tmp_path / "w"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}},
                    resource_pools=[{"pool_name": "default"}, {"pool_name": "aux", "description": "zero-slot work"}])
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    (tmp_path / "w").mkdir(exist_ok=True)
    monkeypatch.setenv("DET_MASTER", m.url)
    monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "config"))
    monkeypatch.delenv("DET_SESSION_TOKEN", raising=False)
    yield m, tmp_path
    m.stop()
    srv.stop()


def run(*argv):
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        try:
            rc = cli.main(list(argv))
        except SystemExit as e:
            rc = e.code if isinstance(e.code, int) else 1
            if not isinstance(e.code, int) and e.code:
                buf.write(str(e.code))
    return rc, buf.getvalue()


def admin(*argv):
    return run("-u", "admin", *argv)


def test_help_exists_at_every_level_and_the_tree_is_as_wide_as_the_reference(cluster):
    import argparse

    rc, out = run("help")
    assert rc == 0 and "experiment" in out and "workspace" in out
    rc, out = run("workspace", "help")
    assert rc == 0 and "list-pools" in out and "archive" in out
    rc, out = run("resource-pool", "bindings", "help")
    assert rc == 0 and "list-workspaces" in out

    def leaves(p, n=0):
        subs = next((a for a in p._actions if isinstance(a, argparse._SubParsersAction)), None)
        if subs is None:
            return 1
        return sum(leaves(c) for c in {id(c): c for c in subs.choices.values()}.values())
    assert leaves(cli.make_parser()) >= 200          # the reference's tree has 200 leaf commands


def test_workspace_project_editing_and_pool_bindings(cluster):
    m, tmp = cluster
    assert run("workspace", "create", "research")[0] == 0 and run("workspace", "create", "other")[0] == 0
    assert run("workspace", "create-project", "research", "vision")[0] == 0
    rc, out = run("project", "edit", "research", "vision", "--new_name", "vision2", "--description", "images", "--json")
    assert rc == 0 and json.loads(out)["name"] == "vision2" and json.loads(out)["description"] == "images"
    assert run("project", "archive", "research", "vision2")[0] == 0
    sess = api.authentication.login(m.url, "determined", "")
    assert sess.get("/api/v1/workspaces/2/projects").json()["projects"][0]["archived"] is True
    assert run("project", "unarchive", "research", "vision2")[0] == 0
    assert run("project", "archive", "Uncategorized", "Uncategorized")[0] == 1           # the built-in one stays
    # bindings: only administrators; a bound pool is closed to other workspaces
    assert run("resource-pool", "bindings", "add", "aux", "research")[0] == 1
    assert admin("resource-pool", "bindings", "add", "aux", "research")[0] == 0
    rc, out = run("resource-pool", "bindings", "list-workspaces", "aux")
    assert "research" in out and "other" not in out
    rc, out = run("workspace", "list-pools", "other")
    assert "default" in out and "aux" not in out
    rc, out = run("workspace", "list-pools", "research")
    assert "default" in out and "aux" in out
    assert run("workspace", "edit", "other", "--default-compute-pool", "aux")[0] == 1    # not bound to it
    assert run("workspace", "edit", "research", "--default-compute-pool", "aux", "--name", "research-lab")[0] == 0
    assert admin("workspace", "edit", "research-lab", "--agent-uid", "1001", "--agent-user", "ml", "--agent-gid", "1001", "--agent-group", "ml",
                 "--checkpoint-storage-config", '{"type": "shared_fs", "host_path": "/ws-ckpt"}')[0] == 0
    rc, out = run("workspace", "describe", "research-lab", "--json")
    w = json.loads(out)
    assert w["defaultComputePool"] == "aux" and w["agentUserGroup"]["agentUser"] == "ml" and w["checkpointStorageConfig"]["host_path"] == "/ws-ckpt"
    # an experiment in that workspace's project lands in the workspace's default pool; one in another workspace may not name it
    cfgf = tmp / "cfg.yaml"
    cfgf.write_text(yaml.safe_dump(CFG))
    pid = sess.get("/api/v1/workspaces/2/projects").json()["projects"][0]["id"]
    assert run("experiment", "create", str(cfgf), str(FIX), "--paused", "--project_id", str(pid))[0] == 0
    assert sess.get("/api/v1/experiments/1").json()["experiment"]["resourcePool"] == "aux"
    cfgf.write_text(yaml.safe_dump(dict(CFG, resources={"resource_pool": "aux"})))
    rc, out = run("experiment", "create", str(cfgf), str(FIX), "--paused")
    assert rc == 1
    assert admin("resource-pool", "bindings", "replace", "aux", "other")[0] == 0 and admin("resource-pool", "bindings", "remove", "aux", "other")[0] == 0
    assert sess.get("/api/v1/resource-pools/aux/workspace-bindings").json()["workspaceIds"] == []
    assert run("workspace", "archive", "other")[0] == 0 and run("workspace", "unarchive", "other")[0] == 0


def test_rbac_reads_users_groups_and_oauth_clients(cluster):
    m, tmp = cluster
    assert admin("user", "create", "alice")[0] == 0
    assert admin("user-group", "create", "team", "--add-user", "alice")[0] == 0
    assert admin("rbac", "assign-role", "Editor", "-g", "team", "-w", "1")[0] == 0
    rc, out = run("rbac", "list-groups-roles", "team", "--json")
    assert [r["Role"] for r in json.loads(out)] == ["Editor"]
    rc, out = run("rbac", "describe-role", "Editor", "--json")
    r = json.loads(out)
    assert "PERMISSION_TYPE_CREATE_EXPERIMENT" in [p["id"] for p in r["role"]["permissions"]] and len(r["groupRoleAssignments"]) == 1
    rc, out = run("-u", "alice", "rbac", "my-permissions", "--json")
    mine = json.loads(out)
    assert mine["assignments"] == [{"roleId": 6, "scopeCluster": False, "scopeWorkspaceIds": [1]}]
    rc, out = admin("rbac", "my-permissions")
    assert "ClusterAdmin" in out
    assert admin("user-group", "change-name", "team", "crew")[0] == 0
    assert "crew" in run("user-group", "list")[1]
    assert admin("user", "edit", "alice", "--display-name", "Alice A.", "--active", "true")[0] == 0
    assert run("-u", "alice", "user", "edit", "alice", "--admin", "true")[0] == 1         # not by herself
    assert admin("user", "link-with-agent-user", "alice", "--agent-uid", "1200", "--agent-user", "alice", "--agent-gid", "1200", "--agent-group", "users")[0] == 0
    sess = api.authentication.login(m.url, "admin", "")
    assert sess.get("/api/v1/users/alice").json()["user"]["agentUserGroup"]["agentUid"] == 1200
    assert sess.get("/api/v1/users/alice").json()["user"]["displayName"] == "Alice A."
    # oauth clients
    rc, out = admin("oauth", "client", "add", "scim-sync", "https://idp.example.com")
    assert rc == 0 and "Client secret" in out
    cid = out.split("Client ID:")[1].split()[0]
    assert "scim-sync" in admin("oauth", "client", "list")[1]
    assert run("oauth", "client", "list")[0] == 1                                       # administrators only
    assert admin("oauth", "client", "remove", cid)[0] == 0 and "scim-sync" not in admin("oauth", "client", "ls")[1]
    assert "No SSO providers" in run("auth", "list-providers")[1]
    assert run("auth", "login", "alice", "--provider", "okta")[0] == 1
    rc, out = run("dev", "auth-token")
    assert rc == 0 and len(out.strip()) >= 32


def test_master_config_logs_and_accounting(cluster):
    m, tmp = cluster
    rc, out = run("master", "config")
    assert rc == 0 and "resource_pools" in out
    rc, out = run("master", "config", "show", "--json")
    assert {p["pool_name"] for p in json.loads(out)["resource_pools"]} == {"default", "aux"}
    assert run("master", "config", "set", "--log.level", "debug")[0] == 1                # administrators only
    assert admin("master", "config", "set", "--log.level", "debug")[0] == 0
    import logging

    assert logging.getLogger("determined").level == logging.DEBUG
    logging.getLogger("determined.master").info("a line for the ring buffer")
    rc, out = run("master", "logs", "--tail", "5")
    assert "a line for the ring buffer" in out
    admin("master", "config", "set", "--log.level", "info")
    # accounting: fabricate two finished allocations, one day apart
    now = time.time()
    day = 86400.0
    cfgf = tmp / "cfg.yaml"
    cfgf.write_text(yaml.safe_dump(CFG))
    assert run("experiment", "create", str(cfgf), str(FIX), "--paused")[0] == 0
    tid = m.db.exec("INSERT INTO trials(experiment_id,request_id,state,hparams,seed,start_time) VALUES(1,'r','COMPLETED','{}',0,?)", (now - 2 * day,))
    m.db.exec("INSERT INTO allocations(id,task_id,trial_id,state,slots,start_time,end_time,resource_pool) VALUES('a1',?,?,'TERMINATED',2,?,?,'default')",
              (f"trial-{tid}", tid, now - 2 * day, now - 2 * day + 100))
    m.db.exec("INSERT INTO allocations(id,task_id,trial_id,state,slots,start_time,end_time,resource_pool) VALUES('a2',?,?,'TERMINATED',1,?,?,'aux')",
              (f"trial-{tid}", tid, now - day, now - day + 50))
    rc, out = run("resources", "raw", str(now - 3 * day), str(now), "--json")
    rows = json.loads(out)
    assert [(r["allocationId"], r["slots"], round(r["seconds"])) for r in rows] == [("a1", 2, 200), ("a2", 1, 50)]
    assert rows[0]["username"] == "determined" and rows[0]["labels"] == ["nightly"] and rows[0]["experimentId"] == 1
    d0 = time.strftime("%Y-%m-%d", time.gmtime(now - 2 * day))
    d1 = time.strftime("%Y-%m-%d", time.gmtime(now))
    rc, out = run("resources", "aggregated", d0, d1, "--json")
    per_day = json.loads(out)
    assert len(per_day) == 3 and round(sum(e["seconds"] for e in per_day)) == 250
    assert round(per_day[0]["byResourcePool"]["default"]) + round(per_day[0]["byResourcePool"].get("aux", 0)) in (200, 250)
    rc, out = run("res", "agg", d0, d1)
    assert out.splitlines()[0] == "aggregation_type,aggregation_key,date,seconds" and any(l.startswith("experiment_label,nightly,") for l in out.splitlines())
    rc, out = run("resources", "aggregated", d0[:7], d1[:7], "--monthly", "--json")
    assert round(sum(e["seconds"] for e in json.loads(out))) == 250


def test_log_retention_and_tensorboard_files(cluster):
    m, tmp = cluster
    cfgf = tmp / "cfg.yaml"
    cfgf.write_text(yaml.safe_dump(CFG))
    assert run("experiment", "create", str(cfgf), str(FIX), "--paused")[0] == 0
    assert m.db.one("SELECT task_id FROM trials WHERE id=1")["task_id"] == "trial-1"      # the single searcher made its trial at creation
    now = time.time()
    m.db.add_logs([("trial-1", "a", now - 5 * 86400, "INFO", 0, "stdout", "task", "old line\n"), ("trial-1", "a", now - 3600, "INFO", 0, "stdout", "task", "fresh line\n")])
    assert run("trial", "set", "log-retention", "1", "--days", "2")[0] == 0
    assert run("task", "cleanup-logs")[0] == 1 and admin("task", "cleanup-logs")[0] == 0
    assert [r["log"] for r in m.db.get_logs("trial-1")] == ["fresh line\n"]
    assert run("experiment", "set", "log-retention", "1", "forever")[0] == 0
    assert run("experiment", "set-log-retention", "1", "--days", "30")[0] == 0
    assert run("trial", "set", "log-retention", "1")[0] == 1                             # neither --days nor --forever
    tb = tmp / "ck" / "tensorboard" / "experiment-1"
    tb.mkdir(parents=True)
    (tb / "events.out.tfevents.1").write_bytes(b"x")
    rc, out = run("experiment", "delete-tb-files", "1")
    assert rc == 0 and "Deleted 1 tensorboard file" in out and not tb.exists()
    # the support bundle holds the records and the logs
    rc, out = run("trial", "support-bundle", "1", "-o", str(tmp / "bundles"))
    assert rc == 0
    with tarfile.open(out.strip()) as tf:
        names = {n.split("/", 1)[1] for n in tf.getnames()}
        assert {"trial.json", "experiment_config.yaml", "metrics.json", "trial_logs.txt", "master.json"} <= names
        logs = tf.extractfile(next(n for n in tf.getnames() if n.endswith("trial_logs.txt"))).read().decode()
        assert "fresh line" in logs


def test_generic_tasks_context_pause_fork_and_ntsc_extras(cluster):
    m, tmp = cluster
    ctxdir = tmp / "ctx"
    ctxdir.mkdir()
    (ctxdir / "hello.txt").write_text("from the context directory")
    taskf = tmp / "task.yaml"
    taskf.write_text(yaml.safe_dump({"entrypoint": [sys.executable, "-c", "import time; print(open('hello.txt').read(), flush=True); time.sleep(60)"],
                                     "resources": {"slots": 0}, "description": "generic"}))
    rc, out = run("task", "create", str(taskf), "--context", str(ctxdir))
    assert rc == 0
    tid = out.split()[-1]
    sess = api.authentication.login(m.url, "determined", "")
    deadline = time.time() + 60
    while "from the context directory" not in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{tid}/logs").json()["logs"]):
        assert time.time() < deadline
        time.sleep(0.2)
    rc, out = run("task", "config", tid, "--json")
    assert json.loads(out)["description"] == "generic"
    assert run("task", "pause", tid)[0] == 0
    proc = m.commands.tasks[tid].proc
    assert open(f"/proc/{proc.pid}/stat").read().split(") ")[1][0] == "T"               # stopped
    assert run("task", "unpause", tid)[0] == 0
    time.sleep(0.2)
    assert open(f"/proc/{proc.pid}/stat").read().split(") ")[1][0] in "SR"
    rc, out = run("command", "config", tid)
    assert "generic" in out
    assert run("command", "set", "priority", tid, "7")[0] == 0
    assert sess.get(f"/api/v1/commands/{tid}").json()["command"]["config"]["resources"]["priority"] == 7
    rc, out = run("task", "fork", tid)
    fork_id = out.split()[2]
    assert rc == 0 and sess.get(f"/api/v1/commands/{fork_id}").json()["command"]["description"] == "generic"
    assert run("task", "pause", "no-such-task")[0] == 1
    for t in (tid, fork_id):
        run("command", "kill", t)
    rc, out = run("job", "update-batch", "nojob.priority=3", "nojob.weight=2.5")
    assert rc == 0 and "2 update" in out
    assert run("job", "update-batch", "garbage")[0] == 1
    tplf = tmp / "tpl.yaml"
    tplf.write_text("max_restarts: 3\n")
    assert run("template", "create", "t1", str(tplf))[0] == 0
    tplf.write_text("max_restarts: 9\n")
    assert run("template", "set-value", "config", "t1", str(tplf))[0] == 0
    assert sess.get("/api/v1/templates/t1").json()["template"]["config"] == {"max_restarts": 9}
    assert run("workspace", "create", "mw")[0] == 0 and run("model", "create", "mm")[0] == 0
    assert run("model", "move", "mm", "-w", "mw")[0] == 0
    assert sess.get("/api/v1/models/mm").json()["model"]["workspaceId"] == 2


def test_logs_filters_head_tail_and_json(cluster):
    m, tmp = cluster
    cfgf = tmp / "cfg.yaml"
    cfgf.write_text(yaml.safe_dump(CFG))
    assert run("experiment", "create", str(cfgf), str(FIX), "--paused")[0] == 0
    now = time.time()
    rows = [("trial-1", "alloc-a", now - 100 + i, "ERROR" if i % 5 == 0 else "INFO", i % 2, "stderr" if i % 5 == 0 else "stdout", "task", f"line {i}\n") for i in range(20)]
    rows.append(("trial-1", "alloc-b", now - 10, "WARNING", None, "stdout", "master", "from the master\n"))
    m.db.add_logs(rows)
    assert run("trial", "logs", "1")[1].count("\n") == 21
    assert run("trial", "logs", "1", "--head", "3")[1] == "line 0\nline 1\nline 2\n"
    assert run("trial", "logs", "1", "--tail", "2")[1] == "line 19\nfrom the master\n"
    assert run("trial", "logs", "1", "--level", "ERROR")[1] == "line 0\nline 5\nline 10\nline 15\n"
    assert run("experiment", "logs", "1", "--level", "ERROR", "--level", "WARNING", "--tail", "2")[1] == "line 15\nfrom the master\n"
    assert run("trial", "logs", "1", "--rank-id", "1", "--head", "2")[1] == "line 1\nline 3\n"
    assert run("trial", "logs", "1", "--source", "master")[1] == "from the master\n"
    assert run("trial", "logs", "1", "--stdtype", "stderr", "--rank-id", "0")[1] == "line 0\nline 10\n"
    assert run("trial", "logs", "1", "--allocation-id", "alloc-b")[1] == "from the master\n"
    assert run("trial", "logs", "1", "--search", "line 1", "--head", "2")[1] == "line 1\nline 10\n"
    assert run("trial", "logs", "1", "--timestamp-after", str(now - 50))[1] == "from the master\n"
    import datetime

    iso = datetime.datetime.fromtimestamp(now - 99.5, datetime.timezone.utc).isoformat().replace("+00:00", "Z")
    assert run("trial", "logs", "1", "--timestamp-before", iso)[1] == "line 0\n"
    rc, out = run("task", "logs", "trial-1", "--json", "--tail", "1")
    entry = json.loads(out)
    assert entry["log"] == "from the master\n" and entry["level"] == "LOG_LEVEL_WARNING" and entry["source"] == "master"


def test_command_run_ships_context_and_includes(cluster):
    m, tmp = cluster
    ctxdir, inc = tmp / "ctx2", tmp / "extra"
    ctxdir.mkdir()
    inc.mkdir()
    (ctxdir / "a.txt").write_text("A")
    (inc / "b.txt").write_text("B")
    (tmp / "c.txt").write_text("C")
    rc, out = run("cmd", "run", "--context", str(ctxdir), "--include", str(inc), "-i", str(tmp / "c.txt"), "-d", "--",
                  sys.executable, "-c", "print(open('a.txt').read() + open('extra/b.txt').read() + open('c.txt').read())")
    assert rc == 0
    tid = out.split("id: ")[1].split(")")[0]
    sess = api.authentication.login(m.url, "determined", "")
    deadline = time.time() + 60
    while sess.get(f"/api/v1/commands/{tid}").json()["command"]["state"] != "STATE_TERMINATED":
        assert time.time() < deadline
        time.sleep(0.2)
    assert "ABC" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{tid}/logs").json()["logs"])
    assert run("cmd", "list", "-q")[1].strip() == tid
    assert run("cmd", "run", "--include", str(tmp / "missing"), "--", "true")[0] == 1
    # a context entry that tries to leave the directory is refused by the master
    r = None
    try:
        sess.post("/api/v1/commands", json={"config": {"entrypoint": ["true"]}, "files": [{"path": "../evil", "type": 48, "content": ""}]})
    except api.errors.APIException as e:
        r = e
    deadline = time.time() + 30
    ts = sess.get("/api/v1/commands").json()["commands"]
    bad = [t for t in ts if t["id"] != tid]
    while bad and sess.get(f"/api/v1/commands/{bad[0]['id']}").json()["command"]["state"] != "STATE_TERMINATED":
        assert time.time() < deadline
        time.sleep(0.2)
    assert r is not None or (bad and not (pathlib.Path(m.work_dir).parent / "evil").exists() and not (pathlib.Path(m.work_dir) / "evil").exists())


def test_the_committed_gap_list_is_fresh():
    import subprocess

    root = pathlib.Path(__file__).resolve().parents[1]
    r = subprocess.run([sys.executable, str(root / "scripts" / "cli_gaps.py"), "--check"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
