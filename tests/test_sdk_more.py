"""The wider SDK object model (``experimental/_sdk_more.py``) against an in-process master without trial processes: metrics
and checkpoints are reported through the REST API the way a harness would."""
import pathlib
import time
import uuid

import pytest
import yaml

from determined_b200.common import api
from determined_b200.experimental import client
from determined_b200.master import api as mapi
from determined_b200.master import core

FIX = pathlib.Path(__file__).parent / "fixtures" / "onevar_exp"
CFG = {"name": "sdk-exp", "entrypoint": "model_def:OneVarTrial", "searcher": {"name": "random", "metric": "val_loss", "max_length": 20, "max_trials": 3, "max_concurrent_trials": 3},
       "hyperparameters": {"learning_rate": {"type": "double", "minval": 0.01, "maxval": 0.1}, "global_batch_size": 8}}


@pytest.fixture()
def det(tmp_path, monkeypatch):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=4, launch=False, work_dir=str(tmp_path / "w"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}},
                    resource_pools=[{"pool_name": "default"}, {"pool_name": "aux"}])
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "config"))
    monkeypatch.delenv("DET_SESSION_TOKEN", raising=False)
    d = client.Determined(m.url, "admin", "")
    yield d, m, tmp_path
    m.stop()
    srv.stop()


def test_experiment_trial_checkpoint_extras(det):
    d, m, tmp = det
    exp = d.create_experiment(CFG, str(FIX), activate=False)
    exp.set_labels(["b", "a", "a"])
    exp.set_notes("remember this")
    exp.reload()
    assert exp.labels == ["a", "b"] and d._session.get(f"/api/v1/experiments/{exp.id}").json()["experiment"]["notes"] == "remember this"
    code = exp.download_code(str(tmp / "code"))
    assert (pathlib.Path(code) / "model_def.py").exists()
    trials = exp.get_trials()
    assert len(trials) == 3 and [t.id for t in exp.iter_trials(limit=2)] == [t.id for t in trials]
    t = trials[0]
    sess = d._session
    for steps, v in ((10, 0.5), (20, 0.25)):
        sess.post(f"/api/v1/trials/{t.id}/metrics", json={"group": "validation", "metrics": {"trialId": t.id, "trialRunId": 0, "stepsCompleted": steps,
                                                                                                "metrics": {"avgMetrics": {"val_loss": v}}}})
    ck_uuid = str(uuid.uuid4())
    sess.post("/api/v1/checkpoints", json={"uuid": ck_uuid, "taskId": f"trial-{t.id}", "allocationId": None, "reportTime": "2024-01-01T00:00:00Z",
                                           "state": "STATE_COMPLETED", "resources": {"w.bin": "3"}, "metadata": {"steps_completed": 20}, "training": {}})
    ck = d.get_checkpoint(ck_uuid)
    got = list(ck.get_metrics("validation"))
    assert [g["steps_completed"] for g in got] == [20] and got[0]["metrics"] == {"val_loss": 0.25}
    assert [x["steps_completed"] for x in t.stream_metrics("validation")] == [10, 20]
    # logs with filters
    now = time.time()
    m.db.add_logs([(f"trial-{t.id}", "a", now - 5 + i, "ERROR" if i == 3 else "INFO", i % 2, "stdout", "task", f"l{i}\n") for i in range(6)])
    assert list(t.iter_logs(head=2)) == ["l0\n", "l1\n"] and list(t.iter_logs(tail=1)) == ["l5\n"]
    assert list(t.iter_logs(min_level="WARNING")) == ["l3\n"] and list(t.iter_logs(rank_ids=[1], search_text="l5")) == ["l5\n"]
    tb = tmp / "ck" / "tensorboard" / f"experiment-{exp.id}"
    tb.mkdir(parents=True)
    (tb / "events").write_bytes(b"x")
    exp.delete_tensorboard_files()
    assert not tb.exists()
    # model registry extras
    model = d.create_model("sdk-model", metadata={"a": 1, "b": 2})
    v = model.register_version(ck_uuid)
    v.set_name("first")
    v.set_notes("notes!")
    v.reload()
    assert (v.name, v.notes) == ("first", "notes!") and [g["steps_completed"] for g in v.iter_metrics("validation")] == [20]
    model.remove_metadata(["a"])
    model.reload()
    assert model.metadata == {"b": 2}
    ws = d.create_workspace("sdk-ws")
    model.move_to_workspace("sdk-ws")
    model.reload()
    assert model.workspace_id == ws.id
    assert [x.name for x in d.list_models(workspace_names=["sdk-ws"])] == ["sdk-model"] and d.list_models(workspace_ids=[1]) == []


def test_projects_workspaces_pools_users_oauth(det):
    d, m, tmp = det
    ws = d.create_workspace("team")
    p = ws.create_project("vision", "images")
    p.set_name("vision2")
    p.set_description("pictures")
    p.add_note("todo", "label more data")
    p.add_note("done", "cleaned set A")
    p.remove_note("todo")
    with pytest.raises(ValueError):
        p.remove_note("nope")
    p.reload()
    assert (p.name, p.description, p.notes) == ("vision2", "pictures", [{"name": "done", "contents": "cleaned set A"}])
    p.archive()
    p.reload()
    assert p.archived
    p.unarchive()
    pool = next(x for x in d.list_resource_pools() if x.name == "aux")
    pool.add_bindings(["team"])
    assert [w.name for w in pool.list_workspaces()] == ["team"]
    other = d.create_workspace("other")
    assert {x.name for x in ws.list_pools()} == {"default", "aux"} and {x.name for x in other.list_pools()} == {"default"}
    pool.replace_bindings(["other"])
    assert [w.name for w in pool.list_workspaces()] == ["other"]
    pool.remove_bindings(["other"])
    assert pool.list_workspaces() == []
    ws.set_default_pools(compute="aux")
    ws.reload()
    assert ws.default_compute_pool == "aux"
    ws.archive()
    ws.reload()
    assert ws.archived
    ws.unarchive()
    ws.delete_project("vision2")
    assert ws.list_projects() == []
    u = d.create_user("bob")
    u.link_with_agent(agent_uid=1500, agent_gid=1500, agent_user="bob", agent_group="staff")
    assert d._session.get("/api/v1/users/bob").json()["user"]["agentUserGroup"] == {"agentUid": 1500, "agentGid": 1500, "agentUser": "bob", "agentGroup": "staff"}
    c = d.add_oauth_client("https://idp.example.com", "idp")
    assert c.secret and [x.id for x in d.list_oauth_clients()] == [c.id]
    d.remove_oauth_client(c.id)
    assert d.list_oauth_clients() == []
    # module-level shortcuts use the default client
    client._determined = d
    assert "team" in [w.name for w in client.list_workspaces()] and client.get_session_username() == "admin"
    assert client.get_user_by_name("bob").username == "bob"


def test_reference_import_paths_and_metric_objects():
    """``determined.common.experimental.*`` modules resolve to the SDK classes; metric reports are objects with the
    reference's attribute names that still read like the dicts this SDK used to yield."""
    import determined_b200

    determined_b200.install_alias()
    from determined.common.experimental import checkpoint, experiment, metrics, model, trial  # type: ignore
    from determined.experimental import client  # type: ignore

    assert checkpoint.Checkpoint is client.Checkpoint and experiment.ExperimentReference is client.Experiment and trial.TrialSortBy is client.TrialSortBy
    assert model.ModelOrderBy is client.OrderBy and checkpoint.DownloadMode.AUTO.value == "auto"
    m = client.TrialMetrics._from_json({"trialId": 3, "trialRunId": 1, "totalBatches": 20, "endTime": "2026-01-01T00:00:00Z", "group": "validation",
                                        "metrics": {"avgMetrics": {"loss": 0.5}, "batchMetrics": None}})
    assert isinstance(m, metrics.ValidationMetrics) and (m.trial_id, m.steps_completed, m.metrics) == (3, 20, {"loss": 0.5})
    assert m["metrics"] == {"loss": 0.5} and m.get("group") == "validation" and m.get("nope", 7) == 7 and "trial_run_id" in m.keys()
    t = client.TrialMetrics._from_json({"trialId": 3, "totalBatches": 10, "group": "training", "metrics": {"avgMetrics": {}}})
    assert isinstance(t, metrics.TrainingMetrics) and not isinstance(t, metrics.ValidationMetrics)


def test_module_paths_of_the_reference_resolve_under_the_alias(tmp_path, monkeypatch):
    """Code written for the reference imports from its module layout (``determined.common.storage.s3``,
    ``determined.tensorboard.metric_writers.pytorch`` ...): those paths resolve to the modules that hold the names here."""
    import importlib
    import pathlib

    import determined_b200

    determined_b200.install_alias()
    ref = pathlib.Path("/root/reference/harness/determined")
    if not ref.is_dir():
        pytest.skip("reference tree not present")
    missing = []
    for f in sorted(ref.rglob("*.py")):
        mod = ".".join(f.relative_to(ref).with_suffix("").parts)
        mod = mod[:-len(".__init__")] if mod.endswith(".__init__") else mod
        if mod == "__init__" or mod.startswith(("keras", "deploy", "cli.")) or "tensorflow" in mod:
            continue
        try:
            importlib.import_module("determined." + mod)
        except ImportError:
            missing.append(mod)
    # private plumbing of the reference's own client stack (its logging setup, its websocket and requests wrappers, the Horovod shim)
    # (tensorboard.build is a function of the tensorboard package here: `from determined.tensorboard import build` is the spelling that works)
    assert set(missing) <= {"common._logging", "common.detlomond", "common.requests", "horovod", "tensorboard.build"}, missing
    from determined.common.storage.s3 import S3StorageManager  # type: ignore  # noqa: F401
    from determined.searcher._search_method import SearchMethod  # type: ignore  # noqa: F401
    from determined.tensorboard.metric_writers.pytorch import TorchWriter  # type: ignore

    monkeypatch.chdir(tmp_path)
    w = TorchWriter()
    w.add_scalar("x", 1.0, 1)
    w.reset()
    from determined_b200 import tensorboard as det_tb

    assert any(p.name.startswith("events.out.tfevents") for p in det_tb.get_base_path({}).rglob("*")), "written where the sync thread looks"


def test_api_logs_module_reads_and_filters(tmp_path):
    import time

    from determined_b200.common import api
    from determined_b200.common.api import bindings, logs
    from determined_b200.master import api as mapi
    from determined_b200.master import core

    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=0, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        sess = api.authentication.login(m.url, "admin", "")
        now = time.time()
        m.db.add_logs([("task-x", "task-x.1", now - 10 + i, "ERROR" if i == 2 else "INFO", i % 2, "stdout", "task", f"l{i}\n") for i in range(5)])
        got = list(logs.task_logs(sess, "task-x"))
        assert [e.log for e in got] == [f"l{i}\n" for i in range(5)] and isinstance(got[0], bindings.v1LogEntry)
        assert [e.log for e in logs.task_logs(sess, "task-x", tail=2)] == ["l3\n", "l4\n"]
        assert [e.log for e in logs.task_logs(sess, "task-x", head=2)] == ["l0\n", "l1\n"]
        assert [e.log for e in logs.task_logs(sess, "task-x", min_level=bindings.v1LogLevel.ERROR)] == ["l2\n"]
        assert [e.log for e in logs.task_logs(sess, "task-x", rank_ids=[1])] == ["l1\n", "l3\n"]
        assert [e.log for e in logs.task_logs(sess, "task-x", timestamp_after=now - 7.5)] == ["l3\n", "l4\n"]
        with pytest.raises(ValueError):
            list(logs.task_logs(sess, "task-x", timestamp_before="2021-10-26 23:17:12Z"))
        assert list(logs.task_logs(sess, "task-x", follow=True))[-1].log == "l4\n", "follow ends when the task is not running"
    finally:
        m.stop()
        srv.stop()


def test_public_helper_names_of_the_reference_exist_and_work(tmp_path, monkeypatch):
    from determined_b200 import constants, errors, experimental, tensorboard
    from determined_b200.common import api, storage
    from determined_b200.master import api as mapi
    from determined_b200.master import core
    from determined_b200.pytorch import experimental as px

    assert storage.normalize_prefix("/a/b//") == "a/b/" and storage.normalize_prefix(None) == ""
    with pytest.raises(ValueError):
        storage.normalize_prefix("a/../b")
    assert issubclass(errors.MultipleDownloadsFailed, Exception) and issubclass(api.errors.EmptyResultException, Exception)
    assert constants.HOROVOD_STARTUP_TIMEOUT_SECONDS > 0 and callable(px.get_default_device) and callable(experimental.test_one_batch)
    assert experimental.TrialState.RUNNING.value == "STATE_RUNNING" and experimental.ExperimentReference is experimental.Experiment
    mgr = tensorboard.build("c", 1, 2, {"type": "directory", "container_path": str(tmp_path / "tb")}, async_upload=False)
    assert isinstance(mgr, tensorboard.DirectoryTensorboardManager) and issubclass(tensorboard.S3TensorboardManager, tensorboard.TensorboardManager)

    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        sess = api.authentication.login(m.url, "admin", "")
        sess.post("/api/v1/users", json={"user": {"username": "zed"}})
        g = sess.post("/api/v1/groups", json={"name": "gg"}).json()["group"]["groupId"]
        w = sess.post("/api/v1/workspaces", json={"name": "ww"}).json()["workspace"]["id"]
        assert api.group_name_to_group_id(sess, "gg") == g and api.workspace_by_name(sess, "ww")["id"] == w
        assert api.usernames_to_user_ids(sess, ["admin", "zed"])[0] == 1 and api.role_name_to_role_id(sess, "Editor") > 0
        with pytest.raises(LookupError):
            api.usernames_to_user_ids(sess, ["nobody-here"])
        req = api.create_user_assignment_request(sess, "zed", "Viewer", "ww")
        assert req[0]["roleAssignment"]["scopeWorkspaceId"] == w
        pages = list(api.read_paginated(lambda off: sess.get("/api/v1/users", params={"offset": off, "limit": 2}).json()))
        assert sum(len(p["users"]) for p in pages) == 3 and len(pages) == 2
        assert len(list(api.read_paginated(lambda off: sess.get("/api/v1/users", params={"offset": off, "limit": 2}).json(), pages=api.PageOpts.single))) == 1
        cmd = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["sleep", "30"], "resources": {"slots": 0}}}).json()["command"]["id"]
        assert api.wait_for_ntsc_state(sess, api.NTSC_Kind.command, cmd, lambda s: s == "STATE_RUNNING", timeout_secs=20) == "STATE_RUNNING"
        api.wait_for_task_ready(sess, cmd, timeout_secs=20)
        assert api.get_ntsc_details(sess, api.NTSC_Kind.command, cmd)["id"] == cmd
        monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "cfg"))
        st = api.authentication.TokenStore(m.url)
        st.set_token("admin", sess.token)
        st.set_token("zed", "whatever")
        api.authentication.logout_all(m.url)
        assert api.authentication.TokenStore(m.url).get_all_users() == []
    finally:
        m.stop()
        srv.stop()


def test_sdk_signatures_of_the_reference_are_accepted(tmp_path):
    """``order_by`` / ``sort_by`` / ``max_results`` on checkpoint and version listings, filtered ``Trial.logs``, ``workspace_name``
    on ``create_model``, ``remote`` on ``create_user``: the reference's keyword arguments."""
    import time
    import uuid as _uuid

    from determined_b200.common import api
    from determined_b200.experimental import client
    from determined_b200.master import api as mapi
    from determined_b200.master import core

    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=0, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        sess = api.authentication.login(m.url, "admin", "")
        d = client.Determined._from_session(sess)
        cfg = {"name": "x", "entrypoint": "y", "searcher": {"name": "single", "metric": "val_loss", "max_length": 1}, "hyperparameters": {}}
        eid = sess.post("/api/v1/experiments", json={"config": __import__("json").dumps(cfg), "unmanaged": True}).json()["experiment"]["id"]
        tid = sess.post(f"/api/v1/experiments/{eid}/trials", json={"hparams": {}}).json()["trial"]["id"]
        for steps, loss in ((10, 0.5), (20, 0.2), (30, 0.4)):
            sess.post(f"/api/v1/trials/{tid}/metrics", json={"group": "validation", "metrics": {"stepsCompleted": steps, "metrics": {"avgMetrics": {"val_loss": loss, "acc": steps}}}})
            sess.post("/api/v1/checkpoints", json={"uuid": str(_uuid.uuid4()), "taskId": f"trial-{tid}", "allocationId": None, "reportTime": "2026-01-01T00:00:00Z",
                                                   "state": "STATE_COMPLETED", "resources": {}, "metadata": {"steps_completed": steps}, "training": {}})
        t, e = d.get_trial(tid), d.get_experiment(eid)
        steps_of = lambda cks: [c.metadata["steps_completed"] for c in cks]  # noqa: E731
        assert steps_of(t.list_checkpoints(sort_by=client.CheckpointSortBy.SEARCHER_METRIC, order_by=client.OrderBy.ASC)) == [20, 30, 10]
        assert steps_of(t.list_checkpoints(sort_by="acc", order_by=client.OrderBy.DESC, max_results=2)) == [30, 20]
        assert steps_of(t.get_checkpoints(sort_by=client.CheckpointSortBy.BATCH_NUMBER, order_by=client.OrderBy.DESC))[0] == 30
        with pytest.raises(AssertionError):
            t.list_checkpoints(sort_by="acc")
        assert steps_of(e.list_checkpoints(order_by=client.OrderBy.ASC, max_results=1)) == [20]
        assert steps_of(e.list_checkpoints(sort_by="acc", order_by=client.OrderBy.DESC, max_results=1)) == [30]
        now = time.time()
        m.db.add_logs([(f"trial-{tid}", "a", now - 5 + i, "ERROR" if i == 3 else "INFO", i % 2, "stdout", "task", f"l{i}\n") for i in range(6)])
        assert list(t.logs(min_level="ERROR")) == ["l3\n"] and list(t.logs(rank_ids=[0], tail=2)) == ["l2\n", "l4\n"]
        assert list(t.logs(search_text="l5")) == ["l5\n"] and list(t.logs(head=1)) == ["l0\n"]
        sess.post("/api/v1/workspaces", json={"name": "regws"})
        mod = d.create_model("sig-model", workspace_name="regws")
        assert mod.workspace_id != 1 if hasattr(mod, "workspace_id") else True
        assert sess.get("/api/v1/models/sig-model").json()["model"]["workspaceId"] != 1
        assert [v.model_version for v in mod.list_versions(order_by=client.OrderBy.ASC)] == []
        u = d.create_user("sso-only", remote=True)
        assert sess.get(f"/api/v1/users/{u.user_id if hasattr(u, 'user_id') else u.id}").json()["user"]["remote"] is True
    finally:
        m.stop()
        srv.stop()
