"""``core_v2``: unmanaged runs land in the named project, ship what they print as trial logs, write TensorBoard event files
next to their checkpoints, resume by external ids, and close the trial (reference ``e2e_tests/tests/cluster/test_unmanaged.py``
runs the same scenarios from example scripts)."""
import pathlib
import time

import pytest

from determined_b200.common import api
from determined_b200.experimental import client, core_v2
from determined_b200.master import api as mapi
from determined_b200.master import core


@pytest.fixture()
def cluster(tmp_path, monkeypatch):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, launch=False, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    monkeypatch.setenv("DET_MASTER", m.url)
    monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "config"))
    monkeypatch.delenv("DET_SESSION_TOKEN", raising=False)
    sess = api.authentication.login(m.url, "determined", "")
    yield m, sess, tmp_path
    core_v2.close()
    m.stop()
    srv.stop()


def test_unmanaged_run_in_a_project_with_logs_and_tensorboard(cluster, capsys):
    m, sess, tmp = cluster
    d = client.Determined._from_session(sess)
    ws = d.create_workspace("detached")
    proj = ws.create_project("runs")
    ck_dir = tmp / "ck"
    core_v2.init(defaults=core_v2.DefaultConfig(name="detached-run", hparams={"lr": 0.1}, labels=["v2"], description="from a script",
                                                searcher={"name": "single", "metric": "loss", "max_length": 4}, checkpoint_storage=str(ck_dir)),
                 unmanaged=core_v2.UnmanagedConfig(workspace="detached", project="runs", external_experiment_id="exp-A", external_trial_id="trial-A"),
                 client=d)
    assert core_v2.info.trial.experiment_id == 1 and core_v2.train is not None and core_v2.checkpoint is not None
    print("hello from the detached script")
    for i in range(4):
        core_v2.train.report_training_metrics(steps_completed=i + 1, metrics={"loss": 1.0 / (i + 1)})
    core_v2.train.report_validation_metrics(steps_completed=4, metrics={"loss": 0.2})
    with core_v2.checkpoint.store_path({"steps_completed": 4}) as (p, uuid_):
        (p / "weights.bin").write_bytes(b"1234")
    eid = core_v2.info.trial.experiment_id
    tid = core_v2.info.trial.trial_id
    assert core_v2.url_reverse_webui_exp_view().endswith(f"/det/experiments/{eid}")
    core_v2.close()
    assert core_v2.train is None
    exp = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]
    assert exp["projectId"] == proj.id and exp["unmanaged"] and exp["labels"] == ["v2"] and exp["description"] == "from a script"
    trial = sess.get(f"/api/v1/trials/{tid}").json()["trial"]
    assert trial["state"] == "STATE_COMPLETED" and trial["totalBatchesProcessed"] == 4 and trial["latestCheckpoint"] == uuid_
    deadline = time.time() + 10
    while True:
        logs = "".join(l["log"] for l in sess.get(f"/api/v1/trials/{tid}/logs").json()["logs"])
        if "hello from the detached script" in logs or time.time() > deadline:
            break
        time.sleep(0.2)
    assert "hello from the detached script" in logs
    # event files were synced next to the checkpoints
    events = [p for p in ck_dir.rglob("*tfevents*")]
    assert events, list(ck_dir.rglob("*"))[:20]
    # running the same script again with the same external ids continues the same trial from its checkpoint
    core_v2.init(defaults=core_v2.DefaultConfig(name="detached-run", hparams={"lr": 0.1}, checkpoint_storage=str(ck_dir)),
                 unmanaged=core_v2.UnmanagedConfig(workspace="detached", project="runs", external_experiment_id="exp-A", external_trial_id="trial-A"), client=d)
    assert core_v2.info.trial.trial_id == tid and core_v2.info.latest_checkpoint == uuid_
    with core_v2.checkpoint.restore_path(uuid_) as p:
        assert (pathlib.Path(p) / "weights.bin").read_bytes() == b"1234"
    core_v2.close()
    # a second trial of the same experiment (a client-side hyperparameter search)
    core_v2.init(defaults=core_v2.DefaultConfig(name="detached-run", hparams={"lr": 0.2}, checkpoint_storage=str(ck_dir)),
                 unmanaged=core_v2.UnmanagedConfig(workspace="detached", project="runs", external_experiment_id="exp-A", external_trial_id="trial-B"), client=d)
    assert core_v2.info.trial.experiment_id == eid and core_v2.info.trial.trial_id != tid
    core_v2.close()
    assert len(sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]) == 2


def test_unmanaged_config_mistakes(cluster):
    m, sess, tmp = cluster
    d = client.Determined._from_session(sess)
    with pytest.raises(ValueError, match="external_experiment_id"):
        core_v2.init(unmanaged=core_v2.UnmanagedConfig(external_trial_id="t"), client=d)
    with pytest.raises(ValueError, match="does not exist"):
        core_v2.init(unmanaged=core_v2.UnmanagedConfig(workspace="nope"), client=d)
    with pytest.raises(ValueError, match="no project named"):
        core_v2.init(unmanaged=core_v2.UnmanagedConfig(workspace="Uncategorized", project="nope"), client=d)
    core_v2.init(defaults=core_v2.DefaultConfig(name="ok"), client=d, checkpoint_storage=str(tmp / "ck2"))
    with pytest.raises(RuntimeError, match="already called"):
        core_v2.init(client=d)
    core_v2.close()
