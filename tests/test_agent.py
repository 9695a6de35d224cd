"""Agent protocol: a master without local slots runs trials and commands on a remote agent (REST long-poll), kills them
through it, survives the agent restarting with a live task, and fails tasks of an agent that disappears."""
import os
import signal
import threading
import time

import pytest

from determined_b200.agent import Agent, detect_slots
from determined_b200.common import api
from determined_b200.master import agents as magents
from determined_b200.master import api as mapi
from determined_b200.master import core

from tests.test_master_e2e import BASE, create, dump_logs, wait_state


@pytest.fixture()
def cluster(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=0, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}},
                    scheduler_type="priority")
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    sess = api.authentication.login(m.url, "determined", "")
    started = []

    def agent(agent_id="agent-1", slots=2, work_dir=None):
        a = Agent(m.url, agent_id, slots=slots, slot_type="cpu", work_dir=work_dir or str(tmp_path / agent_id))
        th = threading.Thread(target=a.run, daemon=True, name=agent_id)
        th.start()
        started.append((a, th))
        return a, th

    yield m, sess, tmp_path, agent
    for a, th in started:
        a.stop()
    m.stop()
    srv.stop()


def _wait(cond, timeout=30.0, what="condition"):
    t0 = time.time()
    while time.time() - t0 < timeout:
        if cond():
            return
        time.sleep(0.1)
    raise AssertionError(f"timed out waiting for {what}")


def test_detect_slots_cpu():
    s = detect_slots(3, "cpu")
    assert len(s) == 3 and all(x["type"] == "DEVICE_TYPE_CPU" for x in s)


def test_trial_runs_on_remote_agent(cluster):
    m, sess, tmp, agent = cluster
    eid = create(sess, BASE)                       # no slots anywhere yet: queued
    time.sleep(1.0)
    assert sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"] in ("STATE_ACTIVE", "STATE_QUEUED", "STATE_RUNNING")
    a, _ = agent()
    _wait(lambda: any(x["id"] == "agent-1" for x in sess.get("/api/v1/agents").json()["agents"]), what="agent registration")
    ag = sess.get("/api/v1/agents/agent-1").json()["agent"]
    assert len(ag["slots"]) == 2 and ag["connected"] and ag["resourcePools"] == ["default"]
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    trials = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
    assert trials[0]["totalBatchesProcessed"] == 24
    logs = "".join(l["log"] for l in sess.get(f"/api/v1/trials/{trials[0]['id']}/logs").json()["logs"])
    assert logs.strip(), "task output must travel agent -> master"
    rp = sess.get("/api/v1/resource-pools").json()["resourcePools"][0]
    assert rp["slotsAvailable"] == 2 and rp["slotsUsed"] == 0
    assert a.runtime.running() == []


def test_kill_goes_through_the_agent(cluster):
    m, sess, tmp, agent = cluster
    a, _ = agent()
    cfg = dict(BASE, searcher={"name": "single", "metric": "val_loss", "max_length": {"batches": 100000}})
    eid = create(sess, cfg)
    _wait(lambda: a.runtime.running(), 60, "task start on the agent")
    sess.post(f"/api/v1/experiments/{eid}/kill")
    assert wait_state(sess, eid, {"STATE_CANCELED", "STATE_ERROR", "STATE_COMPLETED"}, timeout=60) == "STATE_CANCELED"
    _wait(lambda: not a.runtime.running(), 30, "task exit on the agent")


def test_command_runs_on_remote_agent(cluster):
    m, sess, tmp, agent = cluster
    a, _ = agent()
    _wait(lambda: a.registered, 30, "agent registration")
    r = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["/bin/sh", "-c", "echo agent=$DET_AGENT_ID; echo done"], "resources": {"slots": 0}}}).json()
    cid = r["command"]["id"]
    _wait(lambda: sess.get(f"/api/v1/commands/{cid}").json()["command"]["state"] == "STATE_TERMINATED", 30, "command exit")
    get_logs = lambda: "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"])  # noqa: E731
    _wait(lambda: "done" in get_logs(), 10, "the command's output in the task log")
    assert "agent=agent-1" in get_logs()


def test_agent_restart_reattaches_and_disconnect_fails_tasks(cluster, monkeypatch):
    m, sess, tmp, agent = cluster
    wd = str(tmp / "agent-wd")
    a, th = agent(work_dir=wd)
    r = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["/bin/sh", "-c", "sleep 300"], "resources": {"slots": 1}}}).json()
    cid = r["command"]["id"]
    _wait(lambda: a.runtime.running() == [cid], 30, "command start")
    pid = a.runtime.tasks[cid].pid
    # the agent process goes away but leaves its task running ...
    a.stop(kill_tasks=False)
    th.join(timeout=20)
    # ... and a new agent with the same id and work dir picks it up again
    b, _ = agent(work_dir=wd)
    _wait(lambda: b.runtime.running() == [cid] and b.registered, 30, "re-attachment")
    assert b.runtime.tasks[cid].pid == pid
    assert sess.get(f"/api/v1/commands/{cid}").json()["command"]["state"] == "STATE_RUNNING"
    # now the agent vanishes for good: the master notices the silence and fails the task
    monkeypatch.setattr(magents, "DISCONNECT_AFTER_S", 1.0)
    b.stop(kill_tasks=False)
    _wait(lambda: sess.get(f"/api/v1/commands/{cid}").json()["command"]["state"] == "STATE_TERMINATED", 30, "task failed after disconnect")
    assert not sess.get("/api/v1/agents/agent-1").json()["agent"]["connected"]
    os.kill(pid, signal.SIGKILL)


def test_resource_pools_route_work_to_the_agents_registered_in_them(tmp_path):
    """master.yaml `resource_pools`: a second pool with its own scheduler; an agent joins it; `resources.resource_pool` sends an
    experiment and a command there; unknown pools are refused."""
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}},
                    resource_pools=[{"pool_name": "aux", "description": "extra box", "scheduler": {"type": "priority", "fitting_policy": "worst"}}])
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    a = Agent(m.url, "aux-1", slots=2, slot_type="cpu", resource_pool="aux", work_dir=str(tmp_path / "aux-1"))
    th = threading.Thread(target=a.run, daemon=True)
    th.start()
    try:
        sess = api.authentication.login(m.url, "determined", "")
        _wait(lambda: any(x["id"] == "aux-1" for x in sess.get("/api/v1/agents").json()["agents"]), what="agent registration")
        pools = {p["name"]: p for p in sess.get("/api/v1/resource-pools").json()["resourcePools"]}
        assert set(pools) == {"default", "aux"} and pools["aux"]["slotsAvailable"] == 2 and pools["default"]["slotsAvailable"] == 1
        assert pools["aux"]["schedulerType"] == "SCHEDULER_TYPE_PRIORITY" and pools["aux"]["schedulerFittingPolicy"] == "FITTING_POLICY_WORST"
        assert sess.get("/api/v1/agents/aux-1").json()["agent"]["resourcePools"] == ["aux"]
        cfg = dict(BASE, resources={"slots_per_trial": 2, "resource_pool": "aux"})          # 2 slots: only the aux agent can hold it
        eid = create(sess, cfg)
        assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
        row = m.db.one("SELECT resource_pool FROM allocations WHERE trial_id IN (SELECT id FROM trials WHERE experiment_id=?)", (eid,))
        assert row["resource_pool"] == "aux"
        r = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["/bin/sh", "-c", "echo on=$DET_AGENT_ID"], "resources": {"slots": 0, "resource_pool": "aux"}}}).json()
        cid = r["command"]["id"]
        assert r["command"]["resourcePool"] == "aux"
        _wait(lambda: sess.get(f"/api/v1/commands/{cid}").json()["command"]["state"] == "STATE_TERMINATED", 30, "command exit")
        _wait(lambda: "on=aux-1" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"]), 10, "command output")
        with pytest.raises(api.errors.APIException, match="unknown resource pool"):
            create(sess, dict(BASE, resources={"slots_per_trial": 1, "resource_pool": "nope"}))
        with pytest.raises(api.errors.APIException):
            sess.post("/api/v1/commands", json={"config": {"entrypoint": "true", "resources": {"resource_pool": "nope"}}})
        assert {p["pool_name"] for p in sess.get("/api/v1/master/config").json()["config"]["resource_pools"]} == {"default", "aux"}
    finally:
        a.stop()
        m.stop()
        srv.stop()
