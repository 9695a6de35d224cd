"""``security.authz.type: rbac``: what a user may do depends on the roles they hold in a workspace (directly or through a
group); ``basic`` (the default) restricts nothing but the admin-only routes."""
import json

import pytest

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import config as mconfig
from determined_b200.master import core

from tests.test_master_e2e import BASE


@pytest.fixture()
def cluster(tmp_path):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, launch=False, work_dir=str(tmp_path / "w"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}})
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    admin = api.authentication.login(m.url, "admin", "")
    yield m, admin
    m.stop()
    srv.stop()


def _create(sess, project_id, name="e"):
    return sess.post("/api/v1/experiments", json={"config": json.dumps(dict(BASE, name=name)), "activate": False, "projectId": project_id,
                                                  "modelDefinition": []}).json()["experiment"]["id"]


def _role(admin, name):
    return next(r for r in admin.get("/api/v1/roles").json()["roles"] if r["name"] == name)


def _assign(sess, role, workspace_id, user_id=None, group_id=None):
    ra = {"role": {"roleId": role["roleId"]}, "scopeWorkspaceId": workspace_id}
    body = {"userRoleAssignments": [{"userId": user_id, "roleAssignment": ra}] if user_id else [],
            "groupRoleAssignments": [{"groupId": group_id, "roleAssignment": ra}] if group_id else []}
    return sess.post("/api/v1/roles/add-assignments", json=body)


def test_basic_mode_restricts_nothing(cluster):
    m, admin = cluster
    assert admin.get("/api/v1/master").json()["rbacEnabled"] is False
    admin.post("/api/v1/users", json={"user": {"username": "u"}})
    u = api.authentication.login(m.url, "u", "")
    ws = admin.post("/api/v1/workspaces", json={"name": "a"}).json()["workspace"]
    pid = admin.post(f"/api/v1/workspaces/{ws['id']}/projects", json={"name": "p"}).json()["project"]["id"]
    assert _create(u, pid) and u.get(f"/api/v1/workspaces/{ws['id']}").json()["workspace"]["name"] == "a"


def test_roles_decide_under_rbac(cluster):
    m, admin = cluster
    m.apply_config(mconfig.load(None, {}, {"security.authz.type": "rbac"}))
    assert admin.get("/api/v1/master").json()["rbacEnabled"] is True
    ids = {}
    for name in ("viewer", "editor", "wsadmin", "nobody", "member"):
        ids[name] = admin.post("/api/v1/users", json={"user": {"username": name}}).json()["user"]["id"]
    s = {name: api.authentication.login(m.url, name, "") for name in ids}
    wa = admin.post("/api/v1/workspaces", json={"name": "team-a"}).json()["workspace"]["id"]
    wb = admin.post("/api/v1/workspaces", json={"name": "team-b"}).json()["workspace"]["id"]
    pa = admin.post(f"/api/v1/workspaces/{wa}/projects", json={"name": "pa"}).json()["project"]["id"]
    pb = admin.post(f"/api/v1/workspaces/{wb}/projects", json={"name": "pb"}).json()["project"]["id"]
    ea, eb = _create(admin, pa, "in-a"), _create(admin, pb, "in-b")
    _assign(admin, _role(admin, "Viewer"), wa, user_id=ids["viewer"])
    _assign(admin, _role(admin, "Editor"), wa, user_id=ids["editor"])
    _assign(admin, _role(admin, "WorkspaceAdmin"), wa, user_id=ids["wsadmin"])
    g = admin.post("/api/v1/groups", json={"name": "a-team", "addUsers": [ids["member"]]}).json()["group"]["groupId"]
    _assign(admin, _role(admin, "Editor"), wa, group_id=g)

    def denied(fn):
        with pytest.raises(api.errors.APIException) as e:
            fn()
        assert e.value.status_code == 403 and "access denied" in e.value.message, e.value.message
        return True

    # nobody: no role anywhere
    assert denied(lambda: s["nobody"].get(f"/api/v1/experiments/{ea}")) and denied(lambda: _create(s["nobody"], pa))
    assert s["nobody"].get("/api/v1/workspaces").json()["workspaces"] == [] and s["nobody"].get("/api/v1/experiments").json()["experiments"] == []
    assert denied(lambda: s["nobody"].post("/api/v1/workspaces", json={"name": "mine"}))
    assert s["nobody"].get("/api/v1/me").json()["user"]["username"] == "nobody", "routes about oneself stay open"

    # viewer: reads in A, nothing in B, no writes
    assert s["viewer"].get(f"/api/v1/experiments/{ea}").json()["experiment"]["name"] == "in-a"
    assert [w["name"] for w in s["viewer"].get("/api/v1/workspaces").json()["workspaces"]] == ["team-a"]
    assert [e["id"] for e in s["viewer"].get("/api/v1/experiments").json()["experiments"]] == [ea]
    assert [e["experiment"]["id"] for e in s["viewer"].get("/api/v1/experiments-search").json()["experiments"]] == [ea]
    assert denied(lambda: s["viewer"].get(f"/api/v1/experiments/{eb}")) and denied(lambda: s["viewer"].get(f"/api/v1/workspaces/{wb}"))
    assert denied(lambda: _create(s["viewer"], pa)) and denied(lambda: s["viewer"].patch(f"/api/v1/experiments/{ea}", json={"description": "x"}))
    assert denied(lambda: s["viewer"].post("/api/v1/experiments/kill", json={"experimentIds": [ea]}))

    # editor (directly, or through a group): creates and changes in A, not in B; cannot change the workspace itself
    for who in ("editor", "member"):
        mine = _create(s[who], pa, f"by-{who}")
        assert s[who].patch(f"/api/v1/experiments/{mine}", json={"description": "changed"}).json()["experiment"]["description"] == "changed"
        assert s[who].post("/api/v1/experiments/archive", json={"experimentIds": [mine]}).json()["results"][0]["id"] == mine
        assert denied(lambda: _create(s[who], pb)) and denied(lambda: s[who].patch(f"/api/v1/workspaces/{wa}", json={"name": "renamed"}))
        assert denied(lambda: s[who].post("/api/v1/experiments/move", json={"experimentIds": [mine], "destinationProjectId": pb}))
    assert s["editor"].post("/api/v1/models", json={"name": "m-a", "workspaceId": wa}).json()["model"]["workspaceId"] == wa
    assert denied(lambda: s["editor"].post("/api/v1/models", json={"name": "m-b", "workspaceId": wb}))
    assert s["editor"].post(f"/api/v1/workspaces/{wa}/projects", json={"name": "p2"}).json()["project"]["workspaceId"] == wa
    assert denied(lambda: s["editor"].delete(f"/api/v1/workspaces/{wa}"))

    # workspace admin: edits the workspace and hands out roles inside it, and only there
    assert s["wsadmin"].patch(f"/api/v1/workspaces/{wa}", json={"name": "team-a2"}).json()["workspace"]["name"] == "team-a2"
    assert _assign(s["wsadmin"], _role(admin, "Viewer"), wa, user_id=ids["nobody"]).status_code == 200
    assert s["nobody"].get(f"/api/v1/experiments/{ea}").json()["experiment"]["id"] == ea
    with pytest.raises(api.errors.APIException):
        _assign(s["wsadmin"], _role(admin, "Viewer"), wb, user_id=ids["nobody"])
    with pytest.raises(api.errors.APIException):
        _assign(s["wsadmin"], _role(admin, "ClusterAdmin"), None, user_id=ids["wsadmin"])

    # a cluster-wide role lifts every restriction; so does the session of a task
    _assign(admin, _role(admin, "ClusterAdmin"), None, user_id=ids["viewer"])
    assert s["viewer"].get(f"/api/v1/experiments/{eb}").json()["experiment"]["id"] == eb
    m.db.exec("INSERT INTO sessions(token,user_id,expiry,task_id) VALUES('task-token',?,?,?)", (ids["nobody"], 9e12, "trial-1"))
    harness = api.Session(m.url, "nobody", "task-token")
    assert harness.get(f"/api/v1/experiments/{eb}").json()["experiment"]["id"] == eb
    assert s["nobody"].get("/api/v1/permissions/summary").status_code == 200


def test_strict_job_queue_control_reserves_the_queue_for_administrators(cluster):
    m, admin = cluster
    admin.post("/api/v1/users", json={"user": {"username": "u"}})
    u = api.authentication.login(m.url, "u", "")
    eid = _create(u, 1)
    assert u.post("/api/v1/job-queues", json={"updates": [{"jobId": f"exp-{eid}", "priority": 10}]}).status_code == 200
    m.apply_config(mconfig.load(None, {}, {"security.authz.strict_job_queue_control": True}))
    assert admin.get("/api/v1/master").json()["strictJobQueueControl"] is True
    with pytest.raises(api.errors.APIException) as e:
        u.post("/api/v1/job-queues", json={"updates": [{"jobId": f"exp-{eid}", "priority": 1}]})
    assert e.value.status_code == 403
    assert admin.post("/api/v1/job-queues", json={"updates": [{"jobId": f"exp-{eid}", "priority": 1}]}).status_code == 200


def test_api_requests_leave_audit_lines(cluster, caplog):
    import logging

    m, admin = cluster
    with caplog.at_level(logging.DEBUG, logger="determined.master.audit"):
        admin.post("/api/v1/workspaces", json={"name": "audited"})
        admin.get("/api/v1/workspaces")
        with pytest.raises(api.errors.APIException):
            api.Session(m.url, "x", "bad-token").get("/api/v1/me")
        import time

        for _ in range(60):       # a line is written after its answer has left: wait (with the level still lowered) for the last one
            lines = [(r.levelname, r.getMessage()) for r in caplog.records if r.name == "determined.master.audit"]
            if sum("status=401" in msg for _, msg in lines):
                break
            time.sleep(0.05)
    assert ("INFO", "audit: user=admin remote_ip=127.0.0.1 method=POST path=/api/v1/workspaces status=200 unauthorized=False") in lines
    assert ("DEBUG", "audit: user=admin remote_ip=127.0.0.1 method=GET path=/api/v1/workspaces status=200 unauthorized=False") in lines
    assert ("INFO", "audit: user=None remote_ip=127.0.0.1 method=GET path=/api/v1/me status=401 unauthorized=True") in lines
    # and they are part of what `det master logs` shows
    assert any("audit: user=admin" in e["message"] and "POST" in e["message"] for e in admin.get("/api/v1/master/logs", params={"limit": 1000}).json()["logs"])
