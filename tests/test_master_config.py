"""The master's configuration: defaults < file < environment < flags, validation, secrets masking, and what the master does
with it (reference ``master/internal/config/config_test.go`` checks the same layering and ``Printable``)."""
import subprocess
import sys
import time

import pytest
import yaml

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import config, core


def test_layering_and_defaults(tmp_path):
    f = tmp_path / "master.yaml"
    f.write_text(yaml.safe_dump({"port": 9000, "cluster_name": "from-file", "db": {"path": str(tmp_path / "m.db")},
                                 "resource_manager": {"scheduler": {"type": "priority"}},
                                 "checkpoint_storage": {"type": "shared_fs", "host_path": "/ckpt"},
                                 "task_container_defaults": {"environment_variables": ["A=1", "B=two words"]},
                                 "security": {"agent_token": "s3cret", "initial_user_password": "hunter2"}}))
    c = config.load(str(f), environ={"DET_PORT": "9001", "DET_RESOURCE_MANAGER_SCHEDULER_FITTING_POLICY": "worst", "DET_NOTEBOOK_TIMEOUT": "120",
                                     "DET_SECURITY_TLS_NAME": "det.internal", "UNRELATED": "x"},
                    flags={"port": 9002, "host": None, "log.level": "debug"})
    assert c["port"] == 9002 and c["host"] == "127.0.0.1" and c["cluster_name"] == "from-file"          # flag > env > file > default
    assert c["resource_manager.scheduler"] == {"type": "priority", "fitting_policy": "worst", "preemption": {"enabled": True}, "default_priority": 42}
    assert c["notebook_timeout"] == 120 and c["tensorboard_timeout"] == 300 and c["security.tls.name"] == "det.internal"
    assert c["log.level"] == "debug" and c["retention_policy.log_retention_days"] == -1 and c["config_file"] == str(f)
    assert c.db_path == str(tmp_path / "m.db")
    assert c.cluster_defaults() == {"checkpoint_storage": {"type": "shared_fs", "host_path": "/ckpt"},
                                    "environment": {"environment_variables": ["A=1", "B=two words"]}}
    p = c.printable()
    assert p["security"]["agent_token"] == "********" and p["security"]["initial_user_password"] == "********"
    assert c["security.agent_token"] == "s3cret"                 # the live object keeps the values
    assert config.load(environ={})["port"] == 8080 and config.load(environ={}).db_path.endswith("master.db")


@pytest.mark.parametrize("raw,needle", [
    ({"port": "eighty"}, "<config>.port: must be of type integer"),
    ({"port": 70000}, "<config>.port"),
    ({"resource_manager": {"scheduler": {"type": "lottery"}}}, "scheduler.type: must be one of"),
    ({"no_such_key": 1}, "no_such_key"),
    ({"security": {"tls": {"cert": "/c.pem"}}}, "cert and key must be given together"),
    ({"resource_pools": [{"pool_name": "a"}, {"pool_name": "a"}]}, "pool names must be unique"),
    ({"resource_pools": [{"description": "nameless"}]}, "pool_name"),
    ({"task_container_defaults": {"environment_variables": ["NOEQUALS"]}}, "environment_variables"),
])
def test_mistakes_are_reported_with_their_key(raw, needle):
    with pytest.raises(config.MasterConfigError) as e:
        config.MasterConfig(raw)
    assert needle in str(e.value), str(e.value)


def test_print_config_and_bad_config_from_the_command_line(tmp_path):
    f = tmp_path / "master.yaml"
    f.write_text("port: 8123\nwebhooks: {signing_key: abc}\n")
    r = subprocess.run([sys.executable, "-m", "determined_b200.master", "--config-file", str(f), "--print-config", "--scheduler", "round_robin"],
                       capture_output=True, text=True, timeout=120)
    out = yaml.safe_load(r.stdout)
    assert r.returncode == 0 and out["port"] == 8123 and out["webhooks"]["signing_key"] == "********"
    assert out["resource_manager"]["scheduler"]["type"] == "round_robin"
    f.write_text("port: [1, 2]\n")
    r = subprocess.run([sys.executable, "-m", "determined_b200.master", "--config-file", str(f), "--print-config"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 2 and "<config>.port" in r.stderr


def test_master_applies_its_configuration(tmp_path):
    cfg = config.MasterConfig({"cluster_name": "cfg-cluster", "security": {"initial_user_password": "hunter2", "agent_token": "tok"},
                               "notebook_timeout": 77, "retention_policy": {"log_retention_days": 1}, "webhooks": {"signing_key": "k123"},
                               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}})
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"), cluster_defaults=cfg.cluster_defaults(), launch=False)
    m.apply_config(cfg)
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        # the built-in users got the initial password (and only while theirs was blank)
        with pytest.raises(api.errors.UnauthenticatedException):
            api.authentication.login(m.url, "admin", "")
        sess = api.authentication.login(m.url, "admin", "hunter2")
        assert sess.get("/api/v1/master").json()["clusterName"] == "cfg-cluster"
        shown = sess.get("/api/v1/master/config").json()["config"]
        assert shown["security"]["initial_user_password"] == "********" and shown["notebook_timeout"] == 77
        assert shown["resource_pools"][0]["pool_name"] == "default" and shown["checkpoint_storage"]["type"] == "shared_fs"
        assert sess.get("/api/v1/master/webhook_signing_key").json()["key"] == "k123"
        sess.post("/api/v1/users/1/password", json=api.authentication.salt_and_hash("changed"))      # travels hashed, like a login
        m.apply_config(cfg)                                                      # a restart does not reset a changed password
        api.authentication.login(m.url, "admin", "changed")
        assert m.agent_token == "tok" and m.commands.default_idle_timeout["NOTEBOOK"] == 77
        # log retention: lines older than a day go, newer ones stay
        now = time.time()
        m.db.add_logs([("t-old", "a", now - 3 * 86400, "INFO", None, "stdout", "task", "old\n"), ("t-new", "a", now - 60, "INFO", None, "stdout", "task", "new\n")])
        assert m.sweep_old_logs() == 1
        assert [r["log"] for r in m.db.get_logs("t-new")] == ["new\n"] and m.db.get_logs("t-old") == []
    finally:
        m.stop()
        srv.stop()
