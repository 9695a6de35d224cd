"""Container runtimes of the agent: the command lines the drivers build, and a task run end to end through an agent whose
"docker" is a shell script that records its arguments and runs the task on the host."""
import json
import os
import stat
import sys
import threading
import time

import pytest

from determined_b200.agent import runtimes
from determined_b200.agent.agent import Agent
from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import core


def _spec(tmp_path, **kw):
    base = dict(name="det-abc", image="img:1", argv=["python3", "-m", "x"], env={"A": "1", "MULTI": "a\nb"}, cwd="/run/model",
                mounts=[runtimes.Mount("/run", "/run"), runtimes.Mount("/data", "/mnt/data", True, "rshared")], gpu_ids=[2, 3],
                shm_size=1 << 30, devices=[{"host_path": "/dev/infiniband", "container_path": "/dev/infiniband", "mode": "rwm"}],
                add_caps=["IPC_LOCK"], drop_caps=["MKNOD"], user="1000:1000", env_file=str(tmp_path / "task.env"), extra_args=["--ulimit", "memlock=-1"])
    base.update(kw)
    return runtimes.TaskSpec(**base)


def test_docker_command_line(tmp_path):
    cmd, env = runtimes.Docker().run(_spec(tmp_path))
    s = " ".join(cmd)
    assert cmd[:5] == ["docker", "run", "--rm", "--name", "det-abc"] and cmd[-4:] == ["img:1", "python3", "-m", "x"]
    assert "--network host" in s and "--ipc host" in s and "--workdir /run/model" in s
    assert '--gpus "device=2,3"' in s and "--shm-size 1073741824" in s
    assert "--mount type=bind,source=/data,target=/mnt/data,readonly,bind-propagation=rshared" in s
    assert "--device /dev/infiniband:/dev/infiniband:rwm" in s and "--cap-add IPC_LOCK" in s and "--cap-drop MKNOD" in s
    assert "--user 1000:1000" in s and "--ulimit memlock=-1" in s
    # the environment travels in a 0600 file, values with newlines through the client's environment
    assert (tmp_path / "task.env").read_text() == "A=1\n" and stat.S_IMODE(os.stat(tmp_path / "task.env").st_mode) == 0o600
    assert env == {"MULTI": "a\nb"} and "--env MULTI" in s and "A=1" not in s
    assert runtimes.Docker().signal("det-abc", "TERM") == ["docker", "kill", "--signal", "TERM", "det-abc"]
    login, pull = runtimes.Docker().pull("reg.io/team/img:1", False, {"username": "u", "password": "p"})
    assert login == ["docker", "login", "--username", "u", "--password-stdin", "reg.io"] and "p" not in login and pull == ["docker", "pull", "reg.io/team/img:1"]
    assert runtimes.Docker().pull("img", False) == [] and runtimes.Docker().pull("img", True) == [["docker", "pull", "img"]]


def test_podman_and_apptainer_command_lines(tmp_path):
    cmd, _ = runtimes.Podman().run(_spec(tmp_path, gpu_uuids=["GPU-aa", "GPU-bb"]))
    s = " ".join(cmd)
    assert cmd[:2] == ["podman", "run"] and "--device nvidia.com/gpu=GPU-aa --device nvidia.com/gpu=GPU-bb" in s and "--userns keep-id" in s and "--gpus" not in s
    cmd, env = runtimes.Apptainer().run(_spec(tmp_path))
    s = " ".join(cmd)
    assert cmd[:2] == ["apptainer", "exec"] and "--nv" in cmd and "--pwd /run/model" in s and "--bind /data:/mnt/data:ro" in s and "--writable-tmpfs" in cmd
    assert "docker://img:1" in cmd and env["APPTAINERENV_MULTI"] == "a\nb" and env["APPTAINERENV_CUDA_VISIBLE_DEVICES"] == "2,3"
    assert runtimes.Apptainer().signal("det-abc", "TERM") is None, "no daemon: the process group is the kill path"
    cmd, env = runtimes.Singularity().run(_spec(tmp_path, image="/images/torch.sif", gpu_ids=[]))
    assert cmd[0] == "singularity" and "/images/torch.sif" in cmd and "--nv" not in cmd and "SINGULARITYENV_MULTI" in env
    assert runtimes.get("process") is None and isinstance(runtimes.get("podman", "/opt/bin/podman"), runtimes.Podman)
    with pytest.raises(ValueError):
        runtimes.get("lxc")


def test_sizes_and_images():
    assert runtimes.parse_size("2g") == 2 << 30 and runtimes.parse_size("512MiB") == 512 << 20 and runtimes.parse_size(4096) == 4096 and runtimes.parse_size(None) is None
    assert runtimes.pick_image({"cuda": "g", "cpu": "c"}, True, "d") == "g" and runtimes.pick_image({"cuda": "g", "cpu": "c"}, False, "d") == "c"
    assert runtimes.pick_image(None, True, "d") == "d" and runtimes.pick_image("x", False, "d") == "x"


FAKE_DOCKER = """#!/bin/bash
# a stand-in for the docker client: record the call; `run` executes the task on the host with the env file applied
dir="$(dirname "$0")"
printf '%s\\n' "$*" >> "$dir/docker.calls"
case "$1" in
  pull|login|rm) exit 0 ;;
  kill) pid=$(cat "$dir/$4.pid"); kill "-$3" "$pid" 2>/dev/null; exit 0 ;;
  run)
    shift
    name=""; envfile=""; wd=""
    while [ $# -gt 0 ]; do
      case "$1" in
        --rm|--init) shift ;;
        --name) name="$2"; shift 2 ;;
        --env-file) envfile="$2"; shift 2 ;;
        --workdir) wd="$2"; shift 2 ;;
        --network|--ipc|--mount|--shm-size|--gpus|--env|--cap-add|--cap-drop|--user|--device) shift 2 ;;
        *) break ;;
      esac
    done
    image="$1"; shift
    echo "$image" > "$dir/image"
    set -a; . "$envfile"; set +a
    cd "$wd"
    [ "$1" = python3 ] && { shift; set -- "$FAKE_PYTHON" "$@"; }
    "$@" &
    echo $! > "$dir/$name.pid"
    wait $!
    ;;
esac
"""


def test_task_through_an_agent_with_a_container_runtime(tmp_path, monkeypatch):
    fake = tmp_path / "docker"
    fake.write_text(FAKE_DOCKER)
    fake.chmod(0o755)
    monkeypatch.setenv("FAKE_PYTHON", sys.executable)
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=0, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}})
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    ag = Agent(m.url, "box-1", slots=2, slot_type="cpu", work_dir=str(tmp_path / "agent"), container_runtime="docker",
               default_image="fallback:1", container_exe=str(fake), container_args=["--shm-size", "1g"])
    th = threading.Thread(target=ag.run, daemon=True)
    th.start()
    try:
        sess = api.authentication.login(m.url, "determined", "")
        for _ in range(100):
            if sess.get("/api/v1/agents").json()["agents"]:
                break
            time.sleep(0.1)
        assert sess.get("/api/v1/agents").json()["agents"][0]["version"].endswith("+docker")
        data = tmp_path / "data"
        data.mkdir()
        cfg = {"entrypoint": ["/bin/sh", "-c", "echo in $DET_AGENT_ID with $MYVAR; exit 5"], "resources": {"slots": 1, "shm_size": "2g"},
               "environment": {"image": {"cpu": "team/tools:7", "cuda": "team/cuda:7"}, "environment_variables": ["MYVAR=hello"], "force_pull_image": True,
                               "add_capabilities": ["SYS_PTRACE"]},
               "bind_mounts": [{"host_path": str(data), "container_path": "data", "read_only": True}]}
        cid = sess.post("/api/v1/commands", json={"config": cfg}).json()["command"]["id"]
        for _ in range(300):
            c = sess.get(f"/api/v1/commands/{cid}").json()["command"]
            if c["state"] == "STATE_TERMINATED":
                break
            time.sleep(0.1)
        logs = "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"])
        assert c["state"] == "STATE_TERMINATED" and c["exitStatus"] == 5, logs
        assert "in box-1 with hello" in logs and "docker: team/tools:7 as det-" in logs
        calls = (tmp_path / "docker.calls").read_text().splitlines()
        assert calls[0] == "pull team/tools:7"
        run = next(c for c in calls if c.startswith("run "))
        assert "--shm-size 2147483648" in run and "--cap-add SYS_PTRACE" in run and "--shm-size 1g team/tools:7" in run, run
        assert f"source={data},target=" in run and "/model/data,readonly" in run and "--gpus" not in run

        # kill goes through `docker kill`
        cid = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["/bin/sh", "-c", "echo up; exec sleep 600"], "resources": {"slots": 1}}}).json()["command"]["id"]
        for _ in range(200):
            if "up" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"]):
                break
            time.sleep(0.1)
        else:
            raise AssertionError("a lone line of output must reach the master without waiting for the next one")
        assert (tmp_path / "image").read_text().strip() == "fallback:1"
        sess.post(f"/api/v1/commands/{cid}/kill")
        for _ in range(200):
            if sess.get(f"/api/v1/commands/{cid}").json()["command"]["state"] == "STATE_TERMINATED":
                break
            time.sleep(0.1)
        assert sess.get(f"/api/v1/commands/{cid}").json()["command"]["state"] == "STATE_TERMINATED"
        assert any(c.startswith("kill --signal TERM det-") for c in (tmp_path / "docker.calls").read_text().splitlines())
    finally:
        ag.stop()
        m.stop()
        srv.stop()
