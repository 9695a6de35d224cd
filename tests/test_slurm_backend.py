"""Slurm pool: tasks become batch jobs (sbatch / squeue / scancel faked by three shell scripts that run the job script as a
local process), output and exit codes come back through the shared run directory, kill goes through scancel."""
import os
import stat
import time

import pytest

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import core

from tests.test_master_e2e import BASE, create, dump_logs, wait_state

SBATCH = """#!/bin/bash
# sbatch --parsable <script>: run it detached, the "job id" is its pid
script="$2"
echo "$@" >> "$(dirname "$0")/sbatch.calls"
out=$(grep -m1 '^#SBATCH --output=' "$script" | cut -d= -f2)
setsid bash "$script" >> "$out" 2>&1 < /dev/null &
echo $!
"""
SQUEUE = """#!/bin/bash
# squeue -h -j id,id -o "%i %T"
ids=$(echo "$3" | tr ',' ' ')
for i in $ids; do if kill -0 "$i" 2>/dev/null; then echo "$i RUNNING"; fi; done
"""
SCANCEL = """#!/bin/bash
kill -TERM -- "-$1" 2>/dev/null || kill -TERM "$1" 2>/dev/null
exit 0
"""


@pytest.fixture()
def cluster(tmp_path):
    bin_dir = tmp_path / "bin"
    bin_dir.mkdir()
    for name, text in (("sbatch", SBATCH), ("squeue", SQUEUE), ("scancel", SCANCEL)):
        p = bin_dir / name
        p.write_text(text)
        p.chmod(p.stat().st_mode | stat.S_IEXEC)
    prov = {"type": "slurm", "slots": 4, "partition": "gpu", "sbatch_args": ["--time=00:10:00"], "poll_seconds": 0.3, "exit_file_grace_seconds": 1.0,
            "sbatch": str(bin_dir / "sbatch"), "squeue": str(bin_dir / "squeue"), "scancel": str(bin_dir / "scancel")}
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=0, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}},
                    resource_pools=[{"pool_name": "hpc", "provider": prov}])
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    yield m, api.authentication.login(m.url, "determined", ""), tmp_path
    m.stop()
    srv.stop()


def test_trial_runs_as_a_batch_job(cluster):
    m, sess, tmp = cluster
    pools = {p["name"]: p for p in sess.get("/api/v1/resource-pools").json()["resourcePools"]}
    assert pools["hpc"]["slotsAvailable"] == 4
    eid = create(sess, dict(BASE, resources={"slots_per_trial": 2, "resource_pool": "hpc"}, slurm={"sbatch_args": ["--exclusive"]}))
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    t = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]
    assert t["totalBatchesProcessed"] == 24
    logs = "".join(l["log"] for l in sess.get(f"/api/v1/trials/{t['id']}/logs").json()["logs"])
    assert "submitted Slurm job" in logs and len(logs.splitlines()) > 3, "the job's own output must come back through slurm.out"
    job_sh = next((tmp / "work" / "slurm-hpc").glob("job-*/job.sh")).read_text()
    assert "#SBATCH --partition=gpu" in job_sh and "#SBATCH --gres=gpu:2" in job_sh and "#SBATCH --time=00:10:00" in job_sh
    assert "#SBATCH --exclusive" in job_sh, "the experiment's own slurm section adds directives"
    assert "export DET_SLOT_IDS='[0, 1]'" in job_sh and "determined_b200.exec.launch" in job_sh
    assert "--parsable" in (tmp / "bin" / "sbatch.calls").read_text()


def test_command_exit_code_and_kill_through_scancel(cluster):
    m, sess, tmp = cluster
    cfg = {"entrypoint": ["/bin/sh", "-c", "echo hello from $DET_AGENT_ID; exit 7"], "resources": {"slots": 0, "resource_pool": "hpc"}}
    cid = sess.post("/api/v1/commands", json={"config": cfg}).json()["command"]["id"]
    for _ in range(200):
        c = sess.get(f"/api/v1/commands/{cid}").json()["command"]
        if c["state"] == "STATE_TERMINATED":
            break
        time.sleep(0.1)
    assert c["state"] == "STATE_TERMINATED" and c["exitStatus"] == 7
    assert "hello from slurm-hpc" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"])
    long = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["/bin/sh", "-c", "sleep 300"], "resources": {"slots": 1, "resource_pool": "hpc"}}}).json()["command"]["id"]
    be = m.backends["hpc"]
    for _ in range(100):
        if long in be.jobs:
            break
        time.sleep(0.1)
    pid = int(be.jobs[long].job_id)
    sess.post(f"/api/v1/commands/{long}/kill")
    for _ in range(200):
        if sess.get(f"/api/v1/commands/{long}").json()["command"]["state"] == "STATE_TERMINATED":
            break
        time.sleep(0.1)
    assert sess.get(f"/api/v1/commands/{long}").json()["command"]["state"] == "STATE_TERMINATED"
    with pytest.raises(ProcessLookupError):
        for _ in range(50):
            os.kill(pid, 0)
            time.sleep(0.1)
