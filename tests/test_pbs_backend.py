"""PBS pool: tasks become qsub jobs (qsub / qstat / qdel faked by three shell scripts that run the job script as a local
process), output and exit codes come back through the shared run directory, kill goes through qdel."""
import stat
import time

import pytest

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import core

from tests.test_master_e2e import BASE, create, dump_logs, wait_state

QSUB = """#!/bin/bash
# qsub <script>: run it detached; the job id is "<pid>.fakepbs"
script="$1"
echo "$@" >> "$(dirname "$0")/qsub.calls"
setsid bash "$script" > /dev/null 2>&1 < /dev/null &
echo "$!.fakepbs"
"""
QSTAT = """#!/bin/bash
# qstat -x id id ...
shift
echo "Job id            Name             User              Time Use S Queue"
echo "----------------  ---------------- ----------------  -------- - -----"
for j in "$@"; do
  pid="${j%%.*}"
  if kill -0 "$pid" 2>/dev/null; then s=R; else s=F; fi
  echo "$j  det-job  root  00:00:01 $s gpu"
done
"""
QDEL = """#!/bin/bash
pid="${1%%.*}"
echo "$1" >> "$(dirname "$0")/qdel.calls"
kill -TERM -- "-$pid" 2>/dev/null || kill -TERM "$pid" 2>/dev/null
exit 0
"""


@pytest.fixture()
def cluster(tmp_path):
    bin_dir = tmp_path / "bin"
    bin_dir.mkdir()
    for name, text in (("qsub", QSUB), ("qstat", QSTAT), ("qdel", QDEL)):
        p = bin_dir / name
        p.write_text(text)
        p.chmod(p.stat().st_mode | stat.S_IEXEC)
    prov = {"type": "pbs", "slots": 4, "queue": "gpu", "pbsbatch_args": ["-l walltime=00:10:00"], "poll_seconds": 0.3, "exit_file_grace_seconds": 1.0,
            "qsub": str(bin_dir / "qsub"), "qstat": str(bin_dir / "qstat"), "qdel": str(bin_dir / "qdel")}
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=0, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}},
                    resource_pools=[{"pool_name": "hpc", "provider": prov}])
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    yield m, api.authentication.login(m.url, "determined", ""), tmp_path
    m.stop()
    srv.stop()


def test_trial_runs_as_a_pbs_job(cluster):
    m, sess, tmp = cluster
    pools = {p["name"]: p for p in sess.get("/api/v1/resource-pools").json()["resourcePools"]}
    assert pools["hpc"]["slotsAvailable"] == 4 and pools["hpc"]["type"] == "RESOURCE_POOL_TYPE_STATIC"
    assert pools["hpc"]["schedulerType"] == "SCHEDULER_TYPE_PBS"
    eid = create(sess, dict(BASE, resources={"slots_per_trial": 2, "resource_pool": "hpc"}, pbs={"pbsbatch_args": ["-l place=scatter"]}))
    assert wait_state(sess, eid, {"STATE_COMPLETED", "STATE_ERROR"}) == "STATE_COMPLETED", dump_logs(sess, eid)
    t = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][0]
    assert t["totalBatchesProcessed"] == 24
    logs = "".join(l["log"] for l in sess.get(f"/api/v1/trials/{t['id']}/logs").json()["logs"])
    assert "submitted PBS job" in logs and len(logs.splitlines()) > 3, "the job's own output must come back through pbs.out"
    job_sh = next((tmp / "work" / "pbs-hpc").glob("job-*/job.sh")).read_text()
    assert "#PBS -q gpu" in job_sh and "#PBS -l select=1:ngpus=2" in job_sh and "#PBS -l walltime=00:10:00" in job_sh and "#PBS -j oe" in job_sh
    assert "#PBS -l place=scatter" in job_sh, "the experiment's own pbs section adds directives"
    assert "export DET_SLOT_IDS='[0, 1]'" in job_sh and "determined_b200.exec.launch" in job_sh
    assert "#SBATCH" not in job_sh


def test_command_exit_code_and_kill_through_qdel(cluster):
    m, sess, tmp = cluster
    cfg = {"entrypoint": ["/bin/sh", "-c", "echo hello from $DET_AGENT_ID; exit 7"], "resources": {"slots": 0, "resource_pool": "hpc"}}
    cid = sess.post("/api/v1/commands", json={"config": cfg}).json()["command"]["id"]
    for _ in range(200):
        c = sess.get(f"/api/v1/commands/{cid}").json()["command"]
        if c["state"] == "STATE_TERMINATED":
            break
        time.sleep(0.1)
    assert c["state"] == "STATE_TERMINATED" and c["exitStatus"] == 7
    assert "hello from pbs-hpc" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"])

    cfg = {"entrypoint": ["/bin/sh", "-c", "echo up; sleep 600"], "resources": {"slots": 1, "resource_pool": "hpc"}}
    cid = sess.post("/api/v1/commands", json={"config": cfg}).json()["command"]["id"]
    for _ in range(200):
        if "up" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"]):
            break
        time.sleep(0.1)
    sess.post(f"/api/v1/commands/{cid}/kill")
    for _ in range(200):
        c = sess.get(f"/api/v1/commands/{cid}").json()["command"]
        if c["state"] == "STATE_TERMINATED":
            break
        time.sleep(0.1)
    assert c["state"] == "STATE_TERMINATED"
    assert (tmp / "bin" / "qdel.calls").read_text().strip().endswith(".fakepbs")
