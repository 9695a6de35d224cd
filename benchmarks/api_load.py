"""REST load test of the master (reference ``performance/src/api_performance_tests.ts``: a k6 script that walks the read
endpoints the web UI uses with a number of virtual users and checks per-endpoint p95 latency and the failure rate).

Here: an in-process master is seeded (experiments with trials, metrics, checkpoints and logs; a workspace / project / model
/ group; a command task), then ``--vus`` threads walk the same list of requests for ``--seconds``.  Per endpoint: count,
failures, p50 / p95 / max latency; the run fails when any request fails or a p95 exceeds ``--p95-ms``.

    python benchmarks/api_load.py [--vus 8] [--seconds 10] [--experiments 20] [--out profiles/api_load.json]
"""
import argparse
import json
import pathlib
import sys
import tempfile
import threading
import time
import urllib.parse
import uuid
from typing import Any, Dict, List, Tuple

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))

from determined_b200.common import api  # noqa: E402
from determined_b200.master import api as mapi  # noqa: E402
from determined_b200.master import core  # noqa: E402

CFG = {"name": "load", "entrypoint": "model_def:Trial", "searcher": {"name": "single", "metric": "validation_error", "max_length": {"batches": 100}},
       "hyperparameters": {"lr": 0.1, "layers": {"type": "int", "minval": 1, "maxval": 4}}, "resources": {"slots_per_trial": 1}}


def seed(m: core.Master, sess: api.Session, experiments: int, reports: int) -> Dict[str, Any]:
    """Unmanaged experiments are rows without processes: the data the read endpoints page through."""
    ws = sess.post("/api/v1/workspaces", json={"name": "load-ws"}).json()["workspace"]
    proj = sess.post(f"/api/v1/workspaces/{ws['id']}/projects", json={"name": "load-proj"}).json()["project"]
    first_trial = first_exp = None
    for i in range(experiments):
        cfg = dict(CFG, name=f"load-{i}", labels=[f"l{i % 3}"])
        e = sess.post("/api/v1/experiments", json={"config": json.dumps(cfg), "unmanaged": True, "projectId": proj["id"],
                                                   "modelDefinition": [{"path": "model_def.py", "content": "cHJpbnQoMSkK", "type": 48, "mode": 0o644}]}).json()["experiment"]
        t = sess.post(f"/api/v1/experiments/{e['id']}/trials", json={"hparams": {"lr": 0.1 * (i + 1), "layers": 1 + i % 4}}).json()["trial"]
        first_exp, first_trial = first_exp or e, first_trial or t
        for s in range(1, reports + 1):
            sess.post(f"/api/v1/trials/{t['id']}/metrics", json={"group": "training", "metrics": {"stepsCompleted": s * 10, "metrics": {"avgMetrics": {"loss": 1.0 / s}}}})
            if s % 5 == 0:
                sess.post(f"/api/v1/trials/{t['id']}/metrics", json={"group": "validation", "metrics": {"stepsCompleted": s * 10, "metrics": {"avgMetrics": {"validation_error": 1.0 / s}}}})
        m.db.add_logs([(f"trial-{t['id']}", f"trial-{t['id']}.1", time.time(), "INFO", 0, "stdout", "task", f"line {k}\n") for k in range(200)])
    ck = str(uuid.uuid4())
    sess.post("/api/v1/checkpoints", json={"uuid": ck, "taskId": f"trial-{first_trial['id']}", "allocationId": None, "reportTime": "2026-01-01T00:00:00Z",
                                           "state": "STATE_COMPLETED", "resources": {"w.bin": "3"}, "metadata": {"steps_completed": 10}, "training": {}})
    sess.post("/api/v1/models", json={"name": "load-model", "labels": ["prod"], "workspaceId": ws["id"]})
    sess.post("/api/v1/models/load-model/versions", json={"checkpointUuid": ck})
    group = sess.post("/api/v1/groups", json={"name": "load-group"}).json()["group"]
    cmd = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["sleep", "600"], "resources": {"slots": 0}}}).json()["command"]
    return {"ws": ws["id"], "proj": proj["id"], "exp": first_exp["id"], "trial": first_trial["id"], "group": group["groupId"], "task": cmd["id"], "pool": "default"}


def requests_of(sd: Dict[str, Any]) -> List[Tuple[str, str]]:
    flt = urllib.parse.quote(json.dumps({"filterGroup": {"children": [{"columnName": "hp.layers", "kind": "field", "operator": ">=", "value": 2}], "conjunction": "and", "kind": "group"}}))
    e, t, w, p, k = sd["exp"], sd["trial"], sd["ws"], sd["proj"], sd["task"]
    return [
        ("get master", "/api/v1/master"), ("get agents", "/api/v1/agents"), ("get workspaces", "/api/v1/workspaces"), ("get user settings", "/api/v1/users/setting"),
        ("get resource pools", "/api/v1/resource-pools"), ("get workspace pools", f"/api/v1/workspaces/{w}/available-resource-pools"), ("get users", "/api/v1/users"),
        ("get pool bindings", f"/api/v1/resource-pools/{sd['pool']}/workspace-bindings"), ("get model labels in workspace", f"/api/v1/model/labels?workspaceId={w}"),
        ("get models", "/api/v1/models"), ("get telemetry", "/api/v1/master/telemetry"), ("get tensorboards", "/api/v1/tensorboards"), ("get shells", "/api/v1/shells"),
        ("get notebooks", "/api/v1/notebooks"), ("get commands", "/api/v1/commands"), ("get job queue stats", "/api/v1/job-queues/stats"),
        ("get job queue", f"/api/v1/job-queues-v2?resourcePool={sd['pool']}"), ("get project", f"/api/v1/projects/{p}"), ("get user activity", "/api/v1/user/projects/activity"),
        ("get workspace notebooks", f"/api/v1/notebooks?limit=1000&workspaceId={w}"), ("get workspace projects", f"/api/v1/workspaces/{w}/projects"),
        ("get webhooks", "/api/v1/webhooks"), ("get metric ranges", f"/api/v1/projects/{p}/experiments/metric-ranges"), ("get project columns", f"/api/v1/projects/{p}/columns"),
        ("search experiments", f"/api/v1/experiments-search?projectId={p}"), ("get model labels", "/api/v1/model/labels"), ("get model versions", "/api/v1/models/load-model/versions"),
        ("get model version", "/api/v1/models/load-model/versions/1"), ("get trial", f"/api/v1/trials/{t}"), ("get trial workloads", f"/api/v1/trials/{t}/workloads"),
        ("get trial log fields", f"/api/v1/trials/{t}/logs/fields"), ("get experiment", f"/api/v1/experiments/{e}"),
        ("get metric names", f"/api/v1/experiments/metrics-stream/metric-names?ids={e}"), ("get metric batches", f"/api/v1/experiments/{e}/metrics-stream/batches?metricName=validation_error"),
        ("get trials sample", f"/api/v1/experiments/{e}/metrics-stream/trials-sample?metricName=validation_error&metricType=METRIC_TYPE_VALIDATION"),
        ("get trials snapshot", f"/api/v1/experiments/{e}/metrics-stream/trials-snapshot?metricName=validation_error&metricType=METRIC_TYPE_VALIDATION&batchesProcessed=50&batchesMargin=1000"),
        ("get experiment trials", f"/api/v1/experiments/{e}/trials"), ("get trial time series", f"/api/v1/trials/time-series?trialIds={t}&startBatches=0"),
        ("get file tree", f"/api/v1/experiments/{e}/file_tree"), ("get experiments", "/api/v1/experiments?showTrialData=true"), ("get master logs", "/api/v1/master/logs?offset=-1&limit=0"),
        ("get allocation aggregates", "/api/v1/resources/allocation/aggregated?startDate=2000-01-01&endDate=2100-01-01&period=RESOURCE_ALLOCATION_AGGREGATION_PERIOD_MONTHLY"),
        ("get tasks", "/api/v1/tasks"), ("get task count", "/api/v1/tasks/count"), ("get task", f"/api/v1/tasks/{k}"), ("get task log fields", f"/api/v1/tasks/{k}/logs/fields"),
        ("get task logs", f"/api/v1/tasks/{k}/logs"), ("get trial logs", f"/api/v1/trials/{t}/logs?limit=100"), ("get permissions", "/api/v1/permissions/summary"),
        ("search groups", "/api/v1/groups"), ("get group", f"/api/v1/groups/{sd['group']}"), ("roles of group", f"/api/v1/roles/search/by-group/{sd['group']}"),
        ("search runs", f"/api/v1/runs?projectId={p}"), ("search runs with a filter", f"/api/v1/runs?projectId={p}&filter={flt}"),
        ("get validation history", f"/api/v1/experiments/{e}/validation-history"), ("get experiment labels", "/api/v1/experiment/labels"),
    ]


def run(vus: int, seconds: float, experiments: int, reports: int) -> Dict[str, Any]:
    tmp = tempfile.mkdtemp(prefix="detb200-load-")
    m = core.Master(db_path=f"{tmp}/m.db", num_slots=1, work_dir=f"{tmp}/w", cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": f"{tmp}/ck"}})
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        admin = api.authentication.login(m.url, "admin", "")
        t0 = time.time()
        sd = seed(m, admin, experiments, reports)
        seed_s = time.time() - t0
        reqs = requests_of(sd)
        lat: Dict[str, List[float]] = {name: [] for name, _ in reqs}
        fails: Dict[str, List[str]] = {name: [] for name, _ in reqs}
        lock = threading.Lock()
        stop = time.time() + seconds

        def user(i: int) -> None:
            sess = api.authentication.login(m.url, "admin", "")
            k = i
            while time.time() < stop:
                name, path = reqs[k % len(reqs)]
                k += 1
                a = time.perf_counter()
                try:
                    r = sess._client().get(m.url + path, headers={"Authorization": f"Bearer {sess.token}"}, timeout=30)
                    ok, why = r.status_code == 200, f"HTTP {r.status_code}: {r.text[:120]}"
                except Exception as e:  # noqa: BLE001
                    ok, why = False, repr(e)
                d = (time.perf_counter() - a) * 1e3
                with lock:
                    lat[name].append(d)
                    if not ok:
                        fails[name].append(why)

        threads = [threading.Thread(target=user, args=(i,), daemon=True) for i in range(vus)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()

        def pct(v: List[float], q: float) -> float:
            s = sorted(v)
            return round(s[min(len(s) - 1, int(q * len(s)))], 2) if s else 0.0
        rows = {name: {"count": len(v), "failures": len(fails[name]), "first_failure": (fails[name] or [None])[0], "p50_ms": pct(v, 0.5), "p95_ms": pct(v, 0.95),
                       "max_ms": round(max(v), 2) if v else 0.0} for name, v in lat.items()}
        total = sum(r["count"] for r in rows.values())
        allv = [x for v in lat.values() for x in v]
        return {"vus": vus, "seconds": seconds, "seeded": {"experiments": experiments, "metric_reports_per_trial": reports, "seed_seconds": round(seed_s, 2)},
                "requests": total, "requests_per_second": round(total / seconds, 1), "failure_rate": round(sum(r["failures"] for r in rows.values()) / max(total, 1), 4),
                "p50_ms": pct(allv, 0.5), "p95_ms": pct(allv, 0.95), "endpoints": rows}
    finally:
        m.stop()
        srv.stop()


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--vus", type=int, default=8)
    ap.add_argument("--seconds", type=float, default=10.0)
    ap.add_argument("--experiments", type=int, default=20)
    ap.add_argument("--reports", type=int, default=50, help="training metric reports per trial")
    ap.add_argument("--p95-ms", type=float, default=2000.0, help="fail when an endpoint's p95 latency exceeds this")
    ap.add_argument("--out")
    a = ap.parse_args()
    res = run(a.vus, a.seconds, a.experiments, a.reports)
    slow = {n: r["p95_ms"] for n, r in res["endpoints"].items() if r["p95_ms"] > a.p95_ms}
    res["thresholds"] = {"p95_ms": a.p95_ms, "failure_rate": 0.0, "slow_endpoints": slow}
    text = json.dumps(res, indent=1)
    if a.out:
        pathlib.Path(a.out).write_text(text + "\n")
    print(json.dumps({k: v for k, v in res.items() if k != "endpoints"}, indent=1))
    worst = sorted(res["endpoints"].items(), key=lambda kv: -kv[1]["p95_ms"])[:5]
    print("slowest p95:", ", ".join(f"{n} {r['p95_ms']} ms" for n, r in worst))
    bad = {n: r["first_failure"] for n, r in res["endpoints"].items() if r["failures"]}
    if bad:
        print("failures:", json.dumps(bad, indent=1))
    return 1 if bad or slow else 0


if __name__ == "__main__":
    sys.exit(main())
