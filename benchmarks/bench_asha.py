"""BASELINE config #5: adaptive_asha packing concurrent CIFAR-10 ResNet-18 trials onto the slots of one box, through the real
single-box master (REST + scheduler + searcher + trial FSM + launcher) — ``bench.py --config asha_pack [--gpus N]``.

What is measured (wall clock; this config is about the control plane, checkpoints and packing, not about one kernel):
* aggregate samples/s = all samples trained by all trials / time from experiment creation to completion;
* slot idle fraction = 1 - (sum over allocations of slots x run time) / (slots x wall time), from the master's allocation table;
* relaunch gap: for every trial that ran in more than one allocation (promoted by ASHA or restarted), the time from the end
  of one allocation to the first training report of the next (scheduling + process start + checkpoint restore);
* preempt -> checkpoint -> exit latency and activate -> first new batch latency, on a long single trial that is paused
  and re-activated, once with synchronous periodic checkpoints and once with ``optimizations.async_checkpoint``.
Reference: master/pkg/searcher/adaptive_asha.go:71-143 (brackets), master/internal/rm/agentrm/fitting.go:72-246 (packing);
the reference publishes no number for this, and its Go master cannot be built offline.
"""
import base64
import json
import os
import pathlib
import statistics
import sys
import tempfile
import time
from typing import Any, Dict, List

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MODEL_DEF = '''
import sys
sys.path.insert(0, {root!r})
from benchmarks import cifar_trial
from determined_b200 import pytorch

_Base = cifar_trial.make_trial(pytorch, per_slot_batch={batch}, pool=8, on_device=False, use_b200_extras={extras}, timing={{}})


class CIFARTrial(_Base):
    def __init__(self, context):
        super().__init__(context)
        lr = float(context.get_hparams().get("learning_rate", 0.05))
        for g in self.opt.param_groups:
            g["lr"] = lr
'''


def _files(batch: int, extras: bool) -> List[Dict[str, Any]]:
    body = MODEL_DEF.format(root=ROOT, batch=batch, extras=extras).encode()
    return [{"path": "model_def.py", "type": 48, "content": base64.b64encode(body).decode(), "mode": 0o644, "mtime": int(time.time())}]


def _wait(sess: Any, eid: int, states: set, timeout: float) -> str:
    t0 = time.time()
    st = "?"
    while time.time() - t0 < timeout:
        st = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"]
        if st in states:
            return st
        time.sleep(0.2)
    raise TimeoutError(f"experiment {eid} still {st} after {timeout}s")


def main(args: Any) -> None:
    import torch
    import yaml

    from determined_b200.common import api
    from determined_b200.master import api as mapi
    from determined_b200.master import core

    gpu = torch.cuda.is_available()
    slots = max(1, args.gpus if args.gpus > 1 else (torch.cuda.device_count() if gpu else 2))
    batch = 256 if gpu else 8
    max_len = max(args.steps, 8) if gpu else 8
    tmp = pathlib.Path(tempfile.mkdtemp(prefix="asha-pack-"))
    m = core.Master(db_path=str(tmp / "m.db"), num_slots=slots, work_dir=str(tmp / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp / "ckpt")}},
                    scheduler_type="priority")
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    out: Dict[str, Any] = {}
    try:
        sess = api.authentication.login(m.url, "determined", "")
        base = {"name": "asha-pack", "entrypoint": "model_def:CIFARTrial", "resources": {"slots_per_trial": 1},
                "hyperparameters": {"global_batch_size": batch, "learning_rate": {"type": "log", "minval": -2.5, "maxval": -0.5, "base": 10}},
                "scheduling_unit": max(max_len // 8, 4), "min_validation_period": {"batches": 0}, "max_restarts": 0,
                "optimizations": {"aggregation_frequency": 1, "cuda_graphs": bool(gpu)},
                "environment": {"environment_variables": ["CUDA_DEVICE_ORDER=PCI_BUS_ID"]}}
        # ---- 1. the packing run
        cfg = dict(base, searcher={"name": "adaptive_asha", "metric": "validation_loss", "smaller_is_better": True,
                                   "max_length": {"batches": max_len}, "max_trials": 2 * slots, "max_concurrent_trials": slots,
                                   "divisor": 4, "max_rungs": 3})
        t0 = time.time()
        eid = sess.post("/api/v1/experiments", json={"config": yaml.safe_dump(cfg), "modelDefinition": _files(batch, gpu), "activate": True}).json()["experiment"]["id"]
        st = _wait(sess, eid, {"STATE_COMPLETED", "STATE_ERROR", "STATE_CANCELED"}, 3600)
        wall = time.time() - t0
        print(f"[asha_pack] search finished: {st} in {wall:.1f}s", file=sys.stderr, flush=True)
        if st != "STATE_COMPLETED":
            for t in sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"][:2]:
                for ln in sess.get(f"/api/v1/trials/{t['id']}/logs").json()["logs"][-25:]:
                    print(f"[trial {t['id']}] {ln['log'].rstrip()}", file=sys.stderr)
            raise RuntimeError(f"the search ended in {st}")
        trials = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
        total_batches = sum(t["totalBatchesProcessed"] for t in trials)
        allocs = m.db.all("SELECT a.trial_id, a.slots, a.start_time, a.end_time FROM allocations a JOIN trials t ON t.id=a.trial_id "
                          "WHERE t.experiment_id=? ORDER BY a.start_time", (eid,))
        busy = sum((a["end_time"] - a["start_time"]) * a["slots"] for a in allocs if a["end_time"])
        gaps = []
        by_trial: Dict[int, List[Any]] = {}
        for a in allocs:
            by_trial.setdefault(a["trial_id"], []).append(a)
        for tid, runs in by_trial.items():
            for prev, nxt in zip(runs, runs[1:]):
                first = m.db.one("SELECT MIN(report_time) AS t FROM metrics WHERE trial_id=? AND grp='training' AND report_time>?",
                                 (tid, nxt["start_time"]))
                if first is not None and first["t"]:
                    gaps.append({"idle_between_allocations_s": nxt["start_time"] - prev["end_time"],
                                 "allocation_start_to_first_report_s": first["t"] - nxt["start_time"]})
        out["packing"] = {
            "state": st, "slots": slots, "trials": len(trials), "wall_s": wall, "total_batches": total_batches,
            "aggregate_samples_per_s": total_batches * batch / wall,
            "slot_idle_fraction": 1.0 - busy / (slots * wall), "allocations": len(allocs),
            "trials_by_length": sorted(t["totalBatchesProcessed"] for t in trials),
            "relaunches": len(gaps),
            "relaunch_idle_s_median": statistics.median(g["idle_between_allocations_s"] for g in gaps) if gaps else None,
            "relaunch_start_to_first_report_s_median": statistics.median(g["allocation_start_to_first_report_s"] for g in gaps) if gaps else None}
        # ---- 2. preempt -> checkpoint -> resume latency, sync vs async periodic checkpoints
        out["preempt_resume"] = {}
        for mode in ("sync", "async"):
            c2 = dict(base, searcher={"name": "single", "metric": "validation_loss", "max_length": {"batches": 1000000}},
                      min_checkpoint_period={"batches": max(max_len // 4, 4)})
            c2["optimizations"] = dict(base["optimizations"], async_checkpoint=(mode == "async"))
            e2 = sess.post("/api/v1/experiments", json={"config": yaml.safe_dump(c2), "modelDefinition": _files(batch, gpu), "activate": True}).json()["experiment"]["id"]

            def batches() -> int:
                tr = sess.get(f"/api/v1/experiments/{e2}/trials").json()["trials"]
                return tr[0]["totalBatchesProcessed"] if tr else 0

            print(f"[asha_pack] preempt/resume probe ({mode})", file=sys.stderr, flush=True)
            t1 = time.time()
            while batches() < max(max_len // 2, 8) and time.time() - t1 < 600:
                time.sleep(0.05)
            start_to_training = time.time() - t1
            tp = time.time()
            sess.post(f"/api/v1/experiments/{e2}/pause")
            while m.allocations and time.time() - tp < 300:
                time.sleep(0.01)
            preempt_exit = time.time() - tp
            done = batches()
            ta = time.time()
            sess.post(f"/api/v1/experiments/{e2}/activate")
            while batches() <= done and time.time() - ta < 600:
                time.sleep(0.02)
            resume = time.time() - ta
            sess.post(f"/api/v1/experiments/{e2}/kill")
            _wait(sess, e2, {"STATE_CANCELED", "STATE_COMPLETED", "STATE_ERROR"}, 120)
            out["preempt_resume"][mode] = {"create_to_training_s": start_to_training, "pause_to_checkpointed_exit_s": preempt_exit,
                                           "activate_to_first_new_report_s": resume, "batches_at_pause": done}
    finally:
        m.stop()
        srv.stop()
    p = out["packing"]
    print(json.dumps({"impl": "ours", "metric": "aggregate samples/sec over an adaptive_asha search", "value": p["aggregate_samples_per_s"],
                      "unit": "samples/s", "n_gpus": slots if gpu else 0, "steps": p["total_batches"], "warmup": 0, "ms_per_step": None,
                      "higher_is_better": True, "scaling": "n/a (throughput of a search)", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                      "config": {"model": "resnet18-cifar10 PyTorchTrial", "searcher": "adaptive_asha", "max_trials": 2 * slots, "max_length_batches": max_len,
                                 "divisor": 4, "max_rungs": 3, "slots": slots, "slots_per_trial": 1, "per_slot_batch": batch},
                      "harness": "single-box master (REST, scheduler, searcher, launcher) + trial processes; wall-clock, includes process start-up, "
                                 "validation, checkpoints and relaunches" + ("" if gpu else "; CPU smoke run"),
                      "detail": out}))
