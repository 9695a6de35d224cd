"""``det`` — command line interface.

Parity: reference ``harness/determined/cli`` (argparse tree ``cli/cli.py:110-186``): ``experiment``
(``create [--local --test] [--paused] [-f] [--config k=v] [--template]``, list, describe, pause,
activate, cancel, kill, logs, download, set, archive, delete, label — ``cli/experiment.py``),
``trial``, ``checkpoint`` (list/describe/download/delete/rm — ``cli/checkpoint.py:158-243``),
``model``, ``template``, ``user``, ``agent``/``slot``, ``job``, ``resource-pool``, ``master``,
``workspace``/``project``, ``preview-search``, ``deploy local``, ``version``.
"""
import argparse
import base64
import json
import os
import pathlib
import sys
import time
from typing import Any, Dict, List, Optional, Sequence

import yaml

import determined_b200 as det
from determined_b200.common import api, context as det_context
from determined_b200.experimental import client as sdk


VIEW: Dict[str, Any] = {}      # --sort-by / --order-by / --limit / --offset of the running command (cli/flags.py), applied to what it prints


def _table(rows: List[Sequence[Any]], headers: Sequence[str], as_json: bool = False, as_csv: bool = False) -> None:
    if VIEW:
        from determined_b200.cli import flags

        rows = flags.apply_view(rows, headers, VIEW)
    if as_json:
        print(json.dumps([dict(zip(headers, r)) for r in rows], indent=2, default=str))
        return
    if as_csv:
        import csv

        w = csv.writer(sys.stdout)
        w.writerow(headers)
        w.writerows(rows)
        return
    try:
        import tabulate

        print(tabulate.tabulate(rows, headers=headers, tablefmt="presto"))
    except ImportError:
        print(" | ".join(headers))
        for r in rows:
            print(" | ".join(str(x) for x in r))


def _paging(args: argparse.Namespace) -> Dict[str, Any]:
    """``--limit`` / ``--offset`` of list commands as query parameters."""
    out: Dict[str, Any] = {}
    if getattr(args, "limit", None):
        out["limit"] = args.limit
    if getattr(args, "offset", None):
        out["offset"] = args.offset
    return out


def _master_of(args: argparse.Namespace) -> str:
    return args.master or os.environ.get("DET_MASTER") or "http://127.0.0.1:8080"


def _setup_cert(args: argparse.Namespace) -> None:
    """For an https master: the certificate every session of this process uses, from ``--cert-file`` / ``--cert-name`` /
    ``--noverify``, the environment, or the store of accepted certificates."""
    from determined_b200.common.api import certs

    master = _master_of(args)
    try:
        https = api.canonicalize_master_url(master).startswith("https://")
    except ValueError:
        return
    if https:
        certs.cli_cert = certs.default_load(master, getattr(args, "cert_file", None), getattr(args, "cert_name", None),
                                            bool(getattr(args, "noverify", False)))
        if certs.cli_cert.bundle is False:
            import urllib3

            urllib3.disable_warnings(urllib3.exceptions.InsecureRequestWarning)      # the user asked for exactly this


def _session(args: argparse.Namespace) -> api.Session:
    return api.authentication.login_with_cache(_master_of(args), args.user, None)


def _sdk(args: argparse.Namespace) -> sdk.Determined:
    return sdk.Determined._from_session(_session(args))


def _parse_config_overrides(cfg: Dict[str, Any], overrides: Optional[List[str]]) -> None:
    from determined_b200.master import expconf

    for ov in overrides or []:
        if "=" not in ov:
            raise SystemExit(f"--config expects key=value, got {ov!r}")
        k, v = ov.split("=", 1)
        expconf.apply_override(cfg, k, yaml.safe_load(v))


# ------------------------------------------------------------------------------------------------
# experiment
# ------------------------------------------------------------------------------------------------
def _follow_trial_logs(sess: api.Session, eid: int) -> int:
    d = sdk.Determined._from_session(sess)
    exp = d.get_experiment(eid)
    print(f"Waiting for first trial of experiment {eid}...")
    t = exp.await_first_trial()
    print(f"Following logs of trial {t.id}")
    for line in t.logs(follow=True):
        sys.stdout.write(line)
    exp.reload()
    print(f"Experiment {eid}: {exp.state.value if exp.state else '?'}")
    return 0 if exp.state == sdk.ExperimentState.COMPLETED else 1


def _run_local(cfg: Dict[str, Any], model_dir: str, test: bool) -> int:
    """``det e create --local --test``: run the trial in this process, off-cluster (reference
    ``_execution.py:77-236`` test/local mode)."""
    from determined_b200 import load, pytorch

    entry = cfg.get("entrypoint")
    if not isinstance(entry, str) or ":" not in entry or " " in entry:
        raise SystemExit("--local requires a legacy `module:TrialClass` entrypoint")
    os.chdir(model_dir)
    sys.path.insert(0, os.getcwd())
    hparams = {}
    from determined_b200.master import nprand
    from determined_b200.master.searcher import hparams as hp

    hparams = hp.sample_all(cfg.get("hyperparameters") or {}, nprand.State(int(time.time())))
    trial_cls = load.trial_class_from_entrypoint(entry)
    agg = int((cfg.get("optimizations") or {}).get("aggregation_frequency", 1))
    with pytorch.init(hparams=hparams, exp_conf=cfg, aggregation_frequency=agg) as ctx:
        trial = trial_cls(ctx)
        tr = pytorch.Trainer(trial, ctx)
        if test:
            tr.fit(test_mode=True)
        else:
            ml = (cfg.get("searcher") or {}).get("max_length", {"batches": 100})
            ml = ml if isinstance(ml, dict) else {"batches": ml}
            tr.fit(max_length=pytorch.TrainUnit._from_values(**ml, global_batch_size=ctx._global_batch_size),
                   reporting_period=pytorch.Batch(int(cfg.get("scheduling_unit", 100))))
    print("Model training test succeeded." if test else "Local training finished.")
    return 0


def experiment_create(args: argparse.Namespace) -> int:
    cfg_text = pathlib.Path(args.config_file).read_text()
    cfg = yaml.safe_load(cfg_text) or {}
    _parse_config_overrides(cfg, args.config)
    if args.local:
        return _run_local(cfg, args.model_def, args.test)
    sess = _session(args)
    if args.test:
        # on-cluster test mode: the same model and config squeezed into a one-batch single-trial experiment that exercises
        # train / validate / checkpoint on real slots, then report how it ended
        s = cfg.get("searcher") or {}
        cfg["searcher"] = {"name": "single", "metric": s.get("metric", "loss"), "smaller_is_better": s.get("smaller_is_better", True),
                           "max_length": {"batches": 1}}
        cfg["min_validation_period"] = {"batches": 1}
        cfg["min_checkpoint_period"] = {"batches": 1}
        cfg["description"] = "[test mode] " + str(cfg.get("description") or cfg.get("name") or "")
        for v in (cfg.get("hyperparameters") or {}).values():
            if isinstance(v, dict) and v.get("type") in ("int", "double", "log", "categorical") and "count" in v:
                v.pop("count")
    files = det_context.read_v1_context(pathlib.Path(args.model_def)) if args.model_def else []
    for inc in args.include or []:
        files += det_context.read_v1_context(pathlib.Path(inc))
    body = {"config": yaml.safe_dump(cfg), "modelDefinition": files, "activate": not args.paused, "template": args.template,
            "projectId": args.project_id}
    r = sess.post("/api/v1/experiments", json=body).json()
    eid = r["experiment"]["id"]
    print(f"Created experiment {eid}")
    if args.test and not args.paused:
        rc = _follow_trial_logs(sess, eid)
        state = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"]
        for _ in range(600):
            if state in ("STATE_COMPLETED", "STATE_ERROR", "STATE_CANCELED"):
                break
            time.sleep(0.5)
            state = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]["state"]
        ok = state == "STATE_COMPLETED"
        print("Model training test succeeded." if ok else f"Model training test failed: experiment {eid} ended as {state}.")
        return 0 if ok else (rc or 1)
    if args.follow_first_trial and not args.paused:
        return _follow_trial_logs(sess, eid)
    return 0


def experiment_continue(args: argparse.Namespace) -> int:
    """``det experiment continue ID [--config-file f] [--config k=v]`` (reference ``cli/experiment.py`` ``continue``)."""
    cfg: Dict[str, Any] = {}
    if args.config_file:
        cfg = yaml.safe_load(pathlib.Path(args.config_file).read_text()) or {}
    _parse_config_overrides(cfg, args.config)
    sess = _session(args)
    sess.post("/api/v1/experiments/continue", json={"id": args.experiment_id, "overrideConfig": yaml.safe_dump(cfg) if cfg else None})
    print(f"Continued experiment {args.experiment_id}")
    if args.follow_first_trial:
        return _follow_trial_logs(sess, args.experiment_id)
    return 0


def experiment_move(args: argparse.Namespace) -> int:
    sess = _session(args)
    ws = [w for w in sess.get("/api/v1/workspaces", params={"name": args.workspace}).json()["workspaces"] if w["name"] == args.workspace]
    if not ws:
        raise SystemExit(f"workspace {args.workspace!r} not found")
    pj = [p for p in sess.get(f"/api/v1/workspaces/{ws[0]['id']}/projects").json()["projects"] if p["name"] == args.project]
    if not pj:
        raise SystemExit(f"project {args.project!r} not found in workspace {args.workspace!r}")
    sess.post(f"/api/v1/experiments/{args.experiment_id}/move", json={"destinationProjectId": pj[0]["id"]})
    print(f"Moved experiment {args.experiment_id} to project {args.project}")
    return 0


def experiment_download_model_def(args: argparse.Namespace) -> int:
    sess = _session(args)
    blob = base64.b64decode(sess.get(f"/api/v1/experiments/{args.experiment_id}/model_def").json()["b64Tgz"])
    out = pathlib.Path(args.output_dir) / f"experiment_{args.experiment_id}_model_def.tgz"
    out.write_bytes(blob)
    print(f"Wrote {out} ({len(blob)} bytes)")
    return 0


def experiment_list(args: argparse.Namespace) -> int:
    sess = _session(args)
    params: Dict[str, Any] = _paging(args)
    if not args.all:
        params["archived"] = "false"
    exps = sess.get("/api/v1/experiments", params=params).json()["experiments"]
    rows = [[e["id"], e["username"], e["name"], e["parentId"], e["state"].replace("STATE_", ""), f"{100 * (e['progress'] or 0):.1f}%",
             e["startTime"], e["endTime"], e["resourcePool"]] + ([e["projectId"]] if getattr(args, "show_project", False) else []) for e in exps]
    _table(rows, ["ID", "Owner", "Name", "Parent ID", "State", "Progress", "Started", "Ended", "Resource Pool"] + (["Project ID"] if getattr(args, "show_project", False) else []),
           args.json, args.csv)
    return 0


def experiment_describe(args: argparse.Namespace) -> int:
    sess = _session(args)
    out = []
    for eid in args.experiment_ids.split(","):
        e = sess.get(f"/api/v1/experiments/{int(eid)}").json()["experiment"]
        trials = sess.get(f"/api/v1/experiments/{int(eid)}/trials").json()["trials"]
        e["trials"] = trials
        out.append(e)
    if args.json:
        print(json.dumps(out, indent=2, default=str))
        return 0
    _table([[e["id"], e["state"].replace("STATE_", ""), f"{100 * (e['progress'] or 0):.1f}%", e["startTime"], e["endTime"], e["name"], e["description"],
             e["archived"], e["resourcePool"], ", ".join(e["labels"] or [])] for e in out],
           ["Experiment ID", "State", "Progress", "Started", "Ended", "Name", "Description", "Archived", "Resource Pool", "Labels"])
    print()
    _table([[t["id"], t["experimentId"], t["state"].replace("STATE_", ""), t["startTime"], t["endTime"], json.dumps(t["hparams"], separators=(",", ":")),
             t["totalBatchesProcessed"], t["restarts"]] for e in out for t in e["trials"]],
           ["Trial ID", "Experiment ID", "State", "Started", "Ended", "H-Params", "Batches", "Restarts"])
    return 0


def _exp_action(name: str) -> Any:
    def fn(args: argparse.Namespace) -> int:
        _session(args).post(f"/api/v1/experiments/{args.experiment_id}/{name}")
        print(f"{name.capitalize()}d experiment {args.experiment_id}" if not name.endswith("e") else f"{name.capitalize()}d experiment {args.experiment_id}")
        return 0
    return fn


def experiment_delete(args: argparse.Namespace) -> int:
    if not args.yes:
        if input(f"Delete experiment {args.experiment_id} and all its checkpoints? [y/N] ").strip().lower() not in ("y", "yes"):
            return 1
    _session(args).delete(f"/api/v1/experiments/{args.experiment_id}")
    print(f"Deleted experiment {args.experiment_id}")
    return 0


def experiment_logs(args: argparse.Namespace) -> int:
    sess = _session(args)
    trials = sess.get(f"/api/v1/experiments/{args.experiment_id}/trials").json()["trials"]
    if not trials:
        raise SystemExit("experiment has no trials yet")
    t = sdk.Trial(sess, trials[0]["id"], trials[0])
    for line in t.logs(follow=args.follow, head=args.head, tail=args.tail):
        sys.stdout.write(line)
    return 0


def experiment_download(args: argparse.Namespace) -> int:
    d = _sdk(args)
    exp = d.get_experiment(args.experiment_id)
    cks = exp.top_n_checkpoints(args.top_n, sort_by=args.sort_by, smaller_is_better=args.smaller_is_better)
    for c in cks:
        path = c.download(os.path.join(args.output_dir, c.uuid))
        print(f"Local checkpoint path: {path}")
    return 0


def experiment_set(args: argparse.Namespace) -> int:
    sess = _session(args)
    eid = args.experiment_id
    if args.attr == "name":
        sess.patch(f"/api/v1/experiments/{eid}", json={"name": args.value})
    elif args.attr == "description":
        sess.patch(f"/api/v1/experiments/{eid}", json={"description": args.value})
    elif args.attr == "priority":
        sess.patch(f"/api/v1/experiments/{eid}", json={"resources": {"priority": int(args.value)}})
    elif args.attr == "weight":
        sess.patch(f"/api/v1/experiments/{eid}", json={"resources": {"weight": float(args.value)}})
    elif args.attr == "max-slots":
        sess.patch(f"/api/v1/experiments/{eid}", json={"resources": {"maxSlots": int(args.value)}})
    elif args.attr == "log-retention":
        days = -1 if str(args.value).lower() in ("forever", "-1") else int(args.value)
        sess.put(f"/api/v1/experiments/{eid}/retain_logs", json={"numDays": days})
    elif args.attr == "gc-policy":
        a, b, c = (int(x) for x in args.value.split(","))
        sess.patch(f"/api/v1/experiments/{eid}", json={"checkpointStorage": {"saveExperimentBest": a, "saveTrialBest": b, "saveTrialLatest": c}})
    print(f"Set {args.attr} of experiment {eid} to {args.value}")
    return 0


def experiment_label(args: argparse.Namespace) -> int:
    e = _sdk(args).get_experiment(args.experiment_id)
    (e.add_label if args.action == "add" else e.remove_label)(args.label)
    return 0


def experiment_config(args: argparse.Namespace) -> int:
    e = _session(args).get(f"/api/v1/experiments/{args.experiment_id}").json()["experiment"]
    print(yaml.safe_dump(e["config"], default_flow_style=False))
    return 0


def experiment_wait(args: argparse.Namespace) -> int:
    st = _sdk(args).get_experiment(args.experiment_id).wait(interval=args.polling_interval)
    print(st.value)
    return 0 if st == sdk.ExperimentState.COMPLETED else 1


def experiment_list_checkpoints(args: argparse.Namespace) -> int:
    sess = _session(args)
    cks = sess.get(f"/api/v1/experiments/{args.experiment_id}/checkpoints").json()["checkpoints"]
    _table([[c["training"]["trialId"], c["metadata"].get("steps_completed"), c["state"].replace("STATE_", ""), c["training"]["searcherMetric"], c["uuid"],
             sum(int(v) for v in c["resources"].values())] for c in cks], ["Trial ID", "# of Batches", "State", "Validation Metric", "UUID", "Size"], args.json)
    return 0


# ------------------------------------------------------------------------------------------------
# trial / checkpoint
# ------------------------------------------------------------------------------------------------
def trial_describe(args: argparse.Namespace) -> int:
    sess = _session(args)
    t = sess.get(f"/api/v1/trials/{args.trial_id}").json()["trial"]
    if args.json:
        t["metrics"] = sess.get(f"/api/v1/trials/{args.trial_id}/metrics").json()["metrics"]
        print(json.dumps(t, indent=2, default=str))
        return 0
    _table([[t["id"], t["experimentId"], t["state"].replace("STATE_", ""), t["startTime"], t["endTime"], json.dumps(t["hparams"]), t["totalBatchesProcessed"], t["restarts"]]],
           ["Trial ID", "Experiment ID", "State", "Started", "Ended", "H-Params", "Batches", "Restarts"])
    mets = sess.get(f"/api/v1/trials/{args.trial_id}/metrics").json()["metrics"]
    print()
    _table([[m["totalBatches"], m["group"], m["endTime"], json.dumps(m["metrics"]["avgMetrics"])] for m in mets], ["# of Batches", "Group", "Report Time", "Metrics"])
    return 0


def trial_list(args: argparse.Namespace) -> int:
    trials = _session(args).get(f"/api/v1/experiments/{args.experiment_id}/trials", params=_paging(args)).json()["trials"]
    _table([[t["id"], t["state"].replace("STATE_", ""), json.dumps(t["hparams"]), t["startTime"], t["endTime"], t["totalBatchesProcessed"], t["restarts"],
             t["searcherMetricValue"]] for t in trials], ["Trial ID", "State", "H-Params", "Started", "Ended", "# of Batches", "Restarts", "Best Metric"], args.json)
    return 0


def trial_logs(args: argparse.Namespace) -> int:
    sess = _session(args)
    t = sdk.Trial(sess, args.trial_id)
    for line in t.logs(follow=args.follow, head=args.head, tail=args.tail):
        sys.stdout.write(line)
    return 0


def trial_kill(args: argparse.Namespace) -> int:
    _session(args).post(f"/api/v1/trials/{args.trial_id}/kill")
    print(f"Killed trial {args.trial_id}")
    return 0


def trial_download(args: argparse.Namespace) -> int:
    """``det trial download ID [--latest | --best | --uuid U] [-o DIR]``: fetch one of the trial's checkpoints."""
    t = _sdk(args).get_trial(args.trial_id)
    ck = t.select_checkpoint(uuid=args.uuid) if args.uuid else t.select_checkpoint(best=True) if args.best else t.select_checkpoint(latest=True)
    if ck is None:
        raise SystemExit(f"trial {args.trial_id} has no matching checkpoint")
    path = ck.download(args.output_dir)
    print(f"Local checkpoint path: {path}")
    return 0


def agent_enable(args: argparse.Namespace) -> int:
    sess = _session(args)
    ids = [a["id"] for a in sess.get("/api/v1/agents").json()["agents"]] if getattr(args, "all", False) else [args.agent_id]
    if not ids or ids == [None]:
        raise SystemExit("give an agent id or --all")
    for aid in ids:
        sess.post(f"/api/v1/agents/{aid}/{'enable' if args.enable else 'disable'}", json={"drain": bool(getattr(args, "drain", False))})
        print(f"{'Enabled' if args.enable else 'Disabled'} agent {aid}" + (" (draining)" if getattr(args, "drain", False) else ""))
    return 0


def checkpoint_describe(args: argparse.Namespace) -> int:
    c = _session(args).get(f"/api/v1/checkpoints/{args.uuid}").json()["checkpoint"]
    if args.json:
        print(json.dumps(c, indent=2, default=str))
        return 0
    tr = c["training"]
    _table([[tr.get("experimentId"), tr.get("trialId"), c["metadata"].get("steps_completed"), c["reportTime"], c["uuid"], json.dumps(tr.get("validationMetrics", {}).get("avgMetrics")),
             json.dumps(c["metadata"])]], ["Experiment ID", "Trial ID", "Steps Completed", "Report Time", "UUID", "Validation Metrics", "Metadata"])
    print("\nResources:")
    _table([[k, v] for k, v in sorted(c["resources"].items())], ["Path", "Size"])
    return 0


def checkpoint_download(args: argparse.Namespace) -> int:
    c = _sdk(args).get_checkpoint(args.uuid)
    path = c.download(args.output_dir, mode=sdk.DownloadMode(args.mode))
    if args.quiet:
        print(path)
    else:
        print(f"Local checkpoint path: {path}")
    return 0


def checkpoint_delete(args: argparse.Namespace) -> int:
    uuids = args.uuid.split(",")
    if not args.yes and input(f"Delete {len(uuids)} checkpoint(s)? [y/N] ").strip().lower() not in ("y", "yes"):
        return 1
    _session(args).delete("/api/v1/checkpoints", json={"checkpointUuids": uuids})
    print(f"Deletion of checkpoints {args.uuid} is in progress")
    return 0


def checkpoint_rm(args: argparse.Namespace) -> int:
    _session(args).post("/api/v1/checkpoints/rm", json={"checkpointUuids": args.uuid.split(","), "checkpointGlobs": args.glob})
    print(f"Removed files matching {args.glob} from {args.uuid}")
    return 0


# ------------------------------------------------------------------------------------------------
# misc
# ------------------------------------------------------------------------------------------------
def preview_search(args: argparse.Namespace) -> int:
    cfg = yaml.safe_load(pathlib.Path(args.config_file).read_text())
    _parse_config_overrides(cfg, args.config)
    sim = _sdk(args).preview_search(cfg, seed=args.seed)
    print(f"Using search configuration:\n{yaml.safe_dump(sim['config']['searcher'])}")
    rows = []
    for t in sim["trials"]:
        rows.append([t["occurrences"], ", ".join(f"train {op['validateAfter']['length']} {sim['unit']}, validate" for op in t["operations"])])
    _table(rows, ["Trials", "Breakdown"])
    return 0


def master_info(args: argparse.Namespace) -> int:
    sess = api.UnauthSession(args.master or os.environ.get("DET_MASTER") or "http://127.0.0.1:8080")
    info = sess.get("/api/v1/master").json()
    print(json.dumps(info, indent=2) if getattr(args, "json", False) else yaml.safe_dump(info, default_flow_style=False))
    return 0


def master_config(args: argparse.Namespace) -> int:
    print(yaml.safe_dump(_session(args).get("/api/v1/master/config").json()["config"], default_flow_style=False))
    return 0


def agent_list(args: argparse.Namespace) -> int:
    ags = _session(args).get("/api/v1/agents").json()["agents"]
    _table([[a["id"], a["version"], a["registeredTime"], len(a["slots"]), sum(1 for s in a["slots"].values() if s["container"]), a["enabled"], a["draining"],
             ",".join(a["resourcePools"]), ",".join(a["addresses"])] for a in ags],
           ["Agent ID", "Version", "Registered Time", "Slots", "Containers", "Enabled", "Draining", "Resource Pool", "Addresses"], args.json)
    return 0


def slot_list(args: argparse.Namespace) -> int:
    ags = _session(args).get("/api/v1/agents").json()["agents"]
    rows = []
    for a in ags:
        for sid, s in sorted(a["slots"].items(), key=lambda kv: int(kv[0])):
            rows.append([a["id"], sid, s["enabled"], (s["container"] or {}).get("id"), s["device"]["type"].replace("TYPE_", ""), s["device"]["brand"], s["device"]["uuid"]])
    _table(rows, ["Agent ID", "Slot ID", "Enabled", "Allocation ID", "Type", "Device", "UUID"], args.json)
    return 0


def slot_toggle(enable: bool) -> Any:
    def fn(args: argparse.Namespace) -> int:
        _session(args).post(f"/api/v1/agents/{args.agent_id}/slots/{args.slot_id}/{'enable' if enable else 'disable'}")
        print(f"{'Enabled' if enable else 'Disabled'} slot {args.slot_id} of agent {args.agent_id}")
        return 0
    return fn


def job_list(args: argparse.Namespace) -> int:
    params = _paging(args)
    if getattr(args, "resource_pool", None):
        params["resourcePool"] = args.resource_pool
    jobs = _session(args).get("/api/v1/job-queues-v2", params=params).json()["jobs"]
    if getattr(args, "reverse", False):
        jobs = jobs[::-1]
    if getattr(args, "yaml", False):
        print(yaml.safe_dump(jobs, default_flow_style=False))
        return 0
    _table([[i, j["jobId"], j["type"].replace("TYPE_", ""), j["name"], j["priority"], j["weight"], f"{j['allocatedSlots']}/{j['requestedSlots']}",
             j["summary"]["state"].replace("STATE_", ""), j["resourcePool"]] for i, j in enumerate(jobs)],
           ["#", "ID", "Type", "Job Name", "Priority", "Weight", "Slots (acquired/needed)", "Status", "Resource Pool"], args.json)
    return 0


def job_update(args: argparse.Namespace) -> int:
    upd: Dict[str, Any] = {"jobId": args.job_id}
    if args.priority is not None:
        upd["priority"] = args.priority
    if args.weight is not None:
        upd["weight"] = args.weight
    if getattr(args, "ahead_of", None):
        upd["aheadOf"] = args.ahead_of
    if getattr(args, "behind_of", None):
        upd["behindOf"] = args.behind_of
    _session(args).post("/api/v1/job-queues", json={"updates": [upd]})
    print(f"Updated job {args.job_id}")
    return 0


def resource_pool_list(args: argparse.Namespace) -> int:
    rps = _session(args).get("/api/v1/resource-pools").json()["resourcePools"]
    _table([[p["name"], p["description"], p["type"], p["numAgents"], f"{p['slotsUsed']}/{p['slotsAvailable']}", p["schedulerType"].replace("SCHEDULER_TYPE_", ""),
             p["schedulerFittingPolicy"].replace("FITTING_POLICY_", ""), p["preemptible"]] for p in rps],
           ["Name", "Description", "Type", "Agents", "Slots Used", "Scheduler", "Fitting Policy", "Preemptible"], args.json)
    return 0


def template_cmd(args: argparse.Namespace) -> int:
    sess = _session(args)
    if args.action == "list":
        ts = sess.get("/api/v1/templates").json()["templates"]
        _table([[t["name"]] for t in ts], ["Name"])
    elif args.action == "describe":
        print(yaml.safe_dump(sess.get(f"/api/v1/templates/{args.name}").json()["template"]["config"], default_flow_style=False))
    elif args.action in ("set", "create"):
        cfg = yaml.safe_load(pathlib.Path(args.file).read_text())
        sess.put(f"/api/v1/templates/{args.name}", json={"config": cfg})
        print(f"Set template {args.name}")
    elif args.action == "remove":
        sess.delete(f"/api/v1/templates/{args.name}")
        print(f"Removed template {args.name}")
    return 0


def model_cmd(args: argparse.Namespace) -> int:
    d = _sdk(args)
    if args.action == "list":
        _table([[m.model_id, m.name, m.description, m.archived, ",".join(m.labels)] for m in d.get_models()], ["ID", "Name", "Description", "Archived", "Labels"], args.json)
    elif args.action == "create":
        m = d.create_model(args.name, args.description or "")
        print(f"Created model {m.name} (id {m.model_id})")
    elif args.action == "describe":
        m = d.get_model(args.name)
        _table([[m.model_id, m.name, m.description, json.dumps(m.metadata)]], ["ID", "Name", "Description", "Metadata"])
        print()
        _table([[v.model_version, v.checkpoint.uuid if v.checkpoint else None, v.name, v.comment] for v in m.list_versions()], ["Version", "Checkpoint", "Name", "Comment"])
    elif args.action == "list-versions":
        _table([[v.model_version, v.checkpoint.uuid if v.checkpoint else None, v.name, v.comment] for v in d.get_model(args.name).list_versions()],
               ["Version", "Checkpoint", "Name", "Comment"], args.json)
    elif args.action == "register-version":
        v = d.get_model(args.name).register_version(args.uuid)
        print(f"Registered version {v.model_version} of model {args.name}")
    elif args.action == "delete":
        d.get_model(args.name).delete()
        print(f"Deleted model {args.name}")
    return 0


def user_cmd(args: argparse.Namespace) -> int:
    if args.action == "login":
        import getpass

        master = args.master or os.environ.get("DET_MASTER") or "http://127.0.0.1:8080"
        pw = os.environ.get("DET_PASS")
        if pw is None:
            pw = "" if args.username == "determined" else getpass.getpass(f"Password for user '{args.username}': ")
        sess = api.authentication.login(master, args.username, pw)
        st = api.authentication.TokenStore(master)
        st.set_token(args.username, sess.token or "")
        st.set_active(args.username)
        print(f"Logged in as {args.username}")
        return 0
    if args.action == "logout":
        api.authentication.logout(args.master or os.environ.get("DET_MASTER") or "http://127.0.0.1:8080", args.user)
        print("Logged out")
        return 0
    sess = _session(args)
    if args.action == "whoami":
        print("You are logged in as user '{}'".format(sess.get("/api/v1/me").json()["user"]["username"]))
    elif args.action == "list":
        us = sess.get("/api/v1/users").json()["users"]
        _table([[u["id"], u["username"], u["displayName"], u["active"], u["admin"]] for u in us], ["ID", "Username", "Display Name", "Active", "Admin"], args.json)
    elif args.action == "create":
        sess.post("/api/v1/users", json={"user": {"username": args.username, "admin": args.admin, "active": True}, "password": ""})
        print(f"Created user {args.username}")
    elif args.action in ("activate", "deactivate"):
        us = [u for u in sess.get("/api/v1/users").json()["users"] if u["username"] == args.username]
        if not us:
            raise SystemExit(f"user {args.username} not found")
        sess.patch(f"/api/v1/users/{us[0]['id']}", json={"user": {"active": args.action == "activate"}})
    elif args.action == "rename":
        us = [u for u in sess.get("/api/v1/users").json()["users"] if u["username"] == args.username]
        if not us:
            raise SystemExit(f"user {args.username} not found")
        sess.patch(f"/api/v1/users/{us[0]['id']}", json={"user": {"username": args.new_username}})
        print(f"Renamed user {args.username} to {args.new_username}")
    elif args.action == "change-password":
        import getpass

        me = sess.get("/api/v1/me").json()["user"]
        pw = getpass.getpass("New password: ")
        sess.post(f"/api/v1/users/{me['id']}/password", data=json.dumps(api.authentication.salt_and_hash(pw)))
    return 0


def workspace_cmd(args: argparse.Namespace) -> int:
    sess = _session(args)
    if args.action == "list":
        ws = sess.get("/api/v1/workspaces").json()["workspaces"]
        _table([[w["id"], w["name"], w["numProjects"], w["archived"]] for w in ws], ["ID", "Name", "# Projects", "Archived"], args.json)
    elif args.action == "create":
        w = sess.post("/api/v1/workspaces", json={"name": args.name}).json()["workspace"]
        print(f"Created workspace {w['name']} (id {w['id']})")
    elif args.action == "delete":
        ws = [w for w in sess.get("/api/v1/workspaces", params={"name": args.name}).json()["workspaces"]]
        if not ws:
            raise SystemExit("workspace not found")
        sess.delete(f"/api/v1/workspaces/{ws[0]['id']}")
        print(f"Deleted workspace {args.name}")
    elif args.action == "list-projects":
        ws = [w for w in sess.get("/api/v1/workspaces", params={"name": args.name}).json()["workspaces"]]
        if not ws:
            raise SystemExit("workspace not found")
        ps = sess.get(f"/api/v1/workspaces/{ws[0]['id']}/projects").json()["projects"]
        _table([[p["id"], p["name"], p["description"], p["numExperiments"]] for p in ps], ["ID", "Name", "Description", "# Experiments"], args.json)
    elif args.action == "create-project":
        ws = [w for w in sess.get("/api/v1/workspaces", params={"name": args.name}).json()["workspaces"]]
        if not ws:
            raise SystemExit("workspace not found")
        p = sess.post(f"/api/v1/workspaces/{ws[0]['id']}/projects", json={"name": args.project, "description": ""}).json()["project"]
        print(f"Created project {p['name']} (id {p['id']})")
    return 0


def project_cmd(args: argparse.Namespace) -> int:
    """``det project list|create|describe|delete|list-experiments WORKSPACE [PROJECT]`` (reference ``cli/project.py``)."""
    sess = _session(args)
    ws = [w for w in sess.get("/api/v1/workspaces", params={"name": args.workspace}).json()["workspaces"] if w["name"] == args.workspace]
    if not ws:
        raise SystemExit(f"workspace {args.workspace!r} not found")
    wid = ws[0]["id"]
    projects = sess.get(f"/api/v1/workspaces/{wid}/projects").json()["projects"]
    if args.action == "list":
        _table([[p["id"], p["name"], p["description"], p["numExperiments"]] for p in projects], ["ID", "Name", "Description", "# Experiments"], args.json)
        return 0
    if args.action == "create":
        p = sess.post(f"/api/v1/workspaces/{wid}/projects", json={"name": args.project, "description": args.description or ""}).json()["project"]
        print(f"Created project {p['name']} (id {p['id']}) in workspace {args.workspace}")
        return 0
    match = [p for p in projects if p["name"] == args.project]
    if not match:
        raise SystemExit(f"project {args.project!r} not found in workspace {args.workspace!r}")
    pid = match[0]["id"]
    if args.action == "describe":
        p = sess.get(f"/api/v1/projects/{pid}").json()["project"]
        if args.json:
            print(json.dumps(p, indent=2))
        else:
            _table([[p["id"], p["name"], p.get("description", ""), p.get("numExperiments", 0), p.get("workspaceId", wid)]],
                   ["ID", "Name", "Description", "# Experiments", "Workspace"], False)
    elif args.action == "list-experiments":
        es = sess.get("/api/v1/experiments", params={"projectId": pid, "limit": 500}).json()["experiments"]
        _table([[e["id"], e["name"], e["state"].replace("STATE_", ""), f"{100 * (e.get('progress') or 0):.0f}%", e.get("username", "")] for e in es],
               ["ID", "Name", "State", "Progress", "Owner"], args.json)
    elif args.action == "delete":
        sess.delete(f"/api/v1/projects/{pid}")
        print(f"Deleted project {args.project}")
    return 0


def task_cmd(args: argparse.Namespace) -> int:
    """``det task list | logs TASK_ID [-f] | kill TASK_ID`` over every kind of task (reference ``cli/task.py``)."""
    sess = _session(args)
    if args.action == "list":
        rows = []
        for aid, a in sess.get("/api/v1/tasks").json()["allocationIdToSummary"].items():
            devs = [f"{ag}:{','.join(str(d) for d in v['devices'])}" for r in a.get("resources", []) for ag, v in r.get("agentDevices", {}).items()]
            rows.append([a["taskId"], aid, a["name"], a["slotsNeeded"], a["state"], " ".join(devs)])
        _table(rows, ["Task ID", "Allocation ID", "Name", "Slots", "State", "Devices"], args.json)
        return 0
    if args.action == "logs":
        last = 0
        while True:
            logs = sess.get(f"/api/v1/tasks/{args.task_id}/logs", params={"offset": last}).json()["logs"]
            for ln in logs:
                sys.stdout.write(ln["log"])
                last = max(last, int(ln["id"]))
            running = any(a["taskId"] == args.task_id for a in sess.get("/api/v1/tasks").json()["allocationIdToSummary"].values())
            if not logs and (not args.follow or not running):
                return 0
            time.sleep(0.3)
    if args.action == "kill":
        for plural in ("commands", "shells", "notebooks", "tensorboards"):
            try:
                sess.post(f"/api/v1/{plural}/{args.task_id}/kill")
                print(f"Killed {plural[:-1]} {args.task_id}")
                return 0
            except api.errors.APIException:
                continue
        raise SystemExit(f"task {args.task_id} is not a command / shell / notebook / tensorboard (kill trials through their experiment)")
    return 2


def dev_cmd(args: argparse.Namespace) -> int:
    """``det dev curl PATH`` / ``det dev bindings list|call`` (reference ``cli/dev.py``): raw authenticated API access."""
    sess = _session(args)
    if args.action == "curl":
        body = json.loads(args.data) if args.data else None
        r = getattr(sess, args.request.lower())(args.path, **({"json": body} if body is not None else {}))
        try:
            print(json.dumps(r.json(), indent=2))
        except ValueError:
            print(r.text)
        return 0
    from determined_b200.common.api import bindings

    names = sorted(n for n in dir(bindings) if n.split("_")[0] in ("get", "post", "put", "patch", "delete") and callable(getattr(bindings, n)))
    if args.action == "list":
        for n in names:
            print(f"{n:60s} {(getattr(bindings, n).__doc__ or '').strip()}")
        return 0
    if args.name not in names:
        raise SystemExit(f"no binding named {args.name!r}; see `det dev bindings list`")
    kw: Dict[str, Any] = {}
    for item in args.args:
        k, _, v = item.partition("=")
        try:
            kw[k] = json.loads(v)
        except ValueError:
            kw[k] = v
    print(json.dumps(getattr(bindings, args.name)(sess, **kw), indent=2))
    return 0


def rbac_cmd(args: argparse.Namespace) -> int:
    """``det user-group ...`` / ``det rbac ...`` (reference ``harness/determined/cli/{user_groups,rbac}.py``)."""
    sess = _session(args)

    def uid(name: str) -> int:
        for u in sess.get("/api/v1/users").json()["users"]:
            if u["username"] == name:
                return int(u["id"])
        raise SystemExit(f"user {name} not found")

    def gid(name: str) -> int:
        for g in sess.get("/api/v1/groups").json()["groups"]:
            if g["group"]["name"] == name:
                return int(g["group"]["groupId"])
        raise SystemExit(f"group {name} not found")

    def rid(name: str) -> int:
        for r in sess.get("/api/v1/roles").json()["roles"]:
            if r["name"] == name:
                return int(r["roleId"])
        raise SystemExit(f"role {name} not found")

    a = args.action
    if a == "group-create":
        g = sess.post("/api/v1/groups", json={"name": args.name, "addUsers": [uid(u) for u in args.add_user or []]}).json()["group"]
        print(f"Created group {g['name']} (id {g['groupId']})")
    elif a == "group-list":
        gs = sess.get("/api/v1/groups").json()["groups"]
        _table([[g["group"]["groupId"], g["group"]["name"], g["numMembers"]] for g in gs], ["ID", "Name", "# Members"], args.json)
    elif a == "group-describe":
        g = sess.get(f"/api/v1/groups/{gid(args.name)}").json()["group"]
        _table([[u["id"], u["username"]] for u in g["users"]], ["User ID", "Username"], args.json)
    elif a in ("group-add-user", "group-remove-user"):
        key = "addUsers" if a == "group-add-user" else "removeUsers"
        sess.put(f"/api/v1/groups/{gid(args.name)}", json={key: [uid(u) for u in args.users]})
        print("ok")
    elif a == "group-delete":
        sess.delete(f"/api/v1/groups/{gid(args.name)}")
        print(f"Deleted group {args.name}")
    elif a == "list-roles":
        _table([[r["roleId"], r["name"]] for r in sess.get("/api/v1/roles").json()["roles"]], ["ID", "Role"], args.json)
    elif a in ("assign-role", "unassign-role"):
        ra = {"role": {"roleId": rid(args.role)}, "scopeWorkspaceId": args.workspace_id}
        body: Dict[str, Any] = {"userRoleAssignments": [], "groupRoleAssignments": []}
        if args.username:
            body["userRoleAssignments"].append({"userId": uid(args.username), "roleAssignment": ra})
        if args.group_name:
            body["groupRoleAssignments"].append({"groupId": gid(args.group_name), "roleAssignment": ra})
        sess.post("/api/v1/roles/" + ("add-assignments" if a == "assign-role" else "remove-assignments"), json=body)
        print("ok")
    elif a == "list-users-roles":
        rs = sess.get(f"/api/v1/roles/search/by-user/{uid(args.username)}").json()["roles"]
        _table([[r["role"]["name"], r["scopeWorkspaceId"], r["groupId"]] for r in rs], ["Role", "Workspace", "Via group"], args.json)
    return 0


def ntsc_cmd(args: argparse.Namespace) -> int:
    """``det cmd|shell|notebook|tensorboard ...`` (reference ``harness/determined/cli/{command,shell,notebook,tensorboard}.py``)."""
    sess = _session(args)
    kind, plural = args.kind, args.kind + "s"
    if args.action in ("run", "start"):
        cfg: Dict[str, Any] = {}
        if getattr(args, "config_file", None):
            cfg = yaml.safe_load(pathlib.Path(args.config_file).read_text()) or {}
        for kv in getattr(args, "config", None) or []:
            k, v = kv.split("=", 1)
            cur = cfg
            parts = k.split(".")
            for part in parts[:-1]:
                cur = cur.setdefault(part, {})
            cur[parts[-1]] = yaml.safe_load(v)
        body: Dict[str, Any] = {"config": cfg}
        if kind == "command":
            if not args.entrypoint:
                raise SystemExit("det cmd run: a command is required")
            cfg["entrypoint"] = [x for i, x in enumerate(args.entrypoint) if not (i == 0 and x == "--")]
        if kind == "tensorboard":
            body["experimentIds"] = [int(x) for x in (args.experiment_ids or [])]
            body["trialIds"] = [int(x) for x in (args.trial_ids or [])]
        if getattr(args, "slots", None) is not None:
            cfg.setdefault("resources", {})["slots"] = args.slots
        if getattr(args, "template", None):
            body["templateName"] = args.template
        files = _collect_context(getattr(args, "context", None), getattr(args, "include", None))
        if files:
            body["files"] = files
        for vol in getattr(args, "volume", None) or []:
            host, _, cont = vol.partition(":")
            cfg.setdefault("bind_mounts", []).append({"host_path": host, "container_path": cont or host})
        if getattr(args, "workspace_name", None):
            ws = sess.get("/api/v1/workspaces", params={"name": args.workspace_name}).json()["workspaces"]
            if not ws:
                raise SystemExit(f"workspace {args.workspace_name!r} not found")
            body["workspaceId"] = ws[0]["id"]
        if getattr(args, "preview", False):
            print(yaml.safe_dump(cfg, default_flow_style=False))
            return 0
        t = sess.post(f"/api/v1/{plural}", json=body).json()[kind]
        print(f"Launched {kind} (id: {t['id']}).")
        if getattr(args, "detach", True) is False:
            return _follow_task_logs(sess, t["id"], plural, kind)
        return 0
    if args.action == "list":
        ts = sess.get(f"/api/v1/{plural}").json()[plural]
        if not getattr(args, "all", False):
            me = sess.get("/api/v1/me").json()["user"]["username"]
            ts = [t for t in ts if t["username"] == me or t["state"] != "STATE_TERMINATED"] if False else ts
        if getattr(args, "quiet", False):
            print("\n".join(t["id"] for t in ts))
            return 0
        _table([[t["id"], t["username"], t["description"], t["state"].replace("STATE_", ""), t["exitStatus"], t["resourcePool"]] for t in ts],
               ["ID", "Owner", "Description", "State", "Exit Status", "Resource Pool"], args.json)
        return 0
    if args.action == "kill":
        for tid in args.ids:
            sess.post(f"/api/v1/{plural}/{tid}/kill")
            print(f"Killed {kind} {tid}")
        return 0
    if args.action == "logs":
        return _follow_task_logs(sess, args.id, plural, kind, follow=args.follow)
    if args.action == "describe":
        print(json.dumps(sess.get(f"/api/v1/{plural}/{args.id}").json()[kind], indent=2))
        return 0
    return 2


def _collect_context(context: Optional[str], includes: Optional[List[str]]) -> List[Dict[str, Any]]:
    """File entries of ``--context DIR`` plus every ``--include PATH`` (a directory keeps its name as a prefix)."""
    from determined_b200.common import context as det_context

    files: List[Dict[str, Any]] = []
    if context:
        files += det_context.read_v1_context(pathlib.Path(context))
    for inc in includes or []:
        p = pathlib.Path(inc)
        if p.is_dir():
            files += [dict(f, path=f"{p.name}/{f['path']}") for f in det_context.read_v1_context(p)]
        elif p.is_file():
            files += [f for f in det_context.read_v1_context(p.parent) if f["path"] == p.name]
        else:
            raise SystemExit(f"--include {inc}: no such file or directory")
    return files


def _follow_task_logs(sess: api.Session, task_id: str, plural: str, kind: str, follow: bool = True) -> int:
    last = 0
    while True:
        logs = sess.get(f"/api/v1/tasks/{task_id}/logs", params={"offset": last}).json()["logs"]
        for l in logs:
            sys.stdout.write(l["log"])
            last = max(last, int(l["id"]))
        t = sess.get(f"/api/v1/{plural}/{task_id}").json()[kind]
        if t["state"] == "STATE_TERMINATED" and not logs:
            return int(t["exitStatus"] or 0)
        if not follow and not logs:
            return 0
        time.sleep(0.3)


def deploy_local(args: argparse.Namespace) -> int:
    """``det deploy local cluster-up | cluster-down | master-up | master-down | agent-up | agent-down | logs`` (reference
    ``harness/determined/deploy/local``: docker-compose master+db+agent; here plain daemons with pid files)."""
    from determined_b200.deploy import local

    if args.action in ("cluster-up", "master-up"):
        pid = local.master_up(port=args.master_port, cluster_name=args.cluster_name, slots=args.slots, scheduler=args.scheduler,
                              detach=not args.no_detach)
        if not args.no_detach:
            print(f"master started (pid {pid}) on http://127.0.0.1:{args.master_port}; `det deploy local logs` shows its output")
        return 0
    if args.action in ("cluster-down", "master-down"):
        print("master stopped" if local.master_down(args.cluster_name, delete_db=args.delete_db) else "master was not running")
        return 0
    if args.action == "agent-up":
        pid = local.agent_up(master_url=args.master or f"http://127.0.0.1:{args.master_port}", agent_name=args.agent_name, slots=args.slots,
                             slot_type=args.slot_type, detach=not args.no_detach)
        if not args.no_detach:
            print(f"agent {args.agent_name} started (pid {pid})")
        return 0
    if args.action == "agent-down":
        print("agent stopped" if local.agent_down(args.agent_name) else "agent was not running")
        return 0
    if args.action == "logs":
        local.logs(args.cluster_name, follow=args.follow)
        return 0
    return 2


def version(args: argparse.Namespace) -> int:
    out = {"client": {"version": det.__version__}}
    try:
        sess = api.UnauthSession(args.master or os.environ.get("DET_MASTER") or "http://127.0.0.1:8080", max_retries=0)
        out["master"] = sess.get("/api/v1/master").json()
    except Exception as e:
        out["master"] = {"error": str(e)}
    print(yaml.safe_dump(out, default_flow_style=False))
    return 0


def make_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="det", description="determined_b200 command line client")
    p.add_argument("-m", "--master", help="master address (default: $DET_MASTER or localhost:8080)")
    p.add_argument("-u", "--user", help="run as this user")
    p.add_argument("--cert-file", help="PEM certificate (or CA bundle) to trust for an https master (default: $DET_MASTER_CERT_FILE)")
    p.add_argument("--cert-name", help="host name the master's certificate carries, when not the one in the URL")
    p.add_argument("--noverify", action="store_true", help="do not verify the master's certificate")
    p.add_argument("-v", "--version", action="version", version=f"det {det.__version__}")
    sub = p.add_subparsers(dest="cmd")

    def outfmt(sp: argparse.ArgumentParser) -> None:
        sp.add_argument("--json", action="store_true")
        sp.add_argument("--csv", action="store_true")

    # experiment
    e = sub.add_parser("experiment", aliases=["e"], help="manage experiments").add_subparsers(dest="sub")
    c = e.add_parser("create", help="create experiment")
    c.add_argument("config_file")
    c.add_argument("model_def", nargs="?", default=None)
    c.add_argument("--local", action="store_true", help="run locally in this process, no master")
    c.add_argument("--test", "--test-mode", "--test_mode", action="store_true", help="one batch of train/validate/checkpoint")
    c.add_argument("--paused", action="store_true")
    c.add_argument("-f", "--follow-first-trial", action="store_true")
    c.add_argument("--config", action="append", help="override key=value (dotted)")
    c.add_argument("--template")
    c.add_argument("--project_id", "--project-id", type=int, default=1)
    c.add_argument("-i", "--include", action="append")
    c.add_argument("--publish", action="append")
    c.set_defaults(fn=experiment_create)
    l = e.add_parser("list", aliases=["ls"])
    l.add_argument("--all", "-a", action="store_true")
    l.add_argument("--limit", type=int)
    l.add_argument("--offset", type=int)
    l.add_argument("--pages", choices=["1", "all"], default="all", help="accepted for compatibility: one request returns everything asked for")
    l.add_argument("--show_project", "--show-project", action="store_true")
    outfmt(l)
    l.set_defaults(fn=experiment_list)
    d = e.add_parser("describe")
    d.add_argument("experiment_ids")
    outfmt(d)
    d.set_defaults(fn=experiment_describe)
    for name in ("pause", "activate", "cancel", "kill", "archive", "unarchive"):
        a = e.add_parser(name)
        a.add_argument("experiment_id", type=int)
        a.set_defaults(fn=_exp_action(name))
    a = e.add_parser("delete")
    a.add_argument("experiment_id", type=int)
    a.add_argument("--yes", action="store_true")
    a.set_defaults(fn=experiment_delete)
    a = e.add_parser("logs")
    a.add_argument("experiment_id", type=int)
    a.add_argument("-f", "--follow", action="store_true")
    a.add_argument("--head", type=int)
    a.add_argument("--tail", type=int)
    a.set_defaults(fn=experiment_logs)
    a = e.add_parser("download")
    a.add_argument("experiment_id", type=int)
    a.add_argument("-o", "--output-dir", default=".")
    a.add_argument("--top-n", type=int, default=1)
    a.add_argument("--sort-by")
    a.add_argument("--smaller-is-better", type=lambda s: s.lower() in ("1", "true", "yes"), default=None)
    a.set_defaults(fn=experiment_download)
    a = e.add_parser("set")
    a.add_argument("attr", choices=["name", "description", "priority", "weight", "max-slots", "gc-policy", "log-retention"])
    a.add_argument("experiment_id", type=int)
    a.add_argument("value")
    a.set_defaults(fn=experiment_set)
    a = e.add_parser("label")
    a.add_argument("action", choices=["add", "remove"])
    a.add_argument("experiment_id", type=int)
    a.add_argument("label")
    a.set_defaults(fn=experiment_label)
    a = e.add_parser("config")
    a.add_argument("experiment_id", type=int)
    a.set_defaults(fn=experiment_config)
    a = e.add_parser("continue")
    a.add_argument("experiment_id", type=int)
    a.add_argument("--config-file")
    a.add_argument("--config", action="append")
    a.add_argument("-f", "--follow-first-trial", action="store_true")
    a.set_defaults(fn=experiment_continue)
    a = e.add_parser("move")
    a.add_argument("experiment_id", type=int)
    a.add_argument("workspace")
    a.add_argument("project")
    a.set_defaults(fn=experiment_move)
    a = e.add_parser("download-model-def")
    a.add_argument("experiment_id", type=int)
    a.add_argument("--output-dir", default=".")
    a.set_defaults(fn=experiment_download_model_def)
    a = e.add_parser("wait")
    a.add_argument("experiment_id", type=int)
    a.add_argument("--polling-interval", type=float, default=5)
    a.set_defaults(fn=experiment_wait)
    a = e.add_parser("list-checkpoints", aliases=["lc"])
    a.add_argument("experiment_id", type=int)
    outfmt(a)
    a.set_defaults(fn=experiment_list_checkpoints)
    a = e.add_parser("list-trials", aliases=["lt"])
    a.add_argument("experiment_id", type=int)
    a.add_argument("--limit", type=int)
    a.add_argument("--offset", type=int)
    a.add_argument("--pages", choices=["1", "all"], default="all")
    outfmt(a)
    a.set_defaults(fn=trial_list)

    # trial
    t = sub.add_parser("trial", aliases=["t"], help="manage trials").add_subparsers(dest="sub")
    a = t.add_parser("describe")
    a.add_argument("trial_id", type=int)
    outfmt(a)
    a.set_defaults(fn=trial_describe)
    a = t.add_parser("list")
    a.add_argument("experiment_id", type=int)
    outfmt(a)
    a.set_defaults(fn=trial_list)
    a = t.add_parser("logs")
    a.add_argument("trial_id", type=int)
    a.add_argument("-f", "--follow", action="store_true")
    a.add_argument("--head", type=int)
    a.add_argument("--tail", type=int)
    a.set_defaults(fn=trial_logs)
    a = t.add_parser("kill")
    a.add_argument("trial_id", type=int)
    a.set_defaults(fn=trial_kill)
    a = t.add_parser("download")
    a.add_argument("trial_id", type=int)
    g = a.add_mutually_exclusive_group()
    g.add_argument("--latest", action="store_true", default=True)
    g.add_argument("--best", action="store_true")
    g.add_argument("--uuid")
    a.add_argument("-o", "--output-dir")
    a.set_defaults(fn=trial_download)

    # checkpoint
    c = sub.add_parser("checkpoint", aliases=["c"], help="manage checkpoints").add_subparsers(dest="sub")
    a = c.add_parser("describe")
    a.add_argument("uuid")
    outfmt(a)
    a.set_defaults(fn=checkpoint_describe)
    a = c.add_parser("download")
    a.add_argument("uuid")
    a.add_argument("-o", "--output-dir", default=None)
    a.add_argument("-q", "--quiet", action="store_true")
    a.add_argument("--mode", choices=["auto", "direct", "master"], default="auto")
    a.set_defaults(fn=checkpoint_download)
    a = c.add_parser("delete")
    a.add_argument("uuid", help="comma-separated uuids")
    a.add_argument("--yes", action="store_true")
    a.set_defaults(fn=checkpoint_delete)
    a = c.add_parser("rm")
    a.add_argument("uuid", help="comma-separated uuids")
    a.add_argument("--glob", action="append", required=True)
    a.set_defaults(fn=checkpoint_rm)
    a = c.add_parser("list")
    a.add_argument("experiment_id", type=int)
    outfmt(a)
    a.set_defaults(fn=experiment_list_checkpoints)

    a = sub.add_parser("preview-search", help="simulate the searcher for a config")
    a.add_argument("config_file")
    a.add_argument("--config", action="append")
    a.add_argument("--seed", type=int, default=0)
    a.set_defaults(fn=preview_search)

    ms = sub.add_parser("master").add_subparsers(dest="sub")
    a = ms.add_parser("info")
    a.add_argument("--json", action="store_true")
    a.add_argument("--yaml", action="store_true")
    a.set_defaults(fn=master_info)
    a = ms.add_parser("config")
    a.set_defaults(fn=master_config)

    ag = sub.add_parser("agent", aliases=["a"]).add_subparsers(dest="sub")
    a = ag.add_parser("list", aliases=["ls"])
    outfmt(a)
    a.set_defaults(fn=agent_list)
    for name, en in (("enable", True), ("disable", False)):
        a = ag.add_parser(name)
        a.add_argument("agent_id", nargs="?")
        a.add_argument("--all", action="store_true", help="every agent")
        if not en:
            a.add_argument("--drain", action="store_true", help="let running tasks finish")
        outfmt(a)
        a.set_defaults(fn=agent_enable, enable=en)
    sl = sub.add_parser("slot", aliases=["s"]).add_subparsers(dest="sub")
    a = sl.add_parser("list", aliases=["ls"])
    outfmt(a)
    a.set_defaults(fn=slot_list)
    for name, en in (("enable", True), ("disable", False)):
        a = sl.add_parser(name)
        a.add_argument("agent_id")
        a.add_argument("slot_id")
        a.set_defaults(fn=slot_toggle(en))

    jb = sub.add_parser("job", aliases=["j"]).add_subparsers(dest="sub")
    a = jb.add_parser("list", aliases=["ls"])
    a.add_argument("--resource-pool")
    a.add_argument("--limit", type=int)
    a.add_argument("--offset", type=int)
    a.add_argument("--pages", choices=["1", "all"], default="all")
    a.add_argument("--reverse", action="store_true")
    a.add_argument("--yaml", action="store_true")
    a.add_argument("--table", action="store_true", help="the default output")
    outfmt(a)
    a.set_defaults(fn=job_list)
    a = jb.add_parser("update")
    a.add_argument("job_id")
    a.add_argument("--priority", type=int)
    a.add_argument("--weight", type=float)
    a.add_argument("--ahead-of", help="job id to queue just before")
    a.add_argument("--behind-of", help="job id to queue just after")
    a.set_defaults(fn=job_update)
    rp = sub.add_parser("resource-pool", aliases=["rp"]).add_subparsers(dest="sub")
    a = rp.add_parser("list", aliases=["ls"])
    outfmt(a)
    a.set_defaults(fn=resource_pool_list)

    tp = sub.add_parser("template", aliases=["tpl"]).add_subparsers(dest="sub")
    for name in ("list", "describe", "set", "create", "remove"):
        a = tp.add_parser(name)
        if name != "list":
            a.add_argument("name")
        if name in ("set", "create"):
            a.add_argument("file")
        a.set_defaults(fn=template_cmd, action=name)

    md = sub.add_parser("model").add_subparsers(dest="sub")
    for name in ("list", "create", "describe", "list-versions", "register-version", "delete"):
        a = md.add_parser(name)
        if name != "list":
            a.add_argument("name")
        if name == "create":
            a.add_argument("--description")
        if name == "register-version":
            a.add_argument("uuid")
        outfmt(a)
        a.set_defaults(fn=model_cmd, action=name)

    us = sub.add_parser("user", aliases=["u"]).add_subparsers(dest="sub")
    for name in ("login", "logout", "whoami", "list", "create", "activate", "deactivate", "rename", "change-password"):
        a = us.add_parser(name)
        if name in ("login", "create", "activate", "deactivate", "rename"):
            a.add_argument("username")
        if name == "rename":
            a.add_argument("new_username")
        if name == "create":
            a.add_argument("--admin", action="store_true")
        outfmt(a)
        a.set_defaults(fn=user_cmd, action=name)
    a = sub.add_parser("auth").add_subparsers(dest="sub")
    for name in ("login", "logout"):
        b = a.add_parser(name)
        if name == "login":
            b.add_argument("username", nargs="?", default="determined")
        b.set_defaults(fn=user_cmd, action=name)

    ws = sub.add_parser("workspace", aliases=["w"]).add_subparsers(dest="sub")
    for name in ("list", "create", "delete", "list-projects", "create-project"):
        a = ws.add_parser(name)
        if name != "list":
            a.add_argument("name")
        if name == "create-project":
            a.add_argument("project")
        outfmt(a)
        a.set_defaults(fn=workspace_cmd, action=name)

    pj = sub.add_parser("project", aliases=["p"]).add_subparsers(dest="sub")
    for name in ("list", "create", "describe", "delete", "list-experiments"):
        a = pj.add_parser(name)
        a.add_argument("workspace")
        if name != "list":
            a.add_argument("project")
        if name == "create":
            a.add_argument("--description")
        outfmt(a)
        a.set_defaults(fn=project_cmd, action=name)

    tk = sub.add_parser("task").add_subparsers(dest="sub")
    a = tk.add_parser("list", aliases=["ls"])
    outfmt(a)
    a.set_defaults(fn=task_cmd, action="list")
    a = tk.add_parser("logs")
    a.add_argument("task_id")
    a.add_argument("-f", "--follow", action="store_true")
    a.set_defaults(fn=task_cmd, action="logs")
    a = tk.add_parser("kill")
    a.add_argument("task_id")
    a.set_defaults(fn=task_cmd, action="kill")

    dv = sub.add_parser("dev").add_subparsers(dest="sub")
    a = dv.add_parser("curl")
    a.add_argument("path")
    a.add_argument("-X", "--request", default="GET", choices=["GET", "POST", "PUT", "PATCH", "DELETE"])
    a.add_argument("-d", "--data", help="JSON request body")
    a.set_defaults(fn=dev_cmd, action="curl")
    bd = dv.add_parser("bindings").add_subparsers(dest="bsub")
    a = bd.add_parser("list")
    a.set_defaults(fn=dev_cmd, action="list")
    a = bd.add_parser("call")
    a.add_argument("name")
    a.add_argument("args", nargs="*", help="key=value (values are parsed as JSON when possible; body={...} for the payload)")
    a.set_defaults(fn=dev_cmd, action="call")

    for kind, aliases in (("command", ["cmd"]), ("shell", []), ("notebook", []), ("tensorboard", [])):
        cp = sub.add_parser(kind, aliases=aliases).add_subparsers(dest="sub")
        start = cp.add_parser("run" if kind == "command" else "start")
        start.add_argument("--config-file")
        start.add_argument("--config", action="append")
        start.add_argument("--template")
        start.add_argument("--slots", type=int)
        start.add_argument("-c", "--context", help="directory shipped to the task's working directory")
        start.add_argument("-i", "--include", action="append", help="extra file or directory to ship (repeatable)")
        start.add_argument("-v", "--volume", action="append", help="HOST_PATH:TASK_PATH bind mount (recorded in the config; tasks are host processes here)")
        start.add_argument("-w", "--workspace-name")
        if kind in ("notebook", "tensorboard"):
            start.add_argument("--no-browser", action="store_true")
        if kind == "notebook":
            start.add_argument("--preview", action="store_true", help="print the configuration instead of launching")
        if kind == "shell":
            start.add_argument("-p", "--passphrase", action="store_true", help="accepted for compatibility")
            start.add_argument("--show-ssh-command", action="store_true")
            start.add_argument("ssh_opts", nargs="*")
        start.add_argument("-d", "--detach", action="store_true", default=(kind != "command"))
        if kind == "command":
            start.add_argument("entrypoint", nargs=argparse.REMAINDER)
        if kind == "tensorboard":
            start.add_argument("experiment_ids", nargs="*")
            start.add_argument("-t", "--trial-ids", nargs="*")
        start.set_defaults(fn=ntsc_cmd, kind=kind, action="run" if kind == "command" else "start")
        a = cp.add_parser("list", aliases=["ls"])
        a.add_argument("-a", "--all", action="store_true")
        a.add_argument("-q", "--quiet", action="store_true", help="only print the ids")
        a.add_argument("-w", "--workspace-name", action="append")
        a.add_argument("--sort-by")
        a.add_argument("--order-by", choices=["asc", "desc"])
        outfmt(a)
        a.set_defaults(fn=ntsc_cmd, kind=kind, action="list")
        a = cp.add_parser("kill")
        a.add_argument("-f", "--force", action="store_true", help="accepted for compatibility: kill always signals the whole process group")
        a.add_argument("ids", nargs="+")
        a.set_defaults(fn=ntsc_cmd, kind=kind, action="kill")
        a = cp.add_parser("logs")
        a.add_argument("id")
        a.add_argument("-f", "--follow", action="store_true")
        a.set_defaults(fn=ntsc_cmd, kind=kind, action="logs")
        a = cp.add_parser("describe")
        a.add_argument("id")
        a.set_defaults(fn=ntsc_cmd, kind=kind, action="describe")

    ug = sub.add_parser("user-group").add_subparsers(dest="sub")
    for name, act in (("create", "group-create"), ("list", "group-list"), ("describe", "group-describe"), ("add-user", "group-add-user"),
                      ("remove-user", "group-remove-user"), ("delete", "group-delete")):
        a = ug.add_parser(name)
        if act != "group-list":
            a.add_argument("name")
        if act == "group-create":
            a.add_argument("--add-user", action="append")
        if act in ("group-add-user", "group-remove-user"):
            a.add_argument("users", nargs="+")
        outfmt(a)
        a.set_defaults(fn=rbac_cmd, action=act)
    rb = sub.add_parser("rbac").add_subparsers(dest="sub")
    a = rb.add_parser("list-roles")
    outfmt(a)
    a.set_defaults(fn=rbac_cmd, action="list-roles")
    for name in ("assign-role", "unassign-role"):
        a = rb.add_parser(name)
        a.add_argument("role")
        a.add_argument("-u", "--username")
        a.add_argument("-g", "--group-name")
        a.add_argument("-w", "--workspace-id", type=int)
        a.set_defaults(fn=rbac_cmd, action=name)
    a = rb.add_parser("list-users-roles")
    a.add_argument("username")
    outfmt(a)
    a.set_defaults(fn=rbac_cmd, action="list-users-roles")

    dp = sub.add_parser("deploy", aliases=["d"]).add_subparsers(dest="sub")
    lo = dp.add_parser("local")
    lo.add_argument("action", choices=["cluster-up", "cluster-down", "master-up", "master-down", "agent-up", "agent-down", "logs"])
    lo.add_argument("--master-port", type=int, default=8080)
    lo.add_argument("--cluster-name", default="determined")
    lo.add_argument("--agent-name", default="determined-agent")
    lo.add_argument("--slots", type=int, help="master: slots it drives itself (0 = agents only); agent: slots it offers")
    lo.add_argument("--slot-type", choices=["auto", "cuda", "cpu"], default="auto")
    lo.add_argument("--scheduler", choices=["fair_share", "priority", "round_robin"])
    lo.add_argument("--no-detach", action="store_true")
    lo.add_argument("--delete-db", action="store_true")
    lo.add_argument("--follow", "-f", action="store_true")
    lo.set_defaults(fn=deploy_local)

    a = sub.add_parser("version")
    a.set_defaults(fn=version)
    from determined_b200.cli import more

    more.extend(p)
    from determined_b200.cli import flags

    flags.extend(p)
    return p


def main(argv: Optional[List[str]] = None) -> int:
    parser = make_parser()
    args = parser.parse_args(argv)
    fn = getattr(args, "fn", None)
    if fn is None:
        parser.print_help()
        return 2
    import requests

    from determined_b200.common.api import certs

    _setup_cert(args)
    from determined_b200.cli import flags

    VIEW.clear()
    VIEW.update(flags.view_of(args))
    try:
        try:
            return int(fn(args) or 0)
        except requests.exceptions.SSLError as e:
            # a certificate nobody vouches for: show its fingerprint and, if the user agrees, remember it (trust on first use)
            if not sys.stdin or not sys.stdin.isatty():
                print(f"Failed: the certificate of {_master_of(args)} is not trusted ({e}); pass --cert-file, set "
                      "DET_MASTER_CERT_FILE, or run det in a terminal once to accept it", file=sys.stderr)
                return 1
            try:
                certs.cli_cert = certs.trust_on_first_use(_master_of(args))
            except PermissionError as pe:
                print(f"Failed: {pe}", file=sys.stderr)
                return 1
            return int(fn(args) or 0)
    except api.errors.APIException as e:
        print(f"Failed: {e.message}", file=sys.stderr)
        return 1
    except api.errors.MasterNotFoundException as e:
        print(f"Failed: {e}", file=sys.stderr)
        return 1
    except KeyboardInterrupt:
        return 130
