"""Third pass over the ``det`` parser: options of the reference's commands that the first two passes (``cli.py``, ``more.py``)
do not give.  Everything here is either a *view* option (sort / order / limit / offset of a printed list — applied by
``cli._table`` to whatever the command prints), a confirmation prompt, or a small handler of its own.

Parity: reference ``harness/determined/cli/{project,workspace,model,rbac,user_groups,trial,user,task,template,job,experiment,
checkpoint,dev}.py`` argument lists.  ``docs/CLI_GAPS.md`` is the generated list of what is still different.
"""
import argparse
import getpass
import json
import os
import pathlib
import shlex
import subprocess
import sys
from typing import Any, Callable, Dict, List, Optional

import yaml

from determined_b200.common import api


def _cli() -> Any:
    from determined_b200.cli import cli

    return cli


def find(parser: argparse.ArgumentParser, path: str) -> Optional[argparse.ArgumentParser]:
    cur = parser
    for name in path.split():
        subs = next((a for a in cur._actions if isinstance(a, argparse._SubParsersAction)), None)
        if subs is None or name not in subs.choices:
            return None
        cur = subs.choices[name]
    return cur


def _has(sp: argparse.ArgumentParser, flag: str) -> bool:
    return any(flag in a.option_strings or a.dest == flag.lstrip("-").replace("-", "_") for a in sp._actions)


def _add(sp: argparse.ArgumentParser, *flags: str, **kw: Any) -> None:
    if not any(_has(sp, f) for f in flags):
        sp.add_argument(*flags, **kw)


def wrap(sp: argparse.ArgumentParser, around: Callable[[Callable[[argparse.Namespace], int], argparse.Namespace], int]) -> None:
    inner = sp.get_default("fn")
    sp.set_defaults(fn=lambda args, _i=inner: around(_i, args))


# ---------------------------------------------------------------------------------------------------------------- list views
def view_flags(sp: argparse.ArgumentParser, sort: bool = True, paging: bool = True, pages: bool = False) -> None:
    if sort:
        _add(sp, "--sort-by", help="column to sort the list by (its header, lower case, spaces as '_': id, name, ...)")
        _add(sp, "--order-by", choices=["asc", "desc"], default="asc")
    if paging:
        _add(sp, "--limit", type=int, help="print at most this many rows")
        _add(sp, "--offset", type=int, default=0, help="skip this many rows")
    if pages:
        _add(sp, "--pages", choices=["1", "all"], default="all", help="with --limit: one page, or all pages one after another")
    sp.set_defaults(view=True)


def _col(h: str) -> str:
    return h.lower().replace("#", "").replace(" of ", " ").strip().replace(" ", "_").replace("-", "_")


def apply_view(rows: List[Any], headers: Any, view: Dict[str, Any]) -> List[Any]:
    """Sort and cut the rows of a table the way the view options of the command line ask."""
    rows = list(rows)
    key = view.get("sort_by")
    if key:
        cols = [_col(h) for h in headers]
        want = _col(str(key))
        if want not in cols:
            raise SystemExit(f"--sort-by {key}: no such column; one of: {', '.join(cols)}")
        i = cols.index(want)
        rows.sort(key=lambda r: (r[i] is None, r[i] if isinstance(r[i], (int, float)) else str(r[i])), reverse=view.get("order_by") == "desc")
    elif view.get("order_by") == "desc":
        rows.reverse()
    off = int(view.get("offset") or 0)
    lim = view.get("limit")
    if view.get("pages") == "all" and lim and "pages_explicit" not in view:
        lim = None                              # every page, one after another, is the whole list
    return rows[off:off + int(lim)] if lim else rows[off:]


def view_of(args: argparse.Namespace) -> Dict[str, Any]:
    if not getattr(args, "view", False):
        return {}
    v = {k: getattr(args, k, None) for k in ("sort_by", "order_by", "limit", "offset", "pages")}
    if getattr(args, "pages", None) == "1":
        v["pages_explicit"] = True
    elif getattr(args, "pages", None) is None:
        v["pages_explicit"] = True             # no --pages option on this command: --limit means what it says
    return v


# ------------------------------------------------------------------------------------------------------------- confirmations
def confirm(prompt: str) -> Callable[[Callable[[argparse.Namespace], int], argparse.Namespace], int]:
    def around(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
        if not getattr(args, "yes", False) and sys.stdin is not None and sys.stdin.isatty():
            what = prompt.format(**{k: v for k, v in vars(args).items() if isinstance(v, (str, int, list))})
            if input(f"{what} [y/N]: ").strip().lower() not in ("y", "yes"):
                print("Aborted.")
                return 1
        return inner(args)
    return around


# ------------------------------------------------------------------------------------------------------------------ handlers
def _ws_ids(sess: api.Session, names: List[str]) -> List[int]:
    out = []
    for n in names:
        ws = [w for w in sess.get("/api/v1/workspaces", params={"name": n}).json()["workspaces"] if w["name"] == n]
        if not ws:
            raise SystemExit(f"workspace {n!r} not found")
        out.append(int(ws[0]["id"]))
    return out


def model_list(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not args.workspace_names:
        return inner(args)
    c = _cli()
    sess = c._session(args)
    wids = set(_ws_ids(sess, args.workspace_names))
    ms = [m for m in sess.get("/api/v1/models").json()["models"] if int(m.get("workspaceId") or 1) in wids]
    c._table([[m["id"], m["name"], m.get("description", ""), m.get("archived", False), ",".join(m.get("labels") or [])] for m in ms],
             ["ID", "Name", "Description", "Archived", "Labels"], args.json)
    return 0


def model_describe(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if args.version is None:
        return inner(args)
    c = _cli()
    v = c._session(args).get(f"/api/v1/models/{args.name}/versions/{args.version}").json()["modelVersion"]
    if args.json:
        print(json.dumps(v, indent=2, default=str))
    else:
        ck = v.get("checkpoint") or {}
        c._table([[v.get("version"), ck.get("uuid"), v.get("name", ""), v.get("comment", ""), json.dumps(v.get("metadata") or {}), v.get("creationTime", "")]],
                 ["Version", "Checkpoint", "Name", "Comment", "Metadata", "Created"])
    return 0


def rbac_list_roles(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not args.exclude_global_roles:
        return inner(args)
    c = _cli()
    sess = c._session(args)
    rows = []
    for r in sess.get("/api/v1/roles").json()["roles"]:
        perms = (sess.get(f"/api/v1/roles/{r['roleId']}").json().get("role") or {}).get("permissions") or []
        if any(p.get("scopeTypeMask", {}).get("cluster") and not p.get("scopeTypeMask", {}).get("workspace") for p in perms):
            continue                             # a role with a permission that only exists cluster-wide cannot be scoped to a workspace
        rows.append([r["roleId"], r["name"]])
    c._table(rows, ["ID", "Role"], args.json)
    return 0


def rbac_assign(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if getattr(args, "workspace_name", None):
        args.workspace_id = _ws_ids(_cli()._session(args), [args.workspace_name])[0]
    return inner(args)


def group_list(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not args.groups_user_belongs_to:
        return inner(args)
    c = _cli()
    sess = c._session(args)
    us = [u for u in sess.get("/api/v1/users").json()["users"] if u["username"] == args.groups_user_belongs_to]
    if not us:
        raise SystemExit(f"user {args.groups_user_belongs_to} not found")
    rows = []
    for g in sess.get("/api/v1/groups").json()["groups"]:
        full = sess.get(f"/api/v1/groups/{g['group']['groupId']}").json()["group"]
        if any(int(u["id"]) == int(us[0]["id"]) for u in full.get("users") or []):
            rows.append([g["group"]["groupId"], g["group"]["name"], g["numMembers"]])
    c._table(rows, ["ID", "Name", "# Members"], args.json)
    return 0


def trial_describe(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """``--metrics-summary``: the trial's summary metrics (per group: count / last / min / max / mean of every numeric metric)
    instead of the report-by-report table; ``--metrics`` is the default view and accepted for the reference's spelling."""
    if not getattr(args, "metrics_summary", False):
        return inner(args)
    c = _cli()
    sess = c._session(args)
    mets = sess.get(f"/api/v1/trials/{args.trial_id}/metrics").json()["metrics"]
    agg: Dict[Any, List[float]] = {}
    for m in mets:
        for k, v in (m["metrics"].get("avgMetrics") or {}).items():
            if isinstance(v, (int, float)) and not isinstance(v, bool):
                agg.setdefault((m["group"], k), []).append(float(v))
    rows = [[g, k, len(vs), vs[-1], min(vs), max(vs), sum(vs) / len(vs)] for (g, k), vs in sorted(agg.items())]
    if args.json:
        print(json.dumps({f"{g}.{k}": {"count": n, "last": la, "min": lo, "max": hi, "mean": me} for g, k, n, la, lo, hi, me in rows}, indent=2))
    else:
        c._table(rows, ["Group", "Metric", "Count", "Last", "Min", "Max", "Mean"], False, getattr(args, "csv", False))
    return 0


def trial_download(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """``--sort-by METRIC --smaller-is-better BOOL``: the checkpoint whose validation has the best METRIC."""
    if not args.sort_by:
        return _quietly(inner, args)
    if args.smaller_is_better is None:
        raise SystemExit("--sort-by needs --smaller-is-better true|false")
    ck = _cli()._sdk(args).get_trial(args.trial_id).select_checkpoint(best=True, sort_by=args.sort_by, smaller_is_better=args.smaller_is_better)
    args.uuid, args.sort_by = ck.uuid, None
    return _quietly(inner, args)


def _quietly(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """``--quiet``: only the path of what was downloaded reaches stdout."""
    if not getattr(args, "quiet", False):
        return inner(args)
    import contextlib
    import io

    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        rc = inner(args)
    for line in buf.getvalue().splitlines():
        if "path" in line.lower() and ":" in line:
            print(line.split(":", 1)[1].strip())
    return rc


def experiment_describe(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """``--outdir DIR``: write experiments.csv / trials.csv / metrics.csv (or .json) there instead of printing;
    ``--metrics`` adds every reported metric of every trial."""
    if not args.outdir and not args.metrics:
        return inner(args)
    c = _cli()
    sess = c._session(args)
    ids = args.experiment_ids if isinstance(getattr(args, "experiment_ids", None), list) else [getattr(args, "experiment_id", None) or args.experiment_ids]
    ids = [int(x) for i in ids for x in str(i).split(",") if str(x).strip()]
    exps, trials, mets = [], [], []
    for eid in ids:
        e = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]
        exps.append(e)
        for t in sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]:
            trials.append(t)
            if args.metrics:
                for m in sess.get(f"/api/v1/trials/{t['id']}/metrics").json()["metrics"]:
                    mets.append(dict(m, trialId=t["id"], experimentId=eid))
    tables = {"experiments": (exps, ["id", "name", "state", "progress", "startTime", "endTime", "description", "resourcePool"]),
              "trials": (trials, ["id", "experimentId", "state", "hparams", "startTime", "endTime", "totalBatchesProcessed", "restarts", "searcherMetricValue"]),
              "metrics": (mets, ["experimentId", "trialId", "group", "totalBatches", "endTime", "metrics"])}
    if not args.metrics:
        del tables["metrics"]
    as_json = getattr(args, "json", False)
    out = pathlib.Path(args.outdir) if args.outdir else None
    if out is not None:
        out.mkdir(parents=True, exist_ok=True)
    for name, (objs, cols) in tables.items():
        if as_json:
            text = json.dumps(objs, indent=2, default=str)
        else:
            import csv
            import io

            buf = io.StringIO()
            w = csv.writer(buf)
            w.writerow(cols)
            for o in objs:
                w.writerow([json.dumps(o.get(k)) if isinstance(o.get(k), (dict, list)) else o.get(k) for k in cols])
            text = buf.getvalue()
        if out is not None:
            (out / f"{name}.{'json' if as_json else 'csv'}").write_text(text)
        else:
            print(text)
    if out is not None:
        print(f"Wrote {', '.join(tables)} to {out}")
    return 0


def list_checkpoints(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """``--best N``: the N checkpoints with the best searcher metric, best first."""
    if args.best is None:
        return inner(args)
    c = _cli()
    sess = c._session(args)
    exp = sess.get(f"/api/v1/experiments/{args.experiment_id}").json()["experiment"]
    smaller = bool(((exp.get("config") or {}).get("searcher") or {}).get("smaller_is_better", True))
    cks = [k for k in sess.get(f"/api/v1/experiments/{args.experiment_id}/checkpoints").json()["checkpoints"]
           if k["state"] == "STATE_COMPLETED" and k["training"].get("searcherMetric") is not None]
    cks.sort(key=lambda k: float(k["training"]["searcherMetric"]), reverse=not smaller)
    c._table([[k["training"]["trialId"], k["metadata"].get("steps_completed"), k["state"].replace("STATE_", ""), k["training"]["searcherMetric"], k["uuid"],
               sum(int(v) for v in k["resources"].values())] for k in cks[:args.best]],
             ["Trial ID", "# of Batches", "State", "Validation Metric", "UUID", "Size"], getattr(args, "json", False), getattr(args, "csv", False))
    return 0


def workspace_create(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    rc = inner(args)
    patch: Dict[str, Any] = {}
    aug = {k: getattr(args, a) for k, a in (("agentUid", "agent_uid"), ("agentUser", "agent_user"), ("agentGid", "agent_gid"), ("agentGroup", "agent_group"))
           if getattr(args, a, None) is not None}
    if aug:
        if ("agentUid" in aug) != ("agentUser" in aug) or ("agentGid" in aug) != ("agentGroup" in aug):
            raise SystemExit("--agent-uid goes with --agent-user, --agent-gid with --agent-group")
        patch["agentUserGroup"] = aug
    if args.checkpoint_storage_config and args.checkpoint_storage_config_file:
        raise SystemExit("give --checkpoint-storage-config or --checkpoint-storage-config-file, not both")
    if args.checkpoint_storage_config:
        patch["checkpointStorageConfig"] = json.loads(args.checkpoint_storage_config)
    if args.checkpoint_storage_config_file:
        patch["checkpointStorageConfig"] = yaml.safe_load(pathlib.Path(args.checkpoint_storage_config_file).read_text())
    if args.default_compute_pool:
        patch["defaultComputePool"] = args.default_compute_pool
    if args.default_aux_pool:
        patch["defaultAuxPool"] = args.default_aux_pool
    if patch and rc == 0:
        sess = _cli()._session(args)
        wid = _ws_ids(sess, [args.name])[0]
        sess.patch(f"/api/v1/workspaces/{wid}", json=patch)
    return rc


def user_change_password(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not getattr(args, "target_user", None):
        return inner(args)
    sess = _cli()._session(args)
    us = [u for u in sess.get("/api/v1/users").json()["users"] if u["username"] == args.target_user]
    if not us:
        raise SystemExit(f"user {args.target_user} not found")
    pw = os.environ.get("DET_NEW_PASS") or getpass.getpass(f"New password for user '{args.target_user}': ")
    if not os.environ.get("DET_NEW_PASS") and getpass.getpass("Confirm password: ") != pw:
        raise SystemExit("passwords do not match")
    sess.post(f"/api/v1/users/{us[0]['id']}/password", data=json.dumps(api.authentication.salt_and_hash(pw)))
    print(f"Changed the password of {args.target_user}")
    return 0


def user_create(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not getattr(args, "remote", False):
        return inner(args)
    # a remote user signs in through an identity provider only: no password can ever match
    sess = _cli()._session(args)
    sess.post("/api/v1/users", json={"user": {"username": args.username, "admin": args.admin, "active": True, "remote": True}, "password": "", "isHashed": True})
    print(f"Created remote user {args.username} (no password sign-in)")
    return 0


def user_list(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    c = _cli()
    us = c._session(args).get("/api/v1/users").json()["users"]
    if not getattr(args, "all", False):
        us = [u for u in us if u["active"]]
    c._table([[u["id"], u["username"], u["displayName"], u["active"], u["admin"], u.get("remote", False)] for u in us],
             ["ID", "Username", "Display Name", "Active", "Admin", "Remote"], args.json)
    return 0


def user_logout(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not getattr(args, "all", False):
        return inner(args)
    master = _cli()._master_of(args)
    st = api.authentication.TokenStore(master)
    for name in list(st.get_all_users()):
        api.authentication.logout(master, name)
    print("Logged out every user of this master")
    return 0


def project_view(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """``-a``: experiments of every user, archived ones included (the default list is the caller's own, unarchived)."""
    c = _cli()
    sess = c._session(args)
    if args.action not in ("describe", "list-experiments"):
        return inner(args)
    rc = 0
    if args.action == "describe":
        args2 = argparse.Namespace(**vars(args))
        rc = inner(args2)
        if getattr(args, "json", False):
            return rc
        print()
    wid = _ws_ids(sess, [args.workspace])[0]
    match = [p for p in sess.get(f"/api/v1/workspaces/{wid}/projects").json()["projects"] if p["name"] == args.project]
    if not match:
        raise SystemExit(f"project {args.project!r} not found in workspace {args.workspace!r}")
    es = sess.get("/api/v1/experiments", params={"projectId": match[0]["id"], "limit": 1000}).json()["experiments"]
    if not getattr(args, "all", False):
        me = sess.get("/api/v1/me").json()["user"]["username"]
        es = [e for e in es if not e.get("archived") and e.get("username", me) == me]
    c._table([[e["id"], e["name"], e["state"].replace("STATE_", ""), f"{100 * (e.get('progress') or 0):.0f}%", e.get("username", ""), e.get("archived", False)] for e in es],
             ["ID", "Name", "State", "Progress", "Owner", "Archived"], getattr(args, "json", False))
    return rc


def shell_open(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """A shell here is a process group on the host that runs it: ``shell open`` goes in through that host's sshd
    (``--show-ssh-command`` prints the command instead of running it; extra words are options for ssh)."""
    from determined_b200.cli import more

    if getattr(args, "show_ssh_command", False):
        return more.shell_show_ssh(args)
    if getattr(args, "ssh_opts", None):
        import contextlib
        import io

        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            more.shell_show_ssh(args)
        return subprocess.call(shlex.split(buf.getvalue().strip()))
    return inner(args)


def task_kill(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not getattr(args, "root", False):
        return inner(args)
    r = _cli()._session(args).post(f"/api/v1/tasks/{args.task_id}/kill", json={"killFromRoot": True}).json()
    print(f"Killed task tree: {', '.join(r.get('killed') or []) or 'nothing was running'}")
    return 0


def template_list(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not getattr(args, "details", False):
        return inner(args)
    c = _cli()
    ts = c._session(args).get("/api/v1/templates").json()["templates"]
    c._table([[t["name"], yaml.safe_dump(t.get("config") or {}, default_flow_style=False).rstrip()] for t in ts], ["Name", "Config"], getattr(args, "json", False))
    return 0


def job_update(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if getattr(args, "resource_pool", None):
        _cli()._session(args).post("/api/v1/job-queues", json={"updates": [{"jobId": args.job_id, "resourcePool": args.resource_pool}]})
        print(f"Moved job {args.job_id} to resource pool {args.resource_pool}")
        if all(getattr(args, k, None) is None for k in ("priority", "weight", "ahead_of", "behind_of")):
            return 0
    return inner(args)


def template_create(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    if not getattr(args, "workspace_name", None):
        return inner(args)
    sess = _cli()._session(args)
    cfg = yaml.safe_load(pathlib.Path(args.file).read_text())
    sess.put(f"/api/v1/templates/{args.name}", json={"config": cfg, "workspaceId": _ws_ids(sess, [args.workspace_name])[0]})
    print(f"Set template {args.name} in workspace {args.workspace_name}")
    return 0


def model_create(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    rc = inner(args)
    if rc == 0 and getattr(args, "workspace_name", None):
        _cli()._session(args).patch(f"/api/v1/models/{args.name}", json={"model": {"workspaceName": args.workspace_name}})
    return rc


def dev_curl(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """Extra words go to curl itself (``det dev curl /api/v1/me -- -v``): curl is run with the session's token."""
    extra = [a for a in (getattr(args, "curl_args", None) or []) if a != "--"]
    show = getattr(args, "x", False)
    if not extra and not show:
        return inner(args)
    sess = _cli()._session(args)
    cmd = ["curl", sess.master.rstrip("/") + "/" + args.path.lstrip("/"), "-H", f"Authorization: Bearer {sess.token}", "-s", "--globoff"]
    if (getattr(args, "request", None) or "GET") != "GET":
        cmd += ["-X", args.request]
    if getattr(args, "data", None):
        cmd += ["-H", "Content-Type: application/json", "-d", args.data]
    cmd += extra
    if show:
        print(shlex.join(cmd))
    if not extra:
        return inner(args)
    return subprocess.call(cmd)


def bindings_call(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """Calls that change something ask first unless ``--auto-confirm`` / ``-y``."""
    name = str(getattr(args, "name", ""))
    if not name.startswith("get_") and not getattr(args, "auto_confirm", False) and sys.stdin is not None and sys.stdin.isatty():
        if input(f"{name} may change state on the master. Call it? [y/N]: ").strip().lower() not in ("y", "yes"):
            print("Aborted.")
            return 1
    return inner(args)


def bindings_list(inner: Callable[[argparse.Namespace], int], args: argparse.Namespace) -> int:
    """Without ``--show-unusable`` only the calls the CLI can express are listed: those whose parameters are all
    scalars or flat bodies (streams and nested-message bodies need code)."""
    if getattr(args, "show_unusable", False):
        return inner(args)
    import inspect

    from determined_b200.common.api import bindings

    for n, f in sorted(vars(bindings).items()):
        if not callable(f) or not n.split("_", 1)[0] in ("get", "post", "put", "patch", "delete") or "_" not in n:
            continue
        sig = inspect.signature(f)
        params = [p for p in list(sig.parameters.values())[1:]]
        ann = sig.return_annotation
        if "Iterable" in str(ann) or "Iterator" in str(ann):
            continue
        print(f"{n}({', '.join(p.name for p in params)})")
    return 0


# --------------------------------------------------------------------------------------------------------------------- wiring
def extend(parser: argparse.ArgumentParser) -> None:
    def on(path: str) -> Optional[argparse.ArgumentParser]:
        return find(parser, path)

    for path in ("project list", "project list-experiments", "workspace list", "workspace list-projects"):
        sp = on(path)
        if sp is not None:
            view_flags(sp)
    sp = on("model list")
    if sp is not None:
        view_flags(sp, paging=False)
        _add(sp, "--workspace-names", nargs="+", help="only models of these workspaces")
        wrap(sp, model_list)
    sp = on("model describe")
    if sp is not None:
        _add(sp, "--version", type=int, help="describe this version only")
        wrap(sp, model_describe)
    sp = on("rbac list-roles")
    if sp is not None:
        view_flags(sp, sort=False, pages=True)
        _add(sp, "--exclude-global-roles", action="store_true", help="leave out roles that can only be assigned cluster-wide")
        wrap(sp, rbac_list_roles)
    for path in ("rbac assign-role", "rbac unassign-role"):
        sp = on(path)
        if sp is not None:
            _add(sp, "--workspace-name", help="scope of the assignment, by name")
            sp.add_argument("--username-to-assign", dest="username", help="same as -u")
            sp.add_argument("--group-name-to-assign", dest="group_name", help="same as -g")
            wrap(sp, rbac_assign)
    sp = on("user-group list")
    if sp is not None:
        view_flags(sp, sort=False, pages=True)
        _add(sp, "--groups-user-belongs-to", metavar="USERNAME")
        wrap(sp, group_list)
    sp = on("trial describe")
    if sp is not None:
        view_flags(sp, sort=False, pages=True)
        _add(sp, "--metrics", action="store_true", help="list every reported metric (the default)")
        _add(sp, "--metrics-summary", action="store_true", help="per-metric count / last / min / max / mean instead of every report")
        wrap(sp, trial_describe)
    sp = on("trial download")
    if sp is not None:
        _add(sp, "--sort-by", metavar="METRIC", help="pick the checkpoint with the best value of this validation metric")
        _add(sp, "--smaller-is-better", type=lambda s: str(s).lower() in ("1", "true", "yes"), default=None)
        _add(sp, "--quiet", action="store_true", help="print only the path")
        wrap(sp, trial_download)
    sp = on("experiment download")
    if sp is not None:
        _add(sp, "--quiet", action="store_true", help="print only the paths")
        wrap(sp, _quietly)
    sp = on("experiment describe")
    if sp is not None:
        _add(sp, "--metrics", action="store_true", help="include every reported metric of every trial")
        _add(sp, "--outdir", help="write experiments / trials / metrics files here instead of printing")
        wrap(sp, experiment_describe)
    sp = on("experiment list-checkpoints")
    if sp is not None:
        _add(sp, "--best", type=int, metavar="N", help="only the N best by the searcher metric")
        wrap(sp, list_checkpoints)
    for path, prompt in (("checkpoint rm", "Delete the given checkpoints from checkpoint storage? This cannot be undone."),
                         ("project delete", "Delete project {project} and everything in it?"), ("user-group delete", "Delete user group {name}?"),
                         ("workspace delete", "Delete workspace {name} and every project in it?")):
        sp = on(path)
        if sp is not None:
            _add(sp, "--yes", "-y", action="store_true", help="do not ask for confirmation")
            wrap(sp, confirm(prompt))
    sp = on("workspace create")
    if sp is not None:
        _add(sp, "--agent-uid", type=int)
        _add(sp, "--agent-user")
        _add(sp, "--agent-gid", type=int)
        _add(sp, "--agent-group")
        _add(sp, "--checkpoint-storage-config", help="JSON checkpoint storage of the workspace")
        _add(sp, "--checkpoint-storage-config-file", help="the same, from a YAML / JSON file")
        _add(sp, "--default-compute-pool")
        _add(sp, "--default-aux-pool")
        wrap(sp, workspace_create)
    sp = on("user change-password")
    if sp is not None:
        _add(sp, "target_user", nargs="?", help="another user (admins only); default: yourself")
        wrap(sp, user_change_password)
    sp = on("user create")
    if sp is not None:
        _add(sp, "--remote", action="store_true", help="a user who signs in through an identity provider only")
        wrap(sp, user_create)
    sp = on("user list")
    if sp is not None:
        _add(sp, "-a", "--all", action="store_true", help="include deactivated users")
        wrap(sp, user_list)
    sp = on("user logout")
    if sp is not None:
        _add(sp, "-a", "--all", action="store_true", help="every user signed in on this machine")
        wrap(sp, user_logout)
    for path in ("project describe", "project list-experiments"):
        sp = on(path)
        if sp is not None:
            _add(sp, "-a", "--all", action="store_true", help="experiments of every user, archived ones too")
            wrap(sp, project_view)
    sp = on("shell open")
    if sp is not None:
        _add(sp, "ssh_opts", nargs="*", help="options passed to ssh")
        _add(sp, "--show-ssh-command", action="store_true")
        wrap(sp, shell_open)
    sp = on("task kill")
    if sp is not None:
        _add(sp, "--root", action="store_true", help="kill the whole tree of tasks this one belongs to, from its root")
        wrap(sp, task_kill)
    sp = on("template list")
    if sp is not None:
        _add(sp, "--details", "-d", action="store_true", help="show the configurations too")
        wrap(sp, template_list)
    sp = on("job update")
    if sp is not None:
        _add(sp, "--resource-pool", help="move the job to another resource pool")
        wrap(sp, job_update)
    sp = on("template create")
    if sp is not None:
        _add(sp, "--workspace-name", "-w", help="the workspace the template belongs to")
        wrap(sp, template_create)
    sp = on("model create")
    if sp is not None:
        _add(sp, "--workspace-name", "-w", help="the workspace the model is created in")
        wrap(sp, model_create)
    sp = on("dev curl")
    if sp is not None:
        _add(sp, "-x", action="store_true", help="display the curl command that does the same")
        _add(sp, "curl_args", nargs=argparse.REMAINDER, help="passed to curl")
        wrap(sp, dev_curl)
    sp = on("dev bindings call")
    if sp is not None:
        _add(sp, "--auto-confirm", "-y", action="store_true")
        wrap(sp, bindings_call)
    sp = on("dev bindings list")
    if sp is not None:
        _add(sp, "--show-unusable", action="store_true")
        wrap(sp, bindings_list)
