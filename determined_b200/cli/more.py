"""The rest of the ``det`` command tree: edit / archive / bind / account commands, help at every level.

Parity (reference files under ``harness/determined/cli/``): ``sso.py`` (``auth``), ``command.py`` / ``notebook.py`` /
``shell.py`` / ``tensorboard.py`` (``config``, ``open``, ``set priority``, ``--context`` / ``--include``), ``experiment.py``
(``delete-tb-files``, ``set log-retention``), ``job.py`` (``update-batch``), ``master.py`` (``config show|set``, ``logs``),
``model.py`` (``move``), ``oauth.py``, ``project.py`` (``archive``, ``unarchive``, ``edit``), ``rbac.py`` (``describe-role``,
``list-groups-roles``, ``my-permissions``), ``resource_pool.py`` (``bindings``), ``resources.py`` (``raw``, ``aggregated``),
``task.py`` (``config``, ``create``, ``fork``, ``pause``, ``unpause``, ``cleanup-logs``), ``template.py`` (``set-value``),
``trial.py`` (``set log-retention``, ``support-bundle``), ``user.py`` (``edit``, ``link-with-agent-user``),
``user_groups.py`` (``change-name``), ``workspace.py`` (``archive``, ``unarchive``, ``describe``, ``edit``, ``list-pools``).

``extend(parser)`` is called by ``cli.make_parser`` on the finished tree; ``docs/CLI_GAPS.md`` is generated from the diff of
this tree against the reference's.
"""
import argparse
import csv
import io
import json
import os
import pathlib
import sys
import tarfile
import time
import webbrowser
from typing import Any, Callable, Dict, List, Optional

import yaml

from determined_b200.common import api


def _cli() -> Any:
    from determined_b200.cli import cli

    return cli


def _subs(parser: argparse.ArgumentParser) -> Optional[argparse._SubParsersAction]:
    return next((a for a in parser._actions if isinstance(a, argparse._SubParsersAction)), None)


def _group(root: argparse.ArgumentParser, name: str, help: str = "", aliases: Optional[List[str]] = None) -> argparse.ArgumentParser:
    subs = _subs(root)
    assert subs is not None
    if name in subs.choices:
        return subs.choices[name]
    p = subs.add_parser(name, help=help, aliases=aliases or [])
    p.add_subparsers(dest="sub_" + name.replace("-", "_"))
    return p


def _cmd(group: argparse.ArgumentParser, name: str, fn: Callable[[argparse.Namespace], int], help: str = "",
         aliases: Optional[List[str]] = None) -> argparse.ArgumentParser:
    subs = _subs(group)
    if subs is None:
        subs = group.add_subparsers(dest="sub_" + group.prog.split()[-1].replace("-", "_"))
    if name in subs.choices:             # already there: the handler given here takes over, arguments stay
        subs.choices[name].set_defaults(fn=fn)
        return argparse.ArgumentParser(add_help=False)      # a throw-away target for the caller's add_argument calls
    p = subs.add_parser(name, help=help, aliases=[a for a in (aliases or []) if a not in subs.choices])
    p.set_defaults(fn=fn)
    return p


def _add_help_everywhere(parser: argparse.ArgumentParser) -> None:
    subs = _subs(parser)
    if subs is None:
        return
    seen = set()
    for child in list(subs.choices.values()):
        if id(child) not in seen:
            seen.add(id(child))
            _add_help_everywhere(child)
    if "help" not in subs.choices:
        h = subs.add_parser("help", help="show help for this command")
        h.set_defaults(fn=lambda a, p=parser: (p.print_help(), 0)[1])


def _workspace_id(sess: api.Session, name: str) -> int:
    ws = sess.get("/api/v1/workspaces", params={"name": name}).json()["workspaces"]
    if not ws:
        raise SystemExit(f"workspace {name!r} not found")
    return int(ws[0]["id"])


def _project_id(sess: api.Session, workspace: str, project: str) -> int:
    wid = _workspace_id(sess, workspace)
    ps = [p for p in sess.get(f"/api/v1/workspaces/{wid}/projects").json()["projects"] if p["name"] == project]
    if not ps:
        raise SystemExit(f"project {project!r} not found in workspace {workspace!r}")
    return int(ps[0]["id"])


def _out(obj: Any, as_json: bool = False, as_yaml: bool = False) -> None:
    if as_json:
        print(json.dumps(obj, indent=2))
    else:
        print(yaml.safe_dump(obj, default_flow_style=False, sort_keys=True), end="")


# --------------------------------------------------------------------------------------------------------------------- auth
def auth_login(args: argparse.Namespace) -> int:
    if getattr(args, "provider", None):
        return sso_login(args)
    c = _cli()
    user = args.username or args.user or input("Username: ")
    sess = api.authentication.login_with_cache(c._master_of(args), user, None)
    store = api.authentication.TokenStore(c._master_of(args))
    store.set_token(user, sess.token or "")
    store.set_active(user)
    print(f"Logged in as {user}")
    return 0


def sso_login(args: argparse.Namespace, open_url: Optional[Callable[[str], Any]] = None) -> int:
    """Sign in through the master's identity provider: a listener on this machine's loopback is the relay target; the
    browser ends up there with the session token (reference ``cli/sso.py``).  ``--headless`` prints the URL and reads the
    address the browser ended on (or the bare token) from stdin."""
    import http.server
    import threading
    import urllib.parse

    c = _cli()
    master = c._master_of(args).rstrip("/")
    provs = api.UnauthSession(master).get("/api/v1/master").json().get("ssoProviders") or []
    match = [p for p in provs if p.get("name", "").lower() == args.provider.lower()]
    if not match:
        raise SystemExit(f"this master has no single sign-on provider {args.provider!r} (see `det auth list-providers`)")
    got: Dict[str, str] = {}

    class H(http.server.BaseHTTPRequestHandler):
        def do_GET(self) -> None:  # noqa: N802
            q = urllib.parse.parse_qs(urllib.parse.urlparse(self.path).query)
            if q.get("jwt"):
                got["token"], got["user"] = q["jwt"][0], (q.get("user") or [""])[0]
            body = b"Signed in; you can close this window." if got else b"No token in this request."
            self.send_response(200)
            self.send_header("Content-Length", str(len(body)))
            self.end_headers()
            self.wfile.write(body)

        def log_message(self, *a: Any) -> None:
            pass

    if getattr(args, "headless", False):
        url = f"{master}{match[0]['ssoUrl']}?" + urllib.parse.urlencode({"relayState": "http://127.0.0.1:1/"})
        print(f"Open this address in a browser, sign in, then paste the address it ends on (or its jwt value):\n  {url}")
        pasted = input("> ").strip()
        q = urllib.parse.parse_qs(urllib.parse.urlparse(pasted).query)
        got["token"] = (q.get("jwt") or [pasted])[0]
        got["user"] = (q.get("user") or [""])[0]
    else:
        srv = http.server.HTTPServer(("127.0.0.1", 0), H)
        url = f"{master}{match[0]['ssoUrl']}?" + urllib.parse.urlencode({"relayState": f"http://127.0.0.1:{srv.server_address[1]}/"})
        th = threading.Thread(target=srv.serve_forever, daemon=True)
        th.start()
        try:
            print(f"Waiting for the sign-in to finish in the browser:\n  {url}")
            (open_url or (lambda u: webbrowser.open(u) if os.environ.get("DISPLAY") else None))(url)
            t0 = time.time()
            while not got and time.time() - t0 < float(getattr(args, "timeout", None) or 300):
                time.sleep(0.1)
        finally:
            srv.shutdown()
            srv.server_close()
    if not got.get("token"):
        raise SystemExit("the sign-in did not finish")
    sess = api.Session(master, got.get("user") or "", got["token"])
    user = sess.get("/api/v1/me").json()["user"]["username"]
    store = api.authentication.TokenStore(master)
    store.set_token(user, got["token"])
    store.set_active(user)
    print(f"Logged in as {user}")
    return 0


def auth_logout(args: argparse.Namespace) -> int:
    api.authentication.logout(_cli()._master_of(args), args.user)
    return 0


def auth_list_providers(args: argparse.Namespace) -> int:
    info = api.UnauthSession(_cli()._master_of(args)).get("/api/v1/master").json()
    provs = info.get("ssoProviders") or []
    if not provs:
        print("No SSO providers configured.")
    else:
        _cli()._table([[p.get("name"), p.get("ssoUrl")] for p in provs], ["Provider", "URL"])
    return 0


# ------------------------------------------------------------------------------------------------------ commands and friends
def ntsc_config(kind: str) -> Callable[[argparse.Namespace], int]:
    def fn(args: argparse.Namespace) -> int:
        t = _cli()._session(args).get(f"/api/v1/{kind}s/{args.id}").json()[kind]
        _out(t["config"])
        return 0
    return fn


def ntsc_set_priority(kind: str) -> Callable[[argparse.Namespace], int]:
    def fn(args: argparse.Namespace) -> int:
        _cli()._session(args).post(f"/api/v1/{kind}s/{args.id}/set_priority", json={"priority": int(args.priority)})
        print(f"Set priority of {kind} {args.id} to {args.priority}")
        return 0
    return fn


def ntsc_open(kind: str) -> Callable[[argparse.Namespace], int]:
    def fn(args: argparse.Namespace) -> int:
        sess = _cli()._session(args)
        t = sess.get(f"/api/v1/{kind}s/{args.id}").json()[kind]
        if not t.get("serviceAddress"):
            raise SystemExit(f"{kind} {args.id} has no service to open (state {t['state'].replace('STATE_', '')})")
        url = f"{sess.master}{t['serviceAddress']}?token={sess.token}"
        print(url)
        if not getattr(args, "no_browser", False) and os.environ.get("DISPLAY"):
            webbrowser.open(url)
        return 0
    return fn


def shell_show_ssh(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    t = sess.get(f"/api/v1/shells/{args.id}").json()["shell"]
    # shells here are process groups on the master's (or an agent's) host: the way in is the host's own sshd
    host = (t.get("config") or {}).get("host") or sess.master.split("://", 1)[1].rsplit(":", 1)[0]
    print(" ".join(["ssh", *(args.ssh_opts or []), f"{t['username']}@{host}"]))
    return 0


# ------------------------------------------------------------------------------------------------------------------- the rest
def dev_auth_token(args: argparse.Namespace) -> int:
    print(_cli()._session(args).token)
    return 0


def experiment_delete_tb_files(args: argparse.Namespace) -> int:
    r = _cli()._session(args).delete(f"/api/v1/experiments/{args.experiment_id}/tensorboard-files").json()
    print(f"Deleted {r.get('removedFiles', 0)} tensorboard file(s) of experiment {args.experiment_id}")
    return 0


def _days(args: argparse.Namespace) -> int:
    if args.forever:
        return -1
    if args.days is None:
        raise SystemExit("give --days N or --forever")
    return int(args.days)


def experiment_log_retention(args: argparse.Namespace) -> int:
    _cli()._session(args).put(f"/api/v1/experiments/{args.experiment_id}/retain_logs", json={"numDays": _days(args)})
    return 0


def trial_log_retention(args: argparse.Namespace) -> int:
    _cli()._session(args).put(f"/api/v1/trials/{args.trial_id}/retain_logs", json={"numDays": _days(args)})
    return 0


def trial_support_bundle(args: argparse.Namespace) -> int:
    """Everything somebody helping with a trial wants in one tarball: trial and experiment records, the configuration,
    metrics, checkpoints and the logs of the trial's task."""
    sess = _cli()._session(args)
    tid = args.trial_id
    trial = sess.get(f"/api/v1/trials/{tid}").json()["trial"]
    exp = sess.get(f"/api/v1/experiments/{trial['experimentId']}").json()["experiment"]
    parts: Dict[str, bytes] = {
        "trial.json": json.dumps(trial, indent=2).encode(),
        "experiment.json": json.dumps({k: v for k, v in exp.items() if k != "config"}, indent=2).encode(),
        "experiment_config.yaml": yaml.safe_dump(exp["config"]).encode(),
        "metrics.json": json.dumps(sess.get(f"/api/v1/trials/{tid}/metrics").json()["metrics"], indent=1).encode(),
        "checkpoints.json": json.dumps(sess.get(f"/api/v1/trials/{tid}/checkpoints").json()["checkpoints"], indent=1).encode(),
        "master.json": json.dumps(sess.get("/api/v1/master").json(), indent=2).encode(),
    }
    lines, last = [], 0
    while True:
        logs = sess.get(f"/api/v1/trials/{tid}/logs", params={"offset": last, "limit": 5000}).json()["logs"]
        if not logs:
            break
        for l in logs:
            lines.append(f"{l['timestamp']} [{(l.get('level') or '').replace('LOG_LEVEL_', '')}] {l['log']}")
            last = max(last, int(l["id"]))
    parts["trial_logs.txt"] = "".join(x if x.endswith("\n") else x + "\n" for x in lines).encode()
    name = f"det-bundle-trial-{tid}-{time.strftime('%Y%m%dT%H%M%S')}"
    out = pathlib.Path(args.output_dir or ".") / f"{name}.tar.gz"
    out.parent.mkdir(parents=True, exist_ok=True)
    with tarfile.open(out, "w:gz") as tf:
        for fname, data in parts.items():
            ti = tarfile.TarInfo(f"{name}/{fname}")
            ti.size, ti.mtime = len(data), int(time.time())
            tf.addfile(ti, io.BytesIO(data))
    print(str(out))
    return 0


def job_update_batch(args: argparse.Namespace) -> int:
    """``det job update-batch JOB.priority=10 JOB.weight=2 JOB.ahead-of=OTHER JOB.behind-of=OTHER``: one request."""
    updates = []
    for op in args.operation:
        try:
            left, value = op.split("=", 1)
            jid, key = left.rsplit(".", 1)
        except ValueError:
            raise SystemExit(f"cannot parse {op!r}: expected JOB_ID.<priority|weight|ahead-of|behind-of|resource-pool>=VALUE")
        field = {"priority": "priority", "weight": "weight", "ahead-of": "aheadOf", "behind-of": "behindOf", "resource-pool": "resourcePool"}.get(key)
        if field is None:
            raise SystemExit(f"unknown job attribute {key!r} in {op!r}")
        updates.append({"jobId": jid, field: int(value) if field == "priority" else float(value) if field == "weight" else value})
    _cli()._session(args).post("/api/v1/job-queues", json={"updates": updates})
    print(f"Applied {len(updates)} update(s)")
    return 0


def master_config_show(args: argparse.Namespace) -> int:
    _out(_cli()._session(args).get("/api/v1/master/config").json()["config"], getattr(args, "json", False))
    return 0


def master_config_set(args: argparse.Namespace) -> int:
    log: Dict[str, Any] = {}
    if args.log_level is not None:
        log["level"] = args.log_level
    if args.log_color is not None:
        log["color"] = args.log_color
    if not log:
        raise SystemExit("nothing to set: give --log.level and / or --log.color")
    _cli()._session(args).patch("/api/v1/master/config", json={"config": {"log": log}})
    print("Updated the master's log configuration: " + ", ".join(f"{k}={v}" for k, v in log.items()))
    return 0


def master_logs(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    off = -int(args.tail) if args.tail else 0
    while True:
        logs = sess.get("/api/v1/master/logs", params={"offset": off, "follow": "true" if args.follow else "false"}, timeout=90).json()["logs"]
        for l in logs:
            print(f"{l['timestamp']} {l['level'].replace('LOG_LEVEL_', ''):<8} {l['message']}")
            off = max(off, int(l["id"])) if off >= 0 else int(l["id"])
        if off < 0:
            off = 0
        if not args.follow:
            return 0


def model_move(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    wid = _workspace_id(sess, args.workspace_name)
    sess.patch(f"/api/v1/models/{args.name}", json={"model": {"workspaceId": wid, "workspaceName": args.workspace_name}, "workspaceId": wid})
    print(f"Moved model {args.name} to workspace {args.workspace_name}")
    return 0


def oauth_add(args: argparse.Namespace) -> int:
    c = _cli()._session(args).post("/oauth2/clients", json={"name": args.name, "domain": args.domain}).json()
    print(f"Client ID:     {c['id']}\nClient secret: {c['secret']}\n(the secret is not shown again)")
    return 0


def oauth_list(args: argparse.Namespace) -> int:
    cs = _cli()._session(args).get("/oauth2/clients").json()
    _cli()._table([[c["name"], c["id"], c["domain"]] for c in cs], ["Name", "Client ID", "Domain"])
    return 0


def oauth_remove(args: argparse.Namespace) -> int:
    _cli()._session(args).delete(f"/oauth2/clients/{args.client_id}")
    print(f"Removed client {args.client_id}")
    return 0


def project_archive(flag: bool) -> Callable[[argparse.Namespace], int]:
    def fn(args: argparse.Namespace) -> int:
        sess = _cli()._session(args)
        pid = _project_id(sess, args.workspace_name, args.project_name)
        sess.post(f"/api/v1/projects/{pid}/{'archive' if flag else 'unarchive'}")
        print(f"{'Archived' if flag else 'Unarchived'} project {args.project_name}")
        return 0
    return fn


def project_edit(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    pid = _project_id(sess, args.workspace_name, args.project_name)
    body = {k: v for k, v in (("name", args.new_name), ("description", args.description)) if v is not None}
    if not body:
        raise SystemExit("nothing to change: give --new_name and / or --description")
    p = sess.patch(f"/api/v1/projects/{pid}", json=body).json()["project"]
    _out(p, args.json) if args.json else print(f"Updated project {p['name']} (id {p['id']})")
    return 0


def rbac_describe_role(args: argparse.Namespace) -> int:
    r = _cli()._session(args).get(f"/api/v1/roles/{args.role_name}").json()
    if args.json:
        print(json.dumps(r, indent=2))
        return 0
    print(f"Role: {r['role']['name']} (id {r['role']['roleId']})")
    _cli()._table([[p["id"].replace("PERMISSION_TYPE_", ""), "cluster" + (" / workspace" if p["scopeTypeMask"]["workspace"] else "")] for p in r["role"]["permissions"]],
                  ["Permission", "Scopes"])
    print(f"Assigned to {len(r['userRoleAssignments'])} user(s) and {len(r['groupRoleAssignments'])} group(s)")
    return 0


def rbac_list_groups_roles(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    gs = [g for g in sess.get("/api/v1/groups").json()["groups"] if g["group"]["name"] == args.group_name]
    if not gs:
        raise SystemExit(f"group {args.group_name!r} not found")
    roles = sess.get(f"/api/v1/roles/search/by-group/{gs[0]['group']['groupId']}").json()["roles"]
    _cli()._table([[r["role"]["name"], r["role"]["roleId"], r["scopeWorkspaceId"] if r["scopeWorkspaceId"] is not None else "(cluster)"] for r in roles],
                  ["Role", "ID", "Workspace"], args.json)
    return 0


def rbac_my_permissions(args: argparse.Namespace) -> int:
    r = _cli()._session(args).get("/api/v1/permissions/summary").json()
    if args.json:
        print(json.dumps(r, indent=2))
        return 0
    names = {x["roleId"]: x["name"] for x in r["roles"]}
    _cli()._table([[names.get(a["roleId"]), "yes" if a["scopeCluster"] else "", ", ".join(str(w) for w in a["scopeWorkspaceIds"])] for a in r["assignments"]],
                  ["Role", "Cluster-wide", "Workspaces"])
    return 0


def rp_bindings(action: str) -> Callable[[argparse.Namespace], int]:
    def fn(args: argparse.Namespace) -> int:
        sess = _cli()._session(args)
        path = f"/api/v1/resource-pools/{args.pool_name}/workspace-bindings"
        if action == "list":
            ids = sess.get(path).json()["workspaceIds"]
            names = {w["id"]: w["name"] for w in sess.get("/api/v1/workspaces").json()["workspaces"]}
            _cli()._table([[i, names.get(i, "?")] for i in ids], ["Workspace ID", "Name"])
            return 0
        body = {"resourcePoolName": args.pool_name, "workspaceNames": list(args.workspace_names)}
        {"add": sess.post, "remove": sess.delete, "replace": sess.put}[action](path, json=body)
        print(f"{action.capitalize()}d bindings of pool {args.pool_name}: {', '.join(args.workspace_names)}" if action != "add"
              else f"Bound pool {args.pool_name} to: {', '.join(args.workspace_names)}")
        return 0
    return fn


def resources_raw(args: argparse.Namespace) -> int:
    rows = _cli()._session(args).get("/api/v1/resources/allocation/raw",
                                      params={"timestampAfter": args.timestamp_after, "timestampBefore": args.timestamp_before}).json()["resourceEntries"]
    if args.json:
        print(json.dumps(rows, indent=2))
        return 0
    w = csv.writer(sys.stdout)
    w.writerow(["kind", "username", "labels", "slots", "start_time", "end_time", "seconds", "experiment_id", "resource_pool"])
    for r in rows:
        w.writerow([r["kind"], r["username"], ",".join(r["labels"]), r["slots"], r["startTime"], r["endTime"], f"{r['seconds']:.3f}", r["experimentId"], r["resourcePool"]])
    return 0


def resources_aggregated(args: argparse.Namespace) -> int:
    period = "RESOURCE_ALLOCATION_AGGREGATION_PERIOD_MONTHLY" if args.monthly else "RESOURCE_ALLOCATION_AGGREGATION_PERIOD_DAILY"
    rows = _cli()._session(args).get("/api/v1/resources/allocation/aggregated",
                                      params={"startDate": args.start_date, "endDate": args.end_date, "period": period}).json()["resourceEntries"]
    if args.json:
        print(json.dumps(rows, indent=2))
        return 0
    w = csv.writer(sys.stdout)
    w.writerow(["aggregation_type", "aggregation_key", "date", "seconds"])
    for r in rows:
        w.writerow(["total", "total", r["periodStart"], f"{r['seconds']:.3f}"])
        for typ, key in (("username", "byUsername"), ("experiment_label", "byExperimentLabel"), ("resource_pool", "byResourcePool")):
            for k, v in sorted(r[key].items()):
                w.writerow([typ, k, r["periodStart"], f"{v:.3f}"])
    return 0


def task_config(args: argparse.Namespace) -> int:
    _out(_cli()._session(args).get(f"/api/v1/tasks/{args.task_id}/config").json()["config"], args.json)
    return 0


def _follow_task(sess: api.Session, task_id: str) -> int:
    last = 0
    while True:
        logs = sess.get(f"/api/v1/tasks/{task_id}/logs", params={"offset": last}).json()["logs"]
        for l in logs:
            sys.stdout.write(l["log"])
            last = max(last, int(l["id"]))
        t = sess.get(f"/api/v1/commands/{task_id}").json()["command"]
        if t["state"] == "STATE_TERMINATED" and not logs:
            return int(t["exitStatus"] or 0)
        time.sleep(0.3)


def task_create(args: argparse.Namespace) -> int:
    from determined_b200.common import context as det_context

    c = _cli()
    sess = c._session(args)
    cfg = yaml.safe_load(pathlib.Path(args.config_file).read_text()) or {}
    c._parse_config_overrides(cfg, args.config)
    files: List[Dict[str, Any]] = []
    if args.context:
        files += det_context.read_v1_context(pathlib.Path(args.context))
    for inc in args.include or []:
        p = pathlib.Path(inc)
        if p.is_dir():
            for f in det_context.read_v1_context(p):
                files.append(dict(f, path=f"{p.name}/{f['path']}"))
        else:
            files += [f for f in det_context.read_v1_context(p.parent) if f["path"] == p.name]
    body = {"config": cfg, "contextDirectory": files, "projectId": args.project_id, "forkedFrom": args.fork, "parentId": args.parent,
            "inheritContext": bool(args.inherit_context), "noPause": bool(args.no_pause)}
    r = sess.post("/api/v1/generic-tasks", json=body).json()
    print(f"Created task {r['taskId']}")
    return _follow_task(sess, r["taskId"]) if args.follow else 0


def task_fork(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    r = sess.post("/api/v1/generic-tasks", json={"config": {}, "forkedFrom": args.parent_task_id, "parentId": args.parent_task_id,
                                                  "inheritContext": True, "projectId": args.project_id}).json()
    print(f"Created task {r['taskId']} (fork of {args.parent_task_id})")
    return _follow_task(sess, r["taskId"]) if args.follow else 0


def task_forward(args: argparse.Namespace) -> int:
    """``det task forward TASK_ID LOCAL[:REMOTE]``: connections to the local port reach the task's service port (or a port it
    declared in ``proxy_ports``) through the master; runs until interrupted."""
    from determined_b200.cli import proxy

    local, _, remote = str(args.ports).partition(":")
    sess = _cli()._session(args)
    cfg = proxy.ListenerConfig(args.task_id, int(local), remote_port=int(remote) if remote else None)
    with proxy.http_tunnel_listener(sess, [cfg]) as ports:
        print(f"forwarding 127.0.0.1:{ports[0]} to task {args.task_id}" + (f" port {remote}" if remote else ""), flush=True)
        try:
            while getattr(args, "_once", None) is None:
                time.sleep(1)
        except KeyboardInterrupt:
            pass
    return 0


def task_pause(flag: bool) -> Callable[[argparse.Namespace], int]:
    def fn(args: argparse.Namespace) -> int:
        _cli()._session(args).post(f"/api/v1/tasks/{args.task_id}/{'pause' if flag else 'unpause'}")
        print(f"{'Paused' if flag else 'Resumed'} task {args.task_id}")
        return 0
    return fn


def task_cleanup_logs(args: argparse.Namespace) -> int:
    r = _cli()._session(args).post("/api/v1/cleanup_logs").json()
    print(f"Removed {r['removedCount']} log line(s) past their retention period")
    return 0


def template_set_value_config(args: argparse.Namespace) -> int:
    cfg = yaml.safe_load(pathlib.Path(args.template_file).read_text()) or {}
    _cli()._session(args).patch(f"/api/v1/templates/{args.template_name}", json={"config": cfg})
    print(f"Set the configuration of template {args.template_name}")
    return 0


def group_change_name(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    gs = [g for g in sess.get("/api/v1/groups").json()["groups"] if g["group"]["name"] == args.old_group_name]
    if not gs:
        raise SystemExit(f"group {args.old_group_name!r} not found")
    sess.put(f"/api/v1/groups/{gs[0]['group']['groupId']}", json={"name": args.new_group_name})
    print(f"Renamed group {args.old_group_name} to {args.new_group_name}")
    return 0


def _user_id(sess: api.Session, name: str) -> int:
    us = [u for u in sess.get("/api/v1/users").json()["users"] if u["username"] == name]
    if not us:
        raise SystemExit(f"user {name!r} not found")
    return int(us[0]["id"])


def _tristate(s: Optional[str]) -> Optional[bool]:
    return None if s is None else s.lower() in ("1", "true", "yes")


def user_edit(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    patch = {k: v for k, v in (("displayName", args.display_name), ("username", args.username), ("active", _tristate(args.active)),
                               ("admin", _tristate(args.admin))) if v is not None}
    if args.remote is not None:
        patch["remote"] = _tristate(args.remote)
    if not patch:
        raise SystemExit("nothing to change")
    u = sess.patch(f"/api/v1/users/{_user_id(sess, args.target_user)}", json={"user": patch}).json()["user"]
    print(f"Updated user {u['username']}")
    return 0


def user_link_agent_user(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    aug = {"agentUid": args.agent_uid, "agentUser": args.agent_user, "agentGid": args.agent_gid, "agentGroup": args.agent_group}
    sess.put(f"/api/v1/users/{_user_id(sess, args.det_username)}/agent-user-group", json={"agentUserGroup": aug})
    print(f"Tasks of {args.det_username} run as {args.agent_user}:{args.agent_group} ({args.agent_uid}:{args.agent_gid})")
    return 0


def workspace_archive(flag: bool) -> Callable[[argparse.Namespace], int]:
    def fn(args: argparse.Namespace) -> int:
        sess = _cli()._session(args)
        sess.post(f"/api/v1/workspaces/{_workspace_id(sess, args.workspace_name)}/{'archive' if flag else 'unarchive'}")
        print(f"{'Archived' if flag else 'Unarchived'} workspace {args.workspace_name}")
        return 0
    return fn


def workspace_describe(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    w = sess.get(f"/api/v1/workspaces/{_workspace_id(sess, args.workspace_name)}").json()["workspace"]
    _out(w, args.json)
    return 0


def workspace_edit(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    wid = _workspace_id(sess, args.workspace_name)
    body: Dict[str, Any] = {}
    if args.name is not None:
        body["name"] = args.name
    if any(v is not None for v in (args.agent_uid, args.agent_user, args.agent_gid, args.agent_group)):
        if any(v is None for v in (args.agent_uid, args.agent_user, args.agent_gid, args.agent_group)):
            raise SystemExit("--agent-uid, --agent-user, --agent-gid and --agent-group must be given together")
        body["agentUserGroup"] = {"agentUid": args.agent_uid, "agentUser": args.agent_user, "agentGid": args.agent_gid, "agentGroup": args.agent_group}
    if args.checkpoint_storage_config is not None:
        body["checkpointStorageConfig"] = yaml.safe_load(args.checkpoint_storage_config)
    if args.checkpoint_storage_config_file is not None:
        body["checkpointStorageConfig"] = yaml.safe_load(pathlib.Path(args.checkpoint_storage_config_file).read_text())
    if args.default_compute_pool is not None:
        body["defaultComputePool"] = args.default_compute_pool
    if args.default_aux_pool is not None:
        body["defaultAuxPool"] = args.default_aux_pool
    if not body:
        raise SystemExit("nothing to change")
    w = sess.patch(f"/api/v1/workspaces/{wid}", json=body).json()["workspace"]
    _out(w, True) if args.json else print(f"Updated workspace {w['name']} (id {w['id']})")
    return 0


def workspace_list_pools(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    pools = sess.get(f"/api/v1/workspaces/{_workspace_id(sess, args.workspace_name)}/available-resource-pools").json()["resourcePools"]
    _cli()._table([[p["name"]] for p in pools], ["Resource Pool"])
    return 0


# -------------------------------------------------------------------------------------------------------------------- logs
def _opt(p: argparse.ArgumentParser, *names: str, **kw: Any) -> None:
    """``add_argument`` unless the parser has the option already."""
    if not any(n in p._option_string_actions for n in names):
        p.add_argument(*names, **kw)


def add_log_args(p: argparse.ArgumentParser) -> None:
    _opt(p, "-f", "--follow", action="store_true", help="keep printing as lines arrive")
    _opt(p, "--head", type=int, help="only the first N lines")
    _opt(p, "--tail", type=int, help="only the last N lines")
    _opt(p, "--json", action="store_true", help="one JSON object per line")
    _opt(p, "--level", action="append", choices=["TRACE", "DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"], help="only these levels")
    _opt(p, "--rank-id", dest="rank_id", action="append", type=int, help="only these ranks")
    _opt(p, "--source", action="append", help="only these sources (task, master, agent)")
    _opt(p, "--stdtype", action="append", choices=["stdout", "stderr"])
    _opt(p, "--allocation-id", dest="allocation_id", action="append")
    _opt(p, "--container-id", dest="container_id", action="append")
    _opt(p, "--agent-id", dest="agent_id", action="append", help="accepted for compatibility: lines are not tagged with an agent here")
    _opt(p, "--timestamp-after", dest="timestamp_after", help="ISO 8601 time or seconds since the epoch")
    _opt(p, "--timestamp-before", dest="timestamp_before")
    _opt(p, "--search", dest="search_text", help="only lines containing this text")


def stream_logs(sess: api.Session, path: str, args: argparse.Namespace, still_running: Callable[[], bool]) -> int:
    params: Dict[str, Any] = {}
    for key, attr in (("levels", "level"), ("rankIds", "rank_id"), ("sources", "source"), ("stdtypes", "stdtype"), ("allocationIds", "allocation_id"),
                      ("containerIds", "container_id")):
        v = getattr(args, attr, None)
        if v:
            params[key] = [f"LOG_LEVEL_{x}" for x in v] if key == "levels" else v
    for key, attr in (("timestampAfter", "timestamp_after"), ("timestampBefore", "timestamp_before"), ("searchText", "search_text")):
        if getattr(args, attr, None):
            params[key] = getattr(args, attr)
    as_json, head, tail = getattr(args, "json", False), getattr(args, "head", None), getattr(args, "tail", None)
    follow = bool(getattr(args, "follow", False))

    def emit(l: Dict[str, Any]) -> None:
        sys.stdout.write(json.dumps(l) + "\n" if as_json else l["log"])

    last, shown, kept = 0, 0, []
    while True:
        logs = sess.get(path, params=dict(params, offset=last, limit=5000)).json()["logs"]
        for l in logs:
            last = max(last, int(l["id"]))
            if tail is not None and not follow:
                kept.append(l)
                kept = kept[-tail:]
                continue
            emit(l)
            shown += 1
            if head is not None and shown >= head:
                return 0
        if not logs:
            if not follow or not still_running():
                break
            time.sleep(0.3)
    for l in kept:
        emit(l)
    return 0


def _trial_running(sess: api.Session, tid: int) -> Callable[[], bool]:
    return lambda: sess.get(f"/api/v1/trials/{tid}").json()["trial"]["state"] in ("STATE_ACTIVE", "STATE_RUNNING", "STATE_QUEUED", "STATE_PAUSED",
                                                                               "STATE_STOPPING_COMPLETED", "STATE_STOPPING_CANCELED")


def trial_logs(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    return stream_logs(sess, f"/api/v1/trials/{args.trial_id}/logs", args, _trial_running(sess, args.trial_id))


def experiment_logs(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    trials = sess.get(f"/api/v1/experiments/{args.experiment_id}/trials").json()["trials"]
    if not trials:
        raise SystemExit("experiment has no trials yet")
    tid = trials[0]["id"]
    return stream_logs(sess, f"/api/v1/trials/{tid}/logs", args, _trial_running(sess, tid))


def task_logs(args: argparse.Namespace) -> int:
    sess = _cli()._session(args)
    tid = getattr(args, "task_id", None) or args.id
    running = lambda: any(a["taskId"] == tid for a in sess.get("/api/v1/tasks").json()["allocationIdToSummary"].values())   # noqa: E731
    return stream_logs(sess, f"/api/v1/tasks/{tid}/logs", args, running)


# ---------------------------------------------------------------------------------------------------------------- registration
def extend(parser: argparse.ArgumentParser) -> None:
    subs = _subs(parser)
    assert subs is not None
    # auth
    g = _group(parser, "auth", "log in and out, single sign-on providers")
    login = _subs(g).choices.get("login") if _subs(g) is not None else None
    if login is None:
        login = _cmd(g, "login", auth_login, "log in")
        login.add_argument("username", nargs="?")
    else:
        login.set_defaults(fn=auth_login)
    login.add_argument("--provider", help="single sign-on provider to use")
    login.add_argument("--headless", action="store_true", help="do not try to open a browser")
    if "logout" not in _subs(g).choices:
        _cmd(g, "logout", auth_logout, "log out")
    _cmd(g, "list-providers", auth_list_providers, "single sign-on providers of the master")
    # commands, notebooks, shells, tensorboards
    for kind, group in (("command", "command"), ("notebook", "notebook"), ("shell", "shell"), ("tensorboard", "tensorboard")):
        if group not in subs.choices:
            continue
        g = subs.choices[group]
        a = _cmd(g, "config", ntsc_config(kind), f"show the configuration of a {kind}")
        a.add_argument("id", metavar=f"{kind}_id")
        s = _group(g, "set", f"change a {kind}")
        a = _cmd(s, "priority", ntsc_set_priority(kind), "set the scheduling priority")
        a.add_argument("id", metavar=f"{kind}_id")
        a.add_argument("priority", type=int)
        if kind != "command":
            a = _cmd(g, "open", ntsc_open(kind), f"print (and, with a display, open) the URL of a {kind}")
            a.add_argument("id", metavar=f"{kind}_id")
            a.add_argument("--no-browser", action="store_true")
        if kind == "shell":
            for name in ("show-ssh-command", "show_ssh_command"):
                a = _cmd(g, name, shell_show_ssh, "print the ssh command that reaches the shell's host")
                a.add_argument("id", metavar="shell_id")
                a.add_argument("ssh_opts", nargs="*")
    # dev
    if "dev" in subs.choices:
        _cmd(subs.choices["dev"], "auth-token", dev_auth_token, "print the session token in use")
    # experiment
    e = subs.choices["experiment"]
    a = _cmd(e, "delete-tb-files", experiment_delete_tb_files, "delete the tensorboard files of an experiment")
    a.add_argument("experiment_id", type=int)
    a = _cmd(e, "set-log-retention", experiment_log_retention, "how long the logs of the experiment's trials are kept (also: `experiment set log-retention`)")
    a.add_argument("experiment_id", type=int)
    a.add_argument("--days", type=int)
    a.add_argument("--forever", action="store_true")
    # trial
    t = subs.choices["trial"]
    s = _group(t, "set", "change a trial")
    a = _cmd(s, "log-retention", trial_log_retention, "how long the trial's logs are kept")
    a.add_argument("trial_id", type=int)
    a.add_argument("--days", type=int)
    a.add_argument("--forever", action="store_true")
    a = _cmd(t, "support-bundle", trial_support_bundle, "tarball of records, configuration, metrics and logs of a trial")
    a.add_argument("trial_id", type=int)
    a.add_argument("-o", "--output-dir", default=".")
    # job
    a = _cmd(subs.choices["job"], "update-batch", job_update_batch, "several queue changes in one request")
    a.add_argument("operation", nargs="+", help="JOB_ID.<priority|weight|ahead-of|behind-of|resource-pool>=VALUE")
    # master
    m = subs.choices["master"]
    a = _cmd(m, "logs", master_logs, "the master's own log")
    a.add_argument("-f", "--follow", action="store_true")
    a.add_argument("--tail", type=int)
    mc = _subs(m).choices.get("config")
    if mc is not None:
        a = _cmd(mc, "show", master_config_show, "show the configuration (secrets masked)")
        a.add_argument("--json", action="store_true")
        a.add_argument("--yaml", action="store_true")
        a = _cmd(mc, "set", master_config_set, "change what can change while the master runs: the log section")
        a.add_argument("--log.level", dest="log_level", choices=["trace", "debug", "info", "warn", "error", "fatal"])
        a.add_argument("--log.color", dest="log_color", type=lambda s: s.lower() in ("1", "true", "yes", "on"))
        _subs(mc).required = False
    # model
    a = _cmd(subs.choices["model"], "move", model_move, "move a model to another workspace")
    a.add_argument("name")
    a.add_argument("-w", "--workspace-name", required=True)
    # oauth
    g = _group(_group(parser, "oauth", "OAuth client registry"), "client", "manage OAuth clients")
    a = _cmd(g, "add", oauth_add, "register a client")
    a.add_argument("name")
    a.add_argument("domain", help="redirect domain, http(s)://...")
    _cmd(g, "list", oauth_list, "registered clients", aliases=["ls"])
    a = _cmd(g, "remove", oauth_remove, "remove a client")
    a.add_argument("client_id")
    # project
    p = subs.choices["project"]
    for name, flag in (("archive", True), ("unarchive", False)):
        a = _cmd(p, name, project_archive(flag), f"{name} a project")
        a.add_argument("workspace_name")
        a.add_argument("project_name")
    a = _cmd(p, "edit", project_edit, "rename a project or change its description")
    a.add_argument("workspace_name")
    a.add_argument("project_name")
    a.add_argument("--new_name", "--new-name")
    a.add_argument("--description")
    a.add_argument("--json", action="store_true")
    # rbac
    r = subs.choices["rbac"]
    a = _cmd(r, "describe-role", rbac_describe_role, "permissions and assignments of a role")
    a.add_argument("role_name")
    a.add_argument("--json", action="store_true")
    a = _cmd(r, "list-groups-roles", rbac_list_groups_roles, "roles assigned to a group")
    a.add_argument("group_name")
    a.add_argument("--json", action="store_true")
    a = _cmd(r, "my-permissions", rbac_my_permissions, "roles the current user holds, and where")
    a.add_argument("--json", action="store_true")
    # resource pools
    rp = subs.choices["resource-pool"]
    b = _group(rp, "bindings", "which workspaces may use a pool")
    for name, action in (("add", "add"), ("remove", "remove"), ("replace", "replace")):
        a = _cmd(b, name, rp_bindings(action), f"{name} workspace bindings")
        a.add_argument("pool_name")
        a.add_argument("workspace_names", nargs="+")
    a = _cmd(b, "list-workspaces", rp_bindings("list"), "workspaces bound to a pool")
    a.add_argument("pool_name")
    # resources
    g = _group(parser, "resources", "resource usage accounting", aliases=["res"])
    a = _cmd(g, "raw", resources_raw, "one line per allocation in a time range (CSV)")
    a.add_argument("timestamp_after")
    a.add_argument("timestamp_before")
    a.add_argument("--json", action="store_true")
    a = _cmd(g, "aggregated", resources_aggregated, "slot-seconds per day or month (CSV)", aliases=["agg"])
    a.add_argument("start_date")
    a.add_argument("end_date")
    a.add_argument("--monthly", action="store_true")
    a.add_argument("--json", action="store_true")
    # task
    tk = subs.choices["task"]
    a = _cmd(tk, "config", task_config, "configuration of a task")
    a.add_argument("task_id")
    a.add_argument("--json", action="store_true")
    a = _cmd(tk, "create", task_create, "run a generic task from a configuration file")
    a.add_argument("config_file")
    a.add_argument("-c", "--context", help="directory shipped to the task")
    a.add_argument("-i", "--include", action="append")
    a.add_argument("--project_id", "--project-id", type=int)
    a.add_argument("--config", action="append", help="override key=value (dotted)")
    a.add_argument("-f", "--follow", action="store_true")
    a.add_argument("--fork", help="start from the configuration of this task")
    a.add_argument("--parent", help="parent task")
    a.add_argument("--inherit_context", "--inherit-context", action="store_true")
    a.add_argument("--no_pause", "--no-pause", action="store_true")
    a = _cmd(tk, "forward", task_forward, "forward a local port to a port of a running task, through the master")
    a.add_argument("task_id")
    a.add_argument("ports", metavar="LOCAL[:REMOTE]")
    a = _cmd(tk, "fork", task_fork, "run a task again, with its configuration and context")
    a.add_argument("parent_task_id")
    a.add_argument("-f", "--follow", action="store_true")
    a.add_argument("--project_id", "--project-id", type=int)
    for name, flag in (("pause", True), ("unpause", False)):
        a = _cmd(tk, name, task_pause(flag), f"{name} a running task")
        a.add_argument("task_id")
    _cmd(tk, "cleanup-logs", task_cleanup_logs, "delete task logs past their retention period now")
    # template
    sv = _group(subs.choices["template"], "set-value", "change part of a template")
    a = _cmd(sv, "config", template_set_value_config, "replace the configuration of a template")
    a.add_argument("template_name")
    a.add_argument("template_file")
    # user groups and users
    a = _cmd(subs.choices["user-group"], "change-name", group_change_name, "rename a group")
    a.add_argument("old_group_name")
    a.add_argument("new_group_name")
    u = subs.choices["user"]
    a = _cmd(u, "edit", user_edit, "change a user's display name, login name, remote / active / admin status")
    a.add_argument("target_user")
    a.add_argument("--display-name")
    a.add_argument("--username")
    a.add_argument("--remote")
    a.add_argument("--active")
    a.add_argument("--admin")
    a = _cmd(u, "link-with-agent-user", user_link_agent_user, "the Unix identity a user's tasks run as")
    a.add_argument("det_username")
    a.add_argument("--agent-uid", type=int, required=True)
    a.add_argument("--agent-user", required=True)
    a.add_argument("--agent-gid", type=int, required=True)
    a.add_argument("--agent-group", required=True)
    # workspace
    w = subs.choices["workspace"]
    for name, flag in (("archive", True), ("unarchive", False)):
        a = _cmd(w, name, workspace_archive(flag), f"{name} a workspace")
        a.add_argument("workspace_name")
    a = _cmd(w, "describe", workspace_describe, "show a workspace")
    a.add_argument("workspace_name")
    a.add_argument("--json", action="store_true")
    a = _cmd(w, "edit", workspace_edit, "rename a workspace, set its agent user, checkpoint storage or default pools")
    a.add_argument("workspace_name")
    a.add_argument("--name")
    a.add_argument("--agent-uid", type=int)
    a.add_argument("--agent-user")
    a.add_argument("--agent-gid", type=int)
    a.add_argument("--agent-group")
    a.add_argument("--checkpoint-storage-config", help="YAML / JSON text")
    a.add_argument("--checkpoint-storage-config-file")
    a.add_argument("--default-compute-pool")
    a.add_argument("--default-aux-pool")
    a.add_argument("--json", action="store_true")
    a = _cmd(w, "list-pools", workspace_list_pools, "resource pools the workspace may use")
    a.add_argument("workspace_name")
    # logs everywhere: the same filters and output options
    for group, name, fn in (("experiment", "logs", experiment_logs), ("trial", "logs", trial_logs), ("task", "logs", task_logs),
                            ("command", "logs", task_logs), ("shell", "logs", task_logs), ("notebook", "logs", task_logs), ("tensorboard", "logs", task_logs)):
        lp = _subs(subs.choices[group]).choices.get(name) if group in subs.choices else None
        if lp is not None:
            add_log_args(lp)
            lp.set_defaults(fn=fn)
    _add_help_everywhere(parser)
