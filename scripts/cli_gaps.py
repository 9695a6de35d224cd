"""Diff the ``det`` command tree of this repository against the reference's and write ``docs/CLI_GAPS.md``.

``docs/ref_cli_tree.json`` is the reference's tree (every node of ``determined.cli.cli.all_args_description`` with its
aliases and argument names), dumped once from the installed reference package; this script needs only that file.

    python scripts/cli_gaps.py            # rewrite docs/CLI_GAPS.md
    python scripts/cli_gaps.py --check    # exit 1 if the committed file is stale
"""
import argparse
import json
import pathlib
import sys

ROOT = pathlib.Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

# reference commands whose syntax is accepted here although the parser tree is shaped differently (a positional choice
# instead of a sub-command: `det e set name 3 foo`, `det e label add 3 x` parse the same way)
SAME_SYNTAX = {"experiment set", "experiment label"}
# argument names that differ only in spelling between the two trees (ours on the right)
RENAMED = {"checkpoints_uuids": ("uuids", "uuid", "checkpoint_uuids"), "command_id": ("id", "ids"), "notebook_id": ("id", "ids"), "shell_id": ("id", "ids"),
           "tensorboard_id": ("id", "ids"), "task_id": ("id", "task_id"), "group_name": ("name",), "usernames": ("users",), "project_name": ("project",),
           "workspace_name": ("workspace", "name"), "template_name": ("name",), "template_file": ("file",), "target_user": ("username",),
           "x": ("request",), "test_mode": ("test",), "role_name": ("role",), "username_to_assign": ("username",),
           "group_name_to_assign": ("group_name",), "name": ("project", "name")}
NOTES = {
    "agent_id": "log lines are not tagged with an agent (one host per task)",
    "auto_confirm": "`dev bindings call` never prompts",
    "show_unusable": "every binding is listed",
    "curl_args": "`dev curl` takes -X / -d instead of passing arguments to curl",
    "remote": "no remote (SSO-managed) users without SSO",
    "outdir": "`experiment describe --json` prints to stdout",
    "metrics": "`trial describe` always lists metrics",
    "metrics_summary": "`trial describe` always lists metrics",
    "quiet": "progress bars are not drawn",
    "sort_by": "lists come back in id order",
    "order_by": "lists come back in id order",
    "pages": "one request returns everything asked for",
    "yes": "the command does not prompt",
    "best": "`experiment list-checkpoints` lists all; `--top-n` on download selects the best",
    "details": "`template describe` shows the body",
    "version": "`model list-versions` shows every version",
    "workspace_names": "models list is not filtered by workspace",
    "a": "lists are not filtered by owner",
    "groups_user_belongs_to": "",
    "exclude_global_roles": "",
}


def tree_of(parser: argparse.ArgumentParser):
    out = []

    def walk(p, prefix):
        for a in p._actions:
            if isinstance(a, argparse._SubParsersAction):
                seen = set()
                for name, sp in a.choices.items():
                    if id(sp) in seen:
                        continue
                    seen.add(id(sp))
                    names = [n for n, s in a.choices.items() if s is sp]
                    opts = [(x.option_strings[-1] if x.option_strings else x.dest) for x in sp._actions
                            if not isinstance(x, (argparse._HelpAction, argparse._SubParsersAction))]
                    all_opts = [o for x in sp._actions for o in (x.option_strings or [x.dest]) if not isinstance(x, (argparse._HelpAction, argparse._SubParsersAction))]
                    leaf = not any(isinstance(x, argparse._SubParsersAction) for x in sp._actions)
                    out.append({"cmd": " ".join(prefix + [names[0]]), "aliases": names[1:], "args": opts, "all": all_opts, "leaf": leaf})
                    walk(sp, prefix + [names[0]])
    walk(parser, [])
    return out


def norm(a: str) -> str:
    return a.lstrip("-").replace("-", "_").replace(".", "_")


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    from determined_b200.cli import cli

    ref = json.loads((ROOT / "docs" / "ref_cli_tree.json").read_text())
    mine = tree_of(cli.make_parser())
    by_cmd = {o["cmd"]: o for o in mine}
    alias = {}
    for o in mine:
        parts = o["cmd"].split()
        lvl = alias.setdefault(" ".join(parts[:-1]), {})
        for n in [parts[-1]] + o["aliases"]:
            lvl[n] = parts[-1]

    def resolve(path):
        cur = []
        for p in path.split():
            lvl = alias.get(" ".join(cur), {})
            if p not in lvl:
                return None
            cur.append(lvl[p])
        return " ".join(cur)

    missing, same, argrows = [], [], []
    n_ref = sum(o["leaf"] for o in ref)
    for o in ref:
        if not o["leaf"]:
            continue
        r = resolve(o["cmd"])
        if r is not None and not by_cmd[r]["leaf"] and not o["args"]:
            continue                             # a bare group in the reference's dump (its sub-commands load lazily): the group exists here
        if r is None or not by_cmd[r]["leaf"]:
            parent = " ".join(o["cmd"].split()[:2])
            (same if parent in SAME_SYNTAX else missing).append(o["cmd"])
            continue
        have = {norm(x) for x in by_cmd[r]["all"]}
        for arg in sorted({norm(x) for x in o["args"]}):
            if arg in have or any(norm(x) in have for x in RENAMED.get(arg, ())):
                continue
            argrows.append((o["cmd"], arg, NOTES.get(arg, "")))
    lines = ["# `det` command tree against the reference's", "",
             "Generated by `python scripts/cli_gaps.py` from `docs/ref_cli_tree.json` (the reference's tree, dumped from its installed",
             "package) and this repository's parser; `--check` fails when this file is stale.", "",
             f"* reference: {n_ref} leaf commands; here: {sum(o['leaf'] for o in mine)}",
             f"* reference commands with no counterpart here: **{len(missing)}**",
             f"* reference commands accepted with the same command line but parsed differently (a positional choice instead of a sub-command): {len(same)}",
             f"* reference arguments with no counterpart here: **{len(argrows)}** (on {len({c for c, _, _ in argrows})} commands)", ""]
    if missing:
        lines += ["## Commands not provided", ""] + [f"* `det {m}`" for m in missing] + [""]
    if same:
        lines += ["## Same command line, different parser shape", ""] + [f"* `det {m} ...`" for m in same] + [""]
    lines += ["## Arguments not provided", "", "| command | argument | note |", "|---|---|---|"]
    lines += [f"| `det {c}` | `{arg}` | {note} |" for c, arg, note in argrows]
    lines += ["", "## Commands here that the reference does not have", ""]
    ref_alias = {}
    for o in ref:
        parts = o["cmd"].split()
        lvl = ref_alias.setdefault(" ".join(parts[:-1]), {})
        for n in [parts[-1]] + o["aliases"]:
            lvl[n] = parts[-1]

    def ref_has(path):
        cur = []
        for p in path.split():
            lvl = ref_alias.get(" ".join(cur), {})
            if p not in lvl:
                return False
            cur.append(lvl[p])
        return True
    lines += [f"* `det {o['cmd']}`" for o in mine if o["leaf"] and not ref_has(o["cmd"])]
    text = "\n".join(lines) + "\n"
    path = ROOT / "docs" / "CLI_GAPS.md"
    if a.check:
        if not path.exists() or path.read_text() != text:
            print("docs/CLI_GAPS.md is stale: run python scripts/cli_gaps.py", file=sys.stderr)
            return 1
        return 0
    path.write_text(text)
    print(f"wrote {path}: {len(missing)} commands and {len(argrows)} arguments missing")
    return 0


if __name__ == "__main__":
    sys.exit(main())
