"""Diff the public Python surface of this package against the installed reference (``baseline/_ref``) and write
``docs/API_SURFACE.md``: module paths that import, public names per module, public methods per class, keyword arguments per
method.  Each side is inspected in its own interpreter (the two packages cannot share one).

    python scripts/api_surface_diff.py            # rewrite docs/API_SURFACE.md
"""
import importlib
import inspect
import json
import pathlib
import subprocess
import sys
import types

ROOT = pathlib.Path(__file__).resolve().parents[1]
MODULES = ["", "core", "pytorch", "pytorch.samplers", "pytorch.experimental", "pytorch.dsat", "experimental", "experimental.client", "common.api", "common.util",
           "util", "searcher", "common.storage", "tensorboard", "transformers", "errors", "constants", "common.api.authentication", "common.api.certs", "launch",
           "experimental.core_v2", "common.api.errors", "common.check", "common.context", "ipc", "gpu", "workload", "layers"]
CLASSES = [("pytorch", "PyTorchTrialContext"), ("pytorch", "PyTorchTrial"), ("pytorch", "Trainer"), ("pytorch", "PyTorchCallback"), ("pytorch", "DataLoader"),
           ("pytorch", "PyTorchExperimentalContext"), ("core", "Context"), ("core", "DistributedContext"), ("core", "CheckpointContext"), ("core", "TrainContext"),
           ("core", "SearcherContext"), ("core", "SearcherOperation"), ("core", "PreemptContext"), ("core", "ProfilerContext"), ("experimental.client", "Determined"),
           ("experimental.client", "Experiment"), ("experimental.client", "Trial"), ("experimental.client", "Checkpoint"), ("experimental.client", "Model"),
           ("experimental.client", "ModelVersion"), ("experimental.client", "Project"), ("experimental.client", "Workspace"), ("experimental.client", "User"),
           ("experimental.client", "ResourcePool"), ("searcher", "SearchMethod"), ("searcher", "SearchRunner"), ("searcher", "SearcherState"),
           ("common.storage", "StorageManager"), ("common.storage", "SharedFSStorageManager"), ("", "ClusterInfo"), ("", "TrialInfo"), ("common.api", "Session"),
           ("transformers", "DetCallback")]
FUNCTIONS = [("pytorch", "init"), ("core", "init"), ("pytorch", "load_trial_from_checkpoint_path"), ("experimental.client", "login"),
             ("experimental.client", "create_experiment"), ("", "get_cluster_info"), ("experimental.core_v2", "init")]
TYPING = {"Any", "Dict", "List", "Optional", "Tuple", "Union", "Iterator", "Callable", "Sequence", "Set", "Type", "Iterable", "cast", "TypeVar", "Generic", "overload",
          "Mapping", "NamedTuple", "TYPE_CHECKING", "Deque", "IO", "Generator", "ContextManager", "Awaitable", "T", "Literal", "no_type_check", "SupportsFloat", "U", "C"}


def _params(f):  # noqa: ANN001, ANN202
    sig = inspect.signature(f)
    return ([p.name for p in sig.parameters.values() if p.kind not in (p.VAR_POSITIONAL, p.VAR_KEYWORD)], any(p.kind == p.VAR_KEYWORD for p in sig.parameters.values()))


def inspect_side(which: str) -> None:
    if which == "ref":
        sys.path.insert(0, str(ROOT / "baseline" / "stubs"))
        sys.path.insert(0, str(ROOT / "baseline" / "_ref"))
        root = "determined"
    else:
        sys.path.insert(0, str(ROOT))
        root = "determined_b200"
    out = {"names": {}, "methods": {}, "params": {}, "modules": {}}
    for m in MODULES:
        try:
            mod = importlib.import_module(root + ("." + m if m else ""))
            out["names"][m] = sorted(n for n in dir(mod) if not n.startswith("_") and not isinstance(getattr(mod, n), types.ModuleType))
        except Exception as e:  # noqa: BLE001
            out["names"][m] = f"import failed: {e}"[:100]
    for m, c in CLASSES:
        try:
            cls = getattr(importlib.import_module(root + ("." + m if m else "")), c)
        except Exception:  # noqa: BLE001
            continue
        out["methods"][f"{m}.{c}"] = sorted(n for n in dir(cls) if not n.startswith("_"))
        for n in dir(cls):
            f = getattr(cls, n, None)
            if (n.startswith("_") and n != "__init__") or not callable(f):
                continue
            try:
                out["params"][f"{m}.{c}.{n}"] = _params(f)
            except (TypeError, ValueError):
                pass
    for m, fn in FUNCTIONS:
        try:
            out["params"][f"{m}.{fn}"] = _params(getattr(importlib.import_module(root + ("." + m if m else "")), fn))
        except Exception:  # noqa: BLE001
            pass
    if which == "ours":
        import determined_b200

        determined_b200.install_alias()
        ref = ROOT / "baseline" / "_ref" / "determined"
        for f in sorted(ref.rglob("*.py")):
            mod = ".".join(f.relative_to(ref).with_suffix("").parts)
            mod = mod[:-len(".__init__")] if mod.endswith(".__init__") else mod
            if mod == "__init__" or mod.startswith(("keras", "deploy", "cli.")) or "tensorflow" in mod:
                continue
            try:
                importlib.import_module("determined." + mod)
                out["modules"][mod] = True
            except Exception:  # noqa: BLE001
                out["modules"][mod] = False
    print(json.dumps(out))


def main() -> int:
    if len(sys.argv) > 1 and sys.argv[1] in ("ref", "ours"):
        inspect_side(sys.argv[1])
        return 0
    if not (ROOT / "baseline" / "_ref" / "determined").is_dir():
        print("the reference is not installed under baseline/_ref (DESIGN.md §3)")
        return 1
    sides = {}
    for which in ("ref", "ours"):
        r = subprocess.run([sys.executable, __file__, which], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900, cwd="/tmp")
        line = next((ln for ln in reversed(r.stdout.splitlines()) if ln.startswith("{")), None)
        if line is None:
            print(r.stderr[-1000:])
            return 1
        sides[which] = json.loads(line)
    ref, ours = sides["ref"], sides["ours"]
    lines = ["# Public Python surface against the reference", "", "Generated by `python scripts/api_surface_diff.py`: both packages are imported (each in its own interpreter) and their",
             "public modules, names, methods and keyword arguments compared.  Typing re-exports are ignored.", ""]
    mods = ours["modules"]
    bad = sorted(m for m, ok in mods.items() if not ok)
    lines += ["## Module paths", "", f"{sum(mods.values())} of {len(mods)} module paths of `determined.*` (keras, deploy and the CLI's internals aside) import under",
              "`determined_b200.install_alias()`." + (f"  Not importable: {', '.join('`' + b + '`' for b in bad)}." if bad else ""), ""]
    lines += ["## Names per module", "", "| module | names of the reference | missing here |", "|---|---|---|"]
    for m in MODULES:
        r, o = ref["names"].get(m), ours["names"].get(m)
        if isinstance(r, str):
            lines.append(f"| `determined.{m}` | (the reference's module does not import here: {r}) | |")
            continue
        if isinstance(o, str):
            lines.append(f"| `determined.{m}` | {len(r)} | module missing |")
            continue
        miss = [n for n in r if n not in o and n not in TYPING]
        lines.append(f"| `determined{'.' + m if m else ''}` | {len([n for n in r if n not in TYPING])} | {', '.join('`' + n + '`' for n in miss) or '—'} |")
    lines += ["", "## Methods and attributes per class", "", "| class | of the reference | missing here |", "|---|---|---|"]
    for k, r in ref["methods"].items():
        o = ours["methods"].get(k)
        miss = [n for n in r if o is None or n not in o]
        lines.append(f"| `{k.lstrip('.')}` | {len(r)} | {'class missing' if o is None else (', '.join('`' + n + '`' for n in miss) or '—')} |")
    gaps = []
    for k, (ps, _) in ref["params"].items():
        if k in ours["params"]:
            ops, var = ours["params"][k]
            miss = [p for p in ps if p not in ops and p not in ("self", "cls")]
            if miss and not var:
                gaps.append((k.lstrip("."), miss))
    lines += ["", "## Keyword arguments", "", f"{len(ref['params'])} callables of the classes and functions above were compared; {len(gaps)} have parameters of the reference that are not",
              "accepted here" + (" (constructors of objects that the SDK and the contexts build themselves; users get them from `get_*` / `create_*` calls):" if gaps else "."), ""]
    lines += [f"* `{k}`: {', '.join('`' + p + '`' for p in miss)}" for k, miss in gaps]
    text = "\n".join(lines) + "\n"
    (ROOT / "docs" / "API_SURFACE.md").write_text(text)
    print(f"wrote docs/API_SURFACE.md: {len(bad)} module paths, {len(gaps)} signatures differ")
    return 0


if __name__ == "__main__":
    sys.exit(main())
