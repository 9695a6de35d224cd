"""A small undefined-name check (no linters in the image): every name a module loads must be bound somewhere in that
module (assignment, import, def / class, argument, comprehension / with / except / for target, global) or be a builtin.
Scope-insensitive on purpose: it reports typos and forgotten imports, not shadowing.

    python scripts/lint_names.py [paths...]        # default: determined_b200 scripts benchmarks bench.py
"""
import ast
import builtins
import pathlib
import sys


def check(path: pathlib.Path) -> list:
    tree = ast.parse(path.read_text(), str(path))
    bound = set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__path__", "__spec__", "__class__"}
    for node in ast.walk(tree):
        if isinstance(node, (ast.Import, ast.ImportFrom)):
            for a in node.names:
                if a.name == "*":
                    return []                 # star imports: nothing can be said
                bound.add((a.asname or a.name).split(".")[0])
        elif isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef, ast.ClassDef)):
            bound.add(node.name)
            if not isinstance(node, ast.ClassDef):
                args = node.args
                for a in args.posonlyargs + args.args + args.kwonlyargs + ([args.vararg] if args.vararg else []) + ([args.kwarg] if args.kwarg else []):
                    bound.add(a.arg)
        elif isinstance(node, ast.Lambda):
            args = node.args
            for a in args.posonlyargs + args.args + args.kwonlyargs + ([args.vararg] if args.vararg else []) + ([args.kwarg] if args.kwarg else []):
                bound.add(a.arg)
        elif isinstance(node, ast.Name) and isinstance(node.ctx, (ast.Store, ast.Del)):
            bound.add(node.id)
        elif isinstance(node, ast.ExceptHandler) and node.name:
            bound.add(node.name)
        elif isinstance(node, (ast.Global, ast.Nonlocal)):
            bound.update(node.names)
        elif isinstance(node, ast.MatchAs) and node.name:
            bound.add(node.name)
    out = []
    for node in ast.walk(tree):
        if isinstance(node, ast.Name) and isinstance(node.ctx, ast.Load) and node.id not in bound:
            out.append(f"{path}:{node.lineno}: undefined name {node.id!r}")
    return out


def main() -> int:
    root = pathlib.Path(__file__).resolve().parents[1]
    targets = [pathlib.Path(a) for a in sys.argv[1:]] or [root / "determined_b200", root / "scripts", root / "benchmarks", root / "bench.py", root / "__graft_entry__.py"]
    problems = []
    for t in targets:
        for f in ([t] if t.is_file() else sorted(t.rglob("*.py"))):
            problems += check(f)
    print("\n".join(problems) or "no undefined names")
    return 1 if problems else 0


if __name__ == "__main__":
    sys.exit(main())
