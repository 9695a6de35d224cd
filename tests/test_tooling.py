"""Packaging and deployment files stay loadable: pyproject entry points resolve, the helm chart renders to valid
manifests whose master.yaml passes the master's own config validation and whose commands only use flags that exist."""
import importlib
import pathlib
import re
import subprocess
import sys

import yaml

ROOT = pathlib.Path(__file__).resolve().parents[1]


def test_pyproject_entry_points_resolve():
    import tomllib

    proj = tomllib.loads((ROOT / "pyproject.toml").read_text())
    for name, target in proj["project"]["scripts"].items():
        mod, fn = target.split(":")
        assert callable(getattr(importlib.import_module(mod), fn)), name
    attr = proj["tool"]["setuptools"]["dynamic"]["version"]["attr"]
    mod, _, var = attr.rpartition(".")
    assert re.match(r"\d+\.\d+\.\d+", getattr(importlib.import_module(mod), var))


def _render(text, values, release="det"):
    def sub(m):
        path = m.group(1).strip()
        if path == ".Release.Name":
            return release
        cur = values
        for part in path.replace(".Values.", "", 1).split("."):
            cur = cur[part]
        return str(cur).lower() if isinstance(cur, bool) else str(cur)
    return re.sub(r"\{\{\s*([^}]+?)\s*\}\}", sub, text)


def test_helm_chart_renders_to_manifests_the_programs_accept(tmp_path):
    chart = ROOT / "helm" / "determined-b200"
    assert yaml.safe_load((chart / "Chart.yaml").read_text())["name"] == "determined-b200"
    values = yaml.safe_load((chart / "values.yaml").read_text())
    docs = []
    for t in sorted((chart / "templates").glob("*.yaml")):
        docs += [d for d in yaml.safe_load_all(_render(t.read_text(), values)) if d]
    kinds = {(d["kind"], d["metadata"]["name"]) for d in docs}
    assert {("Deployment", "det-master"), ("Service", "det-master"), ("DaemonSet", "det-agent"), ("ConfigMap", "det-master-config"),
            ("Secret", "det-master-secrets"), ("PersistentVolumeClaim", "det-master-db")} <= kinds
    from determined_b200.master import config as mconfig

    cm = next(d for d in docs if d["kind"] == "ConfigMap")
    (tmp_path / "master.yaml").write_text(cm["data"]["master.yaml"])
    cfg = mconfig.load(str(tmp_path / "master.yaml"), {"DET_SECURITY_AGENT_TOKEN": "tok"}, {})
    assert cfg["port"] == 8080 and cfg["cluster_name"] == "determined" and cfg["security.agent_token"] == "tok"
    assert cfg.db_path == "/var/lib/determined/master.db"
    # every flag the manifests pass exists in the program it is passed to
    for kind, module in (("Deployment", "determined_b200.master"), ("DaemonSet", "determined_b200.agent")):
        cmd = next(d for d in docs if d["kind"] == kind)["spec"]["template"]["spec"]["containers"][0]["command"]
        help_text = subprocess.run([sys.executable, "-m", module, "--help"], stdout=subprocess.PIPE, text=True, cwd=ROOT, timeout=120).stdout
        for flag in (a for a in cmd if a.startswith("--")):
            assert flag in help_text, f"{module} has no {flag}"
    svc = next(d for d in docs if d["kind"] == "Service")
    assert svc["spec"]["ports"][0]["port"] == 8080


def test_generated_reference_pages_are_fresh():
    """docs/reference/*.md are generated from the parser, the proto files and the two config schemas."""
    r = subprocess.run([sys.executable, str(ROOT / "scripts" / "gen_docs.py"), "--check"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stdout
    r = subprocess.run([sys.executable, str(ROOT / "scripts" / "lint_names.py")], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stdout
