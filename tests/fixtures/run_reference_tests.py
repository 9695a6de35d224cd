"""Run test files of the REFERENCE's own suite against this package (``import determined`` aliased).

usage: run_reference_tests.py <reference harness dir> [pytest args...]   (paths relative to the harness dir)
"""
import os
import pathlib
import sys

ref = sys.argv[1]
repo = str(pathlib.Path(__file__).resolve().parents[2])
sys.path[:] = [p for p in sys.path if pathlib.Path(p or ".").resolve() != pathlib.Path(repo, "tests", "fixtures").resolve()]
sys.path.insert(0, repo)
os.chdir(ref)                        # the reference's `tests` package is imported from here (rootdir-relative)
sys.path.insert(0, ref)
sys.path.append(str(pathlib.Path(__file__).resolve().parent / "ref_shims"))   # stand-ins for test-only imports the image lacks

import determined_b200  # noqa: E402

determined_b200.install_alias()

import pytest  # noqa: E402

import determined  # noqa: E402

assert determined is determined_b200
sys.exit(pytest.main(["-q", "-p", "no:cacheprovider", "-c", os.devnull, "--rootdir", os.environ.get("TMPDIR", "/tmp"), *sys.argv[2:]]))
