"""The reference's shared experiment-config test cases (``/root/reference/schemas/test_cases/v0/*.yaml``, consumed there by
the Go master's ``pkg/schemas/expconf`` tests) run against this repository's schema engine and schemas.

Each case names schemas by URL and says what must happen: ``sane_as`` / ``sanity_errors`` (first validation phase),
``complete_as`` / ``completeness_errors`` (second phase, where ``eventuallyRequired`` applies), ``default_as`` + ``defaulted``
(defaults filled in; ``'*'`` in the expectation matches anything), ``merge_as`` + ``merge_src`` + ``merged``.
Skipped where the reference checkout is not present.
"""
import json
import pathlib
import re

import pytest
import yaml

from determined_b200.master import expconf_schema as X

CASES_DIR = pathlib.Path("/root/reference/schemas/test_cases/v0")
TEST_SCHEMAS = pathlib.Path("/root/reference/schemas/expconf/v0")
pytestmark = pytest.mark.skipif(not CASES_DIR.is_dir(), reason="reference checkout not available")


def _load_cases():
    out = []
    if not CASES_DIR.is_dir():
        return out
    for f in sorted(CASES_DIR.glob("*.yaml")):
        for i, c in enumerate(yaml.safe_load(f.read_text()) or []):
            out.append(pytest.param(c, id=f"{f.stem}:{i}:{c.get('name', '')[:50]}"))
    return out


@pytest.fixture(scope="module")
def registry():
    reg = dict(X.REGISTRY)
    # the reference's own fixture schemas (test-root / test-sub / test-union*): test data, read from the checkout
    for f in TEST_SCHEMAS.glob("test-*.json"):
        s = json.loads(f.read_text())
        reg[s["$id"]] = s
    return reg


def _strip_nulls(v):
    """The reference's null rule (schemas/README.md): a null value is the same as an absent key."""
    if isinstance(v, dict):
        return {k: _strip_nulls(x) for k, x in v.items() if x is not None}
    if isinstance(v, list):
        return [_strip_nulls(x) for x in v]
    return v


# the reference's fixture struct TestRoot fills this field at run time (master/pkg/schemas/expconf/test_structs.go)
_TEST_HOOKS = {"http://determined.ai/schemas/expconf/v0/test-root.json":
               lambda v: dict(v, runtime_defaultable=10) if isinstance(v, dict) and v.get("runtime_defaultable") is None else v}


def _matches(expected, got) -> bool:
    if expected == "*":
        return True
    if isinstance(expected, dict):
        return isinstance(got, dict) and set(expected) == set(got) and all(_matches(v, got[k]) for k, v in expected.items())
    if isinstance(expected, list):
        return isinstance(got, list) and len(expected) == len(got) and all(_matches(a, b) for a, b in zip(expected, got))
    return expected == got


@pytest.mark.parametrize("case", _load_cases())
def test_reference_case(case, registry):
    inst = case.get("case")
    for u in case.get("sane_as") or []:
        errs = X.validate(inst, u, complete=False, registry=registry)
        assert not errs, f"should be sane as {u}: {errs}"
    for u in case.get("complete_as") or []:
        errs = X.validate(inst, u, complete=True, registry=registry)
        assert not errs, f"should be complete as {u}: {errs}"
    for phase, complete in (("sanity_errors", False), ("completeness_errors", True)):
        for u, patterns in (case.get(phase) or {}).items():
            errs = X.validate(inst, u, complete=complete, registry=registry)
            assert errs, f"{phase}: expected errors as {u}, got none"
            for pat in patterns:
                assert any(re.search(pat, e) for e in errs), f"{phase} as {u}: no error matches {pat!r} in {errs}"
    if "default_as" in case:
        got = _strip_nulls(X.with_defaults(inst, case["default_as"], registry=registry, hooks=_TEST_HOOKS))
        assert _matches(_strip_nulls(case["defaulted"]), got), json.dumps({"want": case["defaulted"], "got": got}, indent=1, default=str)
    if "merge_as" in case:
        got = _strip_nulls(X.merge(inst, case["merge_src"], case["merge_as"], registry=registry))
        assert _matches(_strip_nulls(case["merged"]), got), json.dumps({"want": case["merged"], "got": got}, indent=1, default=str)
