"""A small JSON-Schema evaluator with the experiment-config extensions.

The reference validates experiment configs in the Go master with draft-07 JSON Schema plus seven keywords of its own
(``schemas/README.md:34-52``; implementations in ``master/pkg/schemas/extensions/*.go``):

``union``               a ``oneOf`` whose error message comes from the *one* alternative the instance was trying to be
                        (chosen by ``unionKey``: ``const:k=v``, ``singleproperty:k``, ``type:t``, ``always``, ``never``);
``checks``              ``{message: subschema}``: when the subschema fails, report exactly ``message``;
``compareProperties``   ``{"type": "a<b" | "a<=b" | "a_is_subdir_of_b" | "same_units" | "length_a<length_b", "a": path, "b": path}``;
``disallowProperties``  ``{property: message}``;
``eventuallyRequired``  like ``required``, but only enforced in the second ("completeness") phase, after cluster defaults
                        and templates were merged in;
``eventually``          a subschema enforced only in the completeness phase;
``optionalRef``         ``$ref`` that a ``null`` instance satisfies.

Null handling follows the reference: a property whose value is ``null`` counts as absent.  Errors carry the instance path in
the reference's style (``<config>.hyperparameters.lr: ...``), because its shared test cases match on those strings.
This module knows nothing about experiment configs: schemas are plain dicts looked up by ``$id`` in a registry.
"""
import re
from typing import Any, Callable, Dict, List, Optional, Tuple

Schema = Any          # dict | bool
Registry = Dict[str, Dict[str, Any]]

_TYPES: Dict[str, Callable[[Any], bool]] = {
    "object": lambda v: isinstance(v, dict),
    "array": lambda v: isinstance(v, (list, tuple)),
    "string": lambda v: isinstance(v, str),
    "boolean": lambda v: isinstance(v, bool),
    "integer": lambda v: (isinstance(v, int) and not isinstance(v, bool)) or (isinstance(v, float) and v.is_integer()),
    "number": lambda v: isinstance(v, (int, float)) and not isinstance(v, bool),
    "null": lambda v: v is None,
}


def _path(path: Tuple[Any, ...]) -> str:
    out = "<config>"
    for p in path:
        out += f"[{p}]" if isinstance(p, int) else f".{p}"
    return out


def _get(inst: Any, dotted: str) -> Any:
    for part in dotted.split("."):
        if not isinstance(inst, dict) or part not in inst:
            raise KeyError(dotted)
        inst = inst[part]
    return inst


class Validator:
    def __init__(self, registry: Registry, complete: bool) -> None:
        self.reg = registry
        self.complete = complete

    # ------------------------------------------------------------------ entry
    def errors(self, schema: Schema, inst: Any) -> List[str]:
        out: List[str] = []
        self._v(schema, inst, (), out)
        # stable order, no duplicates
        seen, uniq = set(), []
        for e in out:
            if e not in seen:
                seen.add(e)
                uniq.append(e)
        return uniq

    def ok(self, schema: Schema, inst: Any) -> bool:
        probe: List[str] = []
        self._v(schema, inst, (), probe)
        return not probe

    def _resolve(self, ref: str) -> Schema:
        if ref not in self.reg:
            raise KeyError(f"unknown schema reference {ref}")
        return self.reg[ref]

    # ------------------------------------------------------------------ the walk
    def _v(self, s: Schema, inst: Any, path: Tuple[Any, ...], errs: List[str]) -> None:
        if s is True or s == {}:
            return
        if s is False:
            errs.append(f"{_path(path)}: is not allowed")
            return
        here = _path(path)
        if "$ref" in s:
            self._v(self._resolve(s["$ref"]), inst, path, errs)
        if "optionalRef" in s and inst is not None:
            self._v(self._resolve(s["optionalRef"]), inst, path, errs)
        if "type" in s:
            ts = s["type"] if isinstance(s["type"], list) else [s["type"]]
            if not any(_TYPES[t](inst) for t in ts):
                errs.append(f"{here}: must be of type {' or '.join(t for t in ts if t != 'null')}"
                            if len(ts) > 1 else f"{here}: must be of type {ts[0]}")
                return                          # the remaining keywords assume the type
        if "const" in s and inst != s["const"]:
            errs.append(f"{here}: must be {s['const']!r}")
        if "enum" in s and inst not in s["enum"]:
            errs.append(f"{here}: must be one of {[e for e in s['enum'] if e is not None]}")
        if isinstance(inst, (int, float)) and not isinstance(inst, bool):
            if "minimum" in s and inst < s["minimum"]:
                errs.append(f"{here}: must be >= {s['minimum']}")
            if "maximum" in s and inst > s["maximum"]:
                errs.append(f"{here}: must be <= {s['maximum']}")
            if "exclusiveMinimum" in s and inst <= s["exclusiveMinimum"]:
                errs.append(f"{here}: must be > {s['exclusiveMinimum']}")
            if "exclusiveMaximum" in s and inst >= s["exclusiveMaximum"]:
                errs.append(f"{here}: must be < {s['exclusiveMaximum']}")
        if isinstance(inst, str):
            if "pattern" in s and re.search(s["pattern"], inst) is None:
                errs.append(f"{here}: does not match pattern {s['pattern']!r}")
            if "minLength" in s and len(inst) < s["minLength"]:
                errs.append(f"{here}: must be at least {s['minLength']} characters long")
        if isinstance(inst, dict):
            self._object(s, inst, path, errs)
        if isinstance(inst, (list, tuple)) and "items" in s:
            for i, item in enumerate(inst):
                self._v(s["items"], item, path + (i,), errs)
        if "not" in s and self._passes(s["not"], inst, path):
            errs.append(f"{here}: is not allowed here")
        if "allOf" in s:
            for sub in s["allOf"]:
                self._v(sub, inst, path, errs)
        if "anyOf" in s and not any(self._passes(sub, inst, path) for sub in s["anyOf"]):
            errs.append(f"{here}: does not match any of the allowed forms")
        if "oneOf" in s and sum(1 for sub in s["oneOf"] if self._passes(sub, inst, path)) != 1:
            errs.append(f"{here}: must match exactly one of the allowed forms")
        if "if" in s:
            branch = "then" if self._passes(s["if"], inst, path) else "else"
            if branch in s:
                self._v(s[branch], inst, path, errs)
        # ---- extensions
        if "union" in s:
            self._union(s["union"], inst, path, errs)
        if "checks" in s:
            for msg, sub in s["checks"].items():
                if not self._passes(sub, inst, path):
                    errs.append(f"{here}: {msg}")
        if "compareProperties" in s:
            self._compare(s["compareProperties"], inst, path, errs)
        if "disallowProperties" in s and isinstance(inst, dict):
            for prop, msg in s["disallowProperties"].items():
                if inst.get(prop) is not None:
                    errs.append(f"{here}: {msg}")
        if self.complete:
            if "eventuallyRequired" in s and isinstance(inst, dict):
                for prop in s["eventuallyRequired"]:
                    if inst.get(prop) is None:
                        errs.append(f"{here}: {prop} is a required property")
            if "eventually" in s:
                self._v(s["eventually"], inst, path, errs)

    def _passes(self, s: Schema, inst: Any, path: Tuple[Any, ...]) -> bool:
        probe: List[str] = []
        self._v(s, inst, path, probe)
        return not probe

    def _object(self, s: Dict[str, Any], inst: Dict[str, Any], path: Tuple[Any, ...], errs: List[str]) -> None:
        here = _path(path)
        for prop in s.get("required", []):
            if inst.get(prop) is None:                 # null counts as absent
                errs.append(f"{here}: {prop} is a required property")
        props = s.get("properties", {})
        for k, sub in props.items():
            if k in inst:
                self._v(sub, inst[k], path + (k,), errs)
        extra = s.get("additionalProperties", True)
        if extra is not True:
            for k, v in inst.items():
                if k in props:
                    continue
                if extra is False:
                    errs.append(f"{here}: additional property {k!r} is not allowed")
                else:
                    self._v(extra, v, path + (k,), errs)

    # ------------------------------------------------------------------ union
    @staticmethod
    def _union_key_matches(key: str, inst: Any) -> bool:
        if key == "always":
            return True
        if key == "never":
            return False
        kind, _, arg = key.partition(":")
        if kind == "not":
            return not Validator._union_key_matches(arg, inst)
        if kind == "hasattr":
            return isinstance(inst, dict) and arg in inst
        if kind == "const":
            k, _, v = arg.partition("=")
            return isinstance(inst, dict) and k in inst and str(inst[k]) == v
        if kind == "singleproperty":
            return isinstance(inst, dict) and len(inst) == 1 and arg in inst
        if kind == "type":
            return arg in _TYPES and _TYPES[arg](inst)
        raise ValueError(f"unknown unionKey {key!r}")

    def _union(self, u: Dict[str, Any], inst: Any, path: Tuple[Any, ...], errs: List[str]) -> None:
        items = u["items"]
        passing = [it for it in items if self._passes({k: v for k, v in it.items() if k != "unionKey"}, inst, path)]
        if len(passing) == 1:
            return
        here = _path(path)
        if len(passing) > 1:
            errs.append(f"{here}: matches more than one allowed form")
            return
        for it in items:
            if self._union_key_matches(it["unionKey"], inst):
                # the alternative the instance was trying to be: report ITS errors
                self._v({k: v for k, v in it.items() if k != "unionKey"}, inst, path, errs)
                return
        errs.append(f"{here}: {u.get('defaultMessage', 'does not match any of the allowed forms')}")

    # ------------------------------------------------------------------ compareProperties
    def _compare(self, c: Dict[str, Any], inst: Any, path: Tuple[Any, ...], errs: List[str]) -> None:
        try:
            a, b = _get(inst, c["a"]), _get(inst, c["b"])
        except KeyError:
            return                                  # absent operands: nothing to compare
        if a is None or b is None:
            return
        kind = c["type"]
        here = _path(path)
        try:
            if kind == "a<b":
                ok, msg = a < b, f"{c['a']} must be less than {c['b']}"
            elif kind == "a<=b":
                ok, msg = a <= b, f"{c['a']} must be less than {c['b']}"
            elif kind == "a_is_subdir_of_b":
                import os

                aa = a if os.path.isabs(a) else os.path.join(b, a)
                norm = os.path.normpath(aa)
                base = os.path.normpath(b)
                ok = norm == base or norm.startswith(base.rstrip("/") + "/")
                msg = f"{c['a']} must be a subdirectory of {c['b']}"
            elif kind == "same_units":
                ok = not (isinstance(a, dict) and isinstance(b, dict)) or set(a) == set(b)
                msg = f"{c['a']} must be specified in the same units as {c['b']}"
            elif kind == "length_a<length_b":
                va = next(iter(a.values())) if isinstance(a, dict) else a
                vb = next(iter(b.values())) if isinstance(b, dict) else b
                ok, msg = va < vb, f"{c['a']} must be less than {c['b']}"
            else:
                raise ValueError(f"unknown compareProperties type {kind!r}")
        except TypeError:
            return                                  # wrongly typed operands are reported by the type keywords
        if not ok:
            errs.append(f"{here}: {msg}")


# ---------------------------------------------------------------------------------------------------------------------
# defaults and merging, driven by the same schemas
# ---------------------------------------------------------------------------------------------------------------------
def _deref(reg: Registry, s: Schema) -> Schema:
    while isinstance(s, dict) and ("$ref" in s or "optionalRef" in s) and "properties" not in s and "union" not in s:
        s = reg[s.get("$ref") or s["optionalRef"]]
    return s


def _union_member(reg: Registry, s: Dict[str, Any], inst: Any) -> Optional[Schema]:
    """The alternative of ``s``'s union that ``inst`` selects by its tag (None when it selects none).  The union may sit
    directly in ``s`` or under ``then`` of an ``if``/``then`` pair (a section whose tag may still be missing: checkpoint
    storage without ``type``, searcher without ``name``)."""
    u = s.get("union")
    if u is None and isinstance(s.get("then"), dict) and "union" in s["then"]:
        if not Validator(reg, False).ok(s.get("if", True), inst):
            return None
        u = s["then"]["union"]
    if u is None:
        return None
    for it in u["items"]:
        if it["unionKey"] not in ("always", "never") and Validator._union_key_matches(it["unionKey"], inst):
            return _deref(reg, {k: v for k, v in it.items() if k != "unionKey"})
    for it in u["items"]:
        if it["unionKey"] == "always":
            return _deref(reg, {k: v for k, v in it.items() if k != "unionKey"})
    # untagged alternatives ("never"): the one the instance validates against
    for it in u["items"]:
        sub = {k: v for k, v in it.items() if k != "unionKey"}
        if "not" not in sub and Validator(reg, False).ok(sub, inst):
            return _deref(reg, sub)
    return None


def _has_union(s: Dict[str, Any]) -> bool:
    return "union" in s or (isinstance(s.get("then"), dict) and "union" in s["then"])


def fill_defaults(reg: Registry, schema: Schema, inst: Any, hooks: Optional[Dict[str, Callable[[Any], Any]]] = None) -> Any:
    """``inst`` with every ``default`` of the schema filled in, recursively (a ``null`` or missing property takes its
    default; unions follow the member the instance selects).  ``hooks`` maps a schema ``$id`` to a function applied to the
    defaulted instance of that schema (the reference's runtime defaults, e.g. the environment image map)."""
    import copy

    s = _deref(reg, schema)
    if not isinstance(s, dict):
        return inst
    if isinstance(inst, dict) and "properties" in s:
        out = dict(inst)
        for k, sub in s["properties"].items():
            if not isinstance(sub, dict):
                continue
            subd = _deref(reg, sub)
            cur = out.get(k)
            if cur is None and "default" in sub:
                cur = copy.deepcopy(sub["default"])
            elif cur is None and isinstance(subd, dict) and "default" in subd:
                cur = copy.deepcopy(subd["default"])
            if cur is not None:
                cur = fill_defaults(reg, sub, cur, hooks)
            if cur is not None or k in out or "default" in sub:
                out[k] = cur
        extra = s.get("additionalProperties")
        if isinstance(extra, dict):
            for k in out:
                if k not in s["properties"] and out[k] is not None:
                    out[k] = fill_defaults(reg, extra, out[k], hooks)
        inst = out
    elif isinstance(inst, dict) and isinstance(s.get("additionalProperties"), dict):
        inst = {k: (fill_defaults(reg, s["additionalProperties"], v, hooks) if v is not None else v) for k, v in inst.items()}
    elif isinstance(inst, list) and "items" in s:
        inst = [fill_defaults(reg, s["items"], v, hooks) for v in inst]
    if inst is not None and _has_union(s):
        m = _union_member(reg, s, inst)
        if m is not None:
            inst = fill_defaults(reg, m, inst, hooks)
    hook = (hooks or {}).get(s.get("$id", ""))
    return hook(inst) if hook else inst


def merge(reg: Registry, schema: Schema, obj: Any, src: Any, hooks: Optional[Dict[str, Callable[[Any, Any], Any]]] = None) -> Any:
    """``obj`` completed from ``src`` (reference ``master/pkg/schemas/merge.go``): a value present in ``obj`` wins; objects merge
    property by property along the schema; arrays are replaced unless the schema's ``$id`` has a merge hook (bind mounts,
    devices: append what is not already there; hyperparameters: leaves are atomic).  When both sides have chosen different
    members of a union (an s3 section over a shared_fs one), only the properties that every member shares flow over."""
    import copy

    s = _deref(reg, schema)
    if isinstance(s, dict) and s.get("$id", "") in (hooks or {}):
        return hooks[s["$id"]](obj, src)          # type: ignore[index]
    if obj is None:
        return copy.deepcopy(src)
    if src is None:
        return obj
    if isinstance(s, dict) and _has_union(s) and isinstance(obj, dict) and isinstance(src, dict):
        mo, ms = _union_member(reg, s, obj), _union_member(reg, s, src)
        if mo is not None and ms is not None and mo is not ms:
            common = set(mo.get("properties") or {}) & set(ms.get("properties") or {})
            out = dict(obj)
            for k in common:
                if src.get(k) is not None and "const" not in (mo["properties"][k] if isinstance(mo["properties"][k], dict) else {}):
                    out[k] = merge(reg, mo["properties"][k], obj.get(k), src[k], hooks)
            return out
        member = mo or ms
        if member is not None:
            s = member
    if isinstance(obj, dict) and isinstance(src, dict):
        props = (s.get("properties") or {}) if isinstance(s, dict) else {}
        extra = s.get("additionalProperties") if isinstance(s, dict) else None
        out = dict(obj)
        for k, v in src.items():
            sub = props.get(k, extra if isinstance(extra, dict) else {})
            out[k] = merge(reg, sub if isinstance(sub, dict) else {}, obj.get(k), v, hooks)
        return out
    return obj
