"""Runtime of the generated bindings: one table-driven base class instead of unrolled per-class code.

Every class in ``bindings.py`` is ``class v1Thing(Message)`` plus a ``_SPEC`` table (one row per field: JSON name, kind, type,
required / repeated / map).  Construction, ``from_json`` and ``to_json`` are implemented ONCE here and walk that table.
Surface kept compatible with what callers of the reference's ``determined.common.api.bindings`` rely on:

* keyword-only construction; required fields must be given, optional ones default to "unset";
* an unset optional reads as ``None``; ``to_json(omit_unset=True)`` leaves unset fields out, while a field explicitly set to
  ``None`` is sent as ``null``;
* ``from_json`` is tolerant of unknown keys (a newer master may send more) — ``from_json(obj, strict=True)`` is the
  conformance-test mode that rejects them;
* enums are ``enum.Enum`` classes whose member names drop the common prefix (``experimentv1State.COMPLETED``) and whose
  values are the wire strings (``"STATE_COMPLETED"``).
"""
import enum
import math
import sys
from typing import Any, Dict, Iterable, List, Optional, Tuple, Type, TypeVar

T = TypeVar("T", bound="Message")


class Unset:
    _inst: Optional["Unset"] = None

    def __new__(cls) -> "Unset":
        if cls._inst is None:
            cls._inst = super().__new__(cls)
        return cls._inst

    def __repr__(self) -> str:
        return "<unset>"

    def __bool__(self) -> bool:
        return False


_unset = Unset()


class DetEnum(enum.Enum):
    def to_json(self) -> str:
        return str(self.value)

    @classmethod
    def from_json(cls, v: Any) -> "DetEnum":
        return cls(v)


class APIHttpError(Exception):
    """A binding call got an answer that is not a 200."""

    def __init__(self, operation_name: str, response: Any) -> None:
        self.operation_name = operation_name
        self.response = response
        try:
            body = response.json()
            self.message = body.get("message") or (body.get("error") or {}).get("message") or str(body)
        except Exception:
            self.message = getattr(response, "text", "")
        super().__init__(f"{operation_name}: HTTP {getattr(response, 'status_code', '?')}: {self.message}")


# one row of a _SPEC table
#   (json name, kind, type name, required, repeated, is_map)      kind: "s" scalar / json value, "m" message, "e" enum, "f" float
Spec = Tuple[str, str, str, bool, bool, bool]


def dump_float(v: Any) -> Any:
    """JSON has no inf / nan: they travel as the strings proto3 JSON defines."""
    if isinstance(v, float):
        if math.isnan(v):
            return "NaN"
        if math.isinf(v):
            return "Infinity" if v > 0 else "-Infinity"
    return v


def load_float(v: Any) -> Any:
    if isinstance(v, str):
        return float(v.replace("Infinity", "inf"))
    return v


class Message:
    _SPEC: Tuple[Spec, ...] = ()
    _MODULE: str = ""

    def __init__(self, **kw: Any) -> None:
        names = set()
        for name, _, _, required, _, _ in self._SPEC:
            names.add(name)
            if name in kw:
                v = kw[name]
                if not isinstance(v, Unset):
                    self.__dict__[name] = v
                elif required:
                    raise TypeError(f"{type(self).__name__}: {name} is required")
            elif required:
                raise TypeError(f"{type(self).__name__}() missing required keyword argument {name!r}")
        extra = set(kw) - names
        if extra:
            raise TypeError(f"{type(self).__name__}() got unexpected keyword argument(s) {sorted(extra)}")

    def __getattr__(self, name: str) -> Any:
        # only reached when the instance has no such attribute: an unset optional field reads as None
        for row in type(self)._SPEC:
            if row[0] == name:
                return None
        raise AttributeError(f"{type(self).__name__!r} object has no attribute {name!r}")

    # -- JSON ----------------------------------------------------------------------------------------------------------
    @classmethod
    def _resolve(cls, type_name: str) -> Any:
        return getattr(sys.modules[cls._MODULE or cls.__module__], type_name)

    @classmethod
    def from_json(cls: Type[T], obj: Any, strict: bool = False, partial: bool = False) -> T:
        """``strict``: reject keys the message does not declare.  ``partial``: accept missing required fields (they read as
        None) — for request bodies whose required fields travel in the route instead."""
        if not isinstance(obj, dict):
            raise TypeError(f"{cls.__name__}.from_json: expected an object, got {type(obj).__name__}")
        kw: Dict[str, Any] = {}
        known = set()
        for name, kind, tname, required, repeated, is_map in cls._SPEC:
            known.add(name)
            if name not in obj:
                if required and not partial:
                    raise KeyError(f"{cls.__name__}.from_json: required field {name!r} is missing")
                if required:
                    kw[name] = None
                continue
            v = obj[name]
            if v is None:
                kw[name] = None
                continue
            if kind in ("m", "e"):
                tcls = cls._resolve(tname)
                conv = (lambda x, t=tcls: t.from_json(x, strict=strict, partial=partial)) if kind == "m" else tcls.from_json
            elif kind == "f":
                conv = load_float
            else:
                conv = None
            if conv is None:
                if strict and repeated and not isinstance(v, list):
                    raise TypeError(f"{cls.__name__}.{name}: expected a list")
                kw[name] = v
            elif repeated:
                kw[name] = [conv(x) for x in v]
            elif is_map:
                kw[name] = {k: conv(x) for k, x in v.items()}
            else:
                kw[name] = conv(v)
        if strict:
            extra = set(obj) - known
            if extra:
                raise KeyError(f"{cls.__name__}.from_json: fields {sorted(extra)} are not part of the message")
        return cls(**kw)

    def to_json(self, omit_unset: bool = False) -> Dict[str, Any]:
        out: Dict[str, Any] = {}
        for name, kind, _, _, repeated, is_map in self._SPEC:
            if name not in self.__dict__:
                if not omit_unset:
                    out[name] = None
                continue
            v = self.__dict__[name]
            if v is None:
                out[name] = None
            elif kind in ("m", "e"):
                if repeated:
                    out[name] = [x.to_json(omit_unset) if kind == "m" else x.to_json() for x in v]
                elif is_map:
                    out[name] = {k: (x.to_json(omit_unset) if kind == "m" else x.to_json()) for k, x in v.items()}
                else:
                    out[name] = v.to_json(omit_unset) if kind == "m" else v.to_json()
            elif kind == "f":
                out[name] = [dump_float(x) for x in v] if repeated else dump_float(v)
            else:
                out[name] = v
        return out

    # -- conveniences --------------------------------------------------------------------------------------------------
    def __eq__(self, other: Any) -> bool:
        return type(other) is type(self) and self.__dict__ == other.__dict__

    def __ne__(self, other: Any) -> bool:
        return not self.__eq__(other)

    __hash__ = None  # type: ignore[assignment]

    def __repr__(self) -> str:
        body = ", ".join(f"{k}={v!r}" for k, v in self.__dict__.items())
        return f"{type(self).__name__}({body})"

    __str__ = __repr__


def enum_member_names(values: Iterable[str]) -> List[Tuple[str, str]]:
    """(python member name, wire value) pairs: the longest ``_``-terminated prefix shared by all values is dropped
    (``STATE_ACTIVE, STATE_PAUSED`` -> ``ACTIVE, PAUSED``); a name that would start with a digit keeps the prefix."""
    vals = list(values)
    if len(vals) < 2:
        return [(v, v) for v in vals]
    first = vals[0].split("_")
    n = 0
    for i in range(len(first) - 1):
        pre = "_".join(first[: i + 1]) + "_"
        if all(v.startswith(pre) and len(v) > len(pre) for v in vals):
            n = len(pre)
        else:
            break
    out = []
    for v in vals:
        short = v[n:]
        out.append((short if short and not short[0].isdigit() else v, v))
    return out
