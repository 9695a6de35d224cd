"""A small reader for the ``.proto`` files under ``determined_b200/proto/``: the API definition of this master.

Parity: reference ``proto/src/determined/**.proto`` compiled by ``buf`` + grpc-gateway into swagger, which feeds
``bindings/generate_bindings_py.py``.  There is no gRPC server here (REST is the only wire), so nothing needs ``protoc``: the
files use the proto3 subset below and exist to be the ONE description of messages, enums and routes that

* ``common/api/generate_bindings.py`` turns into typed Python bindings (``common/api/bindings.py``),
* ``master/openapi.py`` turns into ``components.schemas`` + typed operations of ``GET /api/v1/openapi.json``,
* ``tests/test_proto_conformance.py`` holds the master's actual answers against.

Subset: ``syntax`` / ``package`` / ``import`` (ignored: all files share one namespace keyed by package); ``message`` with
scalar, message, enum, ``repeated``, ``optional``, ``map<string, T>`` and ``oneof`` fields, nested messages and enums,
``option (required) = "a, b";`` and the field option ``[json_name = "x"]``; ``enum``; ``service`` with ``rpc X(Req) returns
([stream] Resp) { option (http) = { get: "/path/{field}" [body: "*" | "field"] }; }``.  ``//`` comments directly above a
declaration become its doc string.

JSON mapping is proto3's: field ``request_id`` travels as ``requestId``; 64-bit integers may travel as strings; enums travel
by value name; ``google.protobuf.{Struct,Value,Timestamp,Duration,FieldMask,Empty}`` and the wrapper types are the usual
JSON forms.
"""
import dataclasses
import pathlib
import re
from typing import Any, Dict, Iterator, List, Optional, Tuple

PROTO_DIR = pathlib.Path(__file__).resolve().parents[2] / "proto"

SCALARS = {"double": "number", "float": "number", "int32": "integer", "uint32": "integer", "sint32": "integer",
           "int64": "integer64", "uint64": "integer64", "sint64": "integer64", "bool": "boolean", "string": "string", "bytes": "string"}
WELL_KNOWN = {
    "google.protobuf.Struct": "object", "google.protobuf.Value": "any", "google.protobuf.ListValue": "array",
    "google.protobuf.Timestamp": "string", "google.protobuf.Duration": "string", "google.protobuf.FieldMask": "string",
    "google.protobuf.Empty": "object", "google.protobuf.Int32Value": "integer", "google.protobuf.Int64Value": "integer64",
    "google.protobuf.UInt32Value": "integer", "google.protobuf.DoubleValue": "number", "google.protobuf.FloatValue": "number",
    "google.protobuf.BoolValue": "boolean", "google.protobuf.StringValue": "string",
}


class ProtoError(ValueError):
    pass


def json_name(field: str) -> str:
    """proto3 JSON name: ``validate_after_length`` -> ``validateAfterLength``."""
    parts = field.split("_")
    return parts[0] + "".join(p[:1].upper() + p[1:] for p in parts[1:])


@dataclasses.dataclass
class Field:
    name: str                    # proto name (snake_case)
    type: str                    # scalar name, well-known type, or fully qualified message / enum name
    number: int
    repeated: bool = False
    map_value: bool = False      # map<string, type>
    optional: bool = False       # explicit `optional`
    oneof: Optional[str] = None
    doc: str = ""
    json_override: Optional[str] = None      # [json_name = "..."]

    @property
    def json(self) -> str:
        return self.json_override or json_name(self.name)


def _package_of(full_name: str) -> str:
    return ".".join(p for p in full_name.split(".") if p[:1].islower())


def _local_of(full_name: str) -> str:
    return "".join(p for p in full_name.split(".") if p[:1].isupper())


@dataclasses.dataclass
class Message:
    full_name: str               # determined.experiment.v1.SearcherEvent  (nested: ...TrialExitedEarly.ExitedReason)
    fields: List[Field]
    required: List[str]          # proto names
    doc: str = ""

    @property
    def package(self) -> str:
        return _package_of(self.full_name)

    @property
    def local(self) -> str:
        return _local_of(self.full_name)


@dataclasses.dataclass
class Enum:
    full_name: str
    values: List[Tuple[str, int]]
    doc: str = ""

    @property
    def package(self) -> str:
        return _package_of(self.full_name)

    @property
    def local(self) -> str:
        return _local_of(self.full_name)


@dataclasses.dataclass
class Rpc:
    name: str
    request: str
    response: str
    server_stream: bool
    method: str                  # get / post / put / patch / delete
    path: str                    # /api/v1/experiments/{experiment_id}
    body: Optional[str]          # None, "*" or a field name of the request
    doc: str = ""

    @property
    def path_params(self) -> List[str]:
        """Field paths named in the route (``experiment_id``, or ``experiment.id`` for a field of a sub-message)."""
        return re.findall(r"\{([\w.]+)\}", self.path)


_TOKEN = re.compile(r"""\s*(?:(//[^\n]*)|(/\*.*?\*/)|("(?:[^"\\]|\\.)*")|([A-Za-z_][\w.]*)|(-?\d+)|(.))""", re.S)


def _tokens(text: str) -> Iterator[Tuple[str, str]]:
    pos = 0
    while pos < len(text):
        m = _TOKEN.match(text, pos)
        if m is None:
            break
        pos = m.end()
        if m.group(1) is not None:
            yield "doc", m.group(1)[2:].strip()
        elif m.group(2) is not None:
            continue
        elif m.group(3) is not None:
            yield "str", m.group(3)[1:-1]
        elif m.group(4) is not None:
            yield "id", m.group(4)
        elif m.group(5) is not None:
            yield "num", m.group(5)
        elif m.group(6) is not None and m.group(6).strip():
            yield "sym", m.group(6)


class _Parser:
    def __init__(self, text: str, fname: str) -> None:
        self.toks = list(_tokens(text))
        self.i = 0
        self.fname = fname
        self.package = ""
        self.messages: List[Message] = []
        self.enums: List[Enum] = []
        self.rpcs: List[Rpc] = []

    # token helpers -------------------------------------------------------------------------------------------------
    def _doc(self) -> str:
        """Consume the comment lines in front of the next declaration."""
        lines = []
        while self.i < len(self.toks) and self.toks[self.i][0] == "doc":
            lines.append(self.toks[self.i][1])
            self.i += 1
        return " ".join(lines).strip()

    def _peek(self) -> Tuple[str, str]:
        j = self.i
        while j < len(self.toks) and self.toks[j][0] == "doc":
            j += 1
        return self.toks[j] if j < len(self.toks) else ("eof", "")

    def _next(self) -> Tuple[str, str]:
        while self.i < len(self.toks) and self.toks[self.i][0] == "doc":
            self.i += 1
        if self.i >= len(self.toks):
            raise ProtoError(f"{self.fname}: unexpected end of file")
        t = self.toks[self.i]
        self.i += 1
        return t

    def _expect(self, value: str) -> None:
        t = self._next()
        if t[1] != value:
            raise ProtoError(f"{self.fname}: expected {value!r}, found {t[1]!r} (token {self.i})")

    def _ident(self) -> str:
        t = self._next()
        if t[0] != "id":
            raise ProtoError(f"{self.fname}: expected a name, found {t[1]!r}")
        return t[1]

    # grammar ---------------------------------------------------------------------------------------------------------
    def parse(self) -> None:
        while True:
            doc = self._doc()
            if self.i >= len(self.toks):
                return
            kw = self._ident()
            if kw == "syntax":
                self._expect("=")
                self._next()
                self._expect(";")
            elif kw == "package":
                self.package = self._ident()
                self._expect(";")
            elif kw == "import":
                if self._peek()[1] in ("public", "weak"):
                    self._next()
                self._next()
                self._expect(";")
            elif kw == "option":
                self._skip_statement()
            elif kw == "message":
                self._message(self.package, doc)
            elif kw == "enum":
                self._enum(self.package, doc)
            elif kw == "service":
                self._service()
            else:
                raise ProtoError(f"{self.fname}: unexpected {kw!r} at top level")

    def _skip_statement(self) -> None:
        depth = 0
        while True:
            t = self._next()
            if t[1] == "{":
                depth += 1
            elif t[1] == "}":
                depth -= 1
            elif t[1] == ";" and depth == 0:
                return

    def _field_tail(self) -> Tuple[int, Optional[str]]:
        """``= N [json_name = "x", other = options];`` -> (N, json_name or None)."""
        self._expect("=")
        t = self._next()
        if t[0] != "num":
            raise ProtoError(f"{self.fname}: field number expected, found {t[1]!r}")
        override: Optional[str] = None
        if self._peek()[1] == "[":
            self._next()
            while True:
                k = self._next()
                if k[1] == "]":
                    break
                if k[1] == "json_name":
                    self._expect("=")
                    override = self._next()[1]
        self._expect(";")
        return int(t[1]), override

    def _message(self, scope: str, doc: str) -> None:
        name = self._ident()
        full = f"{scope}.{name}"
        self._expect("{")
        msg = Message(full, [], [], doc)
        self.messages.append(msg)
        while True:
            fdoc = self._doc()
            t = self._peek()
            if t[1] == "}":
                self._next()
                break
            kw = self._ident()
            if kw == "message":
                self._message(full, fdoc)
            elif kw == "enum":
                self._enum(full, fdoc)
            elif kw == "option":
                self._expect("(")
                opt = self._ident()
                self._expect(")")
                self._expect("=")
                val = self._next()
                self._expect(";")
                if opt == "required":
                    msg.required += [x.strip() for x in val[1].split(",") if x.strip()]
            elif kw == "reserved":
                self._skip_statement()
            elif kw == "oneof":
                oname = self._ident()
                self._expect("{")
                while True:
                    odoc = self._doc()
                    if self._peek()[1] == "}":
                        self._next()
                        break
                    ftype = self._ident()
                    fname = self._ident()
                    num, jn = self._field_tail()
                    msg.fields.append(Field(fname, ftype, num, oneof=oname, doc=odoc, json_override=jn))
            elif kw == "map":
                self._expect("<")
                ktype = self._ident()
                self._expect(",")
                vtype = self._ident()
                self._expect(">")
                if ktype not in ("string", "int32", "int64"):
                    raise ProtoError(f"{self.fname}: map key {ktype} not supported")
                fname = self._ident()
                num, jn = self._field_tail()
                msg.fields.append(Field(fname, vtype, num, map_value=True, doc=fdoc, json_override=jn))
            else:
                repeated = optional = False
                if kw == "repeated":
                    repeated, kw = True, self._ident()
                elif kw == "optional":
                    optional, kw = True, self._ident()
                fname = self._ident()
                num, jn = self._field_tail()
                msg.fields.append(Field(fname, kw, num, repeated=repeated, optional=optional, doc=fdoc, json_override=jn))

    def _enum(self, scope: str, doc: str) -> None:
        name = self._ident()
        self._expect("{")
        en = Enum(f"{scope}.{name}", [], doc)
        self.enums.append(en)
        while True:
            self._doc()
            if self._peek()[1] == "}":
                self._next()
                break
            vname = self._ident()
            if vname == "option":
                self._skip_statement()
                continue
            self._expect("=")
            num = self._next()
            if self._peek()[1] == "[":
                while self._next()[1] != "]":
                    pass
            self._expect(";")
            en.values.append((vname, int(num[1])))

    def _service(self) -> None:
        self._ident()
        self._expect("{")
        while True:
            doc = self._doc()
            if self._peek()[1] == "}":
                self._next()
                return
            kw = self._ident()
            if kw == "option":
                self._skip_statement()
                continue
            if kw != "rpc":
                raise ProtoError(f"{self.fname}: expected rpc, found {kw!r}")
            name = self._ident()
            self._expect("(")
            req = self._ident()
            self._expect(")")
            if self._ident() != "returns":
                raise ProtoError(f"{self.fname}: rpc {name}: 'returns' expected")
            self._expect("(")
            stream = False
            resp = self._ident()
            if resp == "stream":
                stream, resp = True, self._ident()
            self._expect(")")
            method = path = ""
            body: Optional[str] = None
            if self._peek()[1] == "{":
                self._next()
                while self._peek()[1] != "}":
                    if self._ident() != "option":
                        raise ProtoError(f"{self.fname}: rpc {name}: only options are allowed in the body")
                    self._expect("(")
                    opt = self._ident()
                    self._expect(")")
                    self._expect("=")
                    if opt.split(".")[-1] == "http":
                        self._expect("{")
                        while self._peek()[1] != "}":
                            key = self._ident()
                            self._expect(":")
                            val = self._next()[1]
                            if key == "body":
                                body = val
                            elif key in ("get", "post", "put", "patch", "delete"):
                                method, path = key, val
                            if self._peek()[1] in (",", ";"):
                                self._next()
                        self._next()
                        if self._peek()[1] == ";":
                            self._next()
                    else:
                        self._skip_statement()
                self._next()
                if self._peek()[1] == ";":
                    self._next()
            else:
                self._expect(";")
            if not method:
                raise ProtoError(f"{self.fname}: rpc {name} has no (http) option")
            self.rpcs.append(Rpc(name, req, resp, stream, method, path, body, doc))


class Registry:
    """Every message, enum and rpc of a directory of proto files, with type names resolved."""

    def __init__(self) -> None:
        self.messages: Dict[str, Message] = {}
        self.enums: Dict[str, Enum] = {}
        self.rpcs: List[Rpc] = []

    def add_file(self, text: str, fname: str = "<string>") -> None:
        p = _Parser(text, fname)
        p.parse()
        for m in p.messages:
            if m.full_name in self.messages:
                raise ProtoError(f"{fname}: message {m.full_name} defined twice")
            self.messages[m.full_name] = m
        for e in p.enums:
            self.enums[e.full_name] = e
        for r in p.rpcs:
            r.request, r.response = self._pending(r.request, p.package), self._pending(r.response, p.package)
            self.rpcs.append(r)

    @staticmethod
    def _pending(name: str, scope: str) -> str:
        return name if name.startswith("google.protobuf.") else f"{scope}\x00{name}"

    def _lookup(self, name: str, scope: str) -> str:
        if name in SCALARS or name in WELL_KNOWN:
            return name
        if name.startswith("."):
            name, scope = name[1:], ""
        parts = scope.split(".") if scope else []
        while True:
            cand = ".".join(parts + [name])
            if cand in self.messages or cand in self.enums:
                return cand
            if not parts:
                raise ProtoError(f"type {name!r} (used in {scope}) is not defined")
            parts.pop()

    def resolve(self) -> "Registry":
        for m in self.messages.values():
            for f in m.fields:
                f.type = self._lookup(f.type, m.full_name)
            known = {f.name for f in m.fields}
            for r in m.required:
                if r not in known:
                    raise ProtoError(f"{m.full_name}: required field {r!r} does not exist")
        for r in self.rpcs:
            for attr in ("request", "response"):
                v = getattr(r, attr)
                if "\x00" in v:
                    scope, name = v.split("\x00")
                    setattr(r, attr, self._lookup(name, scope))
            if r.request in self.messages:
                fields = {f.name for f in self.messages[r.request].fields}
                for p in r.path_params:
                    msg: Optional[Message] = self.messages[r.request]
                    for part in p.split("."):
                        f = next((x for x in msg.fields if x.name == part), None) if msg is not None else None
                        if f is None:
                            raise ProtoError(f"rpc {r.name}: path parameter {p!r} is not a field of {r.request}")
                        msg = self.messages.get(f.type)
                if r.body not in (None, "*") and r.body not in fields:
                    raise ProtoError(f"rpc {r.name}: body field {r.body!r} is not a field of {r.request}")
        return self

    # names in the generated bindings / OpenAPI document -------------------------------------------------------------
    def binding_name(self, full_name: str) -> str:
        """``determined.experiment.v1.SearcherEvent`` -> ``v1SearcherEvent``; when two packages define the same local name,
        both get their package in front (``experimentv1State`` / ``checkpointv1State``), like the swagger the reference's
        generator reads."""
        obj: Any = self.messages.get(full_name) or self.enums[full_name]
        local, pkg = obj.local, obj.package
        clash = [o for o in list(self.messages.values()) + list(self.enums.values()) if o.local == local and o.package != pkg]
        parts = pkg.split(".")
        version = parts[-1] if parts else "v1"
        if clash and len(parts) >= 2:
            return f"{parts[-2]}{version}{local}"
        return f"{version}{local}"

    # OpenAPI ---------------------------------------------------------------------------------------------------------
    def field_schema(self, f: Field) -> Dict[str, Any]:
        if f.type in SCALARS or f.type in WELL_KNOWN:
            kind = SCALARS.get(f.type) or WELL_KNOWN[f.type]
            if kind == "integer64":
                base: Dict[str, Any] = {"type": "string", "format": f.type.split(".")[-1].lower()}
            elif kind == "any":
                base = {}
            elif kind == "string" and f.type.endswith("Timestamp"):
                base = {"type": "string", "format": "date-time"}
            else:
                base = {"type": kind}
        else:
            base = {"$ref": "#/components/schemas/" + self.binding_name(f.type)}
        if f.repeated:
            base = {"type": "array", "items": base}
        elif f.map_value:
            base = {"type": "object", "additionalProperties": base}
        if f.doc:
            base = dict(base, description=f.doc) if "$ref" not in base else {"allOf": [base], "description": f.doc}
        return base

    def schemas(self) -> Dict[str, Any]:
        out: Dict[str, Any] = {}
        for e in self.enums.values():
            out[self.binding_name(e.full_name)] = {"type": "string", "enum": [v for v, _ in e.values], "description": e.doc}
        for m in self.messages.values():
            sch: Dict[str, Any] = {"type": "object", "properties": {f.json: self.field_schema(f) for f in m.fields}}
            if m.required:
                by_name = {f.name: f.json for f in m.fields}
                sch["required"] = [by_name[r] for r in m.required]
            if m.doc:
                sch["description"] = m.doc
            out[self.binding_name(m.full_name)] = sch
        return dict(sorted(out.items()))


_CACHE: Dict[str, Registry] = {}


def load(directory: Optional[pathlib.Path] = None) -> Registry:
    d = pathlib.Path(directory or PROTO_DIR)
    key = str(d)
    if key not in _CACHE:
        reg = Registry()
        for p in sorted(d.rglob("*.proto")):
            reg.add_file(p.read_text(), str(p.relative_to(d)))
        _CACHE[key] = reg.resolve()
    return _CACHE[key]
