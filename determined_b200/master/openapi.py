"""OpenAPI 3 description of the master's REST surface, derived from the route table.

Parity: reference ``proto/src/determined/api/v1/api.proto`` -> grpc-gateway swagger
(``proto/build/swagger/determined/api/v1/api.swagger.json``), which feeds the binding generators in
``bindings/``.  Here the route table of ``master/api.py`` is the single source: ``GET
/api/v1/openapi.json`` serves this document and ``common/api/generate_bindings.py`` turns it into
``common/api/bindings.py``.
"""
import re
from typing import Any, Dict, List

from determined_b200 import __version__


def _operation_id(method: str, path: str) -> str:
    parts = [p for p in path.split("/") if p and p not in ("api", "v1")]
    words: List[str] = []
    for p in parts:
        words.append("by_" + p[1:-1] if p.startswith("{") else re.sub(r"[^0-9a-zA-Z]+", "_", p))
    return (method.lower() + "_" + "_".join(words)).strip("_")


def _shape(path: str) -> str:
    """Route with parameter names blanked, for matching the route table against the rpc paths."""
    return re.sub(r"\{[^}]*\}", "{}", path.rstrip("/"))


def _typed(op: Dict[str, Any], rpc: Any, reg: Any) -> None:
    from determined_b200.common.api import protolite

    op["x-rpc"] = rpc.name
    if rpc.doc:
        op["summary"] = rpc.doc
    ref = lambda full: {"$ref": "#/components/schemas/" + reg.binding_name(full)}  # noqa: E731
    if rpc.response in reg.messages:
        op["responses"]["200"]["content"]["application/json"]["schema"] = ref(rpc.response)
    req = reg.messages.get(rpc.request)
    if req is None:
        return
    top = {p.split(".")[0] for p in rpc.path_params if "." not in p}
    if rpc.body == "*":
        op["requestBody"] = {"required": True, "content": {"application/json": {"schema": ref(rpc.request)}}}
        return
    if rpc.body:
        bf = next(f for f in req.fields if f.name == rpc.body)
        op["requestBody"] = {"required": True, "content": {"application/json": {"schema": reg.field_schema(bf)}}}
    for f in req.fields:
        if f.name in top or f.name == rpc.body or (f.type in reg.messages and not f.repeated):
            continue
        op["parameters"].append({"name": protolite.json_name(f.name), "in": "query", "required": False, "schema": reg.field_schema(f)})


def build(routes: Any) -> Dict[str, Any]:
    from determined_b200.common.api import protolite

    reg = protolite.load()
    rpc_by_route = {(r.method.upper(), _shape(r.path)): r for r in reg.rpcs}
    paths: Dict[str, Dict[str, Any]] = {}
    seen = set()
    for method, rx, fn, need_auth in routes:
        path = re.sub(r"\(\?P<(\w+)>\[\^/\]\+\)", r"{\1}", rx.pattern[1:-3])
        op_id = _operation_id(method, path)
        if op_id in seen:
            continue
        seen.add(op_id)
        doc = (fn.__doc__ or "").strip()
        summary = doc.split("\n")[0] if doc else fn.__name__.replace("_", " ")
        params = [{"name": n, "in": "path", "required": True, "schema": {"type": "string"}} for n in re.findall(r"\{(\w+)\}", path)]
        op: Dict[str, Any] = {"operationId": op_id, "summary": summary, "tags": [path.split("/")[3] if path.count("/") >= 3 else "misc"],
                              "parameters": params, "responses": {"200": {"description": "OK", "content": {"application/json": {"schema": {"type": "object"}}}}}}
        if doc and "\n" in doc:
            op["description"] = " ".join(x.strip() for x in doc.split("\n")[1:]).strip()
        if method in ("POST", "PUT", "PATCH"):
            op["requestBody"] = {"required": False, "content": {"application/json": {"schema": {"type": "object"}}}}
        if need_auth:
            op["security"] = [{"bearer": []}]
        rpc = rpc_by_route.get((method, _shape(path)))
        if rpc is not None:
            _typed(op, rpc, reg)
        paths.setdefault(path, {})[method.lower()] = op
    return {"openapi": "3.0.3", "info": {"title": "determined_b200 master API", "version": __version__},
            "components": {"securitySchemes": {"bearer": {"type": "http", "scheme": "bearer"}}, "schemas": reg.schemas()},
            "paths": dict(sorted(paths.items()))}
