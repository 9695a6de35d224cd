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


def build(routes: Any) -> Dict[str, Any]:
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
        paths.setdefault(path, {})[method.lower()] = op
    return {"openapi": "3.0.3", "info": {"title": "determined_b200 master API", "version": __version__},
            "components": {"securitySchemes": {"bearer": {"type": "http", "scheme": "bearer"}}}, "paths": dict(sorted(paths.items()))}
