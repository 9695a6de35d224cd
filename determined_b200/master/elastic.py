"""Task logs in Elasticsearch instead of the master's own database (reference ``master/internal/elastic``:
``logging: {type: elastic, host, port, security}``).

The master stays the only writer and reader: log lines arrive through the same calls as always (``Database.add_logs`` and
friends delegate here when ``log_backend`` is set), go out as one ``_bulk`` request per batch into a daily index, and come
back through ``_search`` ordered by a sequence number the master assigns (microsecond clock, strictly increasing) — that
number is the ``id`` clients page and follow with.  Retention and per-task deletion are ``_delete_by_query`` calls.
"""
import json
import logging
import threading
import time
from typing import Any, Dict, List, Optional, Tuple

import requests

logger = logging.getLogger("determined.master.elastic")
FIELDS = ("task_id", "allocation_id", "ts", "level", "rank_id", "stdtype", "source", "log")


class ElasticLogs:
    def __init__(self, host: str, port: int = 9200, scheme: str = "http", username: Optional[str] = None, password: Optional[str] = None,
                 verify: Any = True, index_prefix: str = "determined-tasklogs") -> None:
        self.base = f"{scheme}://{host}:{port}"
        self.prefix = index_prefix
        self.http = requests.Session()
        if username:
            self.http.auth = (username, password or "")
        self.verify = verify
        self._seq = 0
        self._lock = threading.Lock()

    @classmethod
    def from_config(cls, c: Dict[str, Any]) -> "ElasticLogs":
        sec = c.get("security") or {}
        tls = sec.get("tls") or {}
        verify: Any = True
        if tls.get("skip_verify"):
            verify = False
        elif tls.get("certificate"):
            verify = tls["certificate"]
        return cls(str(c["host"]), int(c.get("port") or 9200), "https" if tls.get("enabled") else "http", sec.get("username"), sec.get("password"),
                   verify, str(c.get("index_prefix") or "determined-tasklogs"))

    # -- plumbing ------------------------------------------------------------------------------------
    def _call(self, method: str, path: str, body: Any = None, ndjson: bool = False, ok_missing: bool = False) -> Dict[str, Any]:
        data = body if ndjson else (json.dumps(body) if body is not None else None)
        r = self.http.request(method, self.base + path, data=data, timeout=30, verify=self.verify,
                              headers={"Content-Type": "application/x-ndjson" if ndjson else "application/json"})
        if r.status_code == 404 and ok_missing:
            return {}
        if r.status_code >= 400:
            raise RuntimeError(f"elasticsearch {method} {path}: HTTP {r.status_code}: {r.text[:300]}")
        return r.json() if r.content else {}

    def _next_seq(self) -> int:
        with self._lock:
            self._seq = max(self._seq + 1, int(time.time() * 1e6))
            return self._seq

    def _index_for(self, ts: float) -> str:
        return f"{self.prefix}-{time.strftime('%Y.%m.%d', time.gmtime(ts))}"

    # -- the log-store surface of master/db.py ------------------------------------------------------------
    def add_logs(self, rows: List[Tuple[Any, ...]]) -> None:
        lines = []
        for row in rows:
            doc = dict(zip(FIELDS, row))
            doc["seq"] = self._next_seq()
            lines.append(json.dumps({"index": {"_index": self._index_for(float(doc["ts"] or time.time()))}}))
            lines.append(json.dumps(doc))
        if not lines:
            return
        out = self._call("POST", "/_bulk", "\n".join(lines) + "\n", ndjson=True)
        if out.get("errors"):
            bad = [it for it in out.get("items", []) if (it.get("index") or {}).get("error")]
            logger.error("elasticsearch refused %d of %d log lines: %s", len(bad), len(rows), (bad[0]["index"]["error"] if bad else ""))

    @staticmethod
    def _row(hit: Dict[str, Any]) -> Dict[str, Any]:
        src = hit["_source"]
        return {"id": src["seq"], **{k: src.get(k) for k in FIELDS}}

    def get_logs(self, task_id: str, after_id: int = 0, limit: int = 1000) -> List[Dict[str, Any]]:
        out: List[Dict[str, Any]] = []
        while len(out) < limit:               # a search window is 10 000 hits at most: page by sequence number
            n = min(limit - len(out), 10000)
            body = {"query": {"bool": {"filter": [{"term": {"task_id": task_id}}, {"range": {"seq": {"gt": after_id}}}]}}, "sort": [{"seq": "asc"}], "size": n}
            hits = (self._call("POST", f"/{self.prefix}-*/_search", body, ok_missing=True).get("hits") or {}).get("hits") or []
            out += [self._row(h) for h in hits]
            if len(hits) < n:
                break
            after_id = out[-1]["id"]
        return out

    def delete_logs_before(self, ts: float) -> int:
        return int(self._call("POST", f"/{self.prefix}-*/_delete_by_query", {"query": {"range": {"ts": {"lt": ts}}}}, ok_missing=True).get("deleted", 0))

    def delete_task_logs(self, task_id: str, before_ts: Optional[float] = None) -> int:
        flt: List[Dict[str, Any]] = [{"term": {"task_id": task_id}}]
        if before_ts is not None:
            flt.append({"range": {"ts": {"lt": before_ts}}})
        return int(self._call("POST", f"/{self.prefix}-*/_delete_by_query", {"query": {"bool": {"filter": flt}}}, ok_missing=True).get("deleted", 0))

    def log_field_values(self, task_id: str) -> Dict[str, List[Any]]:
        cols = ("allocation_id", "rank_id", "stdtype", "source")
        body = {"query": {"bool": {"filter": [{"term": {"task_id": task_id}}]}}, "size": 0, "aggs": {c: {"terms": {"field": c, "size": 1000}} for c in cols}}
        aggs = self._call("POST", f"/{self.prefix}-*/_search", body, ok_missing=True).get("aggregations") or {}
        return {c: [b["key"] for b in (aggs.get(c) or {}).get("buckets") or []] for c in cols}

    def ensure_template(self) -> None:
        """Exact-match fields as keywords (terms / aggregations need them), the line itself as text."""
        kw = {"type": "keyword"}
        self._call("PUT", f"/_index_template/{self.prefix}", {"index_patterns": [f"{self.prefix}-*"], "template": {"mappings": {"properties": {
            "task_id": kw, "allocation_id": kw, "level": kw, "stdtype": kw, "source": kw, "rank_id": {"type": "integer"}, "ts": {"type": "double"},
            "seq": {"type": "long"}, "log": {"type": "text"}}}}})
