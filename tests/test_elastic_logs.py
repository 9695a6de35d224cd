"""Task logs in Elasticsearch: the master writes with _bulk and reads with _search against a stand-in that implements the
subset of the query language the backend uses (bool/filter with term and range, sort by one field, size, terms
aggregations, _delete_by_query)."""
import http.server
import json
import threading
import time

import pytest

from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import config as mconfig
from determined_b200.master import core


class FakeES(http.server.BaseHTTPRequestHandler):
    docs = []          # (index, source)
    calls = []
    templates = {}

    def _send(self, obj, status=200):
        data = json.dumps(obj).encode()
        self.send_response(status)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        self.wfile.write(data)

    @staticmethod
    def _match(src, q):
        if not q:
            return True
        if "bool" in q:
            return all(FakeES._match(src, f) for f in q["bool"].get("filter", []))
        if "term" in q:
            (k, v), = q["term"].items()
            return src.get(k) == v
        if "range" in q:
            (k, cond), = q["range"].items()
            x = src.get(k)
            return x is not None and all({"gt": x > v, "gte": x >= v, "lt": x < v, "lte": x <= v}[op] for op, v in cond.items())
        raise AssertionError(f"query not supported by the stand-in: {q}")

    def _body(self):
        return self.rfile.read(int(self.headers.get("Content-Length") or 0)).decode()

    def do_PUT(self):  # noqa: N802
        type(self).templates[self.path] = json.loads(self._body())
        self._send({"acknowledged": True})

    def do_POST(self):  # noqa: N802
        cls = type(self)
        raw = self._body()
        cls.calls.append((self.path, self.headers.get("Authorization")))
        if self.path == "/_bulk":
            lines = [json.loads(x) for x in raw.splitlines() if x.strip()]
            for action, doc in zip(lines[0::2], lines[1::2]):
                cls.docs.append((action["index"]["_index"], doc))
            return self._send({"errors": False, "items": [{"index": {"status": 201}} for _ in lines[0::2]]})
        index, op = self.path.strip("/").split("/")
        assert index.endswith("-*")
        body = json.loads(raw)
        hits = [(i, d) for i, d in cls.docs if i.startswith(index[:-1]) and self._match(d, body.get("query"))]
        if op == "_delete_by_query":
            cls.docs[:] = [x for x in cls.docs if x not in hits]
            return self._send({"deleted": len(hits)})
        assert op == "_search"
        for s in reversed(body.get("sort") or []):
            (k, direction), = s.items()
            hits.sort(key=lambda h: h[1][k], reverse=direction == "desc")
        aggs = {}
        for name, a in (body.get("aggs") or {}).items():
            vals = {}
            for _, d in hits:
                v = d.get(a["terms"]["field"])
                if v is not None:
                    vals[v] = vals.get(v, 0) + 1
            aggs[name] = {"buckets": [{"key": k, "doc_count": n} for k, n in vals.items()]}
        out = {"hits": {"hits": [{"_index": i, "_source": d} for i, d in hits[:body.get("size", 10)]]}}
        if aggs:
            out["aggregations"] = aggs
        self._send(out)

    def log_message(self, *a):
        pass


@pytest.fixture()
def es():
    FakeES.docs, FakeES.calls, FakeES.templates = [], [], {}
    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), FakeES)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    yield srv.server_address[1]
    srv.shutdown()
    srv.server_close()


def test_task_logs_live_in_elasticsearch(tmp_path, es):
    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "work"),
                    cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}})
    with pytest.raises(ValueError):
        m.apply_config(mconfig.load(None, {}, {"logging.type": "elastic"}))
    cfg = mconfig.load(None, {}, {"logging.type": "elastic", "logging.host": "127.0.0.1", "logging.port": es,
                                  "logging.security.username": "det", "logging.security.password": "pw"})
    assert cfg.printable()["logging"]["security"]["password"] != "pw"
    m.apply_config(cfg)
    assert "/_index_template/determined-tasklogs" in FakeES.templates
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        sess = api.authentication.login(m.url, "admin", "")
        cid = sess.post("/api/v1/commands", json={"config": {"entrypoint": ["/bin/sh", "-c", "for i in 1 2 3; do echo line $i; done; echo oops >&2"],
                                                             "resources": {"slots": 0}}}).json()["command"]["id"]
        for _ in range(200):
            if sess.get(f"/api/v1/commands/{cid}").json()["command"]["state"] == "STATE_TERMINATED":
                break
            time.sleep(0.1)
        logs = sess.get(f"/api/v1/tasks/{cid}/logs").json()["logs"]
        text = "".join(l["log"] for l in logs)
        assert "line 1" in text and "line 3" in text and "oops" in text
        assert m.db.one("SELECT COUNT(*) c FROM task_logs")["c"] == 0, "nothing is written to the master's own table"
        assert FakeES.docs and all(i.startswith("determined-tasklogs-20") for i, _ in FakeES.docs)
        assert all(auth and auth.startswith("Basic ") for _, auth in FakeES.calls)
        ids = [l["id"] for l in logs]
        assert ids == sorted(ids) and len(set(ids)) == len(ids), "sequence numbers order the lines"
        # paging by id and the filters work on top of the backend
        rest = sess.get(f"/api/v1/tasks/{cid}/logs", params={"offset": ids[1]}).json()["logs"]
        assert [l["id"] for l in rest] == ids[2:]
        only = sess.get(f"/api/v1/tasks/{cid}/logs", params={"searchText": "line 2"}).json()["logs"]
        assert len(only) == 1 and "line 2" in only[0]["log"]
        fields = sess.get(f"/api/v1/tasks/{cid}/logs/fields").json()[0]["result"]
        assert fields["stdtypes"] and fields["sources"]
        # retention and per-task deletion are delete-by-query calls
        before = len(FakeES.docs)
        m.log_retention_days = 0
        assert m.sweep_old_logs(now=time.time() + 5) == before and FakeES.docs == []
        m.db.add_logs([("t1", "t1.1", time.time(), "INFO", 0, "stdout", "task", "a\n"), ("t2", "t2.1", time.time(), "INFO", 0, "stdout", "task", "b\n")])
        assert m.db.delete_task_logs("t1") == 1 and [d["task_id"] for _, d in FakeES.docs] == ["t2"]
        # back to the database
        m.apply_config(mconfig.load(None, {}, {}))
        m.db.add_logs([("t3", "t3.1", time.time(), "INFO", 0, "stdout", "task", "c\n")])
        assert m.db.one("SELECT COUNT(*) c FROM task_logs")["c"] == 1
    finally:
        m.stop()
        srv.stop()
