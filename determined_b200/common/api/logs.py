"""Log readers over a session (reference ``harness/determined/common/api/logs.py``: ``trial_logs``, ``task_logs``,
``pprint_logs``).  Entries are ``bindings.v1LogEntry`` objects (``message`` / ``log``, ``level``, ``timestamp``, ``rankId`` ...);
``follow=True`` keeps reading until the task is over."""
import datetime
import time
from typing import Any, Iterable, Iterator, List, Optional, Union


def _ts(v: Union[None, str, int, float, datetime.datetime]) -> Optional[str]:
    if v is None:
        return None
    if isinstance(v, datetime.datetime):
        return v.isoformat("T").replace("+00:00", "") + "Z" if v.tzinfo is None or v.utcoffset() == datetime.timedelta(0) else v.isoformat("T")
    if isinstance(v, (int, float)):
        return datetime.datetime.fromtimestamp(v).isoformat("T") + "Z"
    try:
        datetime.datetime.strptime(v.split(".")[0].rstrip("Z"), "%Y-%m-%dT%H:%M:%S")
        if not v.endswith("Z"):
            raise ValueError
    except ValueError:
        raise ValueError(f"timestamp {v!r} must look like 2021-10-26T23:17:12Z (UTC)")
    return v


def levels_at_or_above(level: Any) -> List[Any]:
    from determined_b200.common.api import bindings

    order = ["TRACE", "DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"]
    name = getattr(level, "name", str(level)).replace("LOG_LEVEL_", "")
    names = order[order.index(name):] if name in order else order
    return [getattr(bindings.v1LogLevel, n) for n in names if hasattr(bindings.v1LogLevel, n)] if hasattr(bindings, "v1LogLevel") else names


def _read(session: Any, path: str, task_running: Any, head: Optional[int], tail: Optional[int], follow: bool, params: dict) -> Iterator[Any]:
    from determined_b200.common.api import bindings

    entries = session.get(path, params=dict(params, limit=head or 100000)).json()["logs"]
    if tail:
        entries = entries[-tail:]
    last = 0
    for e in entries:
        last = max(last, int(e["id"]))
        yield bindings.v1LogEntry.from_json(e)
    while follow and head is None:
        more = session.get(path, params=dict(params, offset=last, follow="true", timeout_seconds=5)).json()["logs"]
        for e in more:
            last = max(last, int(e["id"]))
            yield bindings.v1LogEntry.from_json(e)
        if not more and not task_running():
            return
        if not more:
            time.sleep(0.2)


def _filters(agent_ids: Any, container_ids: Any, rank_ids: Any, sources: Any, stdtypes: Any, min_level: Any, timestamp_before: Any, timestamp_after: Any) -> dict:
    p: dict = {}
    for key, val in (("agentIds", agent_ids), ("containerIds", container_ids), ("rankIds", rank_ids), ("sources", sources), ("stdtypes", stdtypes)):
        if val:
            p[key] = list(val)
    if min_level is not None:
        p["levels"] = [getattr(x, "value", f"LOG_LEVEL_{x}") for x in levels_at_or_above(min_level)]
    if timestamp_before is not None:
        p["timestampBefore"] = _ts(timestamp_before)
    if timestamp_after is not None:
        p["timestampAfter"] = _ts(timestamp_after)
    return p


def trial_logs(session: Any, trial_id: int, head: Optional[int] = None, tail: Optional[int] = None, follow: bool = False,
               agent_ids: Optional[List[str]] = None, container_ids: Optional[List[str]] = None, rank_ids: Optional[List[int]] = None,
               sources: Optional[List[str]] = None, stdtypes: Optional[List[str]] = None, min_level: Any = None,
               timestamp_before: Any = None, timestamp_after: Any = None) -> Iterable[Any]:
    def running() -> bool:
        st = session.get(f"/api/v1/trials/{trial_id}").json()["trial"]["state"]
        return st in ("STATE_ACTIVE", "STATE_RUNNING", "STATE_QUEUED", "STATE_PULLING", "STATE_STARTING", "STATE_PAUSED", "STATE_STOPPING_COMPLETED",
                      "STATE_STOPPING_CANCELED", "STATE_STOPPING_ERROR", "STATE_STOPPING_KILLED")
    return _read(session, f"/api/v1/trials/{trial_id}/logs", running, head, tail, follow,
                 _filters(agent_ids, container_ids, rank_ids, sources, stdtypes, min_level, timestamp_before, timestamp_after))


def task_logs(session: Any, task_id: str, head: Optional[int] = None, tail: Optional[int] = None, follow: bool = False,
              agent_ids: Optional[List[str]] = None, container_ids: Optional[List[str]] = None, rank_ids: Optional[List[int]] = None,
              sources: Optional[List[str]] = None, stdtypes: Optional[List[str]] = None, min_level: Any = None,
              timestamp_before: Any = None, timestamp_after: Any = None) -> Iterable[Any]:
    def running() -> bool:
        return any(a["taskId"] == task_id for a in session.get("/api/v1/tasks").json()["allocationIdToSummary"].values())
    return _read(session, f"/api/v1/tasks/{task_id}/logs", running, head, tail, follow,
                 _filters(agent_ids, container_ids, rank_ids, sources, stdtypes, min_level, timestamp_before, timestamp_after))


def pprint_logs(logs: Iterable[Any]) -> None:
    for e in logs:
        print(getattr(e, "message", None) or getattr(e, "log", ""), end="")
