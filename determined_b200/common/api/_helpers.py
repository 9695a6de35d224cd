"""Small helpers over a session that the CLI, the SDK and user scripts share (reference ``common/api/_util.py`` and
``common/api/_rbac.py``): paging, NTSC lookups and waits, name -> id resolution."""
import contextlib
import enum
import os
import time
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence


class PageOpts(str, enum.Enum):
    single = "1"
    all = "all"


class NTSC_Kind(enum.Enum):  # noqa: N801  (the reference's spelling)
    notebook = "notebook"
    shell = "shell"
    command = "command"
    tensorboard = "tensorboard"


AnyNTSC = Dict[str, Any]
WARNING_MESSAGE_MAP = {"LAUNCH_WARNING_CURRENT_SLOTS_EXCEEDED": "Warning: The requested job requires more slots than currently available. "
                                                               "You may need to increase cluster resources in order for the job to run."}


def get_default_master_url() -> str:
    return os.environ.get("DET_MASTER", os.environ.get("DET_MASTER_ADDR", "localhost:8080"))


def read_paginated(get_with_offset: Callable[[int], Any], offset: int = 0, pages: PageOpts = PageOpts.all) -> Iterator[Any]:
    """Call ``get_with_offset(offset)`` until its ``pagination`` says there is no more (or once, for ``PageOpts.single``)."""
    while True:
        resp = get_with_offset(offset)
        yield resp
        pg = resp.get("pagination") if isinstance(resp, dict) else getattr(resp, "pagination", None)
        get = (lambda k: pg.get(k)) if isinstance(pg, dict) else (lambda k: getattr(pg, k, None))
        if pg is None or pages == PageOpts.single:
            return
        end, total = get("endIndex"), get("total")
        if end is None or total is None or int(end) >= int(total) or int(end) <= offset:
            return
        offset = int(end)


def get_ntsc_details(session: Any, typ: NTSC_Kind, ntsc_id: str) -> AnyNTSC:
    return session.get(f"/api/v1/{typ.value}s/{ntsc_id}").json()[typ.value]


def wait_for_ntsc_state(session: Any, typ: NTSC_Kind, ntsc_id: str, predicate: Callable[[str], bool], timeout_secs: float = 10.0) -> str:
    """Poll until ``predicate(state)`` holds (states as ``STATE_RUNNING`` ...); raises ``TimeoutError`` otherwise."""
    deadline = time.time() + timeout_secs
    while True:
        state = str(get_ntsc_details(session, typ, ntsc_id)["state"])
        if predicate(state):
            return state
        if time.time() > deadline:
            raise TimeoutError(f"{typ.value} {ntsc_id} stayed in {state}")
        time.sleep(0.5)


def wait_for_task_ready(session: Any, task_id: str, progress_report: Optional[Callable[[], None]] = None, timeout_secs: float = 300.0) -> None:
    """Until the task runs (and, if it serves something, its service answers through the proxy)."""
    deadline = time.time() + timeout_secs
    while time.time() < deadline:
        for kind in NTSC_Kind:
            with contextlib.suppress(Exception):
                t = get_ntsc_details(session, kind, task_id)
                if t["state"] == "STATE_TERMINATED":
                    raise RuntimeError(f"task {task_id} terminated before it was ready")
                if t["state"] == "STATE_RUNNING":
                    return
        if progress_report:
            progress_report()
        time.sleep(0.5)
    raise TimeoutError(f"task {task_id} was not ready after {timeout_secs} s")


def not_found_errs(category: str, name: str, session: Any) -> Exception:
    from determined_b200.common.api import errors

    return errors.NotFoundException(f"{category} {name!r} not found", None) if hasattr(errors, "NotFoundException") else LookupError(f"{category} {name!r} not found")


def role_name_to_role_id(session: Any, name: str) -> int:
    for r in session.get("/api/v1/roles").json()["roles"]:
        if r["name"] == name:
            return int(r["roleId"])
    raise LookupError(f"role {name!r} not found")


def group_name_to_group_id(session: Any, name: str) -> int:
    for g in session.get("/api/v1/groups").json()["groups"]:
        if g["group"]["name"] == name:
            return int(g["group"]["groupId"])
    raise LookupError(f"group {name!r} not found")


def usernames_to_user_ids(session: Any, usernames: Sequence[str]) -> List[int]:
    by_name = {u["username"]: int(u["id"]) for u in session.get("/api/v1/users").json()["users"]}
    missing = [u for u in usernames if u not in by_name]
    if missing:
        raise LookupError(f"user(s) not found: {', '.join(missing)}")
    return [by_name[u] for u in usernames]


def workspace_by_name(session: Any, name: str) -> Dict[str, Any]:
    for w in session.get("/api/v1/workspaces", params={"name": name}).json()["workspaces"]:
        if w["name"] == name:
            return w
    raise LookupError(f"workspace {name!r} not found")


def _assignment(role_id: int, workspace_id: Optional[int]) -> Dict[str, Any]:
    return {"role": {"roleId": role_id}, "scopeWorkspaceId": workspace_id}


def create_user_assignment_request(session: Any, user: str, role: str, workspace: Optional[str] = None) -> List[Dict[str, Any]]:
    ws = workspace_by_name(session, workspace)["id"] if workspace else None
    return [{"userId": usernames_to_user_ids(session, [user])[0], "roleAssignment": _assignment(role_name_to_role_id(session, role), ws)}]


def create_group_assignment_request(session: Any, group: str, role: str, workspace: Optional[str] = None) -> List[Dict[str, Any]]:
    ws = workspace_by_name(session, workspace)["id"] if workspace else None
    return [{"groupId": group_name_to_group_id(session, group), "roleAssignment": _assignment(role_name_to_role_id(session, role), ws)}]
