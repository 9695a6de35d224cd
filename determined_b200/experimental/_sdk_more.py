"""The second half of the SDK object model: mix-ins that ``experimental/client.py``'s classes inherit.

Parity (reference ``harness/determined/common/experimental/``): ``checkpoint/_checkpoint.py`` (``get_metrics``),
``experiment.py`` (``set_labels`` / ``set_notes`` / ``download_code`` / ``delete_tensorboard_files`` / ``iter_trials`` /
``get_trials``), ``trial.py`` (``iter_logs`` with filters, ``stream_metrics``), ``model.py`` (``move_to_workspace``,
``remove_metadata``; ``ModelVersion.set_name`` / ``set_notes`` / ``iter_metrics``), ``project.py`` (notes, archive, edit),
``workspace.py`` (``list_pools``, ``delete_project``), ``resource_pool.py`` (workspace bindings), ``user.py``
(``link_with_agent``), ``determined.py`` (OAuth clients, ``list_models``).

Each mix-in only uses ``self._session`` and the identifying attribute of its class, so the classes stay what they were:
thin handles around REST calls.
"""
import base64
import io
import pathlib
import tarfile
from typing import Any, Dict, Iterator, List, Optional


class CheckpointMore:
    _session: Any
    uuid: str

    def get_metrics(self, group: Optional[str] = None) -> Iterator[Dict[str, Any]]:
        """Metrics reported by the trial at the step count of this checkpoint (reference ``Checkpoint.get_metrics``)."""
        self.reload() if not getattr(self, "training", None) else None          # type: ignore[attr-defined]
        tid = (self.training or {}).get("trialId")                                # type: ignore[attr-defined]
        if tid is None:
            return
        steps = (self.metadata or {}).get("steps_completed")                      # type: ignore[attr-defined]
        params = {"group": group} if group else {}
        for m in self._session.get(f"/api/v1/trials/{tid}/metrics", params=params).json()["metrics"]:
            if steps is None or m["totalBatches"] == steps:
                from determined_b200.experimental import client as _client

                yield _client.TrialMetrics._from_json(m)


class TrialMore:
    _session: Any
    id: int

    def iter_logs(self, follow: bool = False, *, head: Optional[int] = None, tail: Optional[int] = None, levels: Optional[List[str]] = None,
                  rank_ids: Optional[List[int]] = None, sources: Optional[List[str]] = None, stdtypes: Optional[List[str]] = None,
                  allocation_ids: Optional[List[str]] = None, container_ids: Optional[List[str]] = None, min_level: Optional[str] = None,
                  timestamp_before: Optional[str] = None, timestamp_after: Optional[str] = None, search_text: Optional[str] = None) -> Iterator[str]:
        """Log lines of the trial, filtered on the master (reference ``Trial.iter_logs``)."""
        order = ["TRACE", "DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"]
        if min_level is not None:
            lv = str(getattr(min_level, "value", min_level)).replace("LOG_LEVEL_", "").upper()
            levels = order[order.index(lv):]
        params: Dict[str, Any] = {}
        for key, val in (("levels", [f"LOG_LEVEL_{x}" for x in levels] if levels else None), ("rankIds", rank_ids), ("sources", sources), ("stdtypes", stdtypes),
                         ("allocationIds", allocation_ids), ("containerIds", container_ids), ("timestampBefore", timestamp_before),
                         ("timestampAfter", timestamp_after), ("searchText", search_text)):
            if val:
                params[key] = val
        off, shown, kept = 0, 0, []
        while True:
            logs = self._session.get(f"/api/v1/trials/{self.id}/logs", params=dict(params, offset=off, limit=1000, follow=str(follow).lower())).json()["logs"]
            for l in logs:
                off = max(off, int(l["id"]))
                if tail is not None and not follow:
                    kept = (kept + [l["log"]])[-tail:]
                    continue
                yield l["log"]
                shown += 1
                if head is not None and shown >= head:
                    return
            if not logs:
                if not follow:
                    break
                self.reload()                                                      # type: ignore[attr-defined]
                if self.state in ("STATE_COMPLETED", "STATE_CANCELED", "STATE_ERROR"):   # type: ignore[attr-defined]
                    break
        yield from kept

    def stream_metrics(self, group: str) -> Iterator[Dict[str, Any]]:
        return self.iter_metrics(group)                                            # type: ignore[attr-defined]


class ExperimentMore:
    _session: Any
    id: int

    def set_labels(self, labels: List[str]) -> None:
        self._session.patch(f"/api/v1/experiments/{self.id}", json={"labels": sorted(set(labels))})
        self.labels = sorted(set(labels))

    def set_notes(self, notes: str) -> None:
        self._session.patch(f"/api/v1/experiments/{self.id}", json={"notes": notes})
        self.notes = notes

    def delete_tensorboard_files(self) -> None:
        self._session.delete(f"/api/v1/experiments/{self.id}/tensorboard-files")

    def download_code(self, output_dir: Optional[str] = None) -> str:
        """Unpack the model definition the experiment was created with; returns the directory."""
        b64 = self._session.get(f"/api/v1/experiments/{self.id}/model_def").json()["b64Tgz"]
        dst = pathlib.Path(output_dir or ".") / f"experiment_{self.id}_model_def"
        dst.mkdir(parents=True, exist_ok=True)
        data = base64.b64decode(b64)
        if data:
            with tarfile.open(fileobj=io.BytesIO(data), mode="r:*") as tf:
                tf.extractall(dst, filter="data")
        return str(dst)

    def iter_trials(self, sort_by: Any = None, order_by: Any = None, limit: int = 100) -> Iterator[Any]:
        """Trials one page at a time (reference ``Experiment.iter_trials``)."""
        from determined_b200.experimental import client

        off = 0
        params: Dict[str, Any] = {}
        if sort_by is not None:
            params["sortBy"] = getattr(sort_by, "value", sort_by)
        if order_by is not None:
            params["orderBy"] = getattr(order_by, "value", order_by)
        while True:
            r = self._session.get(f"/api/v1/experiments/{self.id}/trials", params=dict(params, offset=off, limit=limit)).json()
            for t in r["trials"]:
                yield client.Trial(self._session, t["id"], t)
            off += len(r["trials"])
            if not r["trials"] or off >= int(r["pagination"].get("total", off)):
                return

    def get_trials(self, sort_by: Any = None, order_by: Any = None) -> List[Any]:
        return list(self.iter_trials(sort_by, order_by))


class ModelVersionMore:
    _session: Any

    def _path(self) -> str:
        return f"/api/v1/models/{self.model_name}/versions/{self.model_version}"   # type: ignore[attr-defined]

    def reload(self) -> None:
        d = self._session.get(self._path()).json()["modelVersion"]
        self.name, self.comment, self.notes, self.metadata = d.get("name"), d.get("comment"), d.get("notes"), d.get("metadata") or {}

    def set_name(self, name: str) -> None:
        self._session.patch(self._path(), json={"modelVersion": {"name": name}})
        self.name = name

    def set_notes(self, notes: str) -> None:
        self._session.patch(self._path(), json={"modelVersion": {"notes": notes}})
        self.notes = notes

    def iter_metrics(self, group: Optional[str] = None) -> Iterator[Dict[str, Any]]:
        ck = getattr(self, "checkpoint", None)
        if ck is not None:
            yield from ck.get_metrics(group)

    get_metrics = iter_metrics


class ModelMore:
    _session: Any
    name: str

    def move_to_workspace(self, workspace_name: str) -> None:
        d = self._session.patch(f"/api/v1/models/{self.name}", json={"model": {"workspaceName": workspace_name}}).json()["model"]
        self.workspace_id = d.get("workspaceId")

    def remove_metadata(self, keys: List[str]) -> None:
        self.reload()                                                              # type: ignore[attr-defined]
        md = {k: v for k, v in (self.metadata or {}).items() if k not in keys}     # type: ignore[attr-defined]
        self._session.patch(f"/api/v1/models/{self.name}", json={"model": {"metadata": md}})
        self.metadata = md


class ProjectMore:
    _session: Any
    id: int

    def reload(self) -> None:
        d = self._session.get(f"/api/v1/projects/{self.id}").json()["project"]
        self.name, self.description, self.archived, self.notes = d["name"], d.get("description", ""), bool(d.get("archived")), d.get("notes") or []

    def _patch(self, **fields: Any) -> None:
        d = self._session.patch(f"/api/v1/projects/{self.id}", json=fields).json()["project"]
        self.name, self.description, self.notes = d["name"], d.get("description", ""), d.get("notes") or []

    def set_name(self, name: str) -> None:
        self._patch(name=name)

    def set_description(self, description: str) -> None:
        self._patch(description=description)

    def add_note(self, name: str, contents: str) -> None:
        self.reload()
        self._patch(notes=list(self.notes) + [{"name": name, "contents": contents}])

    def remove_note(self, name: str) -> None:
        self.reload()
        keep = [n for n in self.notes if n.get("name") != name]
        if len(keep) == len(self.notes):
            raise ValueError(f"project {self.name!r} has no note named {name!r}")
        if len(self.notes) - len(keep) > 1:
            raise ValueError(f"project {self.name!r} has several notes named {name!r}")
        self._patch(notes=keep)

    def archive(self) -> None:
        self._session.post(f"/api/v1/projects/{self.id}/archive")
        self.archived = True

    def unarchive(self) -> None:
        self._session.post(f"/api/v1/projects/{self.id}/unarchive")
        self.archived = False


class WorkspaceMore:
    _session: Any
    id: int

    def reload(self) -> None:
        d = self._session.get(f"/api/v1/workspaces/{self.id}").json()["workspace"]
        self.name, self.archived = d["name"], bool(d.get("archived"))
        self.default_compute_pool, self.default_aux_pool = d.get("defaultComputePool"), d.get("defaultAuxPool")
        self.agent_user_group, self.checkpoint_storage_config = d.get("agentUserGroup"), d.get("checkpointStorageConfig")

    def list_pools(self) -> List[Any]:
        from determined_b200.experimental import client

        names = {p["name"] for p in self._session.get(f"/api/v1/workspaces/{self.id}/available-resource-pools").json()["resourcePools"]}
        return [client.ResourcePool(self._session, p) for p in self._session.get("/api/v1/resource-pools").json()["resourcePools"] if p["name"] in names]

    def delete_project(self, name: str) -> None:
        self.get_project(name).delete()                                            # type: ignore[attr-defined]

    def archive(self) -> None:
        self._session.post(f"/api/v1/workspaces/{self.id}/archive")
        self.archived = True

    def unarchive(self) -> None:
        self._session.post(f"/api/v1/workspaces/{self.id}/unarchive")
        self.archived = False

    def set_default_pools(self, compute: Optional[str] = None, aux: Optional[str] = None) -> None:
        body = {k: v for k, v in (("defaultComputePool", compute), ("defaultAuxPool", aux)) if v is not None}
        self._session.patch(f"/api/v1/workspaces/{self.id}", json=body)


class ResourcePoolMore:
    _session: Any
    name: str

    def _path(self) -> str:
        return f"/api/v1/resource-pools/{self.name}/workspace-bindings"

    def add_bindings(self, workspace_names: List[str]) -> None:
        self._session.post(self._path(), json={"resourcePoolName": self.name, "workspaceNames": list(workspace_names)})

    def remove_bindings(self, workspace_names: List[str]) -> None:
        self._session.delete(self._path(), json={"resourcePoolName": self.name, "workspaceNames": list(workspace_names)})

    def replace_bindings(self, workspace_names: List[str]) -> None:
        self._session.put(self._path(), json={"resourcePoolName": self.name, "workspaceNames": list(workspace_names)})

    def list_workspaces(self) -> List[Any]:
        from determined_b200.experimental import client

        ids = set(self._session.get(self._path()).json()["workspaceIds"])
        return [client.Workspace(self._session, w) for w in self._session.get("/api/v1/workspaces").json()["workspaces"] if w["id"] in ids]


class UserMore:
    _session: Any
    user_id: int

    def link_with_agent(self, agent_gid: Optional[int] = None, agent_uid: Optional[int] = None, agent_group: Optional[str] = None,
                        agent_user: Optional[str] = None) -> None:
        """The Unix identity tasks of this user run as (reference ``User.link_with_agent``)."""
        aug = {"agentUid": agent_uid, "agentGid": agent_gid, "agentUser": agent_user, "agentGroup": agent_group}
        self._session.put(f"/api/v1/users/{self.user_id}/agent-user-group", json={"agentUserGroup": aug})
        self.agent_uid, self.agent_gid, self.agent_user, self.agent_group = agent_uid, agent_gid, agent_user, agent_group


class Oauth2ScimClient:
    """One registered OAuth client (reference ``common/experimental/oauth2_scim_client.py``); ``secret`` only right after creation."""

    def __init__(self, client_id: str, name: str, domain: str, secret: Optional[str] = None) -> None:
        self.id, self.name, self.domain, self.secret = client_id, name, domain, secret

    def __repr__(self) -> str:
        return f"Oauth2ScimClient(id={self.id}, name={self.name!r}, domain={self.domain!r})"


class DeterminedMore:
    _session: Any

    def add_oauth_client(self, domain: str, name: str) -> Oauth2ScimClient:
        c = self._session.post("/oauth2/clients", json={"domain": domain, "name": name}).json()
        return Oauth2ScimClient(c["id"], c["name"], c["domain"], c["secret"])

    def list_oauth_clients(self) -> List[Oauth2ScimClient]:
        return [Oauth2ScimClient(c["id"], c["name"], c["domain"]) for c in self._session.get("/oauth2/clients").json()]

    def remove_oauth_client(self, client_id: str) -> None:
        self._session.delete(f"/oauth2/clients/{client_id}")

    def list_models(self, sort_by: Any = None, order_by: Any = None, name: Optional[str] = None, description: Optional[str] = None,
                    model_id: Optional[int] = None, workspace_names: Optional[List[str]] = None, workspace_ids: Optional[List[int]] = None) -> List[Any]:
        from determined_b200.experimental import client

        models = self.get_models(name=name)                                        # type: ignore[attr-defined]
        if description:
            models = [m for m in models if description in (m.description or "")]
        if model_id is not None:
            models = [m for m in models if m.model_id == model_id]
        wanted = set(workspace_ids or [])
        if workspace_names:
            wanted |= {w.id for w in self.list_workspaces() if w.name in workspace_names}      # type: ignore[attr-defined]
        if wanted:
            models = [m for m in models if getattr(m, "workspace_id", None) in wanted]
        key = getattr(sort_by, "value", sort_by)
        if key in (client.ModelSortBy.NAME.value if hasattr(client.ModelSortBy, "NAME") else "SORT_BY_NAME", "SORT_BY_NAME"):
            models.sort(key=lambda m: m.name)
        if str(getattr(order_by, "value", order_by)).endswith("DESC"):
            models.reverse()
        return models
