"""Wire formats of the streaming-updates subsystem (reference ``common/streams/wire.py``: generated
message classes for projects, models and model versions)."""
from typing import Any, Dict, List, Optional


class ServerMsg:
    FIELDS: List[str] = []

    def __init__(self, **kw: Any) -> None:
        for f in self.FIELDS:
            setattr(self, f, kw.get(f))
        for k, v in kw.items():
            if k not in self.FIELDS:
                setattr(self, k, v)

    @classmethod
    def from_json(cls, obj: Any) -> "ServerMsg":
        return cls(**obj)

    def to_json(self) -> Dict[str, Any]:
        return dict(vars(self))

    def __repr__(self) -> str:
        return f"{type(self).__name__}({', '.join(f'{k}={v}' for k, v in vars(self).items())})"

    def __eq__(self, other: object) -> bool:
        return isinstance(other, type(self)) and vars(self) == vars(other)


class DeleteMsg:
    def __init__(self, keys: str) -> None:
        self.keys = keys

    @classmethod
    def from_json(cls, keys: str) -> "DeleteMsg":
        return cls(keys)

    def to_json(self) -> str:
        return self.keys

    def __repr__(self) -> str:
        return f"{type(self).__name__}({self.keys})"

    def __eq__(self, other: object) -> bool:
        return isinstance(other, type(self)) and self.keys == other.keys


class ClientMsg:
    def to_json(self) -> Dict[str, Any]:
        return {k: v for k, v in vars(self).items() if v is not None}

    def __eq__(self, other: object) -> bool:
        return isinstance(other, type(self)) and self.to_json() == other.to_json()


class ProjectMsg(ServerMsg):
    FIELDS = ["id", "name", "description", "archived", "workspace_id", "user_id", "state", "seq"]


class ProjectsDeleted(DeleteMsg):
    pass


class ProjectSubscriptionSpec(ClientMsg):
    def __init__(self, workspace_ids: Optional[List[int]] = None, project_ids: Optional[List[int]] = None, since: Optional[int] = None) -> None:
        self.workspace_ids, self.project_ids, self.since = workspace_ids, project_ids, since


class ModelMsg(ServerMsg):
    FIELDS = ["id", "name", "description", "archived", "labels", "workspace_id", "user_id", "seq"]


class ModelsDeleted(DeleteMsg):
    pass


class ModelSubscriptionSpec(ClientMsg):
    def __init__(self, workspace_ids: Optional[List[int]] = None, model_ids: Optional[List[int]] = None, user_ids: Optional[List[int]] = None,
                 since: Optional[int] = None) -> None:
        self.workspace_ids, self.model_ids, self.user_ids, self.since = workspace_ids, model_ids, user_ids, since


class ModelVersionMsg(ServerMsg):
    FIELDS = ["id", "name", "version", "checkpoint_uuid", "model_id", "comment", "labels", "user_id", "seq"]


class ModelVersionsDeleted(DeleteMsg):
    pass


class ModelVersionSubscriptionSpec(ClientMsg):
    def __init__(self, model_ids: Optional[List[int]] = None, model_version_ids: Optional[List[int]] = None, user_ids: Optional[List[int]] = None,
                 since: Optional[int] = None) -> None:
        self.model_ids, self.model_version_ids, self.user_ids, self.since = model_ids, model_version_ids, user_ids, since
