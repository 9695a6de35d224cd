"""Task identity contract: ``ClusterInfo`` / ``TrialInfo`` / ``RendezvousInfo`` / ``ResourcesInfo``.

Parity: reference ``harness/determined/_info.py:1-403``.  The environment-variable contract
(``DET_MASTER, DET_CLUSTER_ID, DET_AGENT_ID, DET_SLOT_IDS, DET_TASK_ID, DET_ALLOCATION_ID,
DET_SESSION_TOKEN, DET_TASK_TYPE, DET_LATEST_CHECKPOINT`` + trial ``DET_TRIAL_ID,
DET_EXPERIMENT_ID, DET_TRIAL_SEED, DET_HPARAMS, DET_EXPERIMENT_CONFIG, DET_STEPS_COMPLETED,
DET_TRIAL_RUN_ID``; reference ``:109-124,255-291`` and ``master/pkg/tasks/task_trial.go:113-125``)
is kept verbatim so user launch layers keep working.

Design difference: one box, no containers.  The single-box master execs the launcher directly, so
the info may come *either* from JSON files under ``$DET_INFO_DIR`` (what ``prep_container`` writes)
*or* straight from the environment; ``get_cluster_info`` tries files first, then env.
"""
import dataclasses
import json
import os
from typing import Any, Dict, List, Optional, Union

from determined_b200 import constants, gpu


def _info_dir() -> str:
    return os.environ.get("DET_INFO_DIR", constants.INFO_DIR)


def _path(name: str) -> str:
    return os.path.join(_info_dir(), name)


DEFAULT_RENDEZVOUS_INFO_PATH = _path("rendezvous.json")
DEFAULT_TRIAL_INFO_PATH = _path("trial.json")
DEFAULT_RESOURCES_INFO_PATH = _path("resources.json")
DEFAULT_CLUSTER_INFO_PATH = _path("cluster.json")


def getenv_int(key: str) -> Optional[int]:
    v = os.environ.get(key)
    return None if v is None else int(v)


def _load_json(path: str) -> Optional[Dict[str, Any]]:
    try:
        with open(path) as f:
            return json.load(f)
    except FileNotFoundError:
        return None


def _dump_json(path: str, obj: Dict[str, Any]) -> None:
    os.makedirs(os.path.dirname(path) or ".", exist_ok=True)
    tmp = f"{path}.tmp{os.getpid()}"
    with open(tmp, "w") as f:
        json.dump(obj, f)
    os.replace(tmp, path)


@dataclasses.dataclass
class RendezvousInfo:
    """Machine identity configured by the rendezvous layer and consumed by the launch layer."""

    container_addrs: List[str]
    container_rank: int
    container_slot_counts: List[int]

    def _to_file(self, path: Optional[str] = None) -> None:
        _dump_json(path or _path("rendezvous.json"), dataclasses.asdict(self))

    @classmethod
    def _from_file(cls, path: Optional[str] = None) -> Optional["RendezvousInfo"]:
        d = _load_json(path or _path("rendezvous.json"))
        return None if d is None else cls(**d)

    @classmethod
    def single_box(cls, slots: int) -> "RendezvousInfo":
        return cls(["127.0.0.1"], 0, [slots])


class TrialInfo:
    """Information about the running trial (ids, seed, hparams, private config)."""

    def __init__(self, trial_id: int, experiment_id: int, trial_seed: int, hparams: Dict[str, Any],
                 config: Dict[str, Any], steps_completed: int, trial_run_id: int, debug: bool = False,
                 inter_node_network_interface: Optional[str] = None):
        self.trial_id = trial_id
        self.experiment_id = experiment_id
        self.trial_seed = trial_seed
        self.hparams = hparams
        self._config = config
        self._steps_completed = steps_completed
        self._trial_run_id = trial_run_id
        self._debug = debug
        self._inter_node_network_interface = inter_node_network_interface

    def _as_dict(self) -> Dict[str, Any]:
        return {k.lstrip("_"): v for k, v in vars(self).items()}

    @classmethod
    def _from_env(cls) -> "TrialInfo":
        if "DET_TRIAL_ID" not in os.environ:
            raise RuntimeError("DET_TRIAL_ID is not set; this is not a trial task")
        cfg = json.loads(os.environ.get("DET_EXPERIMENT_CONFIG", "{}"))
        return cls(
            trial_id=int(os.environ["DET_TRIAL_ID"]),
            experiment_id=int(os.environ.get("DET_EXPERIMENT_ID", "0")),
            trial_seed=int(os.environ.get("DET_TRIAL_SEED", "0")),
            hparams=json.loads(os.environ.get("DET_HPARAMS", "{}")),
            config=cfg,
            steps_completed=int(os.environ.get("DET_STEPS_COMPLETED", "0")),
            trial_run_id=int(os.environ.get("DET_TRIAL_RUN_ID", "0")),
            debug=bool(cfg.get("debug", False)),
            inter_node_network_interface=os.environ.get("DET_INTER_NODE_NETWORK_INTERFACE"),
        )

    def _to_file(self, path: Optional[str] = None) -> None:
        _dump_json(path or _path("trial.json"), self._as_dict())

    @classmethod
    def _from_file(cls, path: Optional[str] = None) -> Optional["TrialInfo"]:
        d = _load_json(path or _path("trial.json"))
        return None if d is None else cls(**d)


class ResourcesInfo:
    def __init__(self, gpu_uuids: List[str]) -> None:
        self._gpu_uuids = list(gpu_uuids)

    @property
    def gpu_uuids(self) -> List[str]:
        return self._gpu_uuids

    @classmethod
    def _by_inspection(cls) -> "ResourcesInfo":
        return cls(gpu.get_gpu_uuids())

    def _to_file(self, path: Optional[str] = None) -> None:
        _dump_json(path or _path("resources.json"), {"gpu_uuids": self._gpu_uuids})

    @classmethod
    def _from_file(cls, path: Optional[str] = None) -> Optional["ResourcesInfo"]:
        d = _load_json(path or _path("resources.json"))
        return None if d is None else cls(**d)


_REQUIRED_ENV = ("DET_MASTER", "DET_CLUSTER_ID", "DET_AGENT_ID", "DET_SLOT_IDS", "DET_TASK_ID",
                 "DET_ALLOCATION_ID", "DET_SESSION_TOKEN", "DET_TASK_TYPE")


def _canonical_url(url: str) -> str:
    if "://" not in url:
        url = "http://" + url
    return url.rstrip("/")


class ClusterInfo:
    """Properties set for tasks running under the (single-box) master.

    Depend on this only for launch layers / logging; training code that needs it cannot run
    off-cluster (same caveat as the reference)."""

    def __init__(self, master_url: str, cluster_id: str, agent_id: str, slot_ids: List[int], task_id: str,
                 allocation_id: str, session_token: str, task_type: str,
                 master_cert_name: Optional[str] = None, master_cert_file: Optional[str] = None,
                 latest_checkpoint: Optional[str] = None, trial_info: Optional[TrialInfo] = None,
                 rendezvous_info: Optional[RendezvousInfo] = None,
                 resources_info: Optional[ResourcesInfo] = None):
        self.master_url = _canonical_url(master_url)
        self.cluster_id = cluster_id
        self.agent_id = agent_id
        self.slot_ids = slot_ids
        self.task_id = task_id
        self.allocation_id = allocation_id
        self.session_token = session_token
        self.task_type = task_type
        self.master_cert_name = master_cert_name
        self.master_cert_file = master_cert_file
        self._latest_checkpoint = latest_checkpoint
        self._trial_info = trial_info
        self._rendezvous_info = rendezvous_info
        self._resources_info = resources_info

    @classmethod
    def _from_env(cls) -> "ClusterInfo":
        missing = [k for k in _REQUIRED_ENV if k not in os.environ]
        if missing:
            raise RuntimeError(f"missing environment keys [{', '.join(missing)}], is this running on-cluster?")
        env = os.environ
        slot_ids = json.loads(env["DET_SLOT_IDS"])
        trial = TrialInfo._from_file()
        if trial is None and env["DET_TASK_TYPE"] == "TRIAL" and "DET_TRIAL_ID" in env:
            trial = TrialInfo._from_env()
        rdzv = RendezvousInfo._from_file()
        if rdzv is None:
            rdzv = RendezvousInfo.single_box(max(len(slot_ids), 1))
        res = ResourcesInfo._from_file()
        return cls(
            master_url=env["DET_MASTER"], cluster_id=env["DET_CLUSTER_ID"], agent_id=env["DET_AGENT_ID"],
            slot_ids=slot_ids, task_id=env["DET_TASK_ID"], allocation_id=env["DET_ALLOCATION_ID"],
            session_token=env["DET_SESSION_TOKEN"], task_type=env["DET_TASK_TYPE"],
            master_cert_name=env.get("DET_MASTER_CERT_NAME"), master_cert_file=env.get("DET_MASTER_CERT_FILE"),
            latest_checkpoint=env.get("DET_LATEST_CHECKPOINT") or None,
            trial_info=trial, rendezvous_info=rdzv, resources_info=res,
        )

    def _to_file(self, path: Optional[str] = None) -> None:
        skip = ("_trial_info", "_rendezvous_info", "_resources_info")
        _dump_json(path or _path("cluster.json"), {k.lstrip("_"): v for k, v in vars(self).items() if k not in skip})

    @classmethod
    def _from_file(cls, path: Optional[str] = None) -> Optional["ClusterInfo"]:
        d = _load_json(path or _path("cluster.json"))
        if d is None:
            return None
        return cls(trial_info=TrialInfo._from_file(), rendezvous_info=RendezvousInfo._from_file(),
                   resources_info=ResourcesInfo._from_file(), **d)

    @property
    def latest_checkpoint(self) -> Optional[str]:
        return self._latest_checkpoint if self.task_type == "TRIAL" else None

    @property
    def user_data(self) -> Dict[str, Any]:
        if self.task_type != "TRIAL":
            return {}
        return self.trial._config.get("data", {}) or {}

    @property
    def trial(self) -> TrialInfo:
        if self.task_type != "TRIAL":
            raise RuntimeError(f'.trial is only available when task_type == "TRIAL" (got "{self.task_type}")')
        assert self._trial_info is not None
        return self._trial_info

    @property
    def container_addrs(self) -> List[str]:
        if self.task_type != "TRIAL" or self._rendezvous_info is None:
            return ["127.0.0.1"]
        return self._rendezvous_info.container_addrs

    @property
    def container_slot_counts(self) -> List[int]:
        if self.task_type != "TRIAL" or self._rendezvous_info is None:
            return [len(self.slot_ids)]
        return self._rendezvous_info.container_slot_counts

    @property
    def container_rank(self) -> int:
        if self.task_type != "TRIAL" or self._rendezvous_info is None:
            return 0
        return self._rendezvous_info.container_rank

    @property
    def gpu_uuids(self) -> List[str]:
        if self._resources_info is None:
            self._resources_info = ResourcesInfo._by_inspection()
        return self._resources_info.gpu_uuids


_info: Union[ClusterInfo, str, None] = "unloaded"


def get_cluster_info() -> Optional[ClusterInfo]:
    """The :class:`ClusterInfo` for this task, or ``None`` when not running under the master."""
    global _info
    if isinstance(_info, str):
        loaded = ClusterInfo._from_file()
        if loaded is None and all(k in os.environ for k in _REQUIRED_ENV):
            loaded = ClusterInfo._from_env()
        _info = loaded
    return _info


def _reset_cluster_info_cache() -> None:
    global _info
    _info = "unloaded"
