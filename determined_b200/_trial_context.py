"""Framework-agnostic trial plumbing: ``TrialContext``, ``EnvContext``, ``LegacyTrial``, ``TrialController``.

Parity: reference ``harness/determined/_trial_context.py:11-150``, ``_env_context.py``, ``_trial.py``,
``_trial_controller.py``.  The PyTorch and DeepSpeed contexts of this package are built on ``core.Context`` directly and
expose the same getters; these classes exist for code that type-checks against ``det.TrialContext``, builds a context from
a bare config (``TrialContext.from_config``) in a notebook, or subclasses the legacy bases.
"""
import abc
import dataclasses
import pathlib
from typing import Any, Dict, List, Optional, Type


@dataclasses.dataclass
class EnvContext:
    """What a launched trial process knows about its surroundings (filled from ``det.get_cluster_info()``)."""

    master_url: str = ""
    master_cert_file: Optional[str] = None
    master_cert_name: Optional[str] = None
    experiment_config: Dict[str, Any] = dataclasses.field(default_factory=dict)
    hparams: Dict[str, Any] = dataclasses.field(default_factory=dict)
    latest_checkpoint: Optional[str] = None
    steps_completed: int = 0
    det_trial_id: str = ""
    det_experiment_id: str = ""
    det_agent_id: str = ""
    det_cluster_id: str = ""
    trial_seed: int = 0
    trial_run_id: int = 1
    allocation_id: str = ""
    managed_training: bool = True
    test_mode: bool = False
    on_cluster: bool = False
    slot_ids: List[int] = dataclasses.field(default_factory=list)
    use_gpu: bool = False
    container_gpus: List[str] = dataclasses.field(default_factory=list)
    debug: bool = False

    @classmethod
    def from_cluster_info(cls, info: Any) -> "EnvContext":
        t = info.trial
        return cls(master_url=info.master_url, experiment_config=dict(t._config), hparams=dict(t.hparams), latest_checkpoint=info.latest_checkpoint,
                   steps_completed=getattr(t, "_steps_completed", 0), det_trial_id=str(t.trial_id), det_experiment_id=str(t.experiment_id),
                   det_agent_id=info.agent_id, det_cluster_id=info.cluster_id, trial_seed=t.trial_seed, trial_run_id=getattr(t, "_trial_run_id", 1),
                   allocation_id=info.allocation_id, on_cluster=True, slot_ids=list(info.slot_ids))


class TrialContext:
    """Read-only view of a trial's configuration."""

    def __init__(self, core_context: Any = None, env: Optional[EnvContext] = None) -> None:
        self._core = core_context
        self.env = env or EnvContext()
        self.distributed = core_context.distributed if core_context is not None else None
        self._stop_requested = False

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "TrialContext":
        """A context outside any cluster, from an experiment config dict (its ``hyperparameters`` are taken as constants)."""
        from determined_b200 import core

        hp = {k: (v.get("val") if isinstance(v, dict) and v.get("type") == "const" else v) for k, v in (config.get("hyperparameters") or {}).items()}
        return cls(core._dummy_init(), EnvContext(experiment_config=dict(config), hparams=hp, managed_training=False))

    def get_experiment_config(self) -> Dict[str, Any]:
        return self.env.experiment_config

    def get_data_config(self) -> Dict[str, Any]:
        return self.env.experiment_config.get("data") or {}

    def get_experiment_id(self) -> int:
        return int(self.env.det_experiment_id or 0)

    def get_trial_id(self) -> int:
        return int(self.env.det_trial_id or 0)

    def get_trial_seed(self) -> int:
        return self.env.trial_seed

    def get_hparams(self) -> Dict[str, Any]:
        return self.env.hparams

    def get_hparam(self, name: str) -> Any:
        if name not in self.env.hparams:
            raise ValueError(f"Could not find name '{name}' in experiment hyperparameters. Please check your experiment configuration 'hyperparameters' section.")
        if name == "global_batch_size":
            import logging

            logging.getLogger("determined").warning("use context.get_per_slot_batch_size() / get_global_batch_size() rather than get_hparam('global_batch_size')")
        return self.env.hparams[name]

    def get_global_batch_size(self) -> int:
        return int(self.env.hparams["global_batch_size"])

    def get_per_slot_batch_size(self) -> int:
        size = self.distributed.get_size() if self.distributed is not None else 1
        return self.get_global_batch_size() // max(size, 1)

    def get_stop_requested(self) -> bool:
        return self._stop_requested

    def set_stop_requested(self, stop_requested: bool) -> None:
        if not isinstance(stop_requested, bool):
            raise AssertionError("stop_requested must be a bool")
        self._stop_requested = stop_requested

    def get_initial_batch(self) -> int:
        return self.env.steps_completed

    def get_tensorboard_path(self) -> pathlib.Path:
        from determined_b200 import tensorboard

        return tensorboard.get_base_path({})


class LegacyTrial(metaclass=abc.ABCMeta):
    """Base of framework trial classes that pair with a controller (reference ``_trial.py``)."""

    trial_controller_class: Optional[Type["TrialController"]] = None
    trial_context_class: Type[TrialContext] = TrialContext

    @abc.abstractmethod
    def __init__(self, context: TrialContext) -> None:
        ...


class TrialController(metaclass=abc.ABCMeta):
    """Runs one trial: owns the train / validate / checkpoint loop for a framework (reference ``_trial_controller.py``)."""

    def __init__(self, context: Any, env: Optional[EnvContext] = None) -> None:
        self.context, self.env = context, env or EnvContext()

    @classmethod
    def pre_execute_hook(cls, env: EnvContext, distributed_backend: Any) -> Any:
        return None

    @abc.abstractmethod
    def run(self) -> None:
        ...

    def close(self) -> None:
        pass


Trial = LegacyTrial          # the reference's name for the base class of trial definitions (determined._trial.Trial)
