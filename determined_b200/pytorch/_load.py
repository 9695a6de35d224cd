"""Load a trial (class + weights) back from a checkpoint directory.

Parity: reference ``harness/determined/pytorch/_load.py``: ``CheckpointLoadContext`` ``:15-67``,
``load_trial_from_checkpoint_path`` (reads ``load_data.json`` / ``metadata.json``, imports
``code/``, loads ``state_dict.pth``) ``:70-237``.
"""
import importlib
import json
import logging
import pathlib
import sys
from typing import Any, Dict, Optional, Tuple, Type, cast

import torch

import determined_b200 as det
from determined_b200 import core, errors
from determined_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_b200.pytorch._pytorch_trial import PyTorchTrial

logger = logging.getLogger("determined.pytorch")


class CheckpointLoadContext(PyTorchTrialContext):
    """A context for instantiating a trial outside of training: wrap_* only record the objects,
    nothing is moved into flat buffers and no engine is created."""

    def __init__(self, hparams: Optional[Dict[str, Any]] = None, exp_conf: Optional[Dict[str, Any]] = None,
                 map_location: Any = None) -> None:
        core_context = core._dummy_init()
        super().__init__(core_context=core_context, trial_seed=0, hparams=hparams, slots_per_trial=1,
                         num_gpus=0, exp_conf=exp_conf, aggregation_frequency=1, steps_completed=0,
                         managed_training=False, debug_enabled=False, enable_tensorboard_logging=False)
        self.device = torch.device(map_location) if map_location is not None else torch.device("cpu")

    def wrap_model(self, model: torch.nn.Module) -> torch.nn.Module:
        model = model.to(self.device)
        self.models.append(model)
        return model

    def wrap_optimizer(self, optimizer: Any, backward_passes_per_step: int = 1, fp16_compression: Any = None,
                       average_aggregated_gradients: Any = None) -> Any:
        self._optimizer_configs[id(optimizer)] = {}
        self.optimizers.append(optimizer)
        return optimizer


def _trial_class_from_spec(spec: str, code_dir: pathlib.Path) -> Type[PyTorchTrial]:
    module, _, qual = spec.partition(":")
    with det.import_from_path(code_dir):
        mod = importlib.import_module(module)
        obj: Any = mod
        for part in qual.split("."):
            obj = getattr(obj, part)
    if not (isinstance(obj, type) and issubclass(obj, PyTorchTrial)):
        raise errors.InvalidExperimentException(f"{spec} is not a PyTorchTrial subclass")
    return cast(Type[PyTorchTrial], obj)


def load_trial_from_checkpoint_path(path: str, trial_class: Optional[Type[PyTorchTrial]] = None,
                                    trial_kwargs: Optional[Dict[str, Any]] = None,
                                    torch_load_kwargs: Optional[Dict[str, Any]] = None) -> PyTorchTrial:
    """Rebuild the trial object saved under ``path`` with its weights loaded.

    ``trial_class`` overrides the class recorded in ``load_data.json``; ``trial_kwargs`` are extra
    constructor arguments; ``torch_load_kwargs`` go to ``torch.load`` (e.g. ``map_location``)."""
    ckpt_dir = pathlib.Path(path)
    torch_load_kwargs = dict(torch_load_kwargs or {})
    torch_load_kwargs.setdefault("map_location", "cpu")
    torch_load_kwargs.setdefault("weights_only", False)
    load_data_path = ckpt_dir / "load_data.json"
    hparams: Dict[str, Any] = {}
    exp_conf: Optional[Dict[str, Any]] = None
    spec: Optional[str] = None
    # ``PyTorchTrial.build_model()`` is the pre-0.12.13 API (gone after 0.14.0).  Only checkpoints old enough to lack
    # load_data.json can come from it; newer ones may use that method name for their own purposes (reference
    # ``pytorch/_load.py:118-122,213-232``).
    detect_build_model = not load_data_path.exists()
    if load_data_path.exists():
        ld = json.loads(load_data_path.read_text())
        if ld.get("trial_type") not in (None, "PyTorchTrial"):
            raise errors.InvalidCheckpointException(f"checkpoint holds a {ld.get('trial_type')}, not a PyTorchTrial")
        hparams, exp_conf, spec = ld.get("hparams") or {}, ld.get("experiment_config"), ld.get("trial_cls_spec")
    else:
        md_path = ckpt_dir / "metadata.json"
        if md_path.exists():
            md = json.loads(md_path.read_text())
            hparams = md.get("hparams") or {}
            exp_conf = md.get("experiment_config")
            ep = (exp_conf or {}).get("entrypoint")
            if isinstance(ep, str) and ":" in ep and " " not in ep:
                spec = ep
    if trial_class is None:
        if spec is None:
            raise errors.InvalidCheckpointException("checkpoint does not record its trial class; pass trial_class=")
        trial_class = _trial_class_from_spec(spec, ckpt_dir / "code")
    context = CheckpointLoadContext(hparams=hparams, exp_conf=exp_conf, map_location=torch_load_kwargs["map_location"])
    with det.import_from_path(ckpt_dir / "code") if (ckpt_dir / "code").exists() else _null():
        trial = trial_class(context, **(trial_kwargs or {}))
    sd_path = None
    for parts in (("state_dict.pth",), ("determined", "state_dict.pth"), ("pedl", "state_dict.pth"), ("checkpoint.pt",)):
        if ckpt_dir.joinpath(*parts).exists():
            sd_path = ckpt_dir.joinpath(*parts)
            break
    if sd_path is None:
        raise errors.InvalidCheckpointException(f"no state_dict.pth under {ckpt_dir}")
    checkpoint = torch.load(str(sd_path), **torch_load_kwargs)
    if detect_build_model and callable(getattr(trial, "build_model", None)):
        # with these very old checkpoints the caller gets the bare model back, as the reference does
        model = trial.build_model()
        model.load_state_dict(checkpoint["model_state_dict"] if "model_state_dict" in checkpoint else checkpoint["models_state_dict"][0])
        return model  # type: ignore[return-value]
    if "model_state_dict" in checkpoint:
        states = [checkpoint["model_state_dict"]]
    else:
        states = checkpoint["models_state_dict"]
    for model, sd in zip(context.models, states):
        try:
            model.load_state_dict(sd)
        except Exception:
            sd = dict(sd)
            torch.nn.modules.utils.consume_prefix_in_state_dict_if_present(sd, "module.")
            model.load_state_dict(sd)
    if "optimizer_state_dict" in checkpoint:
        ostates = [checkpoint["optimizer_state_dict"]]
    else:
        ostates = checkpoint.get("optimizers_state_dict", [])
    for opt, sd in zip(context.optimizers, ostates):
        try:
            sd = dict(sd)
            sd.pop("detb200_fused", None)
            opt.load_state_dict(sd)
        except Exception as e:
            logger.warning("could not restore optimizer state: %s", e)
    return trial


import contextlib  # noqa: E402


@contextlib.contextmanager
def _null() -> Any:
    yield
