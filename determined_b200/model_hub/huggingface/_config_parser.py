"""Hyperparameter -> kwargs parsing for HF trials (reference ``model_hub/huggingface/_config_parser.py``)."""
import dataclasses
from typing import Any, Dict, Optional, Tuple, Union

from determined_b200.model_hub import utils


class FlexibleDataclass:
    """Dataclass-like bag whose fields without defaults may stay unset."""
    _fields: Dict[str, Any] = {}

    def __init__(self, **kwargs: Any) -> None:
        for k, default in type(self)._fields.items():
            if k in kwargs:
                setattr(self, k, kwargs[k])
            elif default is not dataclasses.MISSING:
                setattr(self, k, default)
        for k, v in kwargs.items():
            if k not in type(self)._fields:
                raise TypeError(f"{type(self).__name__} got an unexpected field {k!r}")

    def as_dict(self) -> Dict[str, Any]:
        return {k: getattr(self, k) for k in type(self)._fields if hasattr(self, k)}

    def __repr__(self) -> str:
        return f"{type(self).__name__}({', '.join(f'{k}={v!r}' for k, v in self.as_dict().items())})"


M = dataclasses.MISSING


class DatasetKwargs(FlexibleDataclass):
    _fields = {"dataset_name": None, "dataset_config_name": None, "validation_split_percentage": None, "train_file": None, "validation_file": None}


class ConfigKwargs(FlexibleDataclass):
    _fields = {"pretrained_model_name_or_path": M, "cache_dir": None, "revision": "main", "use_auth_token": False, "num_labels": M, "finetuning_task": M}


class TokenizerKwargs(FlexibleDataclass):
    _fields = {"pretrained_model_name_or_path": M, "cache_dir": None, "revision": "main", "use_auth_token": False, "use_fast": True, "do_lower_case": M}


class ModelKwargs(FlexibleDataclass):
    _fields = {"pretrained_model_name_or_path": M, "cache_dir": None, "revision": "main", "use_auth_token": False}


@dataclasses.dataclass
class OptimizerKwargs:
    weight_decay: float = 0.0
    adafactor: bool = False
    learning_rate: float = 5e-5
    max_grad_norm: float = 1.0
    adam_beta1: float = 0.9
    adam_beta2: float = 0.999
    adam_epsilon: float = 1e-8
    scale_parameter: bool = False
    relative_step: bool = False


@dataclasses.dataclass
class LRSchedulerKwargs:
    num_training_steps: int
    lr_scheduler_type: str = "linear"
    num_warmup_steps: int = 0


def parse_dict_to_dataclasses(dataclass_types: Tuple[Any, ...], args: Union[Dict[str, Any], utils.AttrDict], as_dict: bool = False) -> Tuple[Any, ...]:
    outputs = []
    for dt in dataclass_types:
        keys = set(dt._fields) if hasattr(dt, "_fields") else {f.name for f in dataclasses.fields(dt) if f.init}
        obj = dt(**{k: v for k, v in args.items() if k in keys})
        if as_dict:
            obj = obj.as_dict() if hasattr(obj, "as_dict") else dataclasses.asdict(obj)
        outputs.append(obj)
    return tuple(outputs)


def default_parse_config_tokenizer_model_kwargs(hparams: Union[Dict, utils.AttrDict]) -> Tuple[Dict, Dict, Dict]:
    if not isinstance(hparams, utils.AttrDict):
        hparams = utils.AttrDict(hparams)
    config_args, tokenizer_args, model_args = parse_dict_to_dataclasses((ConfigKwargs, TokenizerKwargs, ModelKwargs), hparams, as_dict=True)
    if hparams.get("config_name") is not None:
        config_args["pretrained_model_name_or_path"] = hparams.config_name
    if hparams.get("tokenizer_name") is not None:
        tokenizer_args["pretrained_model_name_or_path"] = hparams.tokenizer_name
    for d in (config_args, tokenizer_args, model_args):
        if d.get("pretrained_model_name_or_path") is None:
            raise AssertionError("pretrained_model_name_or_path (or config_name / tokenizer_name) is required to build the model")
    return config_args, tokenizer_args, model_args


def default_parse_optimizer_lr_scheduler_kwargs(hparams: Union[Dict, utils.AttrDict]) -> Tuple[OptimizerKwargs, LRSchedulerKwargs]:
    return parse_dict_to_dataclasses((OptimizerKwargs, LRSchedulerKwargs), hparams)  # type: ignore[return-value]
