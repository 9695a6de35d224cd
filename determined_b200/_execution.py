"""Local / test-mode execution helpers.

Parity: reference ``harness/determined/_execution.py``: ``InvalidHP`` ``:17-24``, ``_get_gpus``
``:27-36``, ``_catch_sys_exit`` ``:39-47``, ``_generate_random_hparam_values`` ``:50-74``,
``_make_test_experiment_config`` ``:77-116``, ``_make_local_execution_exp_config`` ``:119-160``,
``_make_local_execution_env`` ``:163-207``, ``_local_execution_manager`` ``:210-236``.
The B200 build has no ``EnvContext``: local runs go through ``pytorch.init(hparams=, exp_conf=)``,
so ``_make_local_execution_env`` returns the ``core.Context`` plus the resolved (config, hparams).
"""
import contextlib
import logging
import math
import os
import pathlib
import random
import sys
import uuid
from typing import Any, Dict, Iterator, List, Optional, Tuple

from determined_b200 import constants, core, errors, gpu
from determined_b200.errors import InvalidHP  # noqa: F401  (public at ``det.InvalidHP``)

logger = logging.getLogger("determined")


def _get_gpus(limit_gpus: Optional[int]) -> Tuple[bool, List[str], List[int]]:
    gpus, _ = gpu.get_gpus()
    if limit_gpus is not None:
        gpus = gpus[:limit_gpus]
    return len(gpus) > 0, [g.uuid for g in gpus], [g.id for g in gpus]


@contextlib.contextmanager
def _catch_sys_exit() -> Any:
    try:
        yield
    except SystemExit as e:
        raise errors.InvalidExperimentException(
            "Caught a SystemExit exception. This might be raised by directly calling or using a library calling sys.exit(). "
            "Please remove any calls to sys.exit() from your model code.") from e


def _generate_random_hparam_values(hparam_def: Dict[str, Any]) -> Dict[str, Any]:
    def one(h: Any) -> Any:
        if isinstance(h, dict):
            if "type" not in h:
                return _generate_random_hparam_values(h)
            t = h["type"]
            if t == "const":
                return h["val"]
            if t == "int":
                return random.randint(h["minval"], h["maxval"])
            if t == "double":
                return random.uniform(h["minval"], h["maxval"])
            if t == "categorical":
                return random.choice(h["vals"])
            if t == "log":
                return math.pow(h["base"], random.uniform(h["minval"], h["maxval"]))
            raise ValueError(f"Wrong type of hyperparameter: {t}")
        if isinstance(h, (int, float, str, list, type(None))):
            return h
        raise ValueError(f"Wrong type of hyperparameter: {type(h)}")

    return {name: one(hparam_def[name]) for name in hparam_def}


def _make_test_experiment_config(config: Dict[str, Any]) -> Dict[str, Any]:
    """One batch of training, one validation, no restarts, every checkpoint GC'd."""
    out = dict(config)
    out.update({
        "description": "[test-mode] {}".format(out.get("description", str(uuid.uuid4()))),
        "scheduling_unit": 1,
        "min_validation_period": {"batches": 1},
        "checkpoint_storage": {**out.get("checkpoint_storage", {}), "save_experiment_best": 0, "save_trial_best": 0, "save_trial_latest": 0},
        "searcher": {"name": "single", "metric": out["searcher"]["metric"], "max_length": {"batches": 1}},
        "hyperparameters": _generate_random_hparam_values(out.get("hyperparameters", {})),
        "resources": out.get("resources", {"slots_per_trial": 1}),
        "max_restarts": 0,
    })
    return out


def _make_local_execution_exp_config(input_config: Optional[Dict[str, Any]], checkpoint_dir: str, managed_training: bool,
                                     test_mode: bool) -> Dict[str, Any]:
    cfg = dict(input_config or {})
    for key in ("bind_mounts", "checkpoint_storage", "environment", "resources", "optimizations"):
        if key in cfg:
            if test_mode and not managed_training:
                logger.info("'%s' configuration key is not supported by local test mode and will be ignored", key)
            del cfg[key]
    return {"checkpoint_storage": {"type": "shared_fs", "host_path": os.path.abspath(checkpoint_dir)}, **constants.DEFAULT_EXP_CFG, **cfg}


def _make_local_execution_env(managed_training: bool, test_mode: bool, config: Optional[Dict[str, Any]], checkpoint_dir: str,
                              tensorboard_path: Optional[pathlib.Path] = None, hparams: Optional[Dict[str, Any]] = None,
                              limit_gpus: Optional[int] = None) -> Tuple[core.Context, Dict[str, Any], Dict[str, Any]]:
    cfg = _make_local_execution_exp_config(config, checkpoint_dir, managed_training=managed_training, test_mode=test_mode)
    hp = hparams or _generate_random_hparam_values(cfg.get("hyperparameters", {}))
    use_gpu, _, slot_ids = _get_gpus(limit_gpus)
    if use_gpu and limit_gpus is not None:
        os.environ.setdefault("CUDA_VISIBLE_DEVICES", ",".join(str(i) for i in slot_ids))
    ctx = core._dummy_init(checkpoint_storage=cfg["checkpoint_storage"], tensorboard_path=tensorboard_path)
    return ctx, cfg, hp


@contextlib.contextmanager
def _local_execution_manager(context_dir: pathlib.Path) -> Iterator[None]:
    """chdir into the model directory and make it importable for the duration."""
    cwd = os.getcwd()
    path = str(pathlib.Path(context_dir).resolve())
    try:
        os.chdir(path)
        sys.path.insert(0, path)
        with _catch_sys_exit():
            yield
    finally:
        os.chdir(cwd)
        try:
            sys.path.remove(path)
        except ValueError:
            pass
