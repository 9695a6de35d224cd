"""The on-cluster search-runner task of an autotuning job.

Parity: reference ``harness/determined/pytorch/dsat/_run_dsat.py:1-88``: the ``single``-searcher
"runner" experiment loads the pickled CLI args and the user's experiment config shipped in its
context directory, builds the selected search method, and drives a *custom searcher* experiment
through ``RemoteSearchRunner`` (whose state is checkpointed so the runner survives preemption).
"""
import argparse
import copy
import logging
import os
import pathlib
import pickle
from typing import Any, Dict, Type

import determined_b200 as det
from determined_b200 import core, searcher, util
from determined_b200.pytorch.dsat import _dsat_search_method as sm
from determined_b200.pytorch.dsat import _utils, defaults

SEARCH_METHODS: Dict[str, Type[sm.BaseDSATSearchMethod]] = {
    "binary": sm.BinarySearchDSATSearchMethod, "random": sm.RandomDSATSearchMethod,
    "asha": sm.ASHADSATSearchMethod, "_test": sm.TestDSATSearchMethod}


def get_search_method_class(name: str) -> Type[sm.BaseDSATSearchMethod]:
    try:
        return SEARCH_METHODS[name]
    except KeyError:
        raise ValueError(f"unknown autotune search method {name!r}; expected one of {defaults.ALL_SEARCH_METHOD_NAMES}")


def get_custom_dsat_exp_conf_from_args(args: argparse.Namespace) -> Dict[str, Any]:
    """The trial experiment: the user's config with a ``custom`` searcher whose metric is the
    autotuning metric and ``max_restarts`` 0 so an out-of-memory trial reports instead of retrying."""
    conf = copy.deepcopy(_utils.get_dict_from_yaml_or_json_path(args.config_path))
    conf.setdefault("name", "dsat")
    conf["name"] = f"{conf['name']} (autotune trials)"
    conf["searcher"] = {"name": "custom", "metric": args.metric, "smaller_is_better": _utils.smaller_is_better(args.metric)}
    conf["max_restarts"] = 0
    conf.setdefault("resources", {}).setdefault("slots_per_trial", 1)
    # every trial runs the engine's probe between these steps, and the trial controller takes its autotuning path
    probe = {"autotuning": dict(defaults.AUTOTUNING_DICT["autotuning"], start_profile_step=args.start_profile_step, end_profile_step=args.end_profile_step)}
    conf["hyperparameters"] = util.merge_dicts(conf.get("hyperparameters") or {}, {defaults.OVERWRITE_KEY: probe})
    conf["hyperparameters"][defaults.USE_DSAT_MODE_KEY] = True
    return conf


def main(core_context: core.Context) -> None:
    with pathlib.Path(defaults.ARGS_PKL_PATH).open("rb") as f:
        args = pickle.load(f)
    # inside the task container the shipped files sit in the working directory
    args.config_path = os.path.basename(args.config_path)
    args.model_dir = "."
    if args.include:
        args.include = [os.path.basename(p.rstrip("/")) for p in args.include]
    exp_config = get_custom_dsat_exp_conf_from_args(args)
    method = get_search_method_class(args.search_method)(args=args, exp_config=exp_config)
    runner = searcher.RemoteSearchRunner(method, context=core_context)
    runner.run(exp_config, model_dir=args.model_dir, includes=args.include)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO, format=det.LOG_FORMAT if hasattr(det, "LOG_FORMAT") else None)
    with core.init() as ctx:
        main(ctx)
