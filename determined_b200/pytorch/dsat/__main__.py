"""``python -m determined_b200.pytorch.dsat <method> config.yaml model_dir``: submit an autotuning job.

Parity: reference ``harness/determined/pytorch/dsat/__main__.py:1-78``: build the search-runner
``single`` experiment config, stage the user's config + pickled args next to the model
directory, and create the runner experiment (which in turn creates the custom-search experiment
from ``_run_dsat``).
"""
import argparse
import os
import pathlib
import pickle
import shutil
import tempfile
from typing import Any, Dict, List, Optional

from determined_b200.pytorch.dsat import _utils, defaults


def get_search_runner_config_from_args(args: argparse.Namespace) -> Dict[str, Any]:
    if args.search_runner_config is not None:
        return _utils.get_dict_from_yaml_or_json_path(args.search_runner_config)
    user = _utils.get_dict_from_yaml_or_json_path(args.config_path)
    conf: Dict[str, Any] = {k: (dict(v) if isinstance(v, dict) else v) for k, v in defaults.DEFAULT_SEARCH_RUNNER_CONFIG.items()}
    for key in ("project", "workspace", "checkpoint_storage", "environment"):
        if key in user:
            conf[key] = user[key]
    conf["name"] = f"{user.get('name', 'dsat')} (autotune search runner)"
    conf["hyperparameters"] = {k: v for k, v in vars(args).items() if isinstance(v, (int, float, str, bool, list, type(None)))}
    return conf


def stage_context(args: argparse.Namespace, dest: pathlib.Path) -> List[pathlib.Path]:
    """Copy the model dir, the config, extra includes and ``args.pkl`` into ``dest``."""
    shutil.copytree(args.model_dir, dest, dirs_exist_ok=True)
    shutil.copy(args.config_path, dest / os.path.basename(args.config_path))
    for inc in args.include or []:
        src = pathlib.Path(inc)
        if src.is_dir():
            shutil.copytree(src, dest / src.name, dirs_exist_ok=True)
        else:
            shutil.copy(src, dest / src.name)
    with (dest / defaults.ARGS_PKL_PATH).open("wb") as f:
        pickle.dump(args, f)
    return sorted(dest.rglob("*"))


def run_autotuning(args: argparse.Namespace, master_url: Optional[str] = None) -> int:
    from determined_b200.experimental import client

    conf = get_search_runner_config_from_args(args)
    with tempfile.TemporaryDirectory() as tmp:
        ctx_dir = pathlib.Path(tmp) / "ctx"
        ctx_dir.mkdir()
        stage_context(args, ctx_dir)
        d = client.Determined(master=master_url or args.master_url or os.environ.get("DET_MASTER"))
        exp = d.create_experiment(conf, str(ctx_dir))
    print(f"Created autotune search-runner experiment {exp.id}")
    return int(exp.id)


def parse_args(argv: Optional[List[str]] = None) -> argparse.Namespace:
    args = _utils.get_full_parser().parse_args(argv)
    assert args.start_profile_step <= args.end_profile_step, "start-profile-step must not exceed end-profile-step"
    assert args.max_trials >= 2, "need at least the model-profile trial and one tuned trial"
    return args


if __name__ == "__main__":
    run_autotuning(parse_args())
