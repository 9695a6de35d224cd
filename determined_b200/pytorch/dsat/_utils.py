"""Autotune helpers (reference ``harness/determined/pytorch/dsat/_utils.py``): argument parser
``:23-185``, ``smaller_is_better`` ``:187``, ``dsat_reporting_context`` ``:292-329``,
``report_json_results`` ``:332``, zero search space ``:359-379``, batch config ``:381-399``,
``get_ds_config_from_hparams`` ``:402-430``, HF argument rewriting ``:432-520``."""
import argparse
import contextlib
import copy
import json
import logging
import pathlib
import random
from typing import Any, Dict, Generator, List, Optional, Union

import yaml

from determined_b200 import core, util
from determined_b200.pytorch.dsat import defaults

logger = logging.getLogger("determined.pytorch.dsat")


def get_base_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(add_help=False)
    d = defaults.AUTOTUNING_ARG_DEFAULTS
    p.add_argument("config_path")
    p.add_argument("model_dir")
    p.add_argument("-i", "--include", type=str, nargs="+")
    p.add_argument("-mt", "--max-trials", type=int, default=d["max-trials"])
    p.add_argument("-mct", "--max-concurrent-trials", type=int, default=d["max-concurrent-trials"])
    p.add_argument("-ms", "--max-slots", type=int)
    p.add_argument("-m", "--metric", type=str, default=d["metric"],
                   choices=defaults.SMALLER_IS_BETTER_METRICS + defaults.LARGER_IS_BETTER_METRICS)
    p.add_argument("--run-full-experiment", action="store_true")
    p.add_argument("-z", "--zero-stages", type=int, nargs="+", default=d["zero-stages"], choices=list(range(4)))
    p.add_argument("--start-profile-step", type=int, default=d["start-profile-step"])
    p.add_argument("--end-profile-step", type=int, default=d["end-profile-step"])
    p.add_argument("-r", "--random-seed", type=int, default=d["random-seed"])
    p.add_argument("--search-runner-config", type=str)
    p.add_argument("--master-url", type=str)
    return p


def get_full_parser() -> argparse.ArgumentParser:
    d = defaults.AUTOTUNING_ARG_DEFAULTS
    base = get_base_parser()
    p = argparse.ArgumentParser(prog="python3 -m determined.pytorch.dsat", description="DeepSpeed-style engine autotuning")
    sub = p.add_subparsers(required=True, dest="search_method")
    sub.add_parser("_test", parents=[base])
    r = sub.add_parser("random", parents=[base])
    r.add_argument("--trials-per-random-config", type=int, default=d["trials-per-random-config"])
    r.add_argument("--early-stopping", type=int)
    b = sub.add_parser("binary", parents=[base])
    b.add_argument("--search-range-factor", type=float, default=d["search-range-factor"])
    a = sub.add_parser("asha", parents=[base])
    a.add_argument("--max-rungs", type=int, default=d["max-rungs"])
    a.add_argument("--min-binary-search-trials", type=int, default=d["min-binary-search-trials"])
    a.add_argument("--asha-early-stopping", type=int, default=d["asha-early-stopping"])
    a.add_argument("--divisor", type=int, default=d["divisor"])
    a.add_argument("--search-range-factor", type=float, default=d["search-range-factor"])
    return p


def smaller_is_better(metric: str) -> bool:
    if metric in defaults.SMALLER_IS_BETTER_METRICS:
        return True
    if metric in defaults.LARGER_IS_BETTER_METRICS:
        return False
    raise ValueError(f"metric must be one of {defaults.SMALLER_IS_BETTER_METRICS + defaults.LARGER_IS_BETTER_METRICS}, not {metric}")


def get_dict_from_yaml_or_json_path(path: str) -> Dict[str, Any]:
    p = pathlib.Path(path)
    if p.suffix == ".json":
        return json.loads(p.read_text())
    if p.suffix in (".yaml", ".yml"):
        return yaml.safe_load(p.read_text())
    raise ValueError(f"unexpected file type for {path}: expected .json/.yaml/.yml")


@contextlib.contextmanager
def dsat_reporting_context(core_context: core.Context, op: core.SearcherOperation, steps_completed: Optional[int] = None) -> Generator[None, None, None]:
    """Wrap the training loop of an autotuning trial (``engine.forward`` / ``backward`` / ``step`` inside).  When the
    engine's probe ends it writes its result file and exits the process; the ``SystemExit`` is caught here, the file is
    reported as the trial's validation metrics and its searcher operation completed, and the exit continues.  A crash (out
    of memory, typically) is NOT caught: with ``max_restarts: 0`` it reaches the search method as an early exit, which is
    the signal to lower the micro-batch ceiling."""
    steps_completed = op.length if steps_completed is None else steps_completed
    try:
        yield
    except SystemExit:
        found = [pathlib.Path(x) for x in (defaults.MODEL_INFO_PROFILING_PATH, defaults.AUTOTUNING_RESULTS_PATH) if pathlib.Path(x).exists()]
        if len(found) == 1:          # a properly ended probe leaves exactly one of them
            report_json_results(core_context, op, steps_completed, found[0], add_gpu_info=found[0].name == defaults.MODEL_INFO_PROFILING_PATH)
        raise


def device_memory_bytes() -> int:
    """Memory of the device a trial runs on: HBM of GPU 0, or the host's RAM for CPU trials."""
    import torch

    if torch.cuda.is_available():
        return int(torch.cuda.get_device_properties(0).total_memory)
    try:
        import psutil

        return int(psutil.virtual_memory().total)
    except Exception:
        return 0


def report_json_results(core_context: core.Context, op: core.SearcherOperation, steps_completed: int, path: Union[str, pathlib.Path],
                        add_gpu_info: bool = False) -> None:
    if core_context.distributed.rank == 0:
        results = json.loads(pathlib.Path(path).read_text())
        if add_gpu_info:
            results["gpu_mem"] = device_memory_bytes()
        core_context.train.report_validation_metrics(steps_completed=steps_completed, metrics=results)
        op.report_completed(results)
    try:           # a probe trial has exactly one operation: drain the generator so the master sees it acknowledged
        next(core_context.searcher.operations())
    except StopIteration:
        pass
    except Exception:
        pass


def report_profile(core_context: core.Context, op: core.SearcherOperation, steps_completed: int, *, throughput: float,
                   latency_ms: float, flops_per_gpu: float = 0.0, forward_ms: float = 0.0, backward_ms: float = 0.0) -> None:
    """B200 path: our ZeRO engine has no built-in profiler that exits the process; trials measure a
    few steps (CUDA events) and report the autotune metrics directly."""
    if core_context.distributed.rank == 0:
        m = {"throughput": throughput, "latency": latency_ms, "FLOPS_per_gpu": flops_per_gpu, "forward": forward_ms, "backward": backward_ms}
        core_context.train.report_validation_metrics(steps_completed=steps_completed, metrics=m)
        op.report_completed(m)


def get_zero_stage_search_space(zero_stage: int) -> Dict[str, List[Union[bool, float]]]:
    space: Dict[str, List[Union[bool, float]]] = {}
    for s in range(1, zero_stage + 1):
        space.update(defaults.DEFAULT_ZERO_SEARCH_SPACE[s])
    return space


def get_random_zero_optim_config(zero_stage: int, rng: Optional[random.Random] = None) -> Dict[str, Union[bool, float]]:
    rng = rng or random
    cfg: Dict[str, Union[bool, float]] = {k: rng.choice(v) for k, v in get_zero_stage_search_space(zero_stage).items()}
    cfg["stage"] = zero_stage
    return cfg


def get_batch_config_from_mbs_gas_and_slots(ds_config: Dict[str, Any], slots: int) -> Dict[str, int]:
    mbs = ds_config["train_micro_batch_size_per_gpu"]
    gas = ds_config.get("gradient_accumulation_steps", defaults.GAS_DEFAULT)
    if gas == "auto":
        gas = 1
    return {"train_batch_size": mbs * gas * slots, "train_micro_batch_size_per_gpu": mbs, "gradient_accumulation_steps": gas}


def get_ds_config_from_hparams(hparams: Dict[str, Any], base_dir: Union[pathlib.Path, str] = pathlib.Path(".")) -> Dict[str, Any]:
    """Trials call this: base config file named by ``hparams["deepspeed_config"]`` overlaid with
    ``hparams["overwrite_deepspeed_args"]`` (what the search method varies)."""
    base = get_dict_from_yaml_or_json_path(str(pathlib.Path(base_dir) / hparams[defaults.CONFIG_KEY]))
    return util.merge_dicts(base, hparams.get(defaults.OVERWRITE_KEY, {}))


def get_hf_ds_config_path_from_args(args: List[str]) -> Optional[str]:
    for i, a in enumerate(args):
        if a == "--deepspeed" and i + 1 < len(args):
            return args[i + 1]
        if a.startswith("--deepspeed="):
            return a.split("=", 1)[1]
    return None


def update_hf_args(args: List[str], ds_config_dict: Dict[str, Any]) -> List[str]:
    """Make HF Trainer command-line args consistent with the (overwritten) engine config."""
    mapping = {"train_micro_batch_size_per_gpu": "--per_device_train_batch_size", "gradient_accumulation_steps": "--gradient_accumulation_steps"}
    out = list(args)
    for k, flag in mapping.items():
        if k not in ds_config_dict or ds_config_dict[k] == "auto":
            continue
        val = str(ds_config_dict[k])
        for i, a in enumerate(out):
            if a == flag and i + 1 < len(out):
                out[i + 1] = val
                break
            if a.startswith(flag + "="):
                out[i] = f"{flag}={val}"
                break
        else:
            out += [flag, val]
    return out


def get_hf_args_with_overwrites(args: List[str], hparams: Dict[str, Any]) -> List[str]:
    if defaults.OVERWRITE_KEY not in hparams:
        return args
    path = get_hf_ds_config_path_from_args(args)
    if path is None:
        raise ValueError("--deepspeed <config> not found in the HF args")
    cfg = util.merge_dicts(get_dict_from_yaml_or_json_path(path), hparams[defaults.OVERWRITE_KEY])
    with open(path, "w") as f:
        json.dump(cfg, f)
    return update_hf_args(args, cfg)
