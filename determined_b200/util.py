"""Small shared helpers.  Parity map (reference ``harness/determined/util.py``):
``merge_dicts`` :144, ``make_metrics`` :171, ``json_encode`` :199, ``write_user_code`` :249,
``make_timing_log`` :304, ``calculate_batch_sizes`` :313, ``force_create_symlink`` :394,
``signal_process_tree``/``forward_signals`` :418-465, ``is_numerical_scalar`` :467,
``mask_config_dict`` :505.  Implementations are independent.
"""
import contextlib
import copy
import datetime
import enum
import fnmatch
import inspect
import json
import logging
import math
import os
import pathlib
import shutil
import signal
import subprocess
import uuid
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Set, Tuple

from determined_b200 import constants, errors

logger = logging.getLogger("determined")

_LEGACY_TRIAL_RE = None


def rmtree_nfs_safe(path: str, ignore_errors: bool = False) -> None:
    """rmtree that tolerates NFS ``.nfsXXXX`` silly-rename ghosts."""

    def onerror(func: Any, p: str, exc_info: Any) -> None:
        if os.path.basename(p).startswith(".nfs") or ignore_errors:
            return
        raise exc_info[1]

    shutil.rmtree(path, onerror=onerror)


def is_overridden(method: Any, parent_class: type) -> bool:
    """True if bound ``method`` is not the implementation inherited from ``parent_class``."""
    if method is None:
        return False
    name = method.__name__
    parent_attr = getattr(parent_class, name, None)
    if parent_attr is None:
        return True
    return getattr(method, "__func__", method) is not getattr(parent_attr, "__func__", parent_attr)


def has_param(fn: Callable[..., Any], name: str, pos: Optional[int] = None) -> bool:
    sig = inspect.signature(fn)
    if name in sig.parameters:
        return True
    if any(p.kind is inspect.Parameter.VAR_KEYWORD for p in sig.parameters.values()):
        return True
    if pos is not None:
        if any(p.kind is inspect.Parameter.VAR_POSITIONAL for p in sig.parameters.values()):
            return True
    return False


def get_member_func(obj: Any, func_name: str) -> Any:
    f = getattr(obj, func_name, None)
    return f if callable(f) else None


def _list_to_dict(rows: List[Dict[str, Any]]) -> Dict[str, List[Any]]:
    out: Dict[str, List[Any]] = {}
    for row in rows:
        for k, v in row.items():
            out.setdefault(k, []).append(v)
    return out


def _dict_to_list(cols: Dict[str, List]) -> List[Dict[str, Any]]:
    n = max((len(v) for v in cols.values()), default=0)
    return [{k: v[i] for k, v in cols.items()} for i in range(n)]


def merge_dicts(base: Dict[Any, Any], override: Dict[Any, Any]) -> Dict[Any, Any]:
    """Recursive merge; values in ``override`` win; neither input is mutated."""
    out = copy.deepcopy(base)
    for k, v in override.items():
        if isinstance(v, dict) and isinstance(out.get(k), dict):
            out[k] = merge_dicts(out[k], v)
        else:
            out[k] = copy.deepcopy(v)
    return out


def validate_batch_metrics(batch_metrics: List[Dict[str, Any]]) -> None:
    if not batch_metrics:
        return
    keys = set(batch_metrics[0].keys())
    for i, m in enumerate(batch_metrics):
        if set(m.keys()) != keys:
            raise errors.CheckFailedError(
                f"inconsistent training metrics: index: {i}: {sorted(keys)} != {sorted(m.keys())}")


def make_metrics(num_inputs: Optional[int], batch_metrics: List[Dict[str, Any]]) -> Dict[str, Any]:
    """Build ``{"batch_metrics", "avg_metrics"[, "num_inputs"]}``; non-numeric metrics average to None."""
    import numpy as np

    validate_batch_metrics(batch_metrics)
    avg: Dict[str, Optional[float]] = {}
    for name, values in _list_to_dict(batch_metrics).items():
        vals = [v for v in values if v is not None]
        try:
            arr = np.asarray(vals, dtype=np.float64)
            avg[name] = float(arr.mean()) if arr.size and arr.ndim == 1 else (
                float(arr.mean()) if arr.size and all(np.ndim(v) == 0 for v in vals) else None)
        except (TypeError, ValueError):
            avg[name] = None
    out: Dict[str, Any] = {"batch_metrics": batch_metrics, "avg_metrics": avg}
    if num_inputs is not None:
        out["num_inputs"] = num_inputs
    return out


def _json_default(o: Any) -> Any:
    import numpy as np

    if isinstance(o, np.generic):
        return o.item()
    if isinstance(o, np.ndarray):
        return o.tolist()
    if isinstance(o, (datetime.datetime, datetime.date)):
        return o.isoformat()
    if isinstance(o, enum.Enum):
        return o.name
    if isinstance(o, uuid.UUID):
        return str(o)
    if isinstance(o, (set, frozenset)):
        return sorted(o)
    if isinstance(o, pathlib.PurePath):
        return str(o)
    try:
        import torch

        if isinstance(o, torch.Tensor):
            return o.detach().cpu().tolist()
    except ImportError:
        pass
    if hasattr(o, "__json__"):
        return o.__json__()
    raise TypeError(f"{type(o).__name__} is not JSON serializable")


def json_encode(obj: Any, indent: Optional[str] = None, sort_keys: bool = False) -> str:
    def scrub(x: Any) -> Any:
        if isinstance(x, float) and not math.isfinite(x):
            return "Infinity" if x > 0 else ("-Infinity" if x < 0 else "NaN")
        if isinstance(x, dict):
            return {str(k): scrub(v) for k, v in x.items()}
        if isinstance(x, (list, tuple)):
            return [scrub(v) for v in x]
        return x

    return json.dumps(scrub(obj), default=lambda o: scrub(_json_default(o)), indent=indent, sort_keys=sort_keys)


def read_detignore(root: pathlib.Path) -> List[str]:
    p = root / ".detignore"
    if not p.exists():
        return []
    return [ln.strip() for ln in p.read_text().splitlines() if ln.strip() and not ln.startswith("#")]


def make_shutil_ignore(root: pathlib.Path, exclude: Optional[pathlib.Path] = None) -> Callable[[str, List[str]], Set[str]]:
    """``shutil.copytree`` ignore callback honouring ``.detignore`` plus an ``exclude`` path (the
    copy destination, when it lives inside ``root``, must never be copied into itself)."""
    patterns = read_detignore(root) + ["__pycache__", "*.pyc", ".git"]
    excl = os.path.realpath(exclude) if exclude is not None else None

    def ignore(dirpath: str, names: List[str]) -> Set[str]:
        rel = os.path.relpath(dirpath, root)
        skip = set()
        for n in names:
            if excl is not None:
                full = os.path.realpath(os.path.join(dirpath, n))
                if full == excl or excl.startswith(full + os.sep) and os.path.dirname(excl) == full:
                    # the checkpoint's own parent (storage dir) or the checkpoint itself
                    skip.add(n)
                    continue
            relname = n if rel == "." else os.path.join(rel, n)
            for pat in patterns:
                pat_ = pat.rstrip("/")
                if fnmatch.fnmatch(n, pat_) or fnmatch.fnmatch(relname, pat_):
                    skip.add(n)
                    break
        return skip

    return ignore


MAX_CONTEXT_SIZE = 96 * 1024 * 1024


def write_user_code(path: pathlib.Path, on_cluster: bool) -> None:
    """Copy the model definition into ``<ckpt>/code`` so the checkpoint is self-loading."""
    dst = pathlib.Path(path) / "code"
    if dst.exists():
        rmtree_nfs_safe(str(dst))
    managed = pathlib.Path(constants.MANAGED_TRAINING_MODEL_COPY)
    if on_cluster and managed.is_dir():
        shutil.copytree(managed, dst)
    else:
        src = pathlib.Path(os.environ.get("DET_MODEL_DIR", "."))
        ignore = make_shutil_ignore(src, exclude=pathlib.Path(path))
        # Never recurse into the destination, and refuse to copy giant working dirs.
        total = 0
        for dp, dn, fn in os.walk(src):
            skip = ignore(dp, dn + fn)
            dn[:] = [d for d in dn if d not in skip]
            for f in fn:
                if f in skip:
                    continue
                try:
                    total += os.path.getsize(os.path.join(dp, f))
                except OSError:
                    pass
            if total > MAX_CONTEXT_SIZE:
                logger.warning("model directory exceeds %d bytes; storing no code/ copy", MAX_CONTEXT_SIZE)
                dst.mkdir(parents=True, exist_ok=True)
                return
        shutil.copytree(src, dst, ignore=ignore)
    os.chmod(dst, 0o755)


def filter_duplicates(items: Sequence[Any], sorter: Callable[[List[Any]], List[Any]] = sorted) -> Set[Any]:
    seen, dup = set(), set()
    for it in sorter(list(items)):
        (dup if it in seen else seen).add(it)
    return dup


def humanize_float(n: float) -> float:
    if n == 0 or not math.isfinite(n):
        return n
    digits = int(math.ceil(math.log10(abs(n))))
    return round(n, max(digits, 4) - digits)


def make_timing_log(verb: str, duration: float, num_inputs: int, num_batches: int) -> str:
    rps = humanize_float(num_inputs / duration if duration else math.inf)
    bps = humanize_float(num_batches / duration if duration else math.inf)
    return (f"{verb}: {num_inputs} records in {humanize_float(duration)}s ({rps} records/s), "
            f"in {num_batches} batches ({bps} batches/s)")


def calculate_batch_sizes(hparams: Dict[str, Any], slots_per_trial: int, trialname: str) -> Tuple[int, int]:
    slots = max(int(slots_per_trial), 1)
    if "global_batch_size" not in hparams:
        raise errors.InvalidExperimentException(
            f"Please specify an integer `global_batch_size` hyperparameter in your experiment config: "
            f"`global_batch_size` is a required hyperparameter for running experiments with {trialname}-based training.")
    gbs = hparams["global_batch_size"]
    if not isinstance(gbs, int) or isinstance(gbs, bool):
        raise errors.InvalidExperimentException(
            f"`global_batch_size` hyperparameter must be an integer value, not {type(gbs).__name__}. Use int or const types.")
    if gbs < slots:
        raise errors.InvalidExperimentException(
            f"Please set the `global_batch_size` hyperparameter to be greater or equal to the number of slots. "
            f"Current batch_size: {gbs}, slots_per_trial: {slots}.")
    per_slot = gbs // slots
    eff = per_slot * slots
    if eff != gbs:
        logger.warning("`global_batch_size` changed from %d to %d to divide across %d slots", gbs, eff, slots)
    return per_slot, eff


def match_legacy_trial_class(arg: str) -> bool:
    import re

    global _LEGACY_TRIAL_RE
    if _LEGACY_TRIAL_RE is None:
        _LEGACY_TRIAL_RE = re.compile(r"^[a-zA-Z0-9_.]+:[a-zA-Z0-9_]+$")
    return bool(_LEGACY_TRIAL_RE.match(arg))


def legacy_trial_entrypoint_to_script(trial_entrypoint: str) -> List[str]:
    return ["python3", "-m", "determined_b200.exec.harness", trial_entrypoint]


def force_create_symlink(src: str, dst: str) -> None:
    os.makedirs(src, exist_ok=True)
    try:
        os.symlink(src, dst, target_is_directory=True)
    except FileExistsError:
        if os.path.islink(dst) or os.path.isfile(dst):
            os.unlink(dst)
        else:
            shutil.rmtree(dst)
        os.symlink(src, dst, target_is_directory=True)


def signal_process_tree(p: subprocess.Popen, signum: Any) -> None:
    """Signal ``p`` and all its descendants (children first so they cannot be re-parented)."""
    try:
        import psutil

        try:
            kids = psutil.Process(p.pid).children(recursive=True)
        except psutil.NoSuchProcess:
            kids = []
        for k in kids:
            with contextlib.suppress(psutil.NoSuchProcess):
                k.send_signal(signum)
    except ImportError:
        pass
    with contextlib.suppress(ProcessLookupError):
        p.send_signal(signum)


@contextlib.contextmanager
def forward_signals(p: subprocess.Popen, *signums: Any) -> Iterator[None]:
    signums = signums or (signal.SIGTERM, signal.SIGINT)
    old = {}

    def handler(signum: Any, frame: Any) -> None:
        signal_process_tree(p, signum)

    try:
        for s in signums:
            old[s] = signal.signal(s, handler)
        yield
    finally:
        for s, h in old.items():
            signal.signal(s, h)


def is_numerical_scalar(n: Any) -> bool:
    import numpy as np

    if isinstance(n, bool):
        return True
    if isinstance(n, (int, float, np.number)):
        return True
    if isinstance(n, np.ndarray) and n.ndim == 0 and np.issubdtype(n.dtype, np.number):
        return True
    try:
        import torch

        if isinstance(n, torch.Tensor) and n.numel() == 1:
            return True
    except ImportError:
        pass
    return False


_MASK = "********"


def mask_checkpoint_storage(d: Dict) -> Dict:
    out = copy.deepcopy(d)
    for k in ("access_key", "secret_key", "credentials", "connection_string"):
        if k in out and out[k] is not None:
            out[k] = _MASK
    return out


def mask_config_dict(d: Dict) -> Dict:
    out = copy.deepcopy(d)
    if isinstance(out.get("checkpoint_storage"), dict):
        out["checkpoint_storage"] = mask_checkpoint_storage(out["checkpoint_storage"])
    try:
        if out["environment"]["registry_auth"].get("password") is not None:
            out["environment"]["registry_auth"]["password"] = _MASK
    except (KeyError, AttributeError, TypeError):
        pass
    return out


def chunks(seq: Sequence[Any], n: int) -> Iterator[Sequence[Any]]:
    for i in range(0, len(seq), n):
        yield seq[i:i + n]


def safe_load_yaml_with_exceptions(path: Any) -> Any:
    import yaml

    with open(path) as f:
        return yaml.safe_load(f)
