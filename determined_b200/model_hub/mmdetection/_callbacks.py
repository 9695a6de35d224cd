"""mmdet-style learning-rate schedules as a PyTorchCallback.

Parity: reference ``model_hub/model_hub/mmdetection/_callbacks.py`` drives mmcv's ``LrUpdaterHook`` through a
fake runner.  The schedules themselves are small closed forms, so they are implemented here directly (policies
``fixed`` / ``step`` / ``CosineAnnealing`` / ``poly`` / ``exp``, each with ``constant`` / ``linear`` / ``exp`` warm-up,
by iteration or by epoch) and need no mmcv.
"""
import math
from typing import Any, Callable, Dict, List, Optional, Sequence

from determined_b200 import pytorch


def build_lr_schedule(lr_config: Dict[str, Any], max_epochs: int, iters_per_epoch: int) -> Callable[[int, int, float], float]:
    """``f(iteration, epoch, base_lr) -> lr`` for an mmcv ``lr_config`` dict."""
    cfg = dict(lr_config)
    policy = str(cfg.pop("policy", "fixed")).lower()
    by_epoch = bool(cfg.pop("by_epoch", True))
    warmup = cfg.pop("warmup", None)
    warmup_iters = int(cfg.pop("warmup_iters", 0))
    warmup_ratio = float(cfg.pop("warmup_ratio", 0.1))
    warmup_by_epoch = bool(cfg.pop("warmup_by_epoch", False))
    if warmup is not None and warmup not in ("constant", "linear", "exp"):
        raise ValueError(f"unknown warmup {warmup!r}")
    if warmup_by_epoch:
        warmup_iters *= iters_per_epoch
    total = max_epochs if by_epoch else max_epochs * iters_per_epoch

    def regular(progress: int, base: float) -> float:
        if policy == "fixed":
            return base
        if policy == "step":
            step, gamma, min_lr = cfg.get("step", []), float(cfg.get("gamma", 0.1)), cfg.get("min_lr")
            n = progress // step if isinstance(step, int) else sum(1 for s in step if progress >= s)
            lr = base * gamma ** n
            return max(lr, float(min_lr)) if min_lr is not None else lr
        if policy in ("cosineannealing", "cosine"):
            target = float(cfg["min_lr"]) if cfg.get("min_lr") is not None else base * float(cfg.get("min_lr_ratio", 0.0))
            return target + 0.5 * (base - target) * (1 + math.cos(math.pi * min(progress / max(total, 1), 1.0)))
        if policy == "poly":
            power, min_lr = float(cfg.get("power", 1.0)), float(cfg.get("min_lr", 0.0))
            return (base - min_lr) * (1 - min(progress / max(total, 1), 1.0)) ** power + min_lr
        if policy == "exp":
            return base * float(cfg.get("gamma", 0.1)) ** progress
        raise ValueError(f"unsupported lr policy {policy!r}")

    def f(iteration: int, epoch: int, base: float) -> float:
        lr = regular(epoch if by_epoch else iteration, base)
        if warmup is not None and iteration < warmup_iters:
            if warmup == "constant":
                return lr * warmup_ratio
            if warmup == "linear":
                return lr * (1 - (1 - iteration / warmup_iters) * (1 - warmup_ratio))
            return lr * warmup_ratio ** (1 - iteration / warmup_iters)
        return lr

    return f


class LrUpdaterCallback(pytorch.PyTorchCallback):
    """Sets every param group's ``lr`` before each training batch from an mmcv ``lr_config``."""

    def __init__(self, context: Any, lr_config: Dict[str, Any], max_epochs: int, iters_per_epoch: int) -> None:
        self.context = context
        self.schedule = build_lr_schedule(lr_config, max_epochs, iters_per_epoch)
        self.iters_per_epoch = max(int(iters_per_epoch), 1)
        self.base_lrs: Optional[List[List[float]]] = None
        self.iteration = 0

    def _optimizers(self) -> Sequence[Any]:
        return list(getattr(self.context, "optimizers", []))

    def on_training_start(self) -> None:
        self.base_lrs = [[float(g.setdefault("initial_lr", g["lr"])) for g in o.param_groups] for o in self._optimizers()]

    def on_training_batch_start(self, batch_idx: Optional[int] = None) -> None:
        if self.base_lrs is None:
            self.on_training_start()
        it = int(batch_idx) if batch_idx is not None else self.iteration
        epoch = it // self.iters_per_epoch
        for o, bases in zip(self._optimizers(), self.base_lrs or []):
            for g, b in zip(o.param_groups, bases):
                g["lr"] = self.schedule(it, epoch, b)
        self.iteration = it + 1

    def state_dict(self) -> Dict[str, Any]:
        return {"iteration": self.iteration, "base_lrs": self.base_lrs}

    def load_state_dict(self, state: Dict[str, Any]) -> None:
        self.iteration, self.base_lrs = int(state.get("iteration", 0)), state.get("base_lrs")
