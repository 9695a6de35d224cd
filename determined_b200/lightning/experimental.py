"""PyTorch-Lightning logger on top of ``core_v2`` (reference ``harness/determined/lightning/experimental.py:1-64``)."""
from typing import Any, Dict, Optional

from determined_b200.experimental import core_v2

try:
    from lightning.pytorch.loggers.logger import Logger as _Base, rank_zero_experiment  # type: ignore
except Exception:  # lightning is optional
    _Base = object

    def rank_zero_experiment(fn):  # type: ignore
        return fn


class DetLogger(_Base):  # type: ignore[misc, valid-type]
    def __init__(self, *, defaults: Optional[core_v2.DefaultConfig] = None, unmanaged: Optional[core_v2.UnmanagedConfig] = None,
                 client: Any = None, **init_kwargs: Any) -> None:
        self._kw = dict(defaults=defaults or core_v2.DefaultConfig(), unmanaged=unmanaged, client=client, **init_kwargs)
        self._initialized = False

    @property
    def name(self) -> str:
        return "determined"

    @property
    def version(self) -> str:
        return "1"

    def _ensure(self) -> None:
        if not self._initialized:
            core_v2.init(**self._kw)
            self._initialized = True

    def log_hyperparams(self, params: Any, *a: Any, **k: Any) -> None:
        d = dict(params) if isinstance(params, dict) else dict(vars(params))
        self._kw["defaults"].hparams = {k_: (v if isinstance(v, (int, float, str, bool, type(None))) else str(v)) for k_, v in d.items()}

    def log_metrics(self, metrics: Dict[str, float], step: Optional[int] = None) -> None:
        self._ensure()
        assert core_v2.train is not None
        core_v2.train.report_metrics(group="training", steps_completed=int(step or 0), metrics=dict(metrics))

    def finalize(self, status: str) -> None:
        if self._initialized:
            core_v2.close()
            self._initialized = False
