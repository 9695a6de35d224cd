"""Scalar writer used by the trial controllers (reference ``tensorboard/metric_writers/*``).
Falls back to a no-op when the ``tensorboard`` package is not installed."""
import logging
import pathlib
from typing import Any, Dict, Optional

from determined_b200 import util

logger = logging.getLogger("determined.tensorboard")


class MetricWriter:
    """What a scalar writer offers (reference ``tensorboard/base.py`` ``MetricWriter``)."""

    def add_scalar(self, name: str, value: Any, step: int) -> None:
        raise NotImplementedError

    def reset(self) -> None:
        raise NotImplementedError

    def flush(self) -> None:
        pass


class _NullWriter:
    def add_scalar(self, *a: Any, **k: Any) -> None:
        pass

    def flush(self) -> None:
        pass

    def close(self) -> None:
        pass


class BatchMetricWriter:
    def __init__(self, writer: Any) -> None:
        self.writer = writer

    @classmethod
    def try_build(cls, log_dir: Optional[pathlib.Path] = None) -> "BatchMetricWriter":
        try:
            from torch.utils.tensorboard import SummaryWriter

            return cls(SummaryWriter(log_dir=str(log_dir) if log_dir else None))
        except Exception as e:
            logger.debug("tensorboard writer unavailable (%s)", e)
            return cls(_NullWriter())

    def _write(self, prefix: str, metrics: Dict[str, Any], step: int) -> None:
        for k, v in metrics.items():
            if util.is_numerical_scalar(v):
                self.writer.add_scalar(prefix + k, float(v), step)

    def on_train_step_end(self, steps_completed: int, metrics: Dict[str, Any], batch_metrics: Any = None) -> None:
        self._write("Determined/", metrics, steps_completed)
        self.writer.flush()

    def on_validation_step_end(self, steps_completed: int, metrics: Dict[str, Any]) -> None:
        self._write("Determined/val_", {k[4:] if k.startswith("val_") else k: v for k, v in metrics.items()}, steps_completed)
        self.writer.flush()
