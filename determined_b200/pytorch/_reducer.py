"""Metric reducers: distributed-correct custom metrics.

Parity: reference ``harness/determined/pytorch/_reducer.py``: ``Reducer`` enum ``:9-25``,
``_simple_reduce_metrics`` ``:28-46``, ``MetricReducer`` ABC ``:49-166``, ``_SimpleReducer``
(with interleaved un-sharding) ``:169-208``, ``_PyTorchReducerContext.wrap_reducer`` /
``reduce_metrics`` (ONE all-gather for all reducers) ``:273-518``.
"""
import abc
import enum
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

import numpy as np


class Reducer(enum.Enum):
    """How per-batch validation metrics combine.  AVG weighs by batch size across slots."""

    AVG = 1
    SUM = 2
    MAX = 3
    MIN = 4


def _simple_reduce_metrics(reducer: Reducer, metrics: np.ndarray, num_batches: Optional[List[int]] = None) -> float:
    metrics = np.asarray(metrics, dtype=np.float64)
    if reducer == Reducer.AVG:
        if num_batches:
            w = np.asarray(num_batches, dtype=np.float64)
            if w.shape != metrics.shape:
                raise ValueError("num_batches must match the shape of metrics")
            return float((metrics * w).sum() / w.sum())
        return float(metrics.mean())
    if reducer == Reducer.SUM:
        return float(metrics.sum())
    if reducer == Reducer.MAX:
        return float(metrics.max())
    if reducer == Reducer.MIN:
        return float(metrics.min())
    raise NotImplementedError(f"unknown reducer {reducer}")


class MetricReducer(metaclass=abc.ABCMeta):
    """Three-step custom metric: (1) user gathers values during the workload, (2)
    :meth:`per_slot_reduce` condenses them on each slot, (3) :meth:`cross_slot_reduce` combines the
    per-slot values (received as a list ordered by rank) into the final metric."""

    @abc.abstractmethod
    def reset(self) -> None:
        ...

    @abc.abstractmethod
    def per_slot_reduce(self) -> Any:
        ...

    @abc.abstractmethod
    def cross_slot_reduce(self, per_slot_metrics: List[Any]) -> Any:
        ...


class _SimpleReducer(MetricReducer):
    """Wraps ``fn(list_of_values) -> metric``; ``update(v)`` appends.  Values from different
    slots are re-interleaved into the original batch order before ``fn`` sees them (slot ``r``
    processed batches ``r, r+n, …``)."""

    def __init__(self, fn: Callable[[List[Any]], Any]) -> None:
        self.fn = fn
        self.updates: List[Any] = []

    def reset(self) -> None:
        self.updates = []

    def update(self, value: Any) -> None:
        self.updates.append(value)

    def per_slot_reduce(self) -> Any:
        return self.updates

    def cross_slot_reduce(self, per_slot_metrics: List[Any]) -> Any:
        flat: List[Any] = []
        longest = max((len(m) for m in per_slot_metrics), default=0)
        for i in range(longest):
            for m in per_slot_metrics:
                if i < len(m):
                    flat.append(m[i])
        return self.fn(flat)


class _WrappedReducer:
    def __init__(self, reducer: MetricReducer, name: Optional[str], for_training: bool, for_validation: bool) -> None:
        self.reducer, self.name = reducer, name
        self.for_training, self.for_validation = for_training, for_validation

    def reset(self) -> None:
        self.reducer.reset()

    def per_slot_reduce(self) -> Any:
        return self.reducer.per_slot_reduce()

    def cross_slot_reduce(self, per_slot: Any) -> Any:
        return self.reducer.cross_slot_reduce(list(per_slot))


def default_allgather_fn(x: Any) -> List[Any]:
    return [x]


class _PyTorchReducerContext:
    def __init__(self, allgather_fn: Callable[[Any], List[Any]] = default_allgather_fn) -> None:
        self._wrapped_reducers: List[_WrappedReducer] = []
        self._allgather_fn = allgather_fn

    def reset_reducers(self) -> None:
        for w in self._wrapped_reducers:
            w.reset()

    def wrap_reducer(self, reducer: Union[Callable, MetricReducer], name: Optional[str] = None,
                     for_training: bool = True, for_validation: bool = True) -> MetricReducer:
        """Register a custom reducer.  A plain callable becomes a ``_SimpleReducer`` (call
        ``.update(value)`` on the returned object); a :class:`MetricReducer` is returned as is.
        ``name=None`` means the reducer returns a dict of several named metrics."""
        if isinstance(reducer, type):
            raise TypeError("wrap_reducer() expects an instance or a function, not a class")
        if isinstance(reducer, MetricReducer):
            obj: MetricReducer = reducer
        elif callable(reducer):
            obj = _SimpleReducer(reducer)
        else:
            raise TypeError("wrap_reducer() expects a MetricReducer or a callable")
        self._wrapped_reducers.append(_WrappedReducer(obj, name, for_training, for_validation))
        return obj

    def run_cross_slot_reduction(self, reducables: List[_WrappedReducer], gathered: List[Any]) -> Dict[str, Any]:
        metrics: Dict[str, Any] = {}
        for w, per_slot in zip(reducables, zip(*gathered)):
            out = w.cross_slot_reduce(per_slot)
            what = (f"function {w.reducer.fn.__name__}()" if isinstance(w.reducer, _SimpleReducer)
                    else f"MetricReducer {type(w.reducer).__name__}")
            if w.name is None:
                if not isinstance(out, dict):
                    raise AssertionError(
                        f"The custom reduction {what} was wrapped with name=None but returned a "
                        f"{type(out).__name__}, not a dict. Set name for a single metric, or return a dict of metrics.")
                metrics.update(out)
            else:
                if isinstance(out, dict):
                    raise AssertionError(
                        f"The custom reduction {what} was wrapped by a call to wrap_reducer() with name set but it "
                        f"returned a dict anyway (name={w.name!r}).  Leave name=None to return a dict of several "
                        "metrics, or return a single value (not a dict) for a single named metric.")
                metrics[w.name] = out
        return metrics

    def reduce_metrics(self, for_training: bool) -> Dict[str, Any]:
        todo = [w for w in self._wrapped_reducers
                if (for_training and w.for_training) or (not for_training and w.for_validation)]
        if not todo:
            return {}
        gathered = self._allgather_fn([w.per_slot_reduce() for w in todo])   # one collective for all
        return self.run_cross_slot_reduction(todo, gathered)
