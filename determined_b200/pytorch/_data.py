"""``det.pytorch.DataLoader`` — a deferred ``torch.utils.data.DataLoader`` whose batch sampler the
harness rewrites for repeat / shard / skip.

Parity: reference ``harness/determined/pytorch/_data.py``: ``DataLoader`` ``:52-279``,
``adapt_batch_sampler`` (Repeat → Distributed → Skip) ``:282-307``, ``data_length`` ``:310-341``,
``to_device`` ``:344-378``.

B200 addition: :class:`DevicePrefetcher` — pinned staging + a copy stream so the H2D of batch
``i+1`` overlaps the compute of batch ``i``.
"""
import logging
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Set, Type, Union, cast

import numpy as np
import torch
from torch.utils.data import BatchSampler, Dataset, IterableDataset, RandomSampler, Sampler, SequentialSampler
from torch.utils.data import DataLoader as _TorchDataLoader

from determined_b200.pytorch import samplers

logger = logging.getLogger("determined.pytorch")

_Array = Union[np.ndarray, torch.Tensor]
_Data = Union[Dict[str, Any], Sequence[Any], np.ndarray, torch.Tensor]
TorchData = Union[Dict[str, torch.Tensor], Sequence[torch.Tensor], torch.Tensor]


class DataLoader:
    """Same constructor as ``torch.utils.data.DataLoader``; the real loader is only built by
    :meth:`get_data_loader`, once the harness knows rank/size/skip."""

    def __init__(self, dataset: Dataset, batch_size: Optional[int] = 1, shuffle: bool = False,
                 sampler: Optional[Sampler] = None, batch_sampler: Optional[BatchSampler] = None,
                 num_workers: int = 0, collate_fn: Optional[Callable] = None, pin_memory: bool = False,
                 drop_last: bool = False, timeout: float = 0, worker_init_fn: Optional[Callable] = None,
                 multiprocessing_context: Any = None, generator: Any = None, *,
                 prefetch_factor: Optional[int] = None, persistent_workers: bool = False) -> None:
        if isinstance(dataset, IterableDataset):
            raise ValueError("det.pytorch.DataLoader does not support IterableDataset; "
                             "build a torch DataLoader yourself and shard it by context.distributed rank")
        if num_workers < 0:
            raise ValueError("num_workers option should be non-negative")
        if timeout < 0:
            raise ValueError("timeout option should be non-negative")
        if batch_sampler is not None:
            if batch_size not in (1, None) or shuffle or sampler is not None or drop_last:
                raise ValueError("batch_sampler option is mutually exclusive with batch_size, shuffle, sampler, and drop_last")
            batch_size, drop_last = None, False
        elif batch_size is None:
            raise ValueError("det.pytorch.DataLoader requires automatic batching: give batch_size or batch_sampler")
        if sampler is not None and shuffle:
            raise ValueError("sampler option is mutually exclusive with shuffle")
        if batch_sampler is None:
            if sampler is None:
                sampler = RandomSampler(dataset, generator=generator) if shuffle else SequentialSampler(dataset)  # type: ignore
            batch_sampler = BatchSampler(sampler, cast(int, batch_size), drop_last)
        self.dataset = dataset
        self.batch_sampler = batch_sampler
        self.num_workers = num_workers
        self.collate_fn = collate_fn
        self.pin_memory = pin_memory
        self.timeout = timeout
        self.worker_init_fn = worker_init_fn
        self.multiprocessing_context = multiprocessing_context
        self.generator = generator
        self.prefetch_factor = prefetch_factor
        self.persistent_workers = persistent_workers

    def get_data_loader(self, repeat: bool = False, skip: int = 0, num_replicas: int = 1, rank: int = 0) -> _TorchDataLoader:
        bs = adapt_batch_sampler(self.batch_sampler, repeat=repeat, skip=skip, num_replicas=num_replicas, rank=rank)
        kw: Dict[str, Any] = {}
        if self.num_workers > 0:
            if self.prefetch_factor is not None:
                kw["prefetch_factor"] = self.prefetch_factor
            kw["persistent_workers"] = self.persistent_workers
        return _TorchDataLoader(self.dataset, batch_sampler=bs, num_workers=self.num_workers,
                                collate_fn=self.collate_fn, pin_memory=self.pin_memory, timeout=self.timeout,
                                worker_init_fn=self.worker_init_fn,
                                multiprocessing_context=self.multiprocessing_context,
                                generator=self.generator, **kw)

    def __iter__(self) -> Iterator:
        # Iterating the deferred loader directly is a footgun in the reference (it raises);
        # local experimentation is convenient, so we build the plain loader.
        return iter(self.get_data_loader())

    def __len__(self) -> int:
        return len(self.batch_sampler)


def adapt_batch_sampler(batch_sampler: BatchSampler, repeat: bool = False, skip: int = 0,
                        num_replicas: int = 1, rank: int = 0) -> BatchSampler:
    """Order matters: repeat first (so every rank can always draw), then shard, then skip."""
    if repeat:
        batch_sampler = samplers.RepeatBatchSampler(batch_sampler)
    if num_replicas > 1:
        batch_sampler = samplers.DistributedBatchSampler(batch_sampler, num_replicas, rank)
    if skip > 0:
        batch_sampler = samplers.SkipBatchSampler(batch_sampler, skip)
    return batch_sampler


def data_length(data: _Data) -> int:
    """Number of records in a batch: first dimension of the first array found."""
    if isinstance(data, (np.ndarray, torch.Tensor)):
        return int(data.shape[0])
    if isinstance(data, dict):
        if not data:
            raise ValueError("cannot take the length of an empty dict batch")
        return data_length(next(iter(data.values())))
    if isinstance(data, (list, tuple)):
        if not data:
            raise ValueError("cannot take the length of an empty batch")
        return data_length(data[0])
    raise TypeError(f"data of type {type(data).__name__} has no defined batch length; "
                    "override PyTorchTrial.get_batch_length()")


def to_device(data: _Data, device: torch.device, warned_types: Optional[Set[Type]] = None) -> TorchData:
    """Recursively move tensors (and numpy arrays) to ``device``; other leaves pass through."""
    if isinstance(data, torch.Tensor):
        return data.to(device, non_blocking=True)
    if isinstance(data, np.ndarray):
        return torch.from_numpy(data).to(device, non_blocking=True)
    if isinstance(data, dict):
        return {k: to_device(v, device, warned_types) for k, v in data.items()}  # type: ignore
    if isinstance(data, tuple) and hasattr(data, "_fields"):
        return type(data)(*(to_device(v, device, warned_types) for v in data))  # namedtuple
    if isinstance(data, (list, tuple)):
        return type(data)(to_device(v, device, warned_types) for v in data)  # type: ignore
    if warned_types is not None and type(data) not in warned_types:
        warned_types.add(type(data))
        logger.warning("to_device(): not able to move data of type %s to the device; it stays on the host", type(data).__name__)
    return data  # type: ignore


class DevicePrefetcher:
    """Wrap a host-batch iterator; yields device batches whose H2D copy ran on a side stream.

    Host tensors are staged through pinned memory (unless already pinned) so the copy is truly
    asynchronous; the consumer's stream waits on the copy event, never the host."""

    def __init__(self, it: Iterator, device: torch.device, depth: int = 2) -> None:
        self._it = iter(it)
        self._device = device
        self._cuda = device.type == "cuda"
        self._stream = torch.cuda.Stream(device=device) if self._cuda else None
        self._queue: List[Any] = []
        self._depth = max(1, depth)
        self._done = False

    def _pin(self, x: Any) -> Any:
        if isinstance(x, np.ndarray):
            x = torch.from_numpy(x)
        if isinstance(x, torch.Tensor):
            return x if (x.is_cuda or x.is_pinned()) else x.pin_memory()
        if isinstance(x, dict):
            return {k: self._pin(v) for k, v in x.items()}
        if isinstance(x, (list, tuple)) and not hasattr(x, "_fields"):
            return type(x)(self._pin(v) for v in x)
        return x

    def _fill(self) -> None:
        while not self._done and len(self._queue) < self._depth:
            try:
                host = next(self._it)
            except StopIteration:
                self._done = True
                return
            if not self._cuda:
                self._queue.append((host, None, host))
                continue
            host = self._pin(host)
            with torch.cuda.stream(self._stream):
                dev = to_device(host, self._device)
                ev = torch.cuda.Event()
                ev.record(self._stream)
            self._queue.append((dev, ev, host))   # keep host alive until the copy is consumed

    def __iter__(self) -> "DevicePrefetcher":
        return self

    def __next__(self) -> Any:
        self._fill()
        if not self._queue:
            raise StopIteration
        dev, ev, _host = self._queue.pop(0)
        if ev is not None:
            torch.cuda.current_stream(self._device).wait_event(ev)
            _record_stream(dev, torch.cuda.current_stream(self._device))
        self._fill()
        return dev


def _record_stream(x: Any, stream: Any) -> None:
    if isinstance(x, torch.Tensor) and x.is_cuda:
        x.record_stream(stream)
    elif isinstance(x, dict):
        for v in x.values():
            _record_stream(v, stream)
    elif isinstance(x, (list, tuple)):
        for v in x:
            _record_stream(v, stream)
