"""Batch-sampler adaptors that implement repeat / rank-sharding / skip-on-resume.

Parity: reference ``harness/determined/pytorch/samplers.py``: ``RepeatBatchSampler`` ``:25-40``,
``DistributedBatchSampler`` (rank-strided batches) ``:74-119``, ``DistributedSampler`` ``:43-71``,
``SkipBatchSampler`` ``:160-188``, ``ReproducibleShuffleSampler``/``…BatchSampler`` ``:191-246``.
These are the *only* data-parallel mechanism the harness itself implements.
"""
import math
from typing import Iterator, List, Optional

import numpy as np
import torch
from torch.utils.data import BatchSampler, Sampler


class RepeatBatchSampler(BatchSampler):
    """Loop over ``batch_sampler`` forever; ``len()`` stays the length of one pass."""

    def __init__(self, batch_sampler: BatchSampler) -> None:
        self.batch_sampler = batch_sampler

    def __len__(self) -> int:
        return len(self.batch_sampler)

    def __iter__(self) -> Iterator[List[int]]:
        while True:
            n = 0
            for b in self.batch_sampler:
                n += 1
                yield b
            if n == 0:
                raise RuntimeError("wrapped batch sampler yielded no batches; cannot repeat forever")


class DistributedSampler(Sampler):
    """Element-strided shard: worker ``r`` of ``n`` gets samples ``r, r+n, r+2n, …``.
    Unlike ``torch.utils.data.DistributedSampler`` nothing is padded or dropped."""

    def __init__(self, sampler: Sampler, num_workers: int, rank: int) -> None:
        self.sampler, self.num_workers, self.rank = sampler, num_workers, rank

    def __len__(self) -> int:
        n = len(self.sampler)  # type: ignore
        return (n - self.rank + self.num_workers - 1) // self.num_workers if n > self.rank else 0

    def __iter__(self) -> Iterator[int]:
        for i, s in enumerate(self.sampler):
            if i % self.num_workers == self.rank:
                yield s


class DistributedBatchSampler(BatchSampler):
    """Batch-strided shard: worker ``r`` of ``n`` gets batches ``r, r+n, …``.  A trailing group
    with fewer than ``n`` batches is still dealt out (low ranks get the extras), so no rank sees a
    batch another rank saw and none is dropped."""

    def __init__(self, batch_sampler: BatchSampler, num_workers: int, rank: int) -> None:
        self.batch_sampler, self.num_workers, self.rank = batch_sampler, num_workers, rank

    def __len__(self) -> int:
        full = len(self.batch_sampler)
        return (full - self.rank + self.num_workers - 1) // self.num_workers if full > self.rank else 0

    def __iter__(self) -> Iterator[List[int]]:
        if self.num_workers == 1:
            yield from self.batch_sampler
            return
        for i, b in enumerate(self.batch_sampler):
            if i % self.num_workers == self.rank:
                yield b


class SkipBatchSampler(BatchSampler):
    """Drop the first ``skip`` batches (resume-from-checkpoint positioning)."""

    def __init__(self, batch_sampler: BatchSampler, skip: int) -> None:
        self.batch_sampler, self.skip = batch_sampler, skip

    def __len__(self) -> int:
        return len(self.batch_sampler)          # the length of an epoch does not change, only where the first one starts

    def __iter__(self) -> Iterator[List[int]]:
        it = iter(self.batch_sampler)
        for _ in range(self.skip):
            try:
                next(it)
            except StopIteration:
                return
        yield from it


class SkipSampler(Sampler):
    """Record-level twin of :class:`SkipBatchSampler`: drops the first ``skip`` indices, keeps the epoch length."""

    def __init__(self, sampler: Sampler, skip: int) -> None:
        self.sampler, self.skip = sampler, skip

    def __len__(self) -> int:
        return len(self.sampler)  # type: ignore[arg-type]

    def __iter__(self) -> Iterator[int]:
        it = iter(self.sampler)
        for _ in range(self.skip):
            try:
                next(it)
            except StopIteration:
                return
        yield from it


class RepeatSampler(Sampler):
    """Cycles through ``sampler`` forever (every pass re-iterates it, so a shuffling sampler reshuffles)."""

    def __init__(self, sampler: Sampler) -> None:
        self.sampler = sampler

    def __len__(self) -> int:
        return len(self.sampler)  # type: ignore[arg-type]

    def __iter__(self) -> Iterator[int]:
        while True:
            n = 0
            for i in self.sampler:
                n += 1
                yield i
            if n == 0:
                return


class ReproducibleShuffleSampler(Sampler):
    """Shuffle with a generator seeded by ``seed`` so epoch ``k`` is the same permutation on
    every restart (each new iteration advances the private RNG exactly once)."""

    def __init__(self, sampler: Sampler, seed: int) -> None:
        self._sampler = sampler
        self._rng = np.random.RandomState(seed)

    def __len__(self) -> int:
        return len(self._sampler)  # type: ignore

    def __iter__(self) -> Iterator[int]:
        idx = list(self._sampler)
        self._rng.shuffle(idx)
        return iter(idx)


class ReproducibleShuffleBatchSampler(BatchSampler):
    """Shuffle the *order of batches* reproducibly."""

    def __init__(self, batch_sampler: BatchSampler, seed: int) -> None:
        self._batch_sampler = batch_sampler
        self._rng = np.random.RandomState(seed)

    def __len__(self) -> int:
        return len(self._batch_sampler)

    def __iter__(self) -> Iterator[List[int]]:
        batches = list(self._batch_sampler)
        order = np.arange(len(batches))
        self._rng.shuffle(order)
        return iter([batches[i] for i in order])
