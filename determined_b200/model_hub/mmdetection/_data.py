"""Data plumbing of the MMDetection adapter that does not need mmdet itself.

Parity: reference ``model_hub/model_hub/mmdetection/_data.py`` (``GroupSampler``: batches of images with
the same aspect-ratio group, sharded over data-parallel ranks; ``DatasetWithIndex``; ``build_dataloader``)
and ``_data_backends.py`` (file backends that read images from S3 / GCS, or hand out one fake image).
"""
import math
from typing import Any, Iterator, List, Sequence

import numpy as np
import torch


class GroupSampler(torch.utils.data.Sampler):
    """Yields indices so that every batch of ``samples_per_gpu`` consecutive indices comes from one group
    (``dataset.flag``: 0 = portrait, 1 = landscape in mmdet), every rank sees the same number of batches, and
    rank ``r`` of ``num_replicas`` takes every ``num_replicas``-th batch.  Groups are padded by repetition."""

    def __init__(self, dataset: Any, samples_per_gpu: int, num_replicas: int = 1, rank: int = 0, seed: int = 0, shuffle: bool = True) -> None:
        flag = getattr(dataset, "flag", None)
        self.flag = np.asarray(flag if flag is not None else np.zeros(len(dataset)), dtype=np.int64)
        self.samples_per_gpu, self.num_replicas, self.rank, self.seed, self.shuffle = samples_per_gpu, num_replicas, rank, seed, shuffle
        self.epoch = 0
        chunk = samples_per_gpu * num_replicas
        self.group_sizes = np.bincount(self.flag)
        self.num_batches = int(sum(math.ceil(n / chunk) for n in self.group_sizes if n))     # per rank
        self.num_samples = self.num_batches * samples_per_gpu

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __iter__(self) -> Iterator[int]:
        rng = np.random.RandomState(self.seed + self.epoch)
        chunk = self.samples_per_gpu * self.num_replicas
        chunks: List[np.ndarray] = []
        for g, n in enumerate(self.group_sizes):
            if n == 0:
                continue
            idx = np.where(self.flag == g)[0]
            if self.shuffle:
                rng.shuffle(idx)
            pad = (-len(idx)) % chunk
            if pad:
                idx = np.concatenate([idx, rng.choice(idx, pad) if self.shuffle else np.resize(idx, pad)])
            chunks.extend(idx.reshape(-1, chunk))
        order = rng.permutation(len(chunks)) if self.shuffle else np.arange(len(chunks))
        for c in order:
            mine = chunks[c][self.rank * self.samples_per_gpu:(self.rank + 1) * self.samples_per_gpu]
            yield from (int(i) for i in mine)

    def __len__(self) -> int:
        return self.num_samples


class DatasetWithIndex(torch.utils.data.Dataset):
    """Adds the sample index to every item (``item["idx"]``) so predictions can be matched to annotations after the
    cross-rank gather of the evaluation reducer; everything else is forwarded to the wrapped dataset."""

    def __init__(self, dataset: Any) -> None:
        self.dataset = dataset

    def __getattr__(self, item: str) -> Any:
        return getattr(self.__dict__["dataset"], item)

    def __getitem__(self, idx: int) -> Any:
        sample = self.dataset[idx]
        if isinstance(sample, dict):
            sample["idx"] = idx
        return sample

    def __len__(self) -> int:
        return len(self.dataset)


def collate(batch: Sequence[Any], samples_per_gpu: int = 1) -> Any:
    """mmcv's collate when available (DataContainer aware), else torch's default."""
    try:
        from mmcv.parallel import collate as mm_collate

        return mm_collate(batch, samples_per_gpu=samples_per_gpu)
    except ImportError:
        return torch.utils.data.default_collate(list(batch))


def build_dataloader(dataset: Any, samples_per_gpu: int, workers_per_gpu: int, num_replicas: int = 1, rank: int = 0, seed: int = 0,
                     shuffle: bool = True, test_mode: bool = False) -> Any:
    """A ``determined_b200.pytorch.DataLoader`` over ``dataset``: group-sampled for training, sequentially sharded (with the
    index attached) for evaluation."""
    from determined_b200 import pytorch

    if test_mode:
        ds = DatasetWithIndex(dataset)
        idx = list(range(rank, len(ds), num_replicas))
        sampler: Any = _Fixed(idx)
        bs = samples_per_gpu
    else:
        ds = dataset
        sampler = GroupSampler(dataset, samples_per_gpu, num_replicas, rank, seed, shuffle)
        bs = samples_per_gpu
    return pytorch.DataLoader(ds, batch_sampler=torch.utils.data.BatchSampler(sampler, bs, drop_last=False), num_workers=workers_per_gpu,
                              collate_fn=lambda b: collate(b, samples_per_gpu), pin_memory=False)


class _Fixed(torch.utils.data.Sampler):
    def __init__(self, idx: List[int]) -> None:
        self.idx = idx

    def __iter__(self) -> Iterator[int]:
        return iter(self.idx)

    def __len__(self) -> int:
        return len(self.idx)
