"""Distributed batch inference with checkpointed progress.

Parity: reference ``harness/determined/pytorch/experimental/_torch_batch_process.py``:
``TorchBatchProcessorContext`` ``:31-150``, ``TorchBatchProcessor`` ``:194-230``,
``torch_batch_process(batch_processor_cls, dataset, batch_size, max_batches, checkpoint_interval,
dataloader_kwargs, distributed_context)`` ``:366-541``: the dataset is sharded batch-wise across
workers, every ``checkpoint_interval`` batches all workers synchronise, ``on_checkpoint_start``
runs, and the number of processed batches is persisted so a preempted/restarted task skips what
is already done; custom reducers are reduced across workers at the end and reported as
``inference`` metrics.
"""
import abc
import contextlib
import json
import logging
import math
import os
import pathlib
from typing import Any, ContextManager, Dict, Iterator, Optional, Type

import torch
import torch.distributed as dist
from torch.utils import data

import determined_b200 as det
from determined_b200 import core
from determined_b200.pytorch import _data, _reducer, samplers

logger = logging.getLogger("determined.pytorch")
_PROGRESS_FILE = "metadata.json"


class TorchBatchProcessorContext(_reducer._PyTorchReducerContext):
    def __init__(self, core_context: core.Context, storage_path: str) -> None:
        super().__init__(core_context.distributed.allgather)
        self._core_context = core_context
        self.distributed = core_context.distributed
        self.device = get_default_device(core_context)
        self._storage_path = storage_path
        self._tb_path: Optional[pathlib.Path] = None
        self._use_default_storage = False
        info = det.get_cluster_info()
        self._hparams = info.trial.hparams if info is not None and info.task_type == "TRIAL" else {}

    def get_hparams(self) -> Dict[str, Any]:
        return self._hparams

    def to_device(self, data_: Any, warned_types: Any = None) -> Any:
        return _data.to_device(data_, self.device, warned_types)

    def get_tensorboard_path(self) -> pathlib.Path:
        return self._core_context.train.get_tensorboard_path()

    def prepare_model_for_inference(self, model: torch.nn.Module) -> torch.nn.Module:
        model.eval()
        return model.to(self.device)

    def upload_path(self) -> ContextManager[pathlib.Path]:
        """A directory whose contents are uploaded to checkpoint storage on exit (all workers may
        write; results are merged as a sharded checkpoint)."""
        ctx = self._core_context

        @contextlib.contextmanager
        def cm() -> Iterator[pathlib.Path]:
            with ctx.checkpoint.store_path({"steps_completed": 0, "inference_output": True} if ctx.distributed.rank == 0 else None,
                                           shard=True) as (path, _):
                yield pathlib.Path(path)

        return cm()

    def report_metrics(self, group: str, steps_completed: int, metrics: Dict[str, Any]) -> None:
        if group in ("training", "validation"):
            raise ValueError("cannot report training/validation groups from batch inference; use another group name")
        self._core_context.train.report_metrics(group=group, steps_completed=steps_completed, metrics=metrics)

    def report_task_using_model_version(self, model_version: Any) -> None:
        self._core_context.experimental.report_task_using_model_version(model_version)

    def report_task_using_checkpoint(self, checkpoint: Any) -> None:
        self._core_context.experimental.report_task_using_checkpoint(checkpoint)

    def get_distributed_rank(self) -> int:
        return self.distributed.get_rank()

    def get_distributed_size(self) -> int:
        return self.distributed.get_size()


def get_default_device(core_context: core.Context) -> torch.device:
    if torch.cuda.is_available():
        return torch.device("cuda", core_context.distributed.local_rank % torch.cuda.device_count())
    return torch.device("cpu")


def _initialize_distributed_backend() -> Optional[core.DistributedContext]:
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and "RANK" in os.environ:
        if not dist.is_initialized():
            if torch.cuda.is_available():
                torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count())
            dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
        return core.DistributedContext.from_torch_distributed()
    return None


class TorchBatchProcessor(metaclass=abc.ABCMeta):
    @abc.abstractmethod
    def __init__(self, context: TorchBatchProcessorContext) -> None:
        """Load models, open outputs, wrap reducers (``context.wrap_reducer``)."""

    @abc.abstractmethod
    def process_batch(self, batch: Any, batch_idx: int) -> None:
        """Handle one batch; ``batch_idx`` is this worker's batch counter."""

    def on_checkpoint_start(self) -> None:
        """Called on every worker before progress is recorded: flush outputs here."""

    def on_finish(self) -> None:
        """Called once after the last batch (after the final synchronisation)."""


def _validate_dataloader_kwargs(kw: Dict[str, Any], batch_size: Optional[int]) -> None:
    for bad in ("shuffle", "sampler", "batch_sampler"):
        if kw.get(bad):
            raise ValueError(f"{bad} is not supported: torch_batch_process shards the dataset deterministically itself")
    if "batch_size" in kw and batch_size is not None:
        raise ValueError("batch_size is passed both directly and in dataloader_kwargs")
    if "batch_size" not in kw and batch_size is None:
        logger.info("batch_size not specified: using default batch size 1")


def _validate_iterate_length(iterate_length: Optional[int], times_iterate: int) -> int:
    if iterate_length is None:
        return times_iterate
    if iterate_length <= 0:
        raise ValueError("max_batches should be a positive integer")
    if iterate_length > times_iterate:
        logger.warning("max_batches %d exceeds the %d batches available per worker; using %d", iterate_length, times_iterate, times_iterate)
        return times_iterate
    return iterate_length


def torch_batch_process(batch_processor_cls: Type[TorchBatchProcessor], dataset: data.Dataset, batch_size: Optional[int] = None,
                        max_batches: Optional[int] = None, checkpoint_interval: int = 5,
                        dataloader_kwargs: Optional[Dict[str, Any]] = None,
                        distributed_context: Optional[core.DistributedContext] = None) -> None:
    if checkpoint_interval <= 0:
        raise ValueError("checkpoint_interval should be a positive integer")
    kw = dict(dataloader_kwargs or {})
    _validate_dataloader_kwargs(kw, batch_size)
    bs = int(kw.pop("batch_size", batch_size or 1))
    dctx = distributed_context or _initialize_distributed_backend()
    with core.init(distributed=dctx) as cc:
        rank, size = cc.distributed.rank, cc.distributed.size
        info = det.get_cluster_info()
        latest = info.latest_checkpoint if info is not None else None
        skip = 0
        if latest is not None:
            with cc.checkpoint.restore_path(latest) as p:
                f = pathlib.Path(p) / _PROGRESS_FILE
                if f.exists():
                    skip = int(json.loads(f.read_text()).get("steps_completed", 0))
            logger.info("resuming batch inference after %d batches per worker", skip)
        n_total = math.ceil(len(dataset) / bs)            # type: ignore[arg-type]
        per_worker = len(samplers.DistributedBatchSampler(data.BatchSampler(data.SequentialSampler(dataset), bs, False), size, rank))
        # every worker must take part in every synchronisation: iterate to the max over workers
        times = max(cc.distributed.allgather(per_worker))
        times = _validate_iterate_length(max_batches, times)
        bsampler: Any = data.BatchSampler(data.SequentialSampler(dataset), bs, False)
        bsampler = samplers.DistributedBatchSampler(bsampler, size, rank)
        if skip:
            bsampler = samplers.SkipBatchSampler(bsampler, skip)
        loader = data.DataLoader(dataset, batch_sampler=bsampler, **kw)
        ctx = TorchBatchProcessorContext(cc, str(pathlib.Path.cwd()))
        proc = batch_processor_cls(ctx)
        it = iter(loader)
        last_ckpt = skip
        steps = skip
        for batch_idx in range(skip, times):
            try:
                batch = next(it)
            except StopIteration:
                batch = None
            if batch is not None:
                proc.process_batch(batch=batch, batch_idx=batch_idx)
            steps = batch_idx + 1
            if steps % checkpoint_interval == 0 or steps == times:
                proc.on_checkpoint_start()
                cc.distributed.allgather(None)             # all workers reached this point
                if rank == 0:
                    with cc.checkpoint.store_path({"steps_completed": steps}) as (path, _):
                        (pathlib.Path(path) / "progress.json").write_text(json.dumps({"steps_completed": steps, "workers": size}))
                    try:
                        cc.train.report_metrics(group="inference_progress", steps_completed=steps,
                                                metrics={"progress": steps / max(times, 1)})
                    except Exception:
                        pass
                last_ckpt = steps
                if cc.preempt.should_preempt():
                    return
        reduced = ctx.reduce_metrics(for_training=False)
        if rank == 0 and reduced:
            cc.train.report_metrics(group="inference", steps_completed=steps, metrics=reduced)
        proc.on_finish()
