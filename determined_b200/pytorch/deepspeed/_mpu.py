"""``ModelParallelUnit`` — what the harness needs to know when not every rank is data-parallel.

Parity: reference ``harness/determined/pytorch/deepspeed/_mpu.py:8-50``: ``data_parallel_rank``,
``data_parallel_world_size`` (loader sharding + metric scaling), ``should_report_metrics``,
``should_build_data_loader``; ``make_data_parallel_mpu`` and ``make_deepspeed_mpu`` (pipeline
grids: only first/last stage build loaders, only the last stage reports).  Extended with a
context-parallel dimension so loader sharding stays correct if sequence parallelism
is added."""
from typing import Any

from determined_b200 import core


class ModelParallelUnit:
    def __init__(self, data_parallel_rank: int, data_parallel_world_size: int, should_report_metrics: bool,
                 should_build_data_loader: bool, context_parallel_rank: int = 0, context_parallel_world_size: int = 1) -> None:
        self.data_parallel_rank = data_parallel_rank
        self.data_parallel_world_size = data_parallel_world_size
        self.should_report_metrics = should_report_metrics
        self.should_build_data_loader = should_build_data_loader
        self.context_parallel_rank = context_parallel_rank
        self.context_parallel_world_size = context_parallel_world_size

    # DeepSpeed-style accessors so the unit can be passed as ``mpu=`` to the engine
    def get_data_parallel_rank(self) -> int:
        return self.data_parallel_rank

    def get_data_parallel_world_size(self) -> int:
        return self.data_parallel_world_size


def make_data_parallel_mpu(dist_context: core.DistributedContext) -> ModelParallelUnit:
    return ModelParallelUnit(dist_context.get_rank(), dist_context.get_size(), True, True)


def make_deepspeed_mpu(topology: Any) -> ModelParallelUnit:
    """``topology`` is a pipeline grid exposing ``get_data_parallel_rank/world_size``,
    ``get_pipe_parallel_rank/world_size`` and ``get_slice_parallel_rank`` (DeepSpeed's
    ``PipelineParallelGrid`` or :class:`determined_b200.parallel.pipeline.PipelineGrid`)."""
    pipe_rank, pipe_world = topology.get_pipe_parallel_rank(), topology.get_pipe_parallel_world_size()
    first, last = pipe_rank == 0, pipe_rank == pipe_world - 1
    slice0 = getattr(topology, "get_slice_parallel_rank", lambda: 0)() == 0
    return ModelParallelUnit(topology.get_data_parallel_rank(), topology.get_data_parallel_world_size(),
                             should_report_metrics=last and slice0, should_build_data_loader=(first or last) and slice0)
