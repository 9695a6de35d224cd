"""Core API (reference ``harness/determined/core/__init__.py:1-54``)."""
from determined_b200.core._tensorboard_mode import TensorboardMode
from determined_b200.core._distributed import DistributedContext, DummyDistributedContext
from determined_b200.core._checkpoint import (
    CheckpointContext, DummyCheckpointContext, DownloadMode, merge_metadata, merge_resources,
)
from determined_b200.core._train import TrainContext, DummyTrainContext, EarlyExitReason
from determined_b200.core._searcher import (
    SearcherContext, DummySearcherContext, SearcherMode, SearcherOperation, DummySearcherOperation,
    Unit, _parse_searcher_units,
)
from determined_b200.core._preempt import (
    PreemptContext, DummyPreemptContext, PreemptMode, _PreemptionWatcher,
)
from determined_b200.core._profiler import ProfilerContext, DummyProfilerContext, StepTimer
from determined_b200.core._experimental import ExperimentalCoreContext, DummyExperimentalCoreContext
from determined_b200.core._heartbeat import _Heartbeat
from determined_b200.core._log_shipper import _LogShipper
from determined_b200.core._context import Context, init, _dummy_init, _install_stacktrace_on_sigusr1, _run_prepare
