"""``determined.pytorch`` (reference ``harness/determined/pytorch/__init__.py:1-40``)."""
from determined_b200.pytorch._data import (
    DataLoader, DevicePrefetcher, TorchData, _Data, adapt_batch_sampler, data_length, to_device,
)
from determined_b200.pytorch import samplers
from determined_b200.pytorch._callback import PyTorchCallback
from determined_b200.pytorch._lr_scheduler import LRScheduler
from determined_b200.pytorch._reducer import (
    MetricReducer, Reducer, _PyTorchReducerContext, _SimpleReducer, _simple_reduce_metrics,
)
from determined_b200.pytorch._metric_utils import (
    _combine_and_average_training_metrics, _convert_metrics_to_numpy, _log_tb_metrics,
    _prepare_metrics_reducers, _reduce_metrics, materialize_metrics,
)
from determined_b200.pytorch._experimental import PyTorchExperimentalContext
from determined_b200.pytorch._pytorch_context import PyTorchTrialContext
from determined_b200.pytorch._pytorch_trial import (
    Batch, Epoch, PyTorchTrial, TrainUnit, _PyTorchTrialController, _TrialState,
)
from determined_b200.pytorch._trainer import Trainer, init
from determined_b200.pytorch._load import CheckpointLoadContext, load_trial_from_checkpoint_path
from determined_b200.parallel.ddp import FusedClipGradNorm
