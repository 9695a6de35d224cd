"""``TorchWriter``: a ``torch.utils.tensorboard.SummaryWriter`` that writes into the directory the trial's TensorBoard files are
synced from (reference ``tensorboard/metric_writers/pytorch.py``).

    from determined.tensorboard.metric_writers.pytorch import TorchWriter

    class MyTrial(PyTorchTrial):
        def __init__(self, context):
            self.logger = TorchWriter()
        def train_batch(self, batch, epoch_idx, batch_idx):
            self.logger.writer.add_scalar("my_metric", value, batch_idx)
"""
from typing import Any

from determined_b200 import tensorboard as det_tensorboard
from determined_b200.tensorboard.metric_writers import MetricWriter


class TorchWriter(MetricWriter):
    def __init__(self) -> None:
        from torch.utils import tensorboard

        self.writer: Any = tensorboard.SummaryWriter(log_dir=str(det_tensorboard.get_base_path({})))

    def add_scalar(self, name: str, value: Any, step: int) -> None:
        self.writer.add_scalar(name, value, step)

    def reset(self) -> None:
        self.writer.close()          # flush and close: the next write starts a new event file

    def flush(self) -> None:
        self.writer.flush()
