"""MNIST PyTorchTrial (reference ``examples/tutorials/mnist_pytorch/train.py``: MNistTrial)."""
from typing import Any, Dict

import data
import torch
import torch.nn.functional as F

from determined_b200 import pytorch
from determined_b200.models.mnist_cnn import MNISTNet


class MNistTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext, hparams: Dict[str, Any] = None) -> None:
        self.context = context
        hp = hparams or context.get_hparams()
        self.batch_size = context.get_per_slot_batch_size()
        self.model = context.wrap_model(MNISTNet(hp["n_filters1"], hp["n_filters2"], hp["dropout1"], hp["dropout2"]))
        self.optimizer = context.wrap_optimizer(torch.optim.Adadelta(self.model.parameters(), lr=hp["learning_rate"]))

    def build_training_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(data.get_dataset(train=True), batch_size=self.batch_size, shuffle=True)

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(data.get_dataset(train=False), batch_size=self.batch_size)

    def train_batch(self, batch: pytorch.TorchData, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        x, y = batch
        loss = F.nll_loss(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss}

    def evaluate_batch(self, batch: pytorch.TorchData, batch_idx: int) -> Dict[str, Any]:
        x, y = batch
        out = self.model(x)
        return {"validation_loss": F.nll_loss(out, y), "accuracy": (out.argmax(1) == y).float().mean()}
