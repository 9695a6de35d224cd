"""CIFAR-10 ResNet-18 PyTorchTrial: the configuration `bench.py` measures (BASELINE config #2: bf16 autocast,
aggregation_frequency 2, 256 samples per slot).  Reference: ``examples/computer_vision/cifar10_pytorch/model_def.py``."""
import os
import pathlib
import pickle
from typing import Any, Dict

import numpy as np
import torch
import torch.nn.functional as F

from determined_b200 import pytorch
from determined_b200.models.resnet_cifar import resnet18_cifar as resnet18


def load_cifar(train: bool, n_synthetic: int = 8192) -> torch.utils.data.Dataset:
    root = pathlib.Path(os.environ.get("DET_DATA_DIR", "/data")) / "cifar-10-batches-py"
    files = [root / f"data_batch_{i}" for i in range(1, 6)] if train else [root / "test_batch"]
    if all(f.exists() for f in files):
        xs, ys = [], []
        for f in files:
            d = pickle.load(open(f, "rb"), encoding="bytes")
            xs.append(np.asarray(d[b"data"], dtype=np.uint8).reshape(-1, 3, 32, 32))
            ys += list(d[b"labels"])
        x = torch.from_numpy(np.concatenate(xs)).float().div_(255)
        x = (x - torch.tensor([0.4914, 0.4822, 0.4465]).view(1, 3, 1, 1)) / torch.tensor([0.2470, 0.2435, 0.2616]).view(1, 3, 1, 1)
        y = torch.tensor(ys)
    else:
        g = torch.Generator().manual_seed(0 if train else 1)
        y = torch.randint(0, 10, (n_synthetic,), generator=g)
        x = torch.randn(n_synthetic, 3, 32, 32, generator=g) + y.view(-1, 1, 1, 1).float() * 0.2
    return torch.utils.data.TensorDataset(x, y)


class CIFARTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        context.experimental.use_amp(torch.bfloat16)          # native autocast; no GradScaler needed for bf16
        if hp.get("cuda_graphs", True):
            context.experimental.enable_cuda_graphs()
        self.model = context.wrap_model(resnet18(num_classes=10).to(memory_format=torch.channels_last))
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=hp["learning_rate"], momentum=0.9,
                                                                weight_decay=hp.get("weight_decay", 5e-4), nesterov=True))
        self.scheduler = context.wrap_lr_scheduler(torch.optim.lr_scheduler.OneCycleLR(self.optimizer, max_lr=hp["learning_rate"], total_steps=hp.get("total_steps", 4000)),
                                                   pytorch.LRScheduler.StepMode.STEP_EVERY_OPTIMIZER_STEP)

    def build_training_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(load_cifar(True), batch_size=self.context.get_per_slot_batch_size(), shuffle=True, drop_last=True, num_workers=2, pin_memory=True)

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(load_cifar(False), batch_size=self.context.get_per_slot_batch_size())

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        x, y = batch
        loss = F.cross_entropy(self.model(x.contiguous(memory_format=torch.channels_last)), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss}

    def evaluate_batch(self, batch: Any, batch_idx: int) -> Dict[str, Any]:
        x, y = batch
        out = self.model(x.contiguous(memory_format=torch.channels_last))
        return {"validation_loss": F.cross_entropy(out, y), "accuracy": (out.argmax(1) == y).float().mean()}
