"""A torchvision classifier as a DeepSpeedTrial whose engine config DSAT can tune (ZeRO stage x micro batch size).

Reference counterpart: examples/deepspeed_autotune/torchvision/deepspeed_trial/model_def.py.  The images are synthetic
(no dataset download in this environment); what DSAT measures is throughput and memory, which do not care.

    python -m determined_b200.pytorch.dsat asha deepspeed.yaml . --max-trials 16 -m throughput
"""
from typing import Any, Dict, Iterator

import torch
import torchvision

from determined_b200 import pytorch
from determined_b200.pytorch import deepspeed as det_ds
from determined_b200.pytorch import dsat


class RandomImages(torch.utils.data.Dataset):
    def __init__(self, n: int = 4096, size: int = 64, classes: int = 100) -> None:
        self.n, self.size, self.classes = n, size, classes

    def __len__(self) -> int:
        return self.n

    def __getitem__(self, i: int):
        g = torch.Generator().manual_seed(i)
        return torch.randn(3, self.size, self.size, generator=g), int(torch.randint(0, self.classes, (1,), generator=g))


class VisionTrial(det_ds.DeepSpeedTrial):
    def __init__(self, context: det_ds.DeepSpeedTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        ds_config = dsat.get_ds_config_from_hparams(hp)
        model = getattr(torchvision.models, hp.get("model_name", "resnet18"))(num_classes=100)
        engine, _, _, _ = det_ds.initialize(model=model, config=ds_config)
        self.engine = context.wrap_model_engine(engine)
        self.loss = torch.nn.CrossEntropyLoss()
        self.image_size = int(hp.get("image_size", 64))
        self.dataset_len = int(hp.get("dataset_len", 4096))

    def _loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(RandomImages(n=self.dataset_len, size=self.image_size), batch_size=self.engine.train_micro_batch_size_per_gpu(), num_workers=0)

    def build_training_data_loader(self) -> pytorch.DataLoader:
        return self._loader()

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        return self._loader()

    def train_batch(self, it: Iterator[Any], epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        x, y = self.context.to_device(next(it))
        loss = self.loss(self.engine(x).float(), y)
        self.engine.backward(loss)
        self.engine.step()
        return {"loss": loss}

    def evaluate_batch(self, it: Iterator[Any], batch_idx: int) -> Dict[str, Any]:
        x, y = self.context.to_device(next(it))
        out = self.engine(x).float()
        return {"val_loss": self.loss(out, y), "accuracy": (out.argmax(-1) == y).float().mean()}
