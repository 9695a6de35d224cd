"""CIFAR-10 ResNet-18 PyTorchTrial used by ``bench.py`` for BOTH arms (BASELINE config #2).

The reference ships no CIFAR model, so the trial is written once against the
``determined.pytorch`` API surface and instantiated with either implementation of that API:
``make_trial(pytorch_module)``.  Nothing in here touches this repo's engine directly — whatever
happens underneath ``wrap_model / wrap_optimizer / backward / step_optimizer`` is the harness'
business, which is exactly what is being compared.

Data is synthetic (no network on the box): a pool of pinned host batches (or device batches for
the kernel-only number) of CIFAR shape, cycled forever; the pool is larger than L2.
"""
import importlib
from typing import Any, Dict, Iterator, List, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


def _resnet18() -> nn.Module:
    # plain torch.nn definition, shared by both arms
    mod = importlib.import_module("determined_b200.models.resnet_cifar")
    return mod.resnet18_cifar(10)


class SyntheticCIFAR:
    """Endless iterable of (images, labels).  ``on_device``: batches already live on the GPU."""

    def __init__(self, batch_size: int, pool: int, device: Any = None, seed: int = 0) -> None:
        g = torch.Generator().manual_seed(seed)
        self.batches: List[Tuple[torch.Tensor, torch.Tensor]] = []
        for _ in range(pool):
            x = torch.randn(batch_size, 3, 32, 32, generator=g)
            y = torch.randint(0, 10, (batch_size,), generator=g)
            if device is not None:
                x, y = x.to(device), y.to(device)
            elif torch.cuda.is_available():
                x, y = x.pin_memory(), y.pin_memory()
            self.batches.append((x, y))
        self.batch_size = batch_size

    def __len__(self) -> int:
        return len(self.batches)

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        while True:
            for b in self.batches:
                yield b


def make_trial(pytorch: Any, *, per_slot_batch: int, pool: int, on_device: bool, use_b200_extras: bool,
               timing: Dict[str, Any]) -> Any:
    """Build the trial class against ``pytorch`` (= determined_b200.pytorch or the reference's
    determined.pytorch).  ``timing`` receives CUDA events / counters from train_batch so the bench
    can bracket exactly K steps without modifying either harness."""

    class CIFARTrial(pytorch.PyTorchTrial):  # type: ignore[misc, name-defined]
        def __init__(self, context: Any) -> None:
            self.context = context
            if use_b200_extras:
                context.experimental.enable_cuda_graphs(True)
            context.experimental.disable_dataset_reproducibility_checks()
            model = _resnet18().to(memory_format=torch.channels_last)
            self.model = context.wrap_model(model)
            self.opt = context.wrap_optimizer(
                torch.optim.SGD(self.model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4, nesterov=True))
            self.device = context.device

        def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
            hook = timing.get("hook")
            if hook is not None:
                hook(batch_idx)           # reference arm: brackets K steps (eager, not captured)
            x, y = batch
            x = x.contiguous(memory_format=torch.channels_last)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = self.model(x)
            loss = F.cross_entropy(logits.float(), y)
            self.context.backward(loss)
            self.context.step_optimizer(self.opt)
            return {"loss": loss}

        def evaluate_batch(self, batch: Any, batch_idx: int) -> Dict[str, Any]:
            x, y = batch
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = self.model(x.contiguous(memory_format=torch.channels_last))
            return {"validation_loss": F.cross_entropy(logits.float(), y),
                    "accuracy": (logits.argmax(1) == y).float().mean()}

        def build_training_data_loader(self) -> Any:
            return _Loader(SyntheticCIFAR(per_slot_batch, pool, self.device if on_device else None,
                                          seed=1234 + self.context.distributed.rank))

        def build_validation_data_loader(self) -> Any:
            # a real torch DataLoader (the reference asserts the type) over two pre-built batches
            src = SyntheticCIFAR(per_slot_batch, 2, self.device if on_device else None, seed=99)
            return torch.utils.data.DataLoader(src.batches, batch_size=None, shuffle=False)

        def build_callbacks(self) -> Dict[str, Any]:
            cb = timing.get("callback")
            return {"bench": cb} if cb is not None else {}

    return CIFARTrial


class _Loader:
    """Looks enough like a DataLoader for both harnesses (len + iter)."""

    def __init__(self, src: SyntheticCIFAR) -> None:
        self.src = src

    def __len__(self) -> int:
        return len(self.src)

    def __iter__(self) -> Iterator:
        return iter(self.src)


class _FiniteLoader(_Loader):
    def __iter__(self) -> Iterator:
        return iter(self.src.batches)
