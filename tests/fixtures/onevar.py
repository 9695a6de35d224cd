"""One-variable linear model with a closed-form SGD trajectory — the numerical oracle for the
training loop (idea from the reference's ``harness/tests/experiment/fixtures/pytorch_onevar_model.py``:
y = w*x fit to y = 2x with MSE, so one SGD step is ``w' = w + 2*lr*x*(y - w*x)`` per sample mean)."""
from typing import Any, Dict, List, Tuple

import numpy as np
import torch

from determined_b200 import pytorch


class OnesDataset(torch.utils.data.Dataset):
    """x = i+1 (scaled), y = 2x; deterministic."""

    def __init__(self, n: int = 64) -> None:
        self.x = torch.arange(1, n + 1, dtype=torch.float32).reshape(n, 1) / n
        self.y = 2.0 * self.x

    def __len__(self) -> int:
        return len(self.x)

    def __getitem__(self, i: int) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.x[i], self.y[i]


def expected_weights(w0: float, lr: float, batches: List[Tuple[np.ndarray, np.ndarray]], agg: int = 1,
                     average: bool = True) -> List[float]:
    """Replay plain SGD on MSE analytically, with gradient aggregation over ``agg`` batches."""
    w, acc, out = w0, 0.0, []
    for i, (x, y) in enumerate(batches):
        grad = float(np.mean(2.0 * x * (w * x - y)))
        acc += grad
        if (i + 1) % agg == 0:
            w -= lr * (acc / agg if average else acc)
            acc = 0.0
        out.append(w)
    return out


class OneVarTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.lr = float(hp.get("learning_rate", 0.05))
        self.bs = int(hp.get("batch_size", 4))
        self.n = int(hp.get("dataset_len", 64))
        model = torch.nn.Linear(1, 1, bias=False)
        model.weight.data.fill_(0.0)
        self.model = context.wrap_model(model)
        opt_cls = {"sgd": torch.optim.SGD, "adam": torch.optim.Adam, "adamw": torch.optim.AdamW}[hp.get("optimizer", "sgd")]
        self.opt = context.wrap_optimizer(opt_cls(self.model.parameters(), lr=self.lr))
        self.clip = hp.get("clip")
        self.weights: List[float] = []

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x), y)
        self.context.backward(loss)
        clip = None
        if self.clip is not None:
            clip = lambda ps: torch.nn.utils.clip_grad_norm_(ps, self.clip)  # noqa: E731
        self.context.step_optimizer(self.opt, clip_grads=clip)
        self.weights.append(float(self.model.weight.detach().cpu().item()))
        return {"loss": loss, "w": self.model.weight.detach().reshape(())}

    def evaluate_batch(self, batch: Any, batch_idx: int) -> Dict[str, Any]:
        x, y = batch
        return {"val_loss": torch.nn.functional.mse_loss(self.model(x), y)}

    def build_training_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(OnesDataset(self.n), batch_size=self.bs)

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        return pytorch.DataLoader(OnesDataset(self.n), batch_size=self.bs)
