"""Model definition used by the master end-to-end tests (a no-op-ish trial with tunable metric,
in the spirit of the reference's e2e ``no_op`` fixture, ``e2e_tests/tests/fixtures/no_op``)."""
import os
import random
import sys
from typing import Any, Dict

import torch

import determined_b200 as det
from determined_b200 import pytorch


class DS(torch.utils.data.Dataset):
    def __init__(self, n: int = 64) -> None:
        self.x = torch.arange(1, n + 1, dtype=torch.float32).reshape(n, 1) / n

    def __len__(self) -> int:
        return len(self.x)

    def __getitem__(self, i: int):
        return self.x[i], 2.0 * self.x[i]


class OneVarTrial(pytorch.PyTorchTrial):
    def __init__(self, context: pytorch.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        if hp.get("invalid"):
            raise det.InvalidHP("this sample is invalid")
        self.lr = float(hp.get("learning_rate", 0.05))
        self.crash_on = hp.get("crash_on_batch")
        # reference e2e no_op fixture (e2e_tests/tests/fixtures/no_op/model_def.py:55-63,247-251): every batch, every
        # validation and every checkpoint may kill the process with this probability
        self.chaos = float(hp.get("chaos_probability", 0.0))
        self._rng = random.Random(os.getpid() ^ int.from_bytes(os.urandom(4), "little"))
        model = torch.nn.Linear(1, 1, bias=False)
        model.weight.data.fill_(0.0)
        self.model = context.wrap_model(model)
        self.opt = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=self.lr))

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        if self.crash_on is not None and batch_idx == int(self.crash_on) and not os.path.exists("/tmp/.detb200_crashed_" + str(self.context.get_trial_id())):
            open("/tmp/.detb200_crashed_" + str(self.context.get_trial_id()), "w").close()
            sys.exit(17)
        self._chaos("train_batch")
        x, y = batch
        loss = torch.nn.functional.mse_loss(self.model(x), y)
        self.context.backward(loss)
        self.context.step_optimizer(self.opt)
        return {"loss": loss}

    def _chaos(self, where: str) -> None:
        if self.chaos > 0 and self._rng.random() < self.chaos:
            print(f"CHAOS! injected failure in {where}", flush=True)
            os._exit(101)

    def evaluate_batch(self, batch: Any, batch_idx: int) -> Dict[str, Any]:
        self._chaos("evaluate_batch")
        x, y = batch
        return {"val_loss": torch.nn.functional.mse_loss(self.model(x), y)}

    def build_training_data_loader(self):
        return pytorch.DataLoader(DS(), batch_size=int(self.context.get_hparams().get("batch_size", 8)))

    def build_validation_data_loader(self):
        return pytorch.DataLoader(DS(), batch_size=8)
