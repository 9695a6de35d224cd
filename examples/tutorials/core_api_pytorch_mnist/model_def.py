"""MNIST with plain PyTorch and the Core API: metrics, checkpoints, hyperparameter search and data-parallel training in one
script (reference examples/tutorials/core_api_pytorch_mnist/model_def*.py, one file per stage there).

    python3 model_def.py                                                              # off-cluster, one process
    python3 -m determined_b200.launch.torch_distributed python3 model_def.py          # what distributed.yaml runs

Data: real MNIST from $DET_DATA_DIR/MNIST/raw when present, else a synthetic stand-in of the same shape.
"""
import logging
import os
import pathlib
import sys
from typing import Any, Dict, Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import nn

import determined_b200 as det
from determined_b200 import core


class Net(nn.Module):
    def __init__(self, hp: Dict[str, Any]) -> None:
        super().__init__()
        self.conv1 = nn.Conv2d(1, hp["n_filters1"], 3, 1)
        self.conv2 = nn.Conv2d(hp["n_filters1"], hp["n_filters2"], 3, 1)
        self.drop1, self.drop2 = nn.Dropout(hp["dropout1"]), nn.Dropout(hp["dropout2"])
        self.fc1 = nn.Linear(144 * hp["n_filters2"], 128)
        self.fc2 = nn.Linear(128, 10)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = F.relu(self.conv2(F.relu(self.conv1(x))))
        x = self.drop1(F.max_pool2d(x, 2)).flatten(1)
        return F.log_softmax(self.fc2(self.drop2(F.relu(self.fc1(x)))), dim=1)


def dataset(train: bool) -> torch.utils.data.Dataset:
    sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1] / "mnist_pytorch"))
    try:
        import data          # the tutorial next door has the loader (real files or synthetic)

        return data.get_dataset(train, n_synthetic=int(os.environ.get("MNIST_SYNTHETIC", "4096")))
    except ImportError:
        g = torch.Generator().manual_seed(0 if train else 1)
        y = torch.randint(0, 10, (2048,), generator=g)
        return torch.utils.data.TensorDataset(torch.randn(2048, 1, 28, 28, generator=g), y)


def main(ctx: core.Context, hp: Dict[str, Any], latest_checkpoint: Optional[str], trial_id: int) -> None:
    rank, size = ctx.distributed.rank, ctx.distributed.size
    dev = torch.device(f"cuda:{ctx.distributed.local_rank}" if torch.cuda.is_available() else "cpu")
    torch.manual_seed(1)
    model = Net(hp).to(dev)
    if size > 1:
        model = nn.parallel.DistributedDataParallel(model, device_ids=[dev.index] if dev.type == "cuda" else None)
    opt = torch.optim.Adadelta(model.parameters(), lr=hp["learning_rate"])
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=hp.get("gamma", 0.7))
    epochs_done = 0
    if latest_checkpoint is not None:
        with ctx.checkpoint.restore_path(latest_checkpoint) as path:
            st = torch.load(pathlib.Path(path) / "checkpoint.pt", map_location=dev)
        (model.module if size > 1 else model).load_state_dict(st["model"])
        opt.load_state_dict(st["optimizer"])
        epochs_done = st["epochs"] if st["trial_id"] == trial_id else 0

    per_slot = hp["global_batch_size"] // size
    train_ds, test_ds = dataset(True), dataset(False)
    sampler = torch.utils.data.DistributedSampler(train_ds, size, rank) if size > 1 else None
    train = torch.utils.data.DataLoader(train_ds, batch_size=per_slot, sampler=sampler, shuffle=sampler is None)
    test = torch.utils.data.DataLoader(test_ds, batch_size=1000)
    max_batches = int(os.environ.get("MNIST_MAX_BATCHES", "0"))     # for quick smoke runs

    for op in ctx.searcher.operations():       # op.length is in epochs (searcher.max_length in the yaml files)
        while epochs_done < op.length:
            model.train()
            if sampler is not None:
                sampler.set_epoch(epochs_done)
            for i, (x, y) in enumerate(train):
                x, y = x.to(dev), y.to(dev)
                opt.zero_grad()
                loss = F.nll_loss(model(x), y)
                loss.backward()
                opt.step()
                steps = epochs_done * len(train) + i + 1
                if steps % 50 == 0 and rank == 0:
                    ctx.train.report_training_metrics(steps_completed=steps, metrics={"train_loss": loss.item()})
                if max_batches and i + 1 >= max_batches:
                    break
            sched.step()
            epochs_done += 1
            steps = epochs_done * len(train)
            # validation on the chief (small set); every worker waits for the result through the allgather below
            val = None
            if rank == 0:
                model.eval()
                tot, correct = 0.0, 0
                with torch.no_grad():
                    for x, y in test:
                        out = (model.module if size > 1 else model)(x.to(dev))
                        tot += F.nll_loss(out, y.to(dev), reduction="sum").item()
                        correct += (out.argmax(1).cpu() == y).sum().item()
                val = {"test_loss": tot / len(test_ds), "accuracy": correct / len(test_ds)}
                ctx.train.report_validation_metrics(steps_completed=steps, metrics=val)
                op.report_progress(epochs_done)
                with ctx.checkpoint.store_path({"steps_completed": steps}) as (path, uuid):
                    torch.save({"model": (model.module if size > 1 else model).state_dict(), "optimizer": opt.state_dict(), "epochs": epochs_done,
                                "trial_id": trial_id}, pathlib.Path(path) / "checkpoint.pt")
            val = ctx.distributed.broadcast(val)
            if ctx.preempt.should_preempt():
                return
        if rank == 0:
            op.report_completed(val["test_loss"])


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO, format=det.LOG_FORMAT)
    info = det.get_cluster_info()
    hparams = dict({"learning_rate": 1.0, "global_batch_size": 64, "n_filters1": 32, "n_filters2": 64, "dropout1": 0.25, "dropout2": 0.5},
                   **(info.trial.hparams if info is not None else {}))
    distributed = None
    if "RANK" in os.environ:
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
        distributed = core.DistributedContext.from_torch_distributed()
    with core.init(distributed=distributed) as context:
        main(context, hparams, info.latest_checkpoint if info is not None else None, info.trial.trial_id if info is not None else -1)
    sys.exit(0)
