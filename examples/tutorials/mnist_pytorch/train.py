"""Trainer-API entry point: ``python3 train.py`` locally, or as the experiment ``entrypoint`` on-cluster
(reference ``examples/tutorials/mnist_pytorch/train.py:96-142``)."""
import logging

import determined_b200 as det
from determined_b200 import pytorch
from model_def import MNistTrial

HPARAMS = {"n_filters1": 32, "n_filters2": 64, "dropout1": 0.25, "dropout2": 0.5, "learning_rate": 1.0, "global_batch_size": 64}


def main() -> None:
    info = det.get_cluster_info()
    hparams = info.trial.hparams if info is not None else HPARAMS
    with pytorch.init(hparams=None if info is not None else HPARAMS) as ctx:
        trial = MNistTrial(ctx, hparams)
        trainer = pytorch.Trainer(trial, ctx)
        if info is None:
            trainer.fit(max_length=pytorch.Batch(200), validation_period=pytorch.Batch(100), reporting_period=pytorch.Batch(50))
        else:
            trainer.fit(latest_checkpoint=info.latest_checkpoint)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO, format=det.LOG_FORMAT if hasattr(det, "LOG_FORMAT") else None)
    main()
