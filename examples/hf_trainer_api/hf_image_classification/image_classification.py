"""Image classification with the HF Trainer and a ViT, reporting to the cluster through DetCallback
(reference examples/hf_trainer_api/hf_image_classification/image_classification.py).

The image has no network: without ``--model_name_or_path`` (a local snapshot) a small ViT is initialised from a config, and
without ``--train_dir`` (an image-folder dataset) the data is synthetic class-dependent noise of the right shape.
"""
import argparse
import logging

import torch

import determined_b200 as det
from determined_b200 import core
from determined_b200.transformers import DetCallback


class Blobs(torch.utils.data.Dataset):
    """Synthetic images: each class has a bright patch at its own position, so accuracy can rise above chance."""

    def __init__(self, n: int, classes: int, size: int, seed: int) -> None:
        g = torch.Generator().manual_seed(seed)
        self.y = torch.randint(0, classes, (n,), generator=g)
        self.x = torch.randn(n, 3, size, size, generator=g) * 0.2
        step = max(size // classes, 1)
        for c in range(classes):
            self.x[self.y == c, :, c * step:(c + 1) * step, c * step:(c + 1) * step] += 2.0

    def __len__(self) -> int:
        return len(self.y)

    def __getitem__(self, i: int) -> dict:
        return {"pixel_values": self.x[i], "labels": self.y[i]}


def accuracy(eval_pred) -> dict:  # noqa: ANN001
    logits, labels = eval_pred
    return {"accuracy": float((logits.argmax(-1) == labels).mean())}


def main(core_context: core.Context, a: argparse.Namespace) -> None:
    import transformers

    if a.model_name_or_path:
        model = transformers.AutoModelForImageClassification.from_pretrained(a.model_name_or_path, num_labels=a.num_labels, ignore_mismatched_sizes=True)
        size = model.config.image_size
    else:
        cfg = transformers.ViTConfig(image_size=a.image_size, patch_size=a.patch_size, hidden_size=a.hidden_size, num_hidden_layers=a.num_layers,
                                     num_attention_heads=a.num_heads, intermediate_size=4 * a.hidden_size, num_labels=a.num_labels)
        model = transformers.ViTForImageClassification(cfg)
        size = a.image_size
    args = transformers.TrainingArguments(output_dir=a.output_dir, max_steps=a.max_steps, per_device_train_batch_size=a.per_device_train_batch_size,
                                          per_device_eval_batch_size=a.per_device_train_batch_size, learning_rate=a.learning_rate, eval_strategy="steps",
                                          eval_steps=a.eval_steps, save_steps=a.eval_steps, logging_steps=a.logging_steps, bf16=a.bf16, report_to=[],
                                          remove_unused_columns=False, use_cpu=not torch.cuda.is_available())
    trainer = transformers.Trainer(model=model, args=args, train_dataset=Blobs(a.train_records, a.num_labels, size, 0),
                                   eval_dataset=Blobs(a.eval_records, a.num_labels, size, 1), compute_metrics=accuracy,
                                   callbacks=[DetCallback(core_context, args)])
    trainer.train(resume_from_checkpoint=args.resume_from_checkpoint)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    ap = argparse.ArgumentParser()
    ap.add_argument("--model_name_or_path")
    ap.add_argument("--output_dir", default="/tmp/hf_image_classification")
    ap.add_argument("--max_steps", type=int, default=200)
    ap.add_argument("--eval_steps", type=int, default=100)
    ap.add_argument("--logging_steps", type=int, default=10)
    ap.add_argument("--per_device_train_batch_size", type=int, default=32)
    ap.add_argument("--learning_rate", type=float, default=2e-4)
    ap.add_argument("--num_labels", type=int, default=4)
    ap.add_argument("--image_size", type=int, default=224)
    ap.add_argument("--patch_size", type=int, default=16)
    ap.add_argument("--hidden_size", type=int, default=384)
    ap.add_argument("--num_layers", type=int, default=6)
    ap.add_argument("--num_heads", type=int, default=6)
    ap.add_argument("--train_records", type=int, default=2048)
    ap.add_argument("--eval_records", type=int, default=256)
    ap.add_argument("--bf16", action="store_true")
    ns = ap.parse_args()
    info = det.get_cluster_info()
    if info is not None:       # hyperparameters of the experiment override the command line (searchable from the yaml files)
        for k, v in info.trial.hparams.items():
            if hasattr(ns, k):
                setattr(ns, k, type(getattr(ns, k))(v) if getattr(ns, k) is not None else v)
    distributed = None
    import os

    if "RANK" in os.environ and info is not None:
        distributed = core.DistributedContext.from_torch_distributed()
    with core.init(distributed=distributed) as ctx:
        main(ctx, ns)
