"""Causal-LM fine-tuning with the HF Trainer, reporting to the cluster through DetCallback.

Reference: ``examples/hf_trainer_api/hf_language_modeling/run_clm.py`` (HF's run_clm + ``det.core.init`` +
``DetCallback``).  Needs a local model snapshot (``--model_name_or_path /path``) because the image has no network;
``--from_scratch_config gpt2`` trains a randomly initialised GPT-2-sized model on synthetic tokens instead.
"""
import argparse
import logging

import torch

import determined_b200 as det
from determined_b200 import core
from determined_b200.transformers import DetCallback


class Tokens(torch.utils.data.Dataset):
    def __init__(self, vocab: int, seq: int, n: int, seed: int) -> None:
        g = torch.Generator().manual_seed(seed)
        self.x = torch.randint(0, vocab, (n, seq), generator=g)

    def __len__(self) -> int:
        return len(self.x)

    def __getitem__(self, i: int) -> dict:
        return {"input_ids": self.x[i], "labels": self.x[i]}


def main(core_context: core.Context, a: argparse.Namespace) -> None:
    import transformers

    if a.model_name_or_path:
        model = transformers.AutoModelForCausalLM.from_pretrained(a.model_name_or_path)
    else:
        cfg = transformers.GPT2Config(n_layer=a.n_layer, n_head=a.n_head, n_embd=a.n_embd, n_positions=a.block_size)
        model = transformers.GPT2LMHeadModel(cfg)
    args = transformers.TrainingArguments(output_dir=a.output_dir, max_steps=a.max_steps, per_device_train_batch_size=a.per_device_train_batch_size,
                                          eval_strategy="steps", eval_steps=a.eval_steps, save_steps=a.eval_steps, logging_steps=a.logging_steps, bf16=a.bf16,
                                          report_to=[])
    cb = DetCallback(core_context, args)
    trainer = transformers.Trainer(model=model, args=args, train_dataset=Tokens(model.config.vocab_size, a.block_size, 4096, 0),
                                   eval_dataset=Tokens(model.config.vocab_size, a.block_size, 256, 1), callbacks=[cb])
    trainer.train(resume_from_checkpoint=args.resume_from_checkpoint)


if __name__ == "__main__":
    logging.basicConfig(level=logging.INFO)
    ap = argparse.ArgumentParser()
    ap.add_argument("--model_name_or_path")
    ap.add_argument("--output_dir", default="/tmp/run_clm")
    ap.add_argument("--max_steps", type=int, default=100)
    ap.add_argument("--eval_steps", type=int, default=50)
    ap.add_argument("--logging_steps", type=int, default=10)
    ap.add_argument("--per_device_train_batch_size", type=int, default=8)
    ap.add_argument("--block_size", type=int, default=512)
    ap.add_argument("--n_layer", type=int, default=12)
    ap.add_argument("--n_head", type=int, default=12)
    ap.add_argument("--n_embd", type=int, default=768)
    ap.add_argument("--bf16", action="store_true")
    ns = ap.parse_args()
    distributed = core.DistributedContext.from_torch_distributed() if det.get_cluster_info() and torch.distributed.is_available() and "RANK" in __import__("os").environ else None
    with core.init(distributed=distributed) as ctx:
        main(ctx, ns)
