"""Variant A: ``torch_batch_process`` owns the loop — sharding, the data loader, checkpoints of progress, resumption and the
synchronisation between workers; the processor only says what to do with a batch."""
import json
import pathlib

import torch

import model as m
from determined_b200.pytorch import experimental as px


class Scorer(px.TorchBatchProcessor):
    def __init__(self, context: px.TorchBatchProcessorContext) -> None:
        self.context = context
        self.model = context.prepare_model_for_inference(m.get_model())
        self.rank = context.get_distributed_rank() if hasattr(context, "get_distributed_rank") else 0
        self.out = pathlib.Path("comparison_out/tbp")
        self.out.mkdir(parents=True, exist_ok=True)
        self.pending = {}

    def process_batch(self, batch, batch_idx) -> None:
        with torch.no_grad():
            self.pending[batch_idx] = self.model(self.context.to_device(batch)).argmax(1).cpu().tolist()

    def on_checkpoint_start(self) -> None:
        for idx, pred in self.pending.items():
            (self.out / f"rank{self.rank}-batch{idx:05d}.json").write_text(json.dumps(pred))
        self.pending = {}

    def on_finish(self) -> None:
        if self.rank == 0:
            n = sum(len(json.loads(f.read_text())) for f in self.out.glob("*.json"))
            print(f"torch_batch_process variant scored {n} records")


if __name__ == "__main__":
    px.torch_batch_process(Scorer, m.Images(), batch_size=50, checkpoint_interval=4)
