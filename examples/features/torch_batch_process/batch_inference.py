"""Distributed batch inference with ``torch_batch_process``: every worker gets its shard of the dataset, progress is
checkpointed so a preempted job resumes where it stopped, and rank 0 merges the per-worker outputs at the end.

Off-cluster:  python batch_inference.py            On a cluster: det experiment create config.yaml .
Reference counterpart: examples/features/torch_batch_process_embeddings.
"""
import json
import pathlib

import torch

from determined_b200.pytorch import experimental as px


class Squares(torch.utils.data.Dataset):
    def __len__(self) -> int:
        return 200

    def __getitem__(self, i: int) -> torch.Tensor:
        return torch.tensor([float(i)])


class Embedder(px.TorchBatchProcessor):
    def __init__(self, context: px.TorchBatchProcessorContext) -> None:
        self.context = context
        self.model = context.prepare_model_for_inference(torch.nn.Linear(1, 4))
        self.rank = context.get_distributed_rank() if hasattr(context, "get_distributed_rank") else 0
        self.out_dir = pathlib.Path(context.get_hparams().get("output_dir", "batch_out")) if context.get_hparams() else pathlib.Path("batch_out")
        self.out_dir.mkdir(parents=True, exist_ok=True)
        self.rows = []
        self.batches = 0

    def process_batch(self, batch, batch_idx) -> None:
        with torch.no_grad():
            emb = self.model(self.context.to_device(batch))
        self.rows.extend(emb.cpu().tolist())
        self.batches += 1

    def on_checkpoint_start(self) -> None:
        # flush what this worker produced since the last checkpoint: after a restart these batches are skipped
        if self.rows:
            (self.out_dir / f"rank{self.rank}-upto{self.batches}.json").write_text(json.dumps(self.rows))
            self.rows = []

    def on_finish(self) -> None:
        if self.rank == 0:
            n = sum(len(json.loads(f.read_text())) for f in self.out_dir.glob("rank*-upto*.json"))
            print(f"wrote {n} embeddings under {self.out_dir}/")


if __name__ == "__main__":
    px.torch_batch_process(Embedder, Squares(), batch_size=16, checkpoint_interval=5)
