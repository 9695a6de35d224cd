"""GPT-NeoX-style language model as a DeepSpeedTrial over the native ZeRO engine.

Reference: ``examples/deepspeed/gpt_neox/gpt2_trial.py`` (wraps the external gpt-neox package; Megatron
``mpu`` for model parallelism; the DeepSpeed engine owns the optimizer step).  Here the model is
:mod:`determined_b200.models.gpt` (tcgen05 GEMMs, flash attention fwd/bwd, fused rotary), the engine is
:mod:`determined_b200.parallel.zero` (reduce-scatter + AdamW + parameter push fused in one kernel per bucket)
and the config keeps the DeepSpeed schema so DSAT (`python -m determined_b200.pytorch.dsat ...`) can tune it.
"""
from typing import Any, Dict, Iterator

import torch

from determined_b200 import pytorch
from determined_b200.models import gpt
from determined_b200.pytorch import deepspeed as det_ds
from determined_b200.pytorch import dsat


class SyntheticTokens(torch.utils.data.Dataset):
    """Stand-in for the tokenised pile shards of the reference example (no network here)."""

    def __init__(self, vocab: int, seq: int, n: int = 2048, seed: int = 0) -> None:
        g = torch.Generator().manual_seed(seed)
        self.data = torch.randint(0, vocab, (n, seq + 1), generator=g)

    def __len__(self) -> int:
        return len(self.data)

    def __getitem__(self, i: int) -> torch.Tensor:
        return self.data[i]


class GPT2Trial(det_ds.DeepSpeedTrial):
    def __init__(self, context: det_ds.DeepSpeedTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        ds_config = dsat.get_ds_config_from_hparams(hp) if "deepspeed_config" in hp else dict(hp["ds_config"])
        self.seq = int(hp.get("seq_length", 2048))
        model = gpt.build(hp.get("model", "neox-125m"), block_size=self.seq, **hp.get("model_overrides", {}))
        self.vocab = model.cfg.vocab_size
        engine, _, _, _ = det_ds.initialize(model=model, config=ds_config)
        self.engine = context.wrap_model_engine(engine)

    def _loader(self, seed: int) -> pytorch.DataLoader:
        return pytorch.DataLoader(SyntheticTokens(self.vocab, self.seq, seed=seed), batch_size=self.engine.train_micro_batch_size_per_gpu())

    def build_training_data_loader(self) -> pytorch.DataLoader:
        return self._loader(0)

    def build_validation_data_loader(self) -> pytorch.DataLoader:
        return self._loader(1)

    def train_batch(self, it: Iterator[Any], epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
        tokens = self.context.to_device(next(it))
        _, loss = self.engine(tokens[:, :-1], tokens[:, 1:])
        self.engine.backward(loss)
        self.engine.step()
        return {"lm_loss": loss}

    def evaluate_batch(self, it: Iterator[Any], batch_idx: int) -> Dict[str, Any]:
        tokens = self.context.to_device(next(it))
        _, loss = self.engine(tokens[:, :-1], tokens[:, 1:])
        return {"lm_loss": loss, "lm_loss_ppl": torch.exp(loss.float())}
