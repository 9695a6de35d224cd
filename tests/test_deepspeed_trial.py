"""DeepSpeedTrial + ZeRO engine on CPU (test strategy: reference
``tests/experiment/integrations/test_deepspeed_trial.py`` with a Linear(1,1) model)."""
import json
import pathlib

import numpy as np
import pytest
import torch

from determined_b200 import pytorch
from determined_b200.parallel import zero
from determined_b200.pytorch import deepspeed as det_ds


class DS(torch.utils.data.Dataset):
    def __init__(self, n=64):
        self.x = torch.arange(1, n + 1, dtype=torch.float32).reshape(n, 1) / n

    def __len__(self):
        return len(self.x)

    def __getitem__(self, i):
        return self.x[i], 2 * self.x[i]


class LinearDSTrial(det_ds.DeepSpeedTrial):
    def __init__(self, context):
        self.context = context
        hp = context.get_hparams()
        cfg = det_ds.overwrite_deepspeed_config(
            {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2,
             "optimizer": {"type": "AdamW", "params": {"lr": 0.05, "weight_decay": 0.0}}, "zero_optimization": {"stage": 2}},
            hp.get("overwrite_deepspeed_args", {}))
        model = torch.nn.Linear(1, 1, bias=False)
        model.weight.data.fill_(0.0)
        engine, _, _, _ = zero.initialize(model=model, config=cfg)
        self.engine = context.wrap_model_engine(engine)
        self.calls = 0

    def train_batch(self, it, epoch_idx, batch_idx):
        x, y = self.context.to_device(next(it))
        loss = torch.nn.functional.mse_loss(self.engine(x), y)
        self.engine.backward(loss)
        self.engine.step()
        self.calls += 1
        return {"loss": loss}

    def evaluate_batch(self, it, batch_idx):
        x, y = self.context.to_device(next(it))
        return {"val_loss": torch.nn.functional.mse_loss(self.engine(x), y)}

    def build_training_data_loader(self):
        return pytorch.DataLoader(DS(), batch_size=self.engine.train_micro_batch_size_per_gpu())

    def build_validation_data_loader(self):
        return pytorch.DataLoader(DS(), batch_size=self.engine.train_micro_batch_size_per_gpu())


def test_batch_config_resolution():
    assert zero.resolve_batch_config({"train_batch_size": 64, "train_micro_batch_size_per_gpu": 4}, 8) == (64, 4, 2)
    assert zero.resolve_batch_config({"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 3}, 2) == (24, 4, 3)
    assert zero.resolve_batch_config({"train_batch_size": 32, "gradient_accumulation_steps": "auto"}, 4) == (32, 8, 1)
    with pytest.raises(ValueError):
        zero.resolve_batch_config({"train_batch_size": 30, "train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2}, 2)
    with pytest.raises(ValueError):
        zero.resolve_batch_config({}, 1)


def test_overwrite_deepspeed_config(tmp_path):
    f = tmp_path / "ds.json"
    f.write_text(json.dumps({"a": {"b": 1, "c": 2}, "d": 3}))
    assert det_ds.overwrite_deepspeed_config(str(f), {"a": {"b": 9}}) == {"a": {"b": 9, "c": 2}, "d": 3}
    with pytest.raises(TypeError):
        det_ds.overwrite_deepspeed_config(3, {})


def test_deepspeed_trial_trains_accumulates_checkpoints_and_resumes(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    store = tmp_path / "ck"
    store.mkdir()

    def run(max_batches, latest=None):
        with det_ds.init(hparams={}, exp_conf={}, require_gpu=False) as ctx:
            from determined_b200.common import storage

            ctx._core.checkpoint._storage_manager = storage.SharedFSStorageManager(str(store))
            trial = LinearDSTrial(ctx)
            det_ds.Trainer(trial, ctx).fit(max_length=pytorch.Batch(max_batches), checkpoint_policy="all", latest_checkpoint=latest)
            return trial

    t = run(4)
    assert t.calls == 8, "train_batch must be called gradient_accumulation_steps times per batch"
    assert t.engine.micro_steps == 8 and t.engine.global_steps == 4
    w4 = float(t.engine.module.weight)
    assert 0 < w4 < 2
    ck = sorted(store.iterdir(), key=lambda p: p.stat().st_mtime)[-1]
    names = {p.name for p in ck.iterdir()}
    assert {"det_state_dict_rank0.pth", "model0", "metadata.json", "trial_state.pkl", "load_data.json", "code"} <= names
    assert (ck / "model0" / "mp_rank_00_model_states.pt").exists() and (ck / "model0" / "zero_pp_rank_0_mp_rank_00_optim_states.pt").exists()
    assert json.loads((ck / "metadata.json").read_text())["steps_completed"] == 4
    assert json.loads((ck / "load_data.json").read_text())["trial_type"] == "DeepSpeedTrial"
    # resume for 2 more batches == uninterrupted 6 batches
    t2 = run(6, latest=ck.name)
    assert t2.calls == 4 and t2.engine.global_steps == 6
    ref = run(6)
    np.testing.assert_allclose(float(t2.engine.module.weight), float(ref.engine.module.weight), rtol=1e-6)


def test_context_rejects_expconf_optimizations():
    from determined_b200 import core, errors

    with pytest.raises(errors.InvalidExperimentException):
        det_ds.DeepSpeedTrialContext(core._dummy_init(), 0, {}, 1, 0, {"optimizations": {"aggregation_frequency": 2}}, 0, require_gpu=False)
    with pytest.raises(errors.InvalidExperimentException):
        det_ds.DeepSpeedTrialContext(core._dummy_init(), 0, {}, 1, 0, {}, 0, require_gpu=True) if not torch.cuda.is_available() else (_ for _ in ()).throw(errors.InvalidExperimentException("x"))


def test_mpu_from_pipeline_grid():
    class Grid:
        def __init__(self, pr, pw, dr=0, dw=2):
            self.pr, self.pw, self.dr, self.dw = pr, pw, dr, dw

        def get_pipe_parallel_rank(self):
            return self.pr

        def get_pipe_parallel_world_size(self):
            return self.pw

        def get_data_parallel_rank(self):
            return self.dr

        def get_data_parallel_world_size(self):
            return self.dw

    first, mid, last = (det_ds.make_deepspeed_mpu(Grid(r, 3)) for r in range(3))
    assert (first.should_build_data_loader, first.should_report_metrics) == (True, False)
    assert (mid.should_build_data_loader, mid.should_report_metrics) == (False, False)
    assert (last.should_build_data_loader, last.should_report_metrics) == (True, True)
    assert last.data_parallel_world_size == 2


class Stage3Trial(LinearDSTrial):
    """Same trial surface, a model with one weight the stage-3 pass partitions (host twin of the fused kernels on CPU)."""

    def __init__(self, context):
        from determined_b200.ops import gemm as G

        self.context = context
        cfg = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 1,
               "optimizer": {"type": "AdamW", "params": {"lr": 0.01, "weight_decay": 0.0}},
               "zero_optimization": {"stage": 3, "partition_parameters": True, "stage3_param_persistence_threshold": 1}}
        torch.manual_seed(3)

        class Float(torch.nn.Module):          # the fused layer computes in bf16; the plain layers around it are fp32 on CPU
            def forward(self, x):
                return x.float()

        model = torch.nn.Sequential(torch.nn.Linear(1, 64), G.FusedLinear(64, 256, activation="relu"), Float(), torch.nn.Linear(256, 1))
        engine, _, _, _ = zero.initialize(model=model, config=cfg)
        self.engine = context.wrap_model_engine(engine)
        self.calls = 0

    def train_batch(self, it, epoch_idx, batch_idx):
        x, y = self.context.to_device(next(it))
        loss = torch.nn.functional.mse_loss(self.engine(x).float(), y)
        self.engine.backward(loss)
        self.engine.step()
        self.calls += 1
        return {"loss": loss}

    def evaluate_batch(self, it, batch_idx):
        x, y = self.context.to_device(next(it))
        return {"val_loss": torch.nn.functional.mse_loss(self.engine(x).float(), y)}


def test_deepspeed_trial_stage3_checkpoint_holds_full_weights_and_resumes(tmp_path, monkeypatch):
    from determined_b200.ops import gemm_comm

    monkeypatch.chdir(tmp_path)
    store = tmp_path / "ck"
    store.mkdir()

    def run(max_batches, latest=None):
        with det_ds.init(hparams={}, exp_conf={}, require_gpu=False) as ctx:
            from determined_b200.common import storage

            ctx._core.checkpoint._storage_manager = storage.SharedFSStorageManager(str(store))
            trial = Stage3Trial(ctx)
            det_ds.Trainer(trial, ctx).fit(max_length=pytorch.Batch(max_batches), checkpoint_policy="all", latest_checkpoint=latest)
            return trial

    t = run(3)
    assert t.engine.partitioned == 1 and isinstance(t.engine.module[1], gemm_comm.Zero3Linear)
    ck = sorted(store.iterdir(), key=lambda p: p.stat().st_mtime)[-1]
    ms = torch.load(ck / "model0" / "mp_rank_00_model_states.pt", weights_only=False)
    assert ms["module"]["1.weight"].shape == (256, 64) and "1.weight_shard" not in ms["module"]      # the unpartitioned state dict
    t2 = run(5, latest=ck.name)
    ref = run(5)
    assert t2.engine.global_steps == 5
    for a, b in zip(t2.engine.module.parameters(), ref.engine.module.parameters()):
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
