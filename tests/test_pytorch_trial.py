"""Trial-loop tests on CPU (test strategy: SURVEY §4 "Trial-loop tests"; oracle = analytic
one-variable model)."""
import json
import pathlib
import pickle

import numpy as np
import pytest
import torch

from determined_b200 import core, pytorch
from tests.fixtures import onevar


def _batches(n, bs, count, rank=0, size=1):
    ds = onevar.OnesDataset(n)
    all_b = [(ds.x[i:i + bs].numpy(), ds.y[i:i + bs].numpy()) for i in range(0, n, bs)]
    mine = all_b[rank::size]
    return [mine[i % len(mine)] for i in range(count)]


def _run(tmp_path, hparams, max_batches, agg=1, latest=None, **fit_kw):
    exp_conf = {"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")},
                "optimizations": {"aggregation_frequency": agg}}
    with pytorch.init(hparams=hparams, exp_conf=exp_conf, aggregation_frequency=agg) as ctx:
        ctx._core.checkpoint._storage_manager = __import__("determined_b200").common.storage.SharedFSStorageManager(str(tmp_path / "ckpt"))
        (tmp_path / "ckpt").mkdir(exist_ok=True)
        trial = onevar.OneVarTrial(ctx)
        tr = pytorch.Trainer(trial, ctx)
        tr.fit(max_length=pytorch.Batch(max_batches), latest_checkpoint=latest, **fit_kw)
        return trial, tr


def test_onevar_matches_analytic_sgd(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    hp = {"learning_rate": 0.1, "batch_size": 4, "dataset_len": 32}
    trial, _ = _run(tmp_path, hp, 20)
    exp = onevar.expected_weights(0.0, 0.1, _batches(32, 4, 20))
    np.testing.assert_allclose(trial.weights, exp, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("agg", [2, 4])
def test_gradient_aggregation(tmp_path, monkeypatch, agg):
    monkeypatch.chdir(tmp_path)
    hp = {"learning_rate": 0.1, "batch_size": 4, "dataset_len": 32}
    trial, _ = _run(tmp_path, hp, 16, agg=agg)
    exp = onevar.expected_weights(0.0, 0.1, _batches(32, 4, 16), agg=agg)
    np.testing.assert_allclose(trial.weights, exp, rtol=1e-5, atol=1e-6)


def test_checkpoint_layout_and_resume(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    hp = {"learning_rate": 0.1, "batch_size": 4, "dataset_len": 32}
    t1, tr1 = _run(tmp_path, hp, 10, checkpoint_policy="all")
    ckpts = sorted((tmp_path / "ckpt").iterdir(), key=lambda p: p.stat().st_mtime)
    assert ckpts, "no checkpoint written"
    ck = ckpts[-1]
    names = {p.name for p in ck.iterdir()}
    assert {"state_dict.pth", "trial_state.pkl", "load_data.json", "metadata.json", "code"} <= names
    md = json.loads((ck / "metadata.json").read_text())
    assert md["steps_completed"] == 10 and md["format"] == "pickle" and md["framework"].startswith("torch-")
    st = pickle.loads((ck / "trial_state.pkl").read_bytes())
    assert st["batches_trained"] == 10 and set(st) == {"trial_id", "last_ckpt", "step_id", "last_val", "batches_trained", "epochs_trained"}
    sd = torch.load(ck / "state_dict.pth", weights_only=False)
    assert {"models_state_dict", "optimizers_state_dict", "lr_schedulers_state_dict", "callbacks", "rng_state"} <= set(sd)
    ld = json.loads((ck / "load_data.json").read_text())
    assert ld["trial_type"] == "PyTorchTrial" and ld["trial_cls_spec"].endswith(":OneVarTrial")

    # resume: continues from batch 10 on the right data and lands where an uninterrupted run lands
    t2, _ = _run(tmp_path, hp, 20, latest=ck.name)
    exp = onevar.expected_weights(0.0, 0.1, _batches(32, 4, 20))
    assert len(t2.weights) == 10
    np.testing.assert_allclose(t2.weights, exp[10:], rtol=1e-5, atol=1e-6)


def test_clip_grads_callable(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    hp = {"learning_rate": 0.1, "batch_size": 4, "dataset_len": 32, "clip": 1e-4}
    trial, _ = _run(tmp_path, hp, 5)
    # every step is clipped to norm 1e-4 -> |dw| = lr * 1e-4
    w = np.array([0.0] + trial.weights)
    np.testing.assert_allclose(np.diff(w), 0.1 * 1e-4, rtol=1e-3)


def test_test_mode_runs_one_batch(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    hp = {"learning_rate": 0.1, "batch_size": 4, "dataset_len": 32}
    exp_conf = {}
    with pytorch.init(hparams=hp, exp_conf=exp_conf) as ctx:
        trial = onevar.OneVarTrial(ctx)
        pytorch.Trainer(trial, ctx).fit(test_mode=True)
        assert len(trial.weights) == 1


def test_train_units():
    assert pytorch.Batch(5).should_stop(10) and not pytorch.Batch(5).should_stop(11)
    assert pytorch.Batch([1, 5]).should_stop(5) and not pytorch.Batch([1, 5]).should_stop(2)
    assert isinstance(pytorch.TrainUnit._from_values(records=100, global_batch_size=10), pytorch.Batch)
    assert pytorch.TrainUnit._from_values(records=100, global_batch_size=10).value == 10
    assert pytorch.TrainUnit._from_searcher_unit(3, core.Unit.EPOCHS).value == 3
    with pytest.raises(ValueError):
        pytorch.TrainUnit._from_values(batches=1, epochs=1)


def test_load_trial_from_checkpoint(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    hp = {"learning_rate": 0.1, "batch_size": 4, "dataset_len": 32}
    t1, _ = _run(tmp_path, hp, 6, checkpoint_policy="all")
    ck = sorted((tmp_path / "ckpt").iterdir(), key=lambda p: p.stat().st_mtime)[-1]
    loaded = pytorch.load_trial_from_checkpoint_path(str(ck), trial_class=onevar.OneVarTrial)
    assert abs(float(loaded.model.weight) - t1.weights[-1]) < 1e-7


def test_async_checkpoint_writes_the_same_files_behind_training(tmp_path, monkeypatch):
    """Checkpoint fast path: mid-training checkpoints are snapshotted and written by a background thread; the files match the
    synchronous path, the snapshot is taken at the boundary (later steps do not leak into it), callbacks are delivered, and the
    trial resumes from one of them."""
    from determined_b200.pytorch import _async_checkpoint as ac

    monkeypatch.chdir(tmp_path)
    hp = {"learning_rate": 0.1, "batch_size": 4, "dataset_len": 32}
    events = []

    class Spy(pytorch.PyTorchCallback):
        def on_checkpoint_write_end(self, checkpoint_dir):
            events.append(("write", pathlib.Path(checkpoint_dir).name))

        def on_checkpoint_upload_end(self, uuid):
            events.append(("upload", uuid))

    def run(root, enable, latest=None, upto=12):
        exp_conf = {"checkpoint_storage": {"type": "shared_fs", "host_path": str(root)}}
        root.mkdir(exist_ok=True)
        with pytorch.init(hparams=hp, exp_conf=exp_conf) as ctx:
            ctx._core.checkpoint._storage_manager = __import__("determined_b200").common.storage.SharedFSStorageManager(str(root))
            ctx.experimental.enable_async_checkpoint(enable)
            trial = onevar.OneVarTrial(ctx)
            trial.build_callbacks = lambda: {"spy": Spy()}
            pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(upto), checkpoint_period=pytorch.Batch(4), checkpoint_policy="none",
                                            latest_checkpoint=latest)
            return trial

    run(tmp_path / "sync", False)
    sync_events, events[:] = list(events), []
    t = run(tmp_path / "async", True)
    assert sorted(e for e in events if e[0] == "write") and len([e for e in events if e[0] == "upload"]) == len([e for e in sync_events if e[0] == "upload"])

    def by_step(root):
        out = {}
        for d in root.iterdir():
            md = json.loads((d / "metadata.json").read_text())
            out[md["steps_completed"]] = d
        return out

    a, s = by_step(tmp_path / "async"), by_step(tmp_path / "sync")
    assert set(a) == set(s) == {4, 8, 12}
    for step in (4, 8, 12):
        assert {p.name for p in a[step].iterdir()} == {p.name for p in s[step].iterdir()}
        sa, ss = torch.load(a[step] / "state_dict.pth", weights_only=False), torch.load(s[step] / "state_dict.pth", weights_only=False)
        for ma, ms in zip(sa["models_state_dict"], ss["models_state_dict"]):
            for k in ms:
                assert torch.equal(ma[k], ms[k]), (step, k)
        assert pickle.loads((a[step] / "trial_state.pkl").read_bytes()) == pickle.loads((s[step] / "trial_state.pkl").read_bytes())
    t2 = run(tmp_path / "async", True, latest=a[8].name, upto=16)
    exp = onevar.expected_weights(0.0, 0.1, _batches(32, 4, 16))
    np.testing.assert_allclose(t2.weights, exp[8:], rtol=1e-5, atol=1e-6)

    # the building blocks: snapshots are deep and writer failures surface at join
    live = {"w": torch.ones(3), "nested": [torch.zeros(2), {"k": 1}], "t": (torch.ones(1), "s")}
    snap = ac.snapshot(live).wait()
    live["w"].add_(1)
    assert torch.equal(snap["w"], torch.ones(3)) and snap["nested"][1] == {"k": 1} and isinstance(snap["t"], tuple)
    w = ac.CheckpointWriter()
    w.submit(lambda: (_ for _ in ()).throw(OSError("disk full")))
    with pytest.raises(RuntimeError, match="background checkpoint failed"):
        w.join()
    w.close()
