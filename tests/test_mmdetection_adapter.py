"""MMDetection adapter: the pieces that work without mmdet / mmcv (samplers, LR schedules, file backends, config helpers),
and the trial's import gate."""
import numpy as np
import pytest
import torch

from determined_b200.model_hub import mmdetection as mm


class _DS(torch.utils.data.Dataset):
    def __init__(self, n=37):
        self.flag = np.array([i % 3 == 0 for i in range(n)], dtype=np.uint8)     # two aspect-ratio groups

    def __len__(self):
        return len(self.flag)

    def __getitem__(self, i):
        return {"img": torch.full((1,), float(i)), "flag": int(self.flag[i])}


def test_group_sampler_batches_share_a_group_and_ranks_partition_the_epoch():
    ds, spg, world = _DS(), 4, 2
    per_rank = []
    for r in range(world):
        s = mm.GroupSampler(ds, spg, world, r, seed=3)
        idx = list(s)
        assert len(idx) == len(s) and len(idx) % spg == 0
        for b in range(0, len(idx), spg):
            assert len({int(ds.flag[i]) for i in idx[b:b + spg]}) == 1, "a batch mixes aspect-ratio groups"
        per_rank.append(idx)
    assert len(per_rank[0]) == len(per_rank[1])
    assert set(per_rank[0]) | set(per_rank[1]) == set(range(len(ds)))            # padding repeats, nothing is dropped
    s = mm.GroupSampler(ds, spg, world, 0, seed=3)
    a = list(s)
    s.set_epoch(1)
    assert list(s) != a, "a new epoch reshuffles"


def test_dataset_with_index_and_eval_loader_shards():
    ds = mm.DatasetWithIndex(_DS(10))
    assert ds[7]["idx"] == 7 and len(ds) == 10 and ds.flag.shape == (10,)
    seen = []
    for r in range(3):
        dl = mm.build_dataloader(_DS(10), 2, 0, num_replicas=3, rank=r, test_mode=True)
        for batch in dl.get_data_loader(repeat=False, skip=0, num_replicas=1, rank=0):
            seen.extend(int(i) for i in batch["idx"])
    assert sorted(seen) == list(range(10))


def test_lr_schedules_match_closed_forms():
    step = mm.build_lr_schedule({"policy": "step", "step": [8, 11], "warmup": "linear", "warmup_iters": 500, "warmup_ratio": 0.001}, 12, 1000)
    assert step(0, 0, 0.02) == pytest.approx(0.02 * 0.001)
    assert step(250, 0, 0.02) == pytest.approx(0.02 * (1 - 0.5 * 0.999))
    assert step(500, 0, 0.02) == pytest.approx(0.02)
    assert step(8000, 8, 0.02) == pytest.approx(0.002) and step(11500, 11, 0.02) == pytest.approx(0.0002)
    cos = mm.build_lr_schedule({"policy": "CosineAnnealing", "min_lr": 0.0, "by_epoch": False}, 10, 100)
    assert cos(0, 0, 1.0) == pytest.approx(1.0) and cos(500, 5, 1.0) == pytest.approx(0.5) and cos(1000, 10, 1.0) == pytest.approx(0.0, abs=1e-9)
    poly = mm.build_lr_schedule({"policy": "poly", "power": 0.9, "min_lr": 1e-4, "by_epoch": False, "warmup": "exp", "warmup_iters": 10, "warmup_ratio": 0.1}, 1, 100)
    assert poly(50, 0, 0.01) == pytest.approx((0.01 - 1e-4) * 0.5 ** 0.9 + 1e-4)
    assert poly(5, 0, 0.01) == pytest.approx(((0.01 - 1e-4) * 0.95 ** 0.9 + 1e-4) * 0.1 ** 0.5)
    with pytest.raises(ValueError):
        mm.build_lr_schedule({"policy": "nope"}, 1, 1)(0, 0, 1.0)


def test_lr_callback_drives_optimizer_groups():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([{"params": [p], "lr": 0.1}], lr=0.1)
    ctx = type("Ctx", (), {"optimizers": [opt]})()
    cb = mm.LrUpdaterCallback(ctx, {"policy": "step", "step": [1], "warmup": "constant", "warmup_iters": 2, "warmup_ratio": 0.5}, max_epochs=2, iters_per_epoch=4)
    cb.on_training_start()
    lrs = []
    for it in range(8):
        cb.on_training_batch_start(it)
        lrs.append(opt.param_groups[0]["lr"])
    assert lrs == pytest.approx([0.05, 0.05, 0.1, 0.1, 0.01, 0.01, 0.01, 0.01])
    st = cb.state_dict()
    cb2 = mm.LrUpdaterCallback(ctx, {"policy": "fixed"}, 1, 1)
    cb2.load_state_dict(st)
    assert cb2.iteration == 8 and cb2.base_lrs == [[0.1]]


def test_file_backends_and_config_substitution(tmp_path):
    fb = mm.FakeBackend()
    data = fb.get("s3://whatever/0001.jpg")
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    with fb.get_local_path("x.png") as path:
        assert open(path, "rb").read() == data
    cfg = {"data": {"train": {"pipeline": [{"type": "LoadImageFromFile"}, {"type": "Resize"}]},
                    "val": {"pipeline": [{"type": "LoadImageFromFile", "file_client_args": {"backend": "disk"}}, {"type": "MultiScaleFlipAug", "transforms": [{"type": "LoadImageFromWebcam"}]}]}}}
    mm.sub_backend({"backend": "fake"}, cfg)
    assert cfg["data"]["train"]["pipeline"][0]["file_client_args"] == {"backend": "fake"}
    assert cfg["data"]["val"]["pipeline"][0]["file_client_args"] == {"backend": "fake"}
    assert cfg["data"]["val"]["pipeline"][1]["transforms"][0]["file_client_args"] == {"backend": "fake"}
    assert "file_client_args" not in cfg["data"]["train"]["pipeline"][1]


def test_config_helpers(tmp_path):
    assert mm.flatten_overrides({"optimizer": {"lr": 0.1, "paramwise_cfg": {"norm_decay_mult": 0}}, "data.samples_per_gpu": 4}) == \
        {"optimizer.lr": 0.1, "optimizer.paramwise_cfg.norm_decay_mult": 0, "data.samples_per_gpu": 4}
    zoo = tmp_path / "configs" / "faster_rcnn"
    zoo.mkdir(parents=True)
    (zoo / "metafile.yml").write_text("Models:\n  - Name: frcnn\n    Config: configs/faster_rcnn/faster_rcnn_r50_fpn_1x_coco.py\n    Weights: https://example.invalid/frcnn.pth\n")
    m = mm.get_config_pretrained_url_mapping(str(tmp_path / "configs"))
    assert m == {"faster_rcnn/faster_rcnn_r50_fpn_1x_coco.py": "https://example.invalid/frcnn.pth"}
    assert mm.get_pretrained_ckpt_path(str(tmp_path / "dl"), "configs/faster_rcnn/other.py", str(tmp_path / "configs")) == (None, None)
    assert mm.build_fp16_loss_scaler("dynamic") == {} and mm.build_fp16_loss_scaler(512.0)["init_scale"] == 512.0
    with pytest.raises(ValueError):
        mm.build_fp16_loss_scaler({"bogus": 1})


def test_trial_needs_mmdet():
    try:
        import mmdet  # noqa: F401
        pytest.skip("mmdet is installed")
    except ImportError:
        pass
    with pytest.raises(ImportError, match="mmcv and mmdet"):
        mm.MMDetTrial(object())
