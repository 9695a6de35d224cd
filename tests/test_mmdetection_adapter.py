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


# ---------------------------------------------------------------------------------------------------------------------
# MMDetTrial end to end against stand-ins for mmcv / mmdet that have just the surface the trial touches: a Config with
# fromfile / merge_from_dict, build_optimizer, a detector with mmdet's forward contract (losses dict on return_loss=True,
# per-image results otherwise) and a dataset with `flag`, `CLASSES` and `evaluate`.
# ---------------------------------------------------------------------------------------------------------------------
class _Cfg(dict):
    def __getattr__(self, k):
        try:
            v = self[k]
        except KeyError:
            raise AttributeError(k)
        return v

    def __setattr__(self, k, v):
        self[k] = v

    @staticmethod
    def wrap(o):
        if isinstance(o, dict):
            return _Cfg({k: _Cfg.wrap(v) for k, v in o.items()})
        if isinstance(o, list):
            return [_Cfg.wrap(v) for v in o]
        return o

    @classmethod
    def fromfile(cls, path):
        ns = {}
        exec(open(path).read(), ns)
        return cls.wrap({k: v for k, v in ns.items() if not k.startswith("_")})

    def merge_from_dict(self, d):
        for key, v in d.items():
            cur = self
            parts = key.split(".")
            for p in parts[:-1]:
                cur = cur.setdefault(p, _Cfg())
            cur[parts[-1]] = _Cfg.wrap(v)

    @property
    def pretty_text(self):
        import pprint
        return pprint.pformat(dict(self))


class _ToyDetector(torch.nn.Module):
    def __init__(self, width):
        super().__init__()
        self.head = torch.nn.Linear(4, width)
        self.out = torch.nn.Linear(width, 1)
        self.inited = False

    def init_weights(self):
        self.inited = True

    def forward(self, img, target=None, return_loss=True, rescale=False, **kw):
        pred = self.out(torch.relu(self.head(img))).squeeze(-1)
        if return_loss:
            return {"loss_bbox": ((pred - target) ** 2).mean(), "loss_cls": [0.1 * pred.abs().mean(), 0.0 * pred.sum()], "acc": (pred.detach() > 0).float().mean()}
        return [float(p) for p in pred]

    def _parse_losses(self, losses):
        log = {k: (sum(x for x in v) if isinstance(v, list) else v) for k, v in losses.items()}
        loss = sum(v for k, v in log.items() if "loss" in k)
        log["loss"] = loss
        return loss, {k: v.detach() for k, v in log.items()}


class _ToyDataset(torch.utils.data.Dataset):
    CLASSES = ("thing",)

    def __init__(self, n, test_mode=False, **kw):
        import numpy as np

        g = torch.Generator().manual_seed(7)
        self.x = torch.randn(n, 4, generator=g)
        self.y = self.x.sum(1)
        self.flag = np.array([i % 2 for i in range(n)], dtype=np.uint8)
        self.test_mode = test_mode
        self.seen = None

    def __len__(self):
        return len(self.x)

    def __getitem__(self, i):
        return {"img": self.x[i]} if self.test_mode else {"img": self.x[i], "target": self.y[i]}

    def evaluate(self, results, metric="bbox", **kw):
        self.seen = list(results)
        err = sum((r - float(t)) ** 2 for r, t in zip(results, self.y)) / len(results)
        return {f"{metric}_mAP": 1.0 / (1.0 + err), "n": len(results)}


@pytest.fixture()
def fake_mm(monkeypatch):
    import sys
    import types

    mmcv = types.ModuleType("mmcv")
    mmcv.Config = _Cfg
    runner = types.ModuleType("mmcv.runner")
    runner.build_optimizer = lambda model, cfg: getattr(torch.optim, cfg["type"])(model.parameters(), **{k: v for k, v in cfg.items() if k != "type"})
    runner.loaded = []
    runner.load_checkpoint = lambda model, path, map_location=None: runner.loaded.append(path)
    mmcv.runner = runner
    fileio = types.ModuleType("mmcv.fileio")

    class BaseStorageBackend:
        pass

    class FileClient:
        _backends = {}

        @classmethod
        def register_backend(cls, name, backend, force=False):
            assert issubclass(backend, BaseStorageBackend)
            cls._backends[name] = backend
    fileio.BaseStorageBackend, fileio.FileClient = BaseStorageBackend, FileClient
    mmcv.fileio = fileio
    mmdet = types.ModuleType("mmdet")
    models = types.ModuleType("mmdet.models")
    models.build_detector = lambda cfg: _ToyDetector(int(cfg["width"]))
    datasets = types.ModuleType("mmdet.datasets")
    datasets.built = []

    def build_dataset(cfg, default_args=None):
        ds = _ToyDataset(int(cfg["n"]), **{**{k: v for k, v in cfg.items() if k not in ("n", "type", "pipeline")}, **(default_args or {})})
        datasets.built.append(ds)
        return ds
    datasets.build_dataset = build_dataset
    mmdet.models, mmdet.datasets = models, datasets
    for name, mod in (("mmcv", mmcv), ("mmcv.runner", runner), ("mmdet", mmdet), ("mmdet.models", models), ("mmdet.datasets", datasets)):
        monkeypatch.setitem(sys.modules, name, mod)
    return mmcv, mmdet


CONFIG = """
model = dict(type="Toy", width=8)
optimizer = dict(type="SGD", lr=0.05, momentum=0.9)
optimizer_config = dict(grad_clip=dict(max_norm=5.0, norm_type=2))
lr_config = dict(policy="step", step=[1], warmup="linear", warmup_iters=4, warmup_ratio=0.1)
runner = dict(type="EpochBasedRunner", max_epochs=2)
evaluation = dict(interval=1, metric="bbox", save_best="auto")
data = dict(workers_per_gpu=0,
            train=dict(type="ToyDataset", n=32, pipeline=[dict(type="LoadImageFromFile")]),
            val=dict(type="ToyDataset", n=9, pipeline=[dict(type="LoadImageFromFile")]))
"""


def test_mmdet_trial_trains_and_evaluates_through_the_trainer(tmp_path, monkeypatch, fake_mm):
    from determined_b200 import pytorch

    mmcv, mmdet = fake_mm
    monkeypatch.chdir(tmp_path)
    (tmp_path / "configs").mkdir()
    (tmp_path / "configs" / "toy.py").write_text(CONFIG)
    monkeypatch.setenv("MMDETECTION_CONFIG_DIR", str(tmp_path / "configs"))
    hparams = {"global_batch_size": 4, "config_file": "toy.py", "override_mmdet_config": {"model": {"width": 16}, "optimizer.lr": 0.02},
               "data_backend": {"type": "fake", "fake_img_path": "/nonexistent.jpg"}}
    exp_conf = {"searcher": {"name": "single", "metric": "bbox_mAP", "smaller_is_better": False, "max_length": {"batches": 16}}, "hyperparameters": hparams,
                "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}}
    reported = []
    with pytorch.init(hparams=hparams, exp_conf=exp_conf) as ctx:
        trial = mm.MMDetTrial(ctx)
        assert trial.model.inited and trial.model.head.out_features == 16, "override_mmdet_config (nested form) reaches the model"
        assert trial.optimizer.param_groups[0]["lr"] == 0.02, "override_mmdet_config (dotted form) reaches the optimizer"
        assert trial.cfg.data.val.test_mode is True
        assert trial.cfg.data.train.pipeline[0].file_client_args["backend"] == "fake", "the data backend is swapped into every image-loading step"
        assert trial.clip_grads_fn is not None
        assert {"s3", "gcs", "fake"} <= set(mmcv.fileio.FileClient._backends), "the file backends are registered with mmcv"
        orig = ctx._core.train.report_metrics

        def spy(group, steps_completed, metrics, *a, **kw):
            reported.append((group, steps_completed, dict(metrics)))
            return orig(group, steps_completed, metrics, *a, **kw)
        ctx._core.train.report_metrics = spy
        pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(16), validation_period=pytorch.Batch(8), reporting_period=pytorch.Batch(8))
    train = [m for g, s, m in reported if g == "training"]
    val = [m for g, s, m in reported if g == "validation"]
    assert len(train) == 2 and {"loss", "loss_bbox", "loss_cls", "acc", "lr"} <= set(train[0])
    assert train[1]["loss"] < train[0]["loss"], "the detector learns"
    assert len(val) == 2 and val[1]["bbox_mAP"] > val[0]["bbox_mAP"] and val[0]["n"] == 9
    # evaluate() saw one result per validation image, in dataset order, and none of the schedule keys of `evaluation`
    vds = mmdet.datasets.built[-1]
    assert vds.test_mode and len(vds.seen) == 9
    with torch.no_grad():
        ref = trial.model(vds.x, return_loss=False)
    assert vds.seen == pytest.approx(ref, rel=1e-5)
    # warm-up then the step policy: the last lr is base_lr * 0.1 after epoch 1 (8 iterations per epoch here)
    assert train[0]["lr"] <= 0.02 and train[1]["lr"] == pytest.approx(0.002)
    assert trial.model.CLASSES == ("thing",)


def test_mmdet_trial_loads_the_zoo_checkpoint_when_asked(tmp_path, monkeypatch, fake_mm):
    from determined_b200 import pytorch
    from determined_b200.model_hub.mmdetection import utils as mmutils

    mmcv, mmdet = fake_mm
    monkeypatch.chdir(tmp_path)
    (tmp_path / "toy.py").write_text(CONFIG)
    monkeypatch.setattr(mmutils, "get_pretrained_ckpt_path", lambda download_dir, config_file: (str(tmp_path / "zoo.pth"), "http://zoo/x.pth"))
    hparams = {"global_batch_size": 4, "config_file": str(tmp_path / "toy.py"), "use_pretrained": True}
    with pytorch.init(hparams=hparams, exp_conf={"hyperparameters": hparams}) as ctx:
        mm.MMDetTrial(ctx)
    assert mmcv.runner.loaded == [str(tmp_path / "zoo.pth")]
