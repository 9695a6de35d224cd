"""Autotune search methods driven against a simulated cluster (reference strategy:
``harness/tests/experiment/integrations/test_dsat.py`` drives the handlers directly)."""
import json
import pathlib
import uuid

import pytest

from determined_b200 import searcher
from determined_b200.pytorch import dsat
from determined_b200.pytorch.dsat import defaults


def _args(tmp_path, method, extra=()):
    cfg = tmp_path / "exp.yaml"
    ds = tmp_path / "ds_config.json"
    ds.write_text(json.dumps({"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1,
                              "optimizer": {"type": "AdamW", "params": {"lr": 1e-4}}, "bf16": {"enabled": True}}))
    cfg.write_text("name: t\nentrypoint: python3 train.py\nresources: {slots_per_trial: 8}\n"
                   "hyperparameters: {deepspeed_config: ds_config.json}\nsearcher: {name: single, metric: loss, max_length: 10}\n")
    argv = [method, str(cfg), str(tmp_path), "--max-trials", "24", "--max-concurrent-trials", "2", *extra]
    return dsat.get_full_parser().parse_args(argv)


class FakeCluster:
    """throughput grows with micro batch size and falls slightly with ZeRO stage; OOM above a
    per-stage limit that mirrors the tracker's own memory model."""
    PARAMS = 6_000_000_000
    ACT = 3 * 2**30
    GPU = 180 * 2**30

    def limit(self, stage):
        p, g, o, n = self.PARAMS * 2, self.PARAMS * 4, self.PARAMS * 12, 8
        fixed = {0: p + g + o, 1: p + g + o // n, 2: p + (g + o) // n, 3: (p + g + o) // n}[stage]
        return (self.GPU - fixed) // self.ACT

    def run(self, trial):
        if isinstance(trial, dsat.DSATModelProfileInfoTrial):
            return {"num_params": self.PARAMS, "trainable_num_params": self.PARAMS, "activation_mem_per_gpu": self.ACT, "gpu_mem": self.GPU}
        if trial.mbs > self.limit(trial.stage):
            return None
        return {"throughput": trial.mbs / (1.0 + 0.05 * trial.mbs + 0.1 * trial.stage), "FLOPS_per_gpu": float(trial.mbs),
                "latency": 1.0 / trial.mbs}


def drive(method, cluster, max_events=500):
    st = searcher.SearcherState()
    pending = list(method.initial_operations(st))
    shutdown = None
    created = []
    n = 0
    while pending and n < max_events:
        n += 1
        op = pending.pop(0)
        if isinstance(op, searcher.Create):
            created.append(op.request_id)
            pending += method.on_trial_created(st, op.request_id)
        elif isinstance(op, searcher.ValidateAfter):
            trial = method.trial_tracker[op.request_id]
            res = cluster.run(trial)
            if res is None:
                pending += method.on_trial_exited_early(st, op.request_id, searcher.ExitedReason.ERRORED)
                st.trials_closed.add(op.request_id)
                pending += method.on_trial_closed(st, op.request_id)
            else:
                pending += method.on_validation_completed(st, op.request_id, res, op.length)
        elif isinstance(op, searcher.Close):
            st.trials_closed.add(op.request_id)
            pending += method.on_trial_closed(st, op.request_id)
        elif isinstance(op, searcher.Shutdown):
            shutdown = op
            break
    method.last_state = st
    return shutdown, created


def test_memory_model_orders_stages(tmp_path):
    args = _args(tmp_path, "binary")
    conf = dsat.get_custom_dsat_exp_conf_from_args(args)
    assert conf["searcher"]["name"] == "custom" and conf["max_restarts"] == 0
    m = dsat.BinarySearchDSATSearchMethod(args, conf)
    ops = m.initial_operations(searcher.SearcherState())
    assert isinstance(ops[0], searcher.Create) and isinstance(ops[1], searcher.ValidateAfter)
    hp = ops[0].hparams
    assert hp[defaults.USE_DSAT_MODE_KEY] and hp[defaults.OVERWRITE_KEY]["train_micro_batch_size_per_gpu"] == 1
    c = FakeCluster()
    m.trial_tracker.update_trial_metric(m.trial_tracker.model_profile_info_trial, c.run(m.trial_tracker.model_profile_info_trial))
    mem = m.trial_tracker.mem_per_gpu_per_stage
    assert mem[0] > mem[1] > mem[2] > mem[3]
    mx = m.trial_tracker.approx_max_mbs_per_stage
    assert mx[1] <= mx[2] <= mx[3]
    assert mx[3] >= 1


@pytest.mark.parametrize("name", ["binary", "random", "asha", "_test"])
def test_search_finds_feasible_best_and_shuts_down(tmp_path, name):
    args = _args(tmp_path, name)
    conf = dsat.get_custom_dsat_exp_conf_from_args(args)
    m = dsat.get_search_method_class(name)(args, conf)
    c = FakeCluster()
    shutdown, created = drive(m, c)
    assert shutdown is not None and not shutdown.failure
    tr = m.trial_tracker
    assert len(created) == len(set(created)) <= args.max_trials
    assert tr.num_running_trials == 0
    best = tr.best_trial
    assert best is not None and not best.error and best.mbs <= c.limit(best.stage)
    assert 0.0 < m.progress(m.last_state) <= 1.0
    # every trial's overwrite block is a consistent batch config
    for _, t in tr:
        ow = t.hparams[defaults.OVERWRITE_KEY]
        assert ow["train_batch_size"] == ow["train_micro_batch_size_per_gpu"] * ow["gradient_accumulation_steps"] * 8
    if name == "binary":
        # converges to the top of the searched range (the model's estimate, one below the true limit)
        assert best.mbs >= c.limit(best.stage) - 1 or tr.num_completed_trials == args.max_trials


def test_binary_search_is_monotone_per_lineage(tmp_path):
    args = _args(tmp_path, "binary", ["-z", "3"])
    m = dsat.BinarySearchDSATSearchMethod(args, dsat.get_custom_dsat_exp_conf_from_args(args))
    drive(m, FakeCluster())
    lin = [t for _, t in m.trial_tracker if not isinstance(t, dsat.DSATModelProfileInfoTrial)]
    ok = [t.mbs for t in lin if not t.error]
    bad = [t.mbs for t in lin if t.error]
    assert ok and (not bad or max(ok) < min(bad))


def test_profile_trial_failure_fails_search(tmp_path):
    args = _args(tmp_path, "random")
    m = dsat.RandomDSATSearchMethod(args, dsat.get_custom_dsat_exp_conf_from_args(args))
    st = searcher.SearcherState()
    ops = m.initial_operations(st)
    assert m.on_trial_exited_early(st, ops[0].request_id, searcher.ExitedReason.ERRORED) == []      # a crash: handled when the trial closes
    out = m.on_trial_closed(st, ops[0].request_id)
    assert len(out) == 1 and isinstance(out[0], searcher.Shutdown) and out[0].failure
    # an invalid hyperparameter or a user's cancel ends the job at once
    m2 = dsat.RandomDSATSearchMethod(args, dsat.get_custom_dsat_exp_conf_from_args(args))
    ops = m2.initial_operations(st)
    out = m2.on_trial_exited_early(st, ops[0].request_id, searcher.ExitedReason.INVALID_HP)
    assert len(out) == 1 and isinstance(out[0], searcher.Shutdown) and out[0].failure


def test_state_roundtrip(tmp_path):
    args = _args(tmp_path, "asha")
    conf = dsat.get_custom_dsat_exp_conf_from_args(args)
    m = dsat.ASHADSATSearchMethod(args, conf)
    st = searcher.SearcherState()
    ops = m.initial_operations(st)
    m.on_validation_completed(st, ops[0].request_id, FakeCluster().run(m.trial_tracker[ops[0].request_id]), 1)
    p = pathlib.Path(tmp_path / "state")
    m.save(st, p, experiment_id=7)
    m2 = dsat.ASHADSATSearchMethod(args, conf)
    st2, eid = m2.load(p)
    assert eid == 7 and len(m2.trial_tracker) == len(m.trial_tracker)
    assert [t.hparams for _, t in m2.trial_tracker] == [t.hparams for _, t in m.trial_tracker]
    assert m2.rng.random() == m.rng.random()


def test_hf_args_rewrite(tmp_path):
    ds = tmp_path / "ds.json"
    ds.write_text(json.dumps({"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2}))
    argv = ["--deepspeed", str(ds), "--per_device_train_batch_size", "1", "--other", "x"]
    assert dsat.get_hf_ds_config_path_from_args(argv) == str(ds)
    hp = {defaults.OVERWRITE_KEY: {"train_micro_batch_size_per_gpu": 16}, defaults.CONFIG_KEY: str(ds)}
    out = dsat.get_hf_args_with_overwrites(argv, hp)
    i = out.index("--per_device_train_batch_size")
    assert out[i + 1] == "16"


# ---------------------------------------------------------------------------------------------------------------------
# the engine half: "autotuning": {"enabled": true} makes the ZeRO engine measure, write its result file and exit
# ---------------------------------------------------------------------------------------------------------------------
def _mlp(width=32):
    import torch

    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(8, width), torch.nn.GELU(), torch.nn.Linear(width, 4))


def _steps(engine, n, batch):
    import torch

    for _ in range(n):
        x = torch.randn(batch, 8)
        loss = engine(x).square().mean()
        engine.backward(loss)
        engine.step()


def test_engine_model_info_probe_counts_parameters_and_saved_activations(tmp_path, monkeypatch):
    from determined_b200.parallel import zero

    monkeypatch.chdir(tmp_path)
    sizes = {}
    for batch in (1, 4):
        cfg = {"train_micro_batch_size_per_gpu": batch, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
               "autotuning": {"enabled": True, "model_info_path": f"info{batch}.json", "model_info": {"profile": True}}}
        model = _mlp()
        engine, *_ = zero.initialize(model=model, config=cfg)
        with pytest.raises(SystemExit):
            _steps(engine, 3, batch)
        assert engine.global_steps == 1, "the model-info probe needs exactly one step"
        info = json.loads((tmp_path / f"info{batch}.json").read_text())
        n = sum(p.numel() for p in model.parameters())
        assert info["num_params"] == info["trainable_num_params"] == n and info["rank"] == 0
        sizes[batch] = info["activation_mem_per_gpu"]
    # what autograd keeps between forward and backward: input, pre-activation and activation of the hidden layer, the
    # output (saved by the loss) -- a few hundred bytes per sample here, proportional to the batch, parameters excluded
    assert 4 * (8 + 32 + 32) <= sizes[1] <= 4 * (8 + 3 * 32 + 2 * 4) and sizes[4] == 4 * sizes[1]


def test_engine_metric_probe_times_the_profile_window_and_exits(tmp_path, monkeypatch):
    from determined_b200.parallel import zero

    monkeypatch.chdir(tmp_path)
    cfg = {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
           "autotuning": {"enabled": True, "start_profile_step": 2, "end_profile_step": 4}}
    model = _mlp()
    engine, *_ = zero.initialize(model=model, config=cfg)
    with pytest.raises(SystemExit):
        _steps(engine, 2 * 10, 2)               # two micro-batches per optimizer step
    assert engine.global_steps == 4, "the probe exits when the step counter reaches end_profile_step"
    res = json.loads((tmp_path / defaults.AUTOTUNING_RESULTS_PATH).read_text())
    assert set(res) == {"latency", "throughput", "FLOPS_per_gpu", "forward", "backward"}
    assert res["latency"] > 0 and res["throughput"] == pytest.approx(engine.train_batch_size() / res["latency"])
    n = sum(p.numel() for p in model.parameters())
    assert res["FLOPS_per_gpu"] == pytest.approx(6.0 * n * 4 / res["latency"])       # 2 micro-batches x 2 samples per step
    assert 0 < res["forward"] < res["latency"] and 0 < res["backward"] < res["latency"]
    # an engine without the switch never exits
    engine2, *_ = zero.initialize(model=_mlp(), config={k: v for k, v in cfg.items() if k != "autotuning"})
    _steps(engine2, 12, 2)
    assert engine2.global_steps == 6


def test_trial_controller_runs_a_probe_trial_and_reports_its_result(tmp_path, monkeypatch):
    """The controller's autotuning path end to end on one process: hyperparameters carry the dsat switch and the overwrite
    section, the engine's probe exits the process inside ``dsat_reporting_context``, which reports the result file as
    validation metrics and completes the searcher operation."""
    import torch

    from determined_b200 import pytorch
    from determined_b200.parallel import zero
    from determined_b200.pytorch import deepspeed as det_ds

    monkeypatch.chdir(tmp_path)

    class DS(torch.utils.data.Dataset):
        def __len__(self):
            return 64

        def __getitem__(self, i):
            return torch.full((8,), float(i) / 64)

    class ProbeTrial(det_ds.DeepSpeedTrial):
        def __init__(self, context):
            self.context = context
            base = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}}}
            cfg = det_ds.overwrite_deepspeed_config(base, context.get_hparams()[defaults.OVERWRITE_KEY])
            engine, *_ = zero.initialize(model=_mlp(), config=cfg)
            self.engine = context.wrap_model_engine(engine)

        def train_batch(self, it, epoch_idx, batch_idx):
            loss = self.engine(next(it)).square().mean()
            self.engine.backward(loss)
            self.engine.step()
            return {"loss": loss}

        def evaluate_batch(self, it, batch_idx):
            return {"val_loss": self.engine(next(it)).square().mean()}

        def build_training_data_loader(self):
            return pytorch.DataLoader(DS(), batch_size=self.engine.train_micro_batch_size_per_gpu())

        def build_validation_data_loader(self):
            return pytorch.DataLoader(DS(), batch_size=self.engine.train_micro_batch_size_per_gpu())

    reported = {}

    def run(overwrite, length):
        hp = {defaults.USE_DSAT_MODE_KEY: True, defaults.OVERWRITE_KEY: overwrite}
        with det_ds.init(hparams=hp, exp_conf={}, require_gpu=False) as ctx:
            monkeypatch.setattr(ctx._core.train, "report_validation_metrics", lambda steps_completed, metrics: reported.update(metrics, steps=steps_completed))
            trial = ProbeTrial(ctx)
            with pytest.raises(SystemExit):
                det_ds.Trainer(trial, ctx).fit(max_length=pytorch.Batch(length))
            return trial

    t = run({"train_micro_batch_size_per_gpu": 1, "autotuning": {"enabled": True, "model_info_path": defaults.MODEL_INFO_PROFILING_PATH,
                                                                   "model_info": {"profile": True}}}, 1)
    assert t.engine.global_steps == 1 and t.engine.train_micro_batch_size_per_gpu() == 1
    assert reported["num_params"] > 0 and reported["activation_mem_per_gpu"] > 0 and reported["gpu_mem"] > 0 and reported["steps"] == 1
    (tmp_path / defaults.MODEL_INFO_PROFILING_PATH).unlink()
    reported.clear()
    t = run({"train_micro_batch_size_per_gpu": 8, "autotuning": {"enabled": True, "start_profile_step": 2, "end_profile_step": 3}}, 4)
    assert t.engine.global_steps == 3 and t.engine.train_micro_batch_size_per_gpu() == 8
    assert reported["throughput"] > 0 and reported["FLOPS_per_gpu"] > 0 and reported["steps"] == 4 and "gpu_mem" not in reported
