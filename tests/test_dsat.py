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
