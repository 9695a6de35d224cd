"""Searcher tests by simulation (test strategy: SURVEY §4 "searcher tested by simulation",
reference ``master/pkg/searcher/simulate.go`` + ``asha_test.go``/``adaptive_asha_test.go``)."""
import numpy as np
import pytest

from determined_b200.master import nprand
from determined_b200.master import searcher as S

HP = {"lr": {"type": "log", "minval": -4, "maxval": -1, "base": 10}, "layers": {"type": "int", "minval": 1, "maxval": 8},
      "act": {"type": "categorical", "vals": ["relu", "gelu"]}, "drop": {"type": "double", "minval": 0.0, "maxval": 0.5},
      "nested": {"a": {"type": "const", "val": 3}, "b": 7}}


def test_nprand_matches_numpy_randomstate():
    for seed in (0, 1, 42, 2**31 - 1):
        a, rs = nprand.State(seed), np.random.RandomState(seed)
        assert [a.unit_interval() for _ in range(50)] == [rs.random_sample() for _ in range(50)]
        a, rs = nprand.State(seed), np.random.RandomState(seed)
        assert [a.intn(1000) for _ in range(50)] == [int(rs.randint(0, 1000)) for _ in range(50)]
        a, rs = nprand.State(seed), np.random.RandomState(seed)
        assert [a.uniform(-3, 7) for _ in range(20)] == [rs.uniform(-3, 7) for _ in range(20)]


def test_nprand_snapshot_restore():
    a = nprand.State(5)
    for _ in range(700):
        a.bits32()
    b = nprand.State.restore(a.snapshot())
    assert [a.bits32() for _ in range(10)] == [b.bits32() for _ in range(10)]


def test_sampling_respects_ranges_and_is_seeded():
    r1, r2 = nprand.State(9), nprand.State(9)
    from determined_b200.master.searcher import hparams as hp

    s1, s2 = hp.sample_all(HP, r1), hp.sample_all(HP, r2)
    assert s1 == s2
    assert 1e-4 <= s1["lr"] <= 1e-1 and 1 <= s1["layers"] <= 8 and s1["act"] in ("relu", "gelu")
    assert 0 <= s1["drop"] <= 0.5 and s1["nested"] == {"a": 3, "b": 7}


def test_single_and_random():
    r = S.simulate({"name": "single", "max_length": {"batches": 100}, "metric": "loss"}, HP)
    assert r["results"] == {100: 1} and r["searcher"].shutdown
    r = S.simulate({"name": "random", "max_length": {"epochs": 3}, "max_trials": 7, "max_concurrent_trials": 2, "metric": "loss"}, HP)
    assert r["results"] == {3: 7}
    assert len({str(t["hparams"]) for t in r["trials"].values()}) == 7


def test_random_invalid_hp_is_replaced():
    seen = {"n": 0}

    def early(hp_, units):
        seen["n"] += 1
        return S.ExitedReason.INVALID_HP if seen["n"] <= 2 else None

    r = S.simulate({"name": "random", "max_length": 10, "max_trials": 4, "max_concurrent_trials": 1, "metric": "m"}, HP, early_exit_fn=early)
    done = [t for t in r["trials"].values() if "exited" not in t]
    assert len(done) == 4 and len(r["trials"]) == 6


def test_grid_counts():
    hp = {"a": {"type": "categorical", "vals": [1, 2, 3]}, "b": {"type": "int", "minval": 0, "maxval": 10, "count": 3},
          "c": {"type": "double", "minval": 0, "maxval": 1, "count": 2}, "d": {"type": "log", "minval": -2, "maxval": 0, "count": 3}}
    r = S.simulate({"name": "grid", "max_length": 5, "max_concurrent_trials": 4, "metric": "m"}, hp)
    assert r["results"] == {5: 3 * 3 * 2 * 3}
    bs = sorted({t["hparams"]["b"] for t in r["trials"].values()})
    assert bs == [0, 5, 10]
    ds = sorted({round(t["hparams"]["d"], 6) for t in r["trials"].values()})
    assert ds == [0.01, 0.1, 1.0]


@pytest.mark.parametrize("stop_once", [False, True])
def test_asha_rung_structure(stop_once):
    cfg = {"name": "async_halving", "num_rungs": 3, "max_length": {"batches": 900}, "max_trials": 36, "divisor": 3,
           "max_concurrent_trials": 6, "metric": "loss", "stop_once": stop_once}
    # metric improves with lr so ranking is deterministic
    r = S.simulate(cfg, HP, seed=1, metric_fn=lambda hp_, u: hp_["lr"])
    res = r["results"]
    assert sum(res.values()) == 36
    assert set(res) <= {100, 400, 1300}      # rung lengths: 100, +300, +900
    assert all(t["closed"] for t in r["trials"].values()) and r["searcher"].shutdown
    if not stop_once:
        # promotion ASHA: about 1/3 promoted per rung, at least one reaches the top
        assert res.get(1300, 0) >= 1 and res[100] > res.get(400, 0) >= res.get(1300, 0)
    # the best configuration must have been trained to the top rung
    best = min(r["trials"].values(), key=lambda t: t["hparams"]["lr"])
    assert best["trained"] == 1300


def test_asha_smaller_is_better_false():
    cfg = {"name": "async_halving", "num_rungs": 2, "max_length": 100, "max_trials": 16, "divisor": 4,
           "max_concurrent_trials": 4, "metric": "acc", "smaller_is_better": False}
    r = S.simulate(cfg, HP, seed=2, metric_fn=lambda hp_, u: hp_["drop"])
    best = max(r["trials"].values(), key=lambda t: t["hparams"]["drop"])
    assert best["trained"] == 125


def test_adaptive_asha_brackets_and_budget():
    assert S.adaptive_brackets("standard", 5) == [3, 4, 5]
    assert S.adaptive_brackets("conservative", 3) == [1, 2, 3]
    assert S.adaptive_brackets("aggressive", 4) == [4]
    assert sum(S.bracket_max_trials(100, 4, [5, 4, 3])) == 100
    assert S.bracket_max_concurrent(16, 4, [70, 20, 10]) == [6, 5, 5]
    cfg = {"name": "adaptive_asha", "max_length": {"batches": 1024}, "max_trials": 64, "max_concurrent_trials": 8, "metric": "loss"}
    r = S.simulate(cfg, HP, seed=4)
    assert sum(r["results"].values()) == 64
    assert all(t["closed"] for t in r["trials"].values()) and r["searcher"].shutdown
    total = sum(k * v for k, v in r["results"].items())
    assert total < 64 * 1024 * 0.5, "adaptive ASHA must train far less than max_trials * max_length"


def test_asha_errored_trial_counts_as_worst_and_search_finishes():
    n = {"i": 0}

    def early(hp_, units):
        n["i"] += 1
        return S.ExitedReason.ERRORED if n["i"] % 5 == 0 else None

    cfg = {"name": "async_halving", "num_rungs": 3, "max_length": 90, "max_trials": 20, "divisor": 3, "max_concurrent_trials": 5, "metric": "m"}
    r = S.simulate(cfg, HP, seed=6, early_exit_fn=early)
    assert r["searcher"].shutdown and all(t["closed"] for t in r["trials"].values())


def test_snapshot_restore_mid_search_is_equivalent():
    cfg = {"name": "adaptive_asha", "max_length": 256, "max_trials": 20, "max_concurrent_trials": 4, "metric": "loss"}
    m1 = S.new_search_method(cfg)
    s1 = S.Searcher(11, m1, HP)
    ops = s1.initial_operations()
    creates = [o for o in ops if isinstance(o, S.Create)]
    for c in creates:
        s1.trial_created(c.request_id)
    first = [o for o in ops if isinstance(o, S.ValidateAfter)][0]
    snap = s1.snapshot()
    import json

    snap = json.loads(json.dumps(snap))          # must survive JSON
    s2 = S.Searcher(0, S.new_search_method(cfg), HP)
    s2.restore(snap)
    a = s1.validation_completed(first.request_id, 0.5, first)
    b = s2.validation_completed(first.request_id, 0.5, first)
    assert [o.to_json() for o in a] == [o.to_json() for o in b]


def test_custom_search_event_queue():
    m = S.new_search_method({"name": "custom", "metric": "loss", "unit": "batches"})
    s = S.Searcher(0, m, {})
    assert s.initial_operations() == []
    assert m.events[0]["id"] == 1 and "initialOperations" in m.events[0]
    s.trial_created("r1")
    s.validation_completed("r1", 0.3, S.ValidateAfter("r1", 10))
    assert [list(e)[1] for e in m.events] == ["initialOperations", "trialCreated", "validationCompleted"]
    m.pop_events_through(2)
    assert len(m.events) == 1
