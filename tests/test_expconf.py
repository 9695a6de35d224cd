"""expconf: sanity/completeness phases, defaults, merging (reference ``schemas/test_cases/v0``)."""
import pytest

from determined_b200.master import expconf

BASE = """
name: t
entrypoint: python3 train.py
searcher: {name: single, metric: loss, max_length: {batches: 10}}
checkpoint_storage: {type: shared_fs, host_path: /tmp/x}
"""


def test_defaults_filled():
    c = expconf.finalize(BASE)
    assert c["max_restarts"] == 5 and c["scheduling_unit"] == 100 and c["checkpoint_policy"] == "best"
    assert c["resources"]["slots_per_trial"] == 1 and c["resources"]["weight"] == 1
    o = c["optimizations"]
    assert (o["aggregation_frequency"], o["average_aggregated_gradients"], o["average_training_metrics"], o["mixed_precision"],
            o["tensor_fusion_threshold"], o["tensor_fusion_cycle_time"]) == (1, True, True, "O0", 64, 1)
    cs = c["checkpoint_storage"]
    assert (cs["save_experiment_best"], cs["save_trial_best"], cs["save_trial_latest"]) == (0, 1, 1)
    assert c["searcher"]["smaller_is_better"] is True
    assert c["min_validation_period"] == {"batches": 0} and c["perform_initial_validation"] is False


def test_adaptive_asha_defaults():
    c = expconf.finalize(BASE.replace("{name: single, metric: loss, max_length: {batches: 10}}",
                                      "{name: adaptive_asha, metric: loss, max_length: {epochs: 10}, max_trials: 5}"))
    s = c["searcher"]
    assert (s["mode"], s["divisor"], s["max_rungs"], s["max_concurrent_trials"], s["stop_once"]) == ("standard", 4, 5, 16, False)


@pytest.mark.parametrize("bad,frag", [
    ("resources: {slots_per_trial: two}", "slots_per_trial"),
    ("optimizations: {mixed_precision: O9}", "mixed_precision"),
    ("checkpoint_policy: sometimes", "checkpoint_policy"),
    ("bogus_key: 1", "additional property 'bogus_key'"),
    ("optimizations: {aggregation_frequency: 0}", "aggregation_frequency"),
    ("hyperparameters: {lr: {type: double, minval: 3, maxval: 1}}", "minval"),
    ("hyperparameters: {lr: {type: weird}}", "it must be one of"),
    ("min_validation_period: {batches: 1, epochs: 2}", "must have one attribute"),
])
def test_sanity_errors(bad, frag):
    with pytest.raises(expconf.ConfigError) as e:
        expconf.finalize(BASE + bad + "\n")
    assert any(frag in m for m in e.value.errors), e.value.errors


def test_completeness_errors():
    with pytest.raises(expconf.ConfigError) as e:
        expconf.finalize("searcher: {name: random, metric: m}\ncheckpoint_storage: {type: s3}\n")
    msg = " ".join(e.value.errors)
    for frag in ("entrypoint", "max_length", "max_trials", "bucket"):
        assert frag in msg


def test_grid_requires_count():
    with pytest.raises(expconf.ConfigError):
        expconf.finalize(BASE.replace("name: single", "name: grid") + "hyperparameters: {lr: {type: double, minval: 0, maxval: 1}}\n")


def test_merge_user_over_template_over_cluster():
    tmpl = {"resources": {"slots_per_trial": 4, "priority": 7}, "max_restarts": 1}
    cluster = {"checkpoint_storage": {"type": "shared_fs", "host_path": "/cluster"}, "max_restarts": 9}
    user = "entrypoint: x\nsearcher: {name: single, metric: m, max_length: 5}\nresources: {slots_per_trial: 2}\n"
    c = expconf.finalize(user, cluster_defaults=cluster, template=tmpl)
    assert c["resources"]["slots_per_trial"] == 2 and c["resources"]["priority"] == 7
    assert c["max_restarts"] == 1 and c["checkpoint_storage"]["host_path"] == "/cluster"


def test_global_batch_size_vs_slots():
    with pytest.raises(expconf.ConfigError):
        expconf.finalize(BASE + "resources: {slots_per_trial: 8}\nhyperparameters: {global_batch_size: 4}\n")


def test_apply_override():
    c = {"a": {"b": 1}}
    expconf.apply_override(c, "a.c.d", 5)
    assert c == {"a": {"b": 1, "c": {"d": 5}}}
