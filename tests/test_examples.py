"""The shipped examples run (CPU, local mode): a user of the reference finds working counterparts."""
import importlib.util
import os
import pathlib
import subprocess
import sys

import pytest
import yaml

ROOT = pathlib.Path(__file__).resolve().parent.parent
EX = ROOT / "examples"


def _local_test(example_dir: pathlib.Path, cfg: str) -> None:
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    r = subprocess.run([sys.executable, "-m", "determined_b200.cli", "experiment", "create", cfg, ".", "--local", "--test"], cwd=example_dir, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "Model training test succeeded." in r.stdout, r.stdout[-3000:]


def test_mnist_tutorial_local_test_mode():
    _local_test(EX / "tutorials" / "mnist_pytorch", "const.yaml")


def test_all_example_configs_validate():
    from determined_b200.master import expconf

    n = 0
    for f in EX.rglob("*.yaml"):
        cfg = yaml.safe_load(f.read_text())
        full = expconf.finalize(cfg, {"checkpoint_storage": {"type": "shared_fs", "host_path": "/tmp/ck"}})
        assert full["searcher"]["name"] in ("single", "adaptive_asha", "custom"), f
        n += 1
    assert n >= 7


def test_gpt_neox_deepspeed_example_trains_tiny_model(tmp_path, monkeypatch):
    from determined_b200 import pytorch
    from determined_b200.pytorch import deepspeed as det_ds

    spec = importlib.util.spec_from_file_location("gpt2_trial", EX / "deepspeed" / "gpt_neox" / "gpt2_trial.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    monkeypatch.chdir(tmp_path)
    hp = {"model": "tiny", "seq_length": 32, "model_overrides": {"rotary_pct": 0.25, "parallel_residual": True, "tie_embeddings": False},
          "deepspeed_config": str(EX / "deepspeed" / "gpt_neox" / "ds_config.json"),
          "overwrite_deepspeed_args": {"gradient_accumulation_steps": 2, "bf16": {"enabled": False}, "optimizer": {"params": {"lr": 3e-3}}}}
    with det_ds.init(hparams=hp, exp_conf={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}}, require_gpu=False) as ctx:
        trial = mod.GPT2Trial(ctx)
        det_ds.Trainer(trial, ctx).fit(max_length=pytorch.Batch(6), validation_period=pytorch.Batch(6))
    assert trial.engine.global_steps == 6 and trial.engine.micro_steps == 12


def test_core_api_example_runs_off_cluster():
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    r = subprocess.run([sys.executable, "train.py"], cwd=EX / "tutorials" / "core_api", env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0, r.stdout[-2000:]


def test_batch_inference_example_runs_off_cluster(tmp_path):
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    r = subprocess.run([sys.executable, str(EX / "features" / "torch_batch_process" / "batch_inference.py")], cwd=tmp_path, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and "wrote 200 embeddings" in r.stdout, r.stdout[-2000:]


def test_unmanaged_example_reports_to_a_master(tmp_path):
    from determined_b200.common import api
    from determined_b200.master import api as mapi
    from determined_b200.master import core

    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, launch=False, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        env = dict(os.environ, PYTHONPATH=str(ROOT), DET_MASTER=m.url, HOME=str(tmp_path))
        r = subprocess.run([sys.executable, str(EX / "features" / "unmanaged" / "1_singleton.py")], cwd=tmp_path, env=env,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
        assert r.returncode == 0 and "experiment page:" in r.stdout, r.stdout[-2000:]
        sess = api.authentication.login(m.url, "determined", "")
        exps = sess.get("/api/v1/experiments").json()["experiments"]
        assert [e["name"] for e in exps] == ["unmanaged-1-singleton"]
        tid = sess.get(f"/api/v1/experiments/{exps[0]['id']}/trials").json()["trials"][0]["id"]
        mets = sess.get(f"/api/v1/trials/{tid}/metrics").json()["metrics"]
        assert len([x for x in mets if x["group"] == "training"]) == 10 and len([x for x in mets if x["group"] == "validation"]) == 2
    finally:
        m.stop()
        srv.stop()


def test_ports_example_config_launches_a_proxied_service(tmp_path):
    import time

    import requests

    from determined_b200.common import api
    from determined_b200.master import api as mapi
    from determined_b200.master import core

    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        sess = api.authentication.login(m.url, "determined", "")
        cfg = yaml.safe_load((EX / "features" / "ports" / "server.yml").read_text())
        cmd = sess.post("/api/v1/commands", json={"config": cfg}).json()["command"]
        ok = False
        for _ in range(100):
            c = sess.get(f"/api/v1/commands/{cmd['id']}").json()["command"]
            if c.get("serviceAddress") and c["state"] == "STATE_RUNNING":
                r = requests.get(m.url + c["serviceAddress"], params={"token": sess.token}, timeout=5)
                if r.status_code == 200 and "Directory listing" in r.text:
                    ok = True
                    break
            time.sleep(0.2)
        assert ok
        sess.post(f"/api/v1/commands/{cmd['id']}/kill")
    finally:
        m.stop()
        srv.stop()


def test_dsat_torchvision_examples(tmp_path, monkeypatch):
    """The DeepSpeedTrial variant trains a few batches on CPU with an overlaid engine config (what DSAT does per trial); the
    Core API variant runs off-cluster and reports its profile."""
    from determined_b200 import pytorch
    from determined_b200.pytorch import deepspeed as det_ds

    d = EX / "deepspeed_autotune" / "torchvision"
    spec = importlib.util.spec_from_file_location("dsat_vision_model_def", d / "deepspeed_trial" / "model_def.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    monkeypatch.chdir(tmp_path)
    hp = {"model_name": "resnet18", "image_size": 32, "dataset_len": 16, "deepspeed_config": str(d / "deepspeed_trial" / "ds_config.json"),
          "overwrite_deepspeed_args": {"train_micro_batch_size_per_gpu": 4, "bf16": {"enabled": False}, "zero_optimization": {"stage": 2}}}
    with det_ds.init(hparams=hp, exp_conf={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}}, require_gpu=False) as ctx:
        trial = mod.VisionTrial(ctx)
        assert trial.engine.train_micro_batch_size_per_gpu() == 4 and trial.engine.zero_optimization_stage() == 2
        det_ds.Trainer(trial, ctx).fit(max_length=pytorch.Batch(3), validation_period=pytorch.Batch(3))
    assert trial.engine.global_steps == 3
    env = dict(os.environ, PYTHONPATH=str(ROOT), CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, "script.py"], cwd=d / "core_api", env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]


@pytest.mark.parametrize("script", ["1_metrics.py", "2_checkpoints.py", "3_hpsearch.py", "4_distributed.py"])
def test_core_api_tutorial_stages_run_off_cluster(script, tmp_path):
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    r = subprocess.run([sys.executable, str(EX / "tutorials" / "core_api" / script)], cwd=tmp_path, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                       text=True, timeout=120)
    assert r.returncode == 0 and "report_metrics(group=validation" in r.stdout, r.stdout[-2000:]


def test_core_api_pytorch_mnist_trains_checkpoints_and_reports(tmp_path):
    env = dict(os.environ, PYTHONPATH=str(ROOT), MNIST_MAX_BATCHES="5", MNIST_SYNTHETIC="512")
    r = subprocess.run([sys.executable, str(EX / "tutorials" / "core_api_pytorch_mnist" / "model_def.py")], cwd=tmp_path, env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and "saved checkpoint" in r.stdout and "SearcherOperation Complete" in r.stdout and "'accuracy'" in r.stdout, r.stdout[-2000:]


def test_inference_examples_score_every_record(tmp_path):
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    r = subprocess.run([sys.executable, str(EX / "features" / "inference_mnist_pytorch" / "inference.py")], cwd=tmp_path, env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0 and "scored 1000 records" in r.stdout, r.stdout[-2000:]
    comp = EX / "features" / "torch_batch_process_core_api_comparison"
    for script, words in (("torch_batch_process_inference.py", "torch_batch_process variant scored 1000 records"), ("core_api_inference.py", "core api variant scored 1000 records")):
        r = subprocess.run([sys.executable, str(comp / script)], cwd=tmp_path, env=dict(env, PYTHONPATH=f"{ROOT}{os.pathsep}{comp}"), stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True, timeout=300)
        assert r.returncode == 0 and words in r.stdout, r.stdout[-2000:]
    import json

    a = sorted((tmp_path / "comparison_out" / "tbp").glob("*.json"))
    b = sorted((tmp_path / "comparison_out" / "core_api").glob("*.json"))
    assert sum(len(json.loads(f.read_text())) for f in a) == sum(len(json.loads(f.read_text())) for f in b) == 1000
    assert sorted(sum((json.loads(f.read_text()) for f in a), [])) == sorted(sum((json.loads(f.read_text()) for f in b), [])), "both variants make the same predictions"


def test_unmanaged_checkpoint_example_resumes_in_the_same_trial(tmp_path):
    from determined_b200.common import api
    from determined_b200.master import api as mapi
    from determined_b200.master import core

    m = core.Master(db_path=str(tmp_path / "m.db"), num_slots=1, launch=False, work_dir=str(tmp_path / "w"))
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    try:
        env = dict(os.environ, PYTHONPATH=str(ROOT), DET_MASTER=m.url, HOME=str(tmp_path))
        script = str(EX / "features" / "unmanaged" / "2_checkpoints.py")
        r1 = subprocess.run([sys.executable, script, "--steps", "40", "--checkpoint-dir", str(tmp_path / "ck")], cwd=tmp_path, env=env,
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
        assert r1.returncode == 0 and "resumed" not in r1.stdout, r1.stdout[-2000:]
        r2 = subprocess.run([sys.executable, script, "--steps", "80", "--checkpoint-dir", str(tmp_path / "ck")], cwd=tmp_path, env=env,
                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
        assert r2.returncode == 0 and "at step 40" in r2.stdout, r2.stdout[-2000:]
        sess = api.authentication.login(m.url, "determined", "")
        exps = sess.get("/api/v1/experiments").json()["experiments"]
        assert len(exps) == 1, "the second run finds the first one's experiment by its external id"
        trials = sess.get(f"/api/v1/experiments/{exps[0]['id']}/trials").json()["trials"]
        assert len(trials) == 1
        steps = [x["totalBatches"] for x in sess.get(f"/api/v1/trials/{trials[0]['id']}/metrics").json()["metrics"] if x["group"] == "training"]
        assert steps == [10, 20, 30, 40, 50, 60, 70, 80]
    finally:
        m.stop()
        srv.stop()
