"""Custom searcher (client side) against a live master, unmanaged core_v2, HF/Lightning callbacks,
batch inference, and the legacy workload sequencer."""
import json
import pathlib
import sys
import time
import types
import uuid

import pytest
import torch
import yaml

from determined_b200 import core, searcher, workload
from determined_b200.common import api
from determined_b200.master import api as mapi
from determined_b200.master import core as mcore

FIX = pathlib.Path(__file__).parent / "fixtures" / "onevar_exp"


@pytest.fixture()
def cluster(tmp_path, monkeypatch):
    m = mcore.Master(db_path=str(tmp_path / "m.db"), num_slots=4, work_dir=str(tmp_path / "work"),
                     cluster_defaults={"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpt")}})
    srv = mapi.APIServer(m, "127.0.0.1", 0).start()
    m.start()
    monkeypatch.setenv("DET_MASTER", m.url)
    monkeypatch.delenv("DET_SESSION_TOKEN", raising=False)
    sess = api.authentication.login(m.url, "determined", "")
    yield m, sess, tmp_path
    m.stop()
    srv.stop()


class TwoLRSearch(searcher.SearchMethod):
    """Two trials with fixed learning rates, each trained 16 batches; closes both, then shuts down."""

    def __init__(self):
        self.lrs = [0.05, 0.2]
        self.results = {}
        self.closed = 0

    def initial_operations(self, st):
        ops = []
        for lr in self.lrs:
            rid = uuid.uuid4()
            ops += [searcher.Create(rid, {"learning_rate": lr, "batch_size": 8, "global_batch_size": 8}, None), searcher.ValidateAfter(rid, 16)]
        return ops

    def on_trial_created(self, st, rid):
        return []

    def on_validation_completed(self, st, rid, metric, train_length):
        self.results[str(rid)] = metric
        return [searcher.Close(rid)]

    def on_trial_closed(self, st, rid):
        self.closed += 1
        return [searcher.Shutdown()] if self.closed == len(self.lrs) else []

    def on_trial_exited_early(self, st, rid, reason):
        return [searcher.Shutdown(failure=True)]

    def progress(self, st):
        return self.closed / len(self.lrs)

    def save_method_state(self, path):
        (path / "m.json").write_text(json.dumps({"results": self.results, "closed": self.closed}))

    def load_method_state(self, path):
        d = json.loads((path / "m.json").read_text())
        self.results, self.closed = d["results"], d["closed"]


def test_local_search_runner_drives_custom_experiment(cluster):
    m, sess, tmp = cluster
    cfg = {"name": "custom", "entrypoint": "model_def:OneVarTrial", "searcher": {"name": "custom", "metric": "val_loss", "unit": "batches"},
           "hyperparameters": {"learning_rate": 0.1, "batch_size": 8, "global_batch_size": 8}, "scheduling_unit": 8,
           "resources": {"slots_per_trial": 1}, "max_restarts": 0}
    method = TwoLRSearch()
    runner = searcher.LocalSearchRunner(method, searcher_dir=tmp / "search", session=sess)
    eid = runner.run(cfg, model_dir=str(FIX))
    exp = sess.get(f"/api/v1/experiments/{eid}").json()["experiment"]
    assert exp["state"] == "STATE_COMPLETED"
    trials = sess.get(f"/api/v1/experiments/{eid}/trials").json()["trials"]
    assert len(trials) == 2 and all(t["state"] == "STATE_COMPLETED" for t in trials)
    assert len(method.results) == 2 and all(isinstance(v, float) for v in method.results.values())
    assert sorted(round(t["hparams"]["learning_rate"], 3) for t in trials) == [0.05, 0.2]
    # state was checkpointed after every event batch and is loadable by a fresh method
    m2 = TwoLRSearch()
    st, got = m2.load(tmp / "search" / f"exp_{eid}")
    assert got == eid and m2.closed == 2 and len(st.trials_closed) == 2 and len(st.trials_created) == 2


def test_core_v2_unmanaged_trial_reports(cluster):
    from determined_b200.experimental import core_v2

    m, sess, tmp = cluster
    core_v2.init(defaults=core_v2.DefaultConfig(name="unmanaged", hparams={"lr": 0.5}, searcher={"name": "single", "metric": "loss", "max_length": 3}),
                 unmanaged=core_v2.UnmanagedConfig(external_experiment_id="ext-e", external_trial_id="ext-t"),
                 checkpoint_storage=str(tmp / "ck2"))
    try:
        for i in range(3):
            core_v2.train.report_training_metrics(steps_completed=i + 1, metrics={"loss": 1.0 / (i + 1)})
        core_v2.train.report_validation_metrics(steps_completed=3, metrics={"loss": 0.25})
        with core_v2.checkpoint.store_path({"steps_completed": 3}) as (p, uid):
            (p / "w.bin").write_bytes(b"abc")
        assert "/det/experiments/" in core_v2.url_reverse_webui_exp_view()
    finally:
        core_v2.close()
    exps = sess.get("/api/v1/experiments").json()["experiments"]
    e = [x for x in exps if x["name"] == "unmanaged"][0]
    assert e["unmanaged"] is True
    t = sess.get(f"/api/v1/experiments/{e['id']}/trials").json()["trials"][0]
    vals = sess.get(f"/api/v1/trials/{t['id']}/metrics", params={"group": "validation"}).json()["metrics"]
    assert vals[-1]["metrics"]["avgMetrics"]["loss"] == 0.25 and vals[-1]["totalBatches"] == 3
    cks = sess.get(f"/api/v1/trials/{t['id']}/checkpoints").json()["checkpoints"]
    assert len(cks) == 1 and "w.bin" in cks[0]["resources"]
    # same external ids -> the same trial is continued, not duplicated
    core_v2.init(defaults=core_v2.DefaultConfig(name="unmanaged"), unmanaged=core_v2.UnmanagedConfig(external_experiment_id="ext-e", external_trial_id="ext-t"),
                 checkpoint_storage=str(tmp / "ck2"))
    try:
        assert core_v2.info.trial.trial_id == t["id"]
        assert core_v2.info.latest_checkpoint == cks[0]["uuid"]
    finally:
        core_v2.close()


class _FakeTrain:
    def __init__(self):
        self.training, self.validation, self.progress = [], [], []

    def report_training_metrics(self, steps_completed, metrics):
        self.training.append((steps_completed, metrics))

    def report_validation_metrics(self, steps_completed, metrics):
        self.validation.append((steps_completed, metrics))

    def report_progress(self, p):
        self.progress.append(p)


def test_hf_det_callback_reports_and_stops(tmp_path):
    from determined_b200 import transformers as det_tf

    with core._dummy_init(checkpoint_storage=str(tmp_path / "ck"), searcher_length=6) as ctx:
        args = types.SimpleNamespace(max_steps=6, num_train_epochs=1, eval_steps=3, save_steps=3, output_dir=str(tmp_path / "out"),
                                     overwrite_output_dir=True, resume_from_checkpoint=None, deepspeed=None, logging_steps=1)
        cb = det_tf.DetCallback(ctx, args)
        state = types.SimpleNamespace(global_step=0, epoch=0.0, is_world_process_zero=True, max_steps=6, best_model_checkpoint=None)
        control = types.SimpleNamespace(should_training_stop=False, should_save=False, should_evaluate=False)
        seen = []
        orig_t, orig_v = ctx.train.report_training_metrics, ctx.train.report_validation_metrics
        ctx.train.report_training_metrics = lambda steps_completed, metrics: (seen.append(("t", steps_completed, metrics)), orig_t(steps_completed, metrics))
        ctx.train.report_validation_metrics = lambda steps_completed, metrics: (seen.append(("v", steps_completed, metrics)), orig_v(steps_completed, metrics))
        for step in range(1, 7):
            state.global_step = step
            cb.on_log(args, state, control, logs={"loss": 1.0 / step, "learning_rate": 1e-3})
            if step % 3 == 0:
                cb.on_log(args, state, control, logs={"eval_loss": 0.5 / step, "eval_runtime": 1.0})
                out = pathlib.Path(args.output_dir) / f"checkpoint-{step}"
                out.mkdir(parents=True, exist_ok=True)
                (out / "model.bin").write_bytes(b"x")
                cb.on_save(args, state, control)
            cb.on_step_end(args, state, control)
        kinds = [k for k, *_ in seen]
        assert kinds.count("t") == 6 and kinds.count("v") == 2
        assert control.should_training_stop          # the single searcher op (6 batches) is complete
        assert det_tf.get_metric_type({"eval_loss": 1}) == "eval" and det_tf.get_metric_type({"loss": 1}) == "train_progress"


def test_torch_batch_process(tmp_path):
    from determined_b200.pytorch import experimental as px

    seen = []

    class Proc(px.TorchBatchProcessor):
        def __init__(self, context):
            self.context = context
            self.model = context.prepare_model_for_inference(torch.nn.Linear(2, 1))
            self.n = 0

        def process_batch(self, batch, batch_idx):
            x = self.context.to_device(batch)
            with torch.no_grad():
                self.model(x)
            seen.append(batch_idx)
            self.n += 1

        def on_checkpoint_start(self):
            seen.append("ckpt")

        def on_finish(self):
            seen.append("done")

    ds = torch.utils.data.TensorDataset(torch.randn(20, 2))

    class DS(torch.utils.data.Dataset):
        def __len__(self):
            return 20

        def __getitem__(self, i):
            return ds[i][0]

    px.torch_batch_process(Proc, DS(), batch_size=4, checkpoint_interval=2, distributed_context=core.DummyDistributedContext())
    assert [s for s in seen if isinstance(s, int)] == [0, 1, 2, 3, 4]
    assert seen.count("ckpt") >= 2 and seen[-1] == "done"


def test_workload_sequencer_orders_train_validate_checkpoint(tmp_path):
    from determined_b200.layers import WorkloadSequencer

    exp_conf = {"searcher": {"name": "single", "metric": "loss", "smaller_is_better": True, "max_length": {"batches": 8}},
                "scheduling_unit": 2, "min_validation_period": {"batches": 4}, "min_checkpoint_period": {"batches": 4},
                "checkpoint_policy": "best", "records_per_epoch": 0}
    with core._dummy_init(checkpoint_storage=str(tmp_path / "ck"), searcher_length=8) as ctx:
        ctx.searcher._units = core.Unit.BATCHES
        seq = WorkloadSequencer(ctx, exp_conf, trial_id=1, exp_id=1, global_batch_size=4)
        kinds = []
        for wkld, respond in seq:
            kinds.append((wkld.kind.name, wkld.num_batches, wkld.total_batches_processed))
            if wkld.kind == workload.Workload.Kind.RUN_STEP:
                respond({"metrics": {"avg_metrics": {"loss": 1.0}, "batch_metrics": [{"loss": 1.0}] * wkld.num_batches}, "stop_requested": False})
            elif wkld.kind == workload.Workload.Kind.COMPUTE_VALIDATION_METRICS:
                respond({"metrics": {"validation_metrics": {"loss": 1.0 / (1 + wkld.total_batches_processed)}}, "stop_requested": False})
            else:
                respond({"uuid": str(uuid.uuid4())})
    names = [k for k, *_ in kinds]
    assert names[:2] == ["RUN_STEP", "RUN_STEP"]
    assert names.count("RUN_STEP") == 4 and names.count("COMPUTE_VALIDATION_METRICS") == 2
    assert names.count("CHECKPOINT_MODEL") >= 2
    i4 = names.index("COMPUTE_VALIDATION_METRICS")
    assert kinds[i4][2] == 4
    st = seq.get_state()
    assert st["steps_completed"] == 8 if "steps_completed" in st else True


def test_commands_run_through_scheduler_logs_and_kill(cluster, capsys):
    """NTSC: `det cmd run` schedules a generic task on slots, ships its output, reports the exit code;
    a long-running shell task holds its slots until killed."""
    import time

    from determined_b200.cli import cli

    m, sess, tmp = cluster
    rc = cli.main(["cmd", "run", "--slots", "1", "--", "python", "-c", "import os; print('hello from', os.environ['DET_TASK_TYPE'], os.environ['DET_SLOT_IDS'])"])
    out = capsys.readouterr().out
    assert rc == 0 and "hello from COMMAND [" in out and "Launched command" in out
    failing = sess.post("/api/v1/commands", json={"config": {"entrypoint": "exit 3"}}).json()["command"]
    deadline = time.time() + 30
    while time.time() < deadline:
        c = sess.get(f"/api/v1/commands/{failing['id']}").json()["command"]
        if c["state"] == "STATE_TERMINATED":
            break
        time.sleep(0.2)
    assert c["exitStatus"] == 3
    sh = sess.post("/api/v1/shells", json={"config": {"resources": {"slots": 2}}}).json()["shell"]
    deadline = time.time() + 30
    while time.time() < deadline and sess.get(f"/api/v1/shells/{sh['id']}").json()["shell"]["state"] != "STATE_RUNNING":
        time.sleep(0.2)
    used = [s for a in sess.get("/api/v1/agents").json()["agents"] for s in a["slots"].values() if s.get("container")]
    assert len(used) == 2
    assert [x["id"] for x in sess.get("/api/v1/shells").json()["shells"]] == [sh["id"]]
    assert cli.main(["shell", "kill", sh["id"]]) == 0
    deadline = time.time() + 30
    while time.time() < deadline and sess.get(f"/api/v1/shells/{sh['id']}").json()["shell"]["state"] != "STATE_TERMINATED":
        time.sleep(0.2)
    assert sess.get(f"/api/v1/shells/{sh['id']}").json()["shell"]["state"] == "STATE_TERMINATED"
    m.schedule_once()
    used = [s for a in sess.get("/api/v1/agents").json()["agents"] for s in a["slots"].values() if s.get("container")]
    assert used == []


def test_streams_subscribe_sync_upsert_delete(cluster):
    from determined_b200.common import streams

    m, sess, tmp = cluster
    w = sess.post("/api/v1/workspaces", json={"name": "ws-stream"}).json()["workspace"]
    p1 = sess.post(f"/api/v1/workspaces/{w['id']}/projects", json={"name": "p1", "description": "d"}).json()["project"]
    st = streams.Stream(streams.PollingStreamTransport(sess, poll_interval=0.1))
    st.subscribe(sync_id="a", projects=streams.ProjectSpec(workspace_id=w["id"]))
    assert next(st) == streams.Sync("a", False)
    first = next(st)
    assert isinstance(first, streams.wire.ProjectMsg) and first.id == p1["id"] and first.name == "p1"
    assert next(st) == streams.Sync("a", True)
    # online event: a new project in the subscribed workspace shows up without resubscribing
    p2 = sess.post(f"/api/v1/workspaces/{w['id']}/projects", json={"name": "p2", "description": ""}).json()["project"]
    ev = next(st)
    assert isinstance(ev, streams.wire.ProjectMsg) and ev.id == p2["id"]
    # resubscribe after a deletion: the cache's known keys produce a deletion message
    sess.delete(f"/api/v1/projects/{p1['id']}")
    st.subscribe(sync_id="b", projects=streams.ProjectSpec(workspace_id=w["id"]))
    got = [next(st) for _ in range(3)]
    assert got[0] == streams.Sync("b", False) and got[-1] == streams.Sync("b", True)
    assert isinstance(got[1], streams.wire.ProjectsDeleted) and got[1].keys == str(p1["id"])
    assert streams.encode_keys({1, 2, 3, 7}) == "1-3,7" and list(streams.range_encoded_keys("1-3,7")) == [1, 2, 3, 7]


def test_model_hub_base_transformer_trial_native_gpt(tmp_path):
    """model_hub's BaseTransformerTrial drives the native GPT through PyTorchTrial / Trainer on CPU."""
    from determined_b200 import pytorch
    from determined_b200.model_hub import huggingface as hf, utils as mh_utils

    class LMTrial(hf.BaseTransformerTrial):
        def _loader(self):
            g = torch.Generator().manual_seed(0)
            data = torch.randint(0, 512, (64, 33), generator=g)

            class DS(torch.utils.data.Dataset):
                def __len__(self):
                    return 64

                def __getitem__(self, i):
                    return {"input_ids": data[i]}

            return pytorch.DataLoader(DS(), batch_size=self.context.get_per_slot_batch_size())

        def build_training_data_loader(self):
            return self._loader()

        def build_validation_data_loader(self):
            return self._loader()

        def evaluate_batch(self, batch, batch_idx):
            ids = batch["input_ids"]
            _, loss = self.model(ids[:, :-1], ids[:, 1:])
            return {"val_loss": loss}

    hparams = {"model_mode": "native-gpt", "gpt_preset": "tiny", "use_apex_amp": False, "use_pretrained_weights": False, "global_batch_size": 8,
               "learning_rate": 3e-3, "weight_decay": 0.01, "num_warmup_steps": 2, "max_grad_norm": 1.0}
    exp_conf = {"searcher": {"name": "single", "metric": "val_loss", "max_length": {"batches": 12}}, "hyperparameters": hparams,
                "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ck")}}
    with pytorch.init(hparams=hparams, exp_conf=exp_conf) as ctx:
        trial = LMTrial(ctx)
        assert trial.hparams.num_training_steps == 12
        pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(12), validation_period=pytorch.Batch(6), reporting_period=pytorch.Batch(6))
    groups = hf.group_parameters_for_optimizer(trial.model, 0.1)
    assert groups[0]["weight_decay"] == 0.1 and groups[1]["weight_decay"] == 0.0 and len(groups[1]["params"]) > 0
    assert mh_utils.compute_num_training_steps({"searcher": {"max_length": {"epochs": 2}}, "records_per_epoch": 80}, 8) == 20
    d = mh_utils.AttrDict({"a": {"b": 1}})
    assert d.a.b == 1
    opt_kw, sched_kw = hf.default_parse_optimizer_lr_scheduler_kwargs({"learning_rate": 0.1, "num_training_steps": 5, "lr_scheduler_type": "cosine"})
    assert opt_kw.learning_rate == 0.1 and sched_kw.lr_scheduler_type == "cosine"


def test_execution_helpers(tmp_path):
    from determined_b200 import _execution as ex

    hp = ex._generate_random_hparam_values({"lr": {"type": "double", "minval": 0.1, "maxval": 0.2}, "n": {"type": "int", "minval": 1, "maxval": 3},
                                            "c": {"type": "categorical", "vals": ["a", "b"]}, "k": {"type": "const", "val": 7}, "lg": {"type": "log", "base": 10, "minval": -3, "maxval": -1},
                                            "nested": {"x": {"type": "const", "val": 1}}, "plain": 5})
    assert 0.1 <= hp["lr"] <= 0.2 and hp["n"] in (1, 2, 3) and hp["c"] in ("a", "b") and hp["k"] == 7 and 1e-3 <= hp["lg"] <= 1e-1
    assert hp["nested"] == {"x": 1} and hp["plain"] == 5
    t = ex._make_test_experiment_config({"searcher": {"name": "adaptive_asha", "metric": "m", "max_trials": 9}, "hyperparameters": {"k": {"type": "const", "val": 2}}})
    assert t["searcher"] == {"name": "single", "metric": "m", "max_length": {"batches": 1}} and t["max_restarts"] == 0 and t["hyperparameters"] == {"k": 2}
    cfg = ex._make_local_execution_exp_config({"resources": {"slots_per_trial": 8}, "name": "x"}, str(tmp_path), managed_training=False, test_mode=True)
    assert cfg["resources"] == {"slots_per_trial": 1, "native_parallel": False} and cfg["checkpoint_storage"]["host_path"] == str(tmp_path)
    ctx, cfg2, hp2 = ex._make_local_execution_env(False, True, {"hyperparameters": {"k": {"type": "const", "val": 3}}}, str(tmp_path))
    with ctx:
        assert hp2 == {"k": 3}
    with pytest.raises(BaseException, match="SystemExit"):
        with ex._local_execution_manager(tmp_path):
            raise SystemExit(1)


def test_task_proxy_and_port_registry(cluster):
    """A task that serves HTTP gets a port from the registry and is reachable through /proxy/<task id>/ (reference
    master/internal/proxy + portregistry)."""
    import requests as rq

    from determined_b200.master import portregistry

    reg = portregistry.PortRegistry(23500, 23510)
    a, b = reg.acquire(), reg.acquire()
    assert a != b and reg.in_use() == {a, b}
    reg.release(a)
    assert reg.in_use() == {b}

    m, sess = cluster[0], cluster[1]
    (pathlib.Path(m.work_dir)).mkdir(parents=True, exist_ok=True)
    serve = "import http.server,os;os.makedirs('pub',exist_ok=True);open('pub/hello.txt','w').write('hi from task');os.chdir('pub');" \
            "http.server.ThreadingHTTPServer(('127.0.0.1',int(os.environ['DET_SERVICE_PORT'])),http.server.SimpleHTTPRequestHandler).serve_forever()"
    cmd = sess.post("/api/v1/commands", json={"config": {"entrypoint": [sys.executable, "-c", serve], "proxy_ports": True}}).json()["command"]
    url = None
    for _ in range(100):
        c = sess.get(f"/api/v1/commands/{cmd['id']}").json()["command"]
        if c["state"] == "STATE_RUNNING" and c.get("serviceAddress"):
            url = m.url + c["serviceAddress"]
            r = rq.get(url + "hello.txt", headers={"Authorization": f"Bearer {sess.token}"}, timeout=5)
            if r.status_code == 200:
                break
        time.sleep(0.2)
    assert url is not None and r.status_code == 200 and r.text == "hi from task", (r.status_code, r.text)
    assert rq.get(url + "hello.txt", timeout=5).status_code == 401                      # the proxy is behind the master's auth
    assert rq.get(url + "hello.txt", params={"token": sess.token}, timeout=5).status_code == 200
    assert rq.get(m.url + "/proxy/nope/x", headers={"Authorization": f"Bearer {sess.token}"}, timeout=5).status_code == 404
    sess.post(f"/api/v1/commands/{cmd['id']}/kill")
    for _ in range(100):
        if sess.get(f"/api/v1/commands/{cmd['id']}").json()["command"]["state"] == "STATE_TERMINATED":
            break
        time.sleep(0.1)
    assert m.ports.in_use() == set()


def test_tracing_spans_nest_and_export_otlp(cluster, tmp_path):
    """Request spans of the master are children of the caller's span (W3C traceparent) and land in an OTLP JSON export."""
    from determined_b200.common import tracing

    out = tmp_path / "traces.jsonl"
    assert not tracing.enabled()
    with tracing.span("noop") as s0:          # disabled: no-op
        assert s0 is None
    tr = tracing.configure("test-service", f"file://{out}", flush_s=60)
    try:
        m, sess = cluster[0], cluster[1]
        with tracing.span("client.list", purpose="test") as root:
            sess.get("/api/v1/experiments")
            with pytest.raises(ValueError):
                with tracing.span("failing"):
                    raise ValueError("boom")
        tr.flush()
    finally:
        tracing.shutdown()
    spans = [s for line in out.read_text().splitlines() for rs in json.loads(line)["resourceSpans"] for ss in rs["scopeSpans"] for s in ss["spans"]]
    by = {s["name"]: s for s in spans}
    assert {"client.list", "failing", "GET /api/v1/experiments"} <= set(by)
    assert by["GET /api/v1/experiments"]["traceId"] == root.trace_id and by["GET /api/v1/experiments"]["parentSpanId"] == root.span_id
    assert by["failing"]["status"]["code"] == 2 and by["failing"]["parentSpanId"] == root.span_id
    assert {"key": "purpose", "value": {"stringValue": "test"}} in by["client.list"]["attributes"]


def test_top_level_trial_context_and_bases():
    import determined_b200 as det

    c = det.TrialContext.from_config({"hyperparameters": {"lr": {"type": "const", "val": 0.1}, "global_batch_size": 32}, "data": {"path": "/x"}})
    assert c.get_hparam("lr") == 0.1 and c.get_global_batch_size() == 32 and c.get_per_slot_batch_size() == 32
    assert c.get_data_config() == {"path": "/x"} and c.get_trial_id() == 0 and not c.get_stop_requested()
    c.set_stop_requested(True)
    assert c.get_stop_requested()
    with pytest.raises(ValueError, match="Could not find name"):
        c.get_hparam("missing")
    with pytest.raises(TypeError):
        det.LegacyTrial(c)                   # abstract
    with pytest.raises(TypeError):
        det.TrialController(c)
    assert det.EnvContext().managed_training is True


def test_idle_timeout_kills_a_quiet_service_and_proxy_traffic_keeps_it_alive(cluster):
    import requests as rq

    from determined_b200.master import commands

    assert commands.parse_duration("1h30m") == 5400 and commands.parse_duration(45) == 45 and commands.parse_duration("500ms") == 0.5
    assert commands.parse_duration(None) is None
    with pytest.raises(ValueError):
        commands.parse_duration("soon")
    m, sess = cluster[0], cluster[1]
    serve = "import http.server,os;http.server.ThreadingHTTPServer(('127.0.0.1',int(os.environ['DET_SERVICE_PORT'])),http.server.SimpleHTTPRequestHandler).serve_forever()"
    cmd = sess.post("/api/v1/commands", json={"config": {"entrypoint": [sys.executable, "-c", serve], "proxy_ports": True, "idle_timeout": "3s"}}).json()["command"]
    t = m.commands.tasks[cmd["id"]]
    for _ in range(100):
        if t.state == "RUNNING":
            break
        time.sleep(0.1)
    assert t.state == "RUNNING" and t.idle_timeout == 3.0
    # traffic through the proxy counts as activity
    for _ in range(4):
        time.sleep(1.0)
        rq.get(m.url + f"/proxy/{t.id}/", headers={"Authorization": f"Bearer {sess.token}"}, timeout=5)
        m.commands.reap_idle()
    assert t.state == "RUNNING", "a service in use must not be reaped"
    for _ in range(100):                       # then silence
        m.commands.reap_idle()
        if t.state == "TERMINATED":
            break
        time.sleep(0.2)
    assert t.state == "TERMINATED" and t.killed
    assert "idle for more than 3s" in "".join(l["log"] for l in sess.get(f"/api/v1/tasks/{t.id}/logs").json()["logs"])


def test_install_alias_never_executes_a_module_twice(tmp_path):
    """`import determined.x` must hand back the very module object of `determined_b200.x` (a second execution of a package
    under its real name would duplicate every class: isinstance checks and mock.patch targets would silently miss)."""
    import subprocess

    code = (
        "import sys, determined_b200\n"
        "import determined_b200.pytorch as rp, determined_b200.common as rc\n"
        "determined_b200.install_alias()\n"
        "import determined.pytorch as p, determined.common as c\n"
        "from determined.common import storage, api\n"
        "from determined import core\n"
        "assert p is rp and c is rc and sys.modules['determined_b200.common'] is rc and sys.modules['determined_b200.pytorch'] is rp\n"
        "assert p.PyTorchTrial is rp.PyTorchTrial and core is determined_b200.core\n"
        "import determined.common.storage as s2\n"
        "assert s2 is storage and s2 is sys.modules['determined_b200.common.storage']\n"
        "from unittest import mock\n"
        "for _ in range(2):\n"
        "    with mock.patch('determined.common.storage.SharedFSStorageManager') as m:\n"
        "        storage.from_string('/tmp/x')\n"
        "    m.assert_called_once_with(base_path='/tmp/x')\n"
        "print('ALIAS_OK')\n")
    r = subprocess.run([sys.executable, "-c", code], cwd=tmp_path, env=dict(__import__("os").environ, PYTHONPATH=str(pathlib.Path(__file__).resolve().parents[1])),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert "ALIAS_OK" in r.stdout, r.stdout[-2000:]


def test_model_hub_dataset_helpers_and_download(tmp_path):
    """default_load_dataset from local files without the `datasets` package, column pruning against forward(), download_url
    from a local HTTP server, and the reference's own model_hub unit tests through the `model_hub` alias."""
    import http.server
    import subprocess
    import sys
    import threading

    import torch

    from determined_b200.model_hub import huggingface as hf, utils as mh_utils

    (tmp_path / "train.csv").write_text("text,label\n" + "".join(f"row {i},{i % 2}\n" for i in range(10)))
    ds = hf.default_load_dataset({"train_file": str(tmp_path / "train.csv"), "validation_split_percentage": 20})
    assert len(ds["train"]) == 8 and len(ds["validation"]) == 2 and ds["train"].column_names == ["text", "label"]
    (tmp_path / "v.jsonl").write_text('{"text": "a", "label": 1}\n{"text": "b", "label": 0}\n')
    (tmp_path / "t.jsonl").write_text('{"text": "c", "label": 1}\n')
    ds = hf.default_load_dataset({"train_file": str(tmp_path / "t.jsonl"), "validation_file": str(tmp_path / "v.jsonl")})
    assert len(ds["validation"]) == 2 and ds["validation"][1] == {"text": "b", "label": 0}
    # the built-in column store (what loads the files when the `datasets` package is absent)
    val = hf.ColumnDataset.from_file(str(tmp_path / "v.jsonl"))
    assert len(val) == 2 and val[1] == {"text": "b", "label": 0} and val.select([1])[0]["text"] == "b"
    assert hf.ColumnDataset.from_file(str(tmp_path / "train.csv"))[3] == {"text": "row 3", "label": "1"}
    tok = val.map(lambda r: {"input_ids": [ord(c) for c in r["text"]], "labels": int(r["label"])})

    class M(torch.nn.Module):
        def forward(self, input_ids, attention_mask=None, labels=None):
            return input_ids

    tok.set_format("torch", tok.column_names)
    hf.remove_unused_columns(M(), tok)
    assert set(tok[0]) == {"input_ids", "labels", "label"} and torch.is_tensor(tok[0]["input_ids"]) and tok.format["type"] == "torch"
    grouped = val.map(lambda b: {"chunk": ["".join(b["text"])]}, batched=True)
    assert grouped.column_names == ["chunk"] and len(grouped) == 1

    (tmp_path / "www").mkdir()
    (tmp_path / "www" / "weights.bin").write_bytes(b"x" * 3000)
    handler = lambda *a, **kw: http.server.SimpleHTTPRequestHandler(*a, directory=str(tmp_path / "www"), **kw)  # noqa: E731
    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), handler)
    threading.Thread(target=srv.serve_forever, daemon=True).start()
    try:
        p = mh_utils.download_url(str(tmp_path / "dl"), f"http://127.0.0.1:{srv.server_address[1]}/weights.bin?sig=1")
        assert p.endswith("/dl/weights.bin") and open(p, "rb").read() == b"x" * 3000
        srv.shutdown()
        assert mh_utils.download_url(str(tmp_path / "dl"), f"http://127.0.0.1:{srv.server_address[1]}/weights.bin") == p, "a file already there is reused"
    finally:
        srv.server_close()

    runner = pathlib.Path(__file__).parent / "fixtures" / "run_reference_tests.py"
    r = subprocess.run([sys.executable, str(runner), "/root/reference/model_hub", "tests/test_hf.py", "tests/test_utils.py", "-k", "not download_url"],
                       cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "4 passed" in r.stdout, r.stdout[-2000:]


def test_import_from_path_refuses_nesting_and_writes_no_bytecode(tmp_path):
    import determined_b200 as det

    for name in ("a", "b"):
        (tmp_path / name).mkdir()
        (tmp_path / name / "model_def_x.py").write_text(f"val = {name!r}\n")
    with det.import_from_path(tmp_path / "a"):
        import model_def_x as a
        with pytest.raises(RuntimeError, match="does not support nesting"):
            with det.import_from_path(tmp_path / "b"):
                pass
    with det.import_from_path(tmp_path / "b"):
        import model_def_x as b
    assert (a.val, b.val) == ("a", "b") and "model_def_x" not in sys.modules
    assert not (tmp_path / "a" / "__pycache__").exists() and not (tmp_path / "b" / "__pycache__").exists()
