"""Unmodified user code of the reference runs on this harness: the trial classes of the reference's own test fixtures
(``/root/reference/harness/tests/experiment/fixtures``) are imported as they are, with ``import determined`` aliased to this
package by ``determined_b200.install_alias()``, and trained through ``pytorch.init`` + ``Trainer.fit``.  Skipped where the
reference checkout is not present."""
import json
import pathlib
import subprocess
import sys

import pytest

REF = pathlib.Path("/root/reference/harness")
RUNNER = pathlib.Path(__file__).parent / "fixtures" / "run_reference_trial.py"
pytestmark = pytest.mark.skipif(not (REF / "tests" / "experiment" / "fixtures" / "pytorch_onevar_model.py").exists(), reason="reference checkout not available")

CASES = [
    ("pytorch_onevar_model", "OneVarTrial", 8, {}),                              # custom reducers (class + function), callbacks, analytic checks inside train_batch
    ("pytorch_onevar_model", "OneVarTrialWithMultiValidation", 8, {}),
    ("pytorch_onevar_model", "OneVarTrialWithTrainingMetrics", 8, {}),
    ("pytorch_onevar_model", "OneVarTrialCustomEval", 8, {}),                    # evaluate_full_dataset
    ("pytorch_onevar_model", "OneVarTrialGradClipping", 8, {}),                  # clip_grads callables
    ("pytorch_onevar_model", "OneVarTrialWithNonScalarValidation", 8, {}),
    ("pytorch_onevar_model", "OneVarTrialWithLRScheduler", 8, {"lr_scheduler_step_mode": 1}),
    ("pytorch_onevar_model", "OneVarTrialAccessContext", 1, {"lr_scheduler_step_mode": 1}),      # two models / optimizers / schedulers on one context
    ("pytorch_onevar_model", "OneVarTrialCallbacks", 8, {}),                     # includes a callback with the pre-0.17.8 epoch-start signature
    ("pytorch_onevar_model", "OneVarAutoAMPTrial", 4, {}),
    ("pytorch_onevar_model", "OneVarManualAMPTrial", 4, {}),
]


def test_reference_fixture_trials_run_unmodified(tmp_path):
    r = subprocess.run([sys.executable, str(RUNNER), str(REF), "--cases", json.dumps(CASES)], cwd=tmp_path,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    for _, cls, _, _ in CASES:
        assert f"REFERENCE_TRIAL_OK {cls}" in r.stdout, f"{cls} failed:\n" + r.stdout[-4000:]
    assert r.returncode == 0, r.stdout[-3000:]


REF_TESTS = [
    # (file, -k expression, minimum number of tests that must pass)
    ("tests/core/test_searcher.py", "", 5),
    ("tests/core/test_preempt.py", "", 14),
    ("tests/test_util.py", "", 11),
    ("tests/test_ipc.py", "PIDServer or distributed_context", 14),      # the one left out drives ZeroMQ classes directly
    ("tests/experiment/pytorch/test_pytorch_data.py", "", 24),          # samplers, adapt_batch_sampler, to_device
    ("tests/experiment/pytorch/test_reducer.py", "", 5),
    ("tests/storage/test_storage_base.py", "not shared_fs_shortcut_dict", 21),   # the one left out calls a private helper of core._context
    ("tests/storage/test_shared_fs.py", "", 8),          # whole file: path rules, lifecycle + selectors, sharded upload / download / store / restore
                                                          # from 4 thread-ranks, selector-aware copytree (with symlinks), incremental TensorBoard fetcher
    # the reference's storage conformance helpers (tests/storage/util.py) against the S3 / GCS / Azure managers over in-memory SDK stand-ins
    (str(pathlib.Path(__file__).resolve().parent.parent / "conformance" / "ref_cloud_conformance.py"), "", 10),
    # the reference's PyTorchTrial suite: checkpoint/restore, reproducibility, clipping, reducers, callbacks, aggregation, AMP, Trainer,
    # checkpoint policies, test mode, RNG restore, and every ancient-checkpoints/*-pytorch* fixture through
    # load_trial_from_checkpoint_path + the 0.20.0 legacy controller load.  Left out: gpu_parallel (run in tests/test_multi_gpu.py's
    # own form) and five tests that call private methods of the reference controller (_train_with_boundaries, _log_tb_metrics).
    ("tests/experiment/pytorch/test_pytorch_trial.py",
     "not gpu_parallel and not test_onevar_single and not test_training_metrics and not test_variable_workload_size and not "
     "test_reject_unnamed_nondict_metric and not test_trainer_disable_tb_logging", 36),
    # client-side custom search: the runner loop against the reference's mock master (typed searcher events from the bindings)
    ("tests/test_custom_searcher.py", "", 2),
    # the DSAT simulation suite: trial / tracker bookkeeping, random / binary / ASHA methods end to end against a mock master,
    # lineage pruning rules, rung promotion, HF argument overwriting
    ("tests/experiment/pytorch/test_deepspeed_autotuning.py", "", 46),
    ("tests/common/api/test_certs.py", "", 1),                                  # the one-certificate layout moves into the per-master store
    ("tests/common/api/test_authentication.py", "", 2),                         # auth.json of versions 0 and 1 are read and merged
    ("tests/launch/test_launch.py", "not launch_trial", 3),                     # that one pins the Horovod command line; the launch script runs through runpy
    ("tests/test_sdk.py", "", 6),                                               # SDK enums cover exactly the values of the bindings' enums
    ("tests/common/test_yaml.py", "", 4),
    ("tests/common/test_api.py", "", 15),                                       # master URL canonicalisation
    ("tests/launch/test_wrap_rank.py", "", 1),
    ("tests/tensorboard/test_util.py", "not nonexistent_directory", 9),        # that one needs a path that does not exist on the test machine
    ("tests/tensorboard/test_base.py", "not upload_thread", 7),                # that one constructs a private upload-thread class
    ("tests/core/test_checkpoint.py", "not test_checkpoint_context", 16),      # the one left out asserts on private call sites (mocks)
]


@pytest.mark.parametrize("path,expr,n", REF_TESTS, ids=[t[0].split("/")[-1] for t in REF_TESTS])
def test_reference_unit_tests_pass_against_this_package(path, expr, n, tmp_path):
    """The reference's own unit tests of the searcher / preemption contexts, utilities, DistributedContext and the PID
    watchdog, run as they are against ``determined_b200`` under the ``determined`` alias."""
    import re

    args = [sys.executable, str(RUNNER.with_name("run_reference_tests.py")), str(REF), path] + (["-k", expr] if expr else [])
    r = subprocess.run(args, cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1200)
    m = re.search(r"(\d+) passed", r.stdout)
    assert r.returncode == 0 and m and int(m.group(1)) >= n and " failed" not in r.stdout.splitlines()[-1], r.stdout[-3000:]
