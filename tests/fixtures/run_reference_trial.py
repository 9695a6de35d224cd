"""Run an UNMODIFIED trial class of the reference's own test fixtures through this harness (``import determined`` is aliased).

usage: run_reference_trial.py <reference harness dir> <module under tests.experiment.fixtures> <TrialClass> <batches> [hparams json]
   or: run_reference_trial.py <reference harness dir> --cases '<json list of [module, TrialClass, batches, hparams]>'
"""
import json
import pathlib
import sys
import tempfile

ref_harness = sys.argv[1]
if sys.argv[2] == "--cases":
    cases = [tuple(c) for c in json.loads(sys.argv[3])]
else:
    cases = [(sys.argv[2], sys.argv[3], int(sys.argv[4]), json.loads(sys.argv[5]) if len(sys.argv) > 5 else {})]
repo = str(pathlib.Path(__file__).resolve().parents[2])
sys.path[:] = [p for p in sys.path if pathlib.Path(p or ".").resolve() != pathlib.Path(repo, "tests", "fixtures").resolve()]
sys.path.insert(0, ref_harness)          # for the reference's `tests.experiment.fixtures` package only
sys.path.insert(1, repo)

import determined_b200  # noqa: E402

determined_b200.install_alias()           # from here on `import determined` is this package

import importlib  # noqa: E402

import determined as det  # noqa: E402
from determined import pytorch  # noqa: E402

assert det is determined_b200 and pytorch.__name__.startswith("determined_b200"), "alias did not take"
def run_case(module, cls_name, batches, extra):
    mod = importlib.import_module(f"tests.experiment.fixtures.{module}")
    assert mod.__file__.startswith(ref_harness), mod.__file__
    trial_cls = getattr(mod, cls_name)
    hp = {"hidden_size": 2, "learning_rate": 0.5, "global_batch_size": 4, "dataloader_type": "determined"}
    hp.update(extra)
    tmp = tempfile.mkdtemp()
    with pytorch.init(hparams=hp, exp_conf={"checkpoint_storage": {"type": "shared_fs", "host_path": tmp}, "hyperparameters": hp,
                                            "searcher": {"name": "single", "metric": "val_loss", "smaller_is_better": True}}) as ctx:
        trial = trial_cls(ctx)
        pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(batches), validation_period=pytorch.Batch(max(batches // 2, 1)), checkpoint_policy="none")
    print("REFERENCE_TRIAL_OK", cls_name, float(trial.model.weight.data.flatten()[0]) if hasattr(trial, "model") else "")



failed = 0
for module, cls_name, batches, extra in cases:
    try:
        run_case(module, cls_name, int(batches), extra)
    except Exception:  # noqa: BLE001 - report every case, fail at the end
        import traceback

        traceback.print_exc()
        print("REFERENCE_TRIAL_FAILED", cls_name)
        failed += 1
sys.exit(1 if failed else 0)
