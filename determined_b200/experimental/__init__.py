from typing import Any

from determined_b200.experimental import client  # noqa: F401
from determined_b200.experimental.client import (  # noqa: F401
    Checkpoint, CheckpointOrderBy, CheckpointSortBy, CheckpointState, Determined, DownloadMode, Experiment, ExperimentOrderBy, ExperimentSortBy,
    ExperimentState, Model, ModelOrderBy, ModelSortBy, ModelVersion, OrderBy, Project, ResourcePool, TrainingMetrics, Trial, TrialMetrics, TrialOrderBy,
    TrialSortBy, TrialState, User, ValidationMetrics, Workspace,
)
from determined_b200.experimental.client import ExperimentReference, TrialReference  # noqa: E402,F401



def test_one_batch(trial_class: Any, config: Any = None) -> Any:
    """Train and validate one batch of ``trial_class`` in this process (reference ``experimental/_native.py``): the quickest
    check that a trial definition runs at all."""
    import logging

    from determined_b200 import pytorch

    config = {**(config or {}), "scheduling_unit": 1}
    if not (isinstance(trial_class, type) and issubclass(trial_class, pytorch.PyTorchTrial)):
        raise TypeError("test_one_batch runs PyTorchTrial (and DeepSpeedTrial) subclasses")
    with pytorch.init(hparams=config.get("hyperparameters", {}), exp_conf=config) as ctx:
        trial = trial_class(ctx)
        pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(1), test_mode=True)
    logging.getLogger("determined").info("The test experiment passed.")
    return trial


test_one_batch.__test__ = False      # a library function, not a test, whatever its name says to pytest
