"""Import paths of the reference's ``determined.common.experimental`` package: the SDK classes live in
``determined_b200.experimental.client``; these modules re-export them under the names user code and type annotations
written for the reference use (``from determined.common.experimental import checkpoint`` ... ``checkpoint.Checkpoint``)."""
from determined_b200.experimental.client import (  # noqa: F401
    Checkpoint, Determined, Experiment, ExperimentReference, Model, ModelOrderBy, ModelSortBy, ModelVersion, OrderBy, Project, TrainingMetrics, Trial,
    TrialMetrics, TrialOrderBy, TrialReference, TrialSortBy, ValidationMetrics)
