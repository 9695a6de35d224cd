from determined_b200.experimental import client  # noqa: F401
from determined_b200.experimental.client import (  # noqa: F401
    Checkpoint, CheckpointOrderBy, CheckpointSortBy, CheckpointState, Determined, DownloadMode, Experiment, ExperimentOrderBy, ExperimentSortBy,
    ExperimentState, Model, ModelOrderBy, ModelSortBy, ModelVersion, OrderBy, Project, ResourcePool, TrainingMetrics, Trial, TrialMetrics, TrialOrderBy,
    TrialSortBy, User, ValidationMetrics, Workspace,
)
