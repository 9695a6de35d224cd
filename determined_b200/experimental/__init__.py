from determined_b200.experimental import client  # noqa: F401
from determined_b200.experimental.client import (  # noqa: F401
    Checkpoint, CheckpointState, Determined, DownloadMode, Experiment, ExperimentState, Model, ModelSortBy, ModelVersion, Trial,
)
