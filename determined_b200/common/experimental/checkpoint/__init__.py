from determined_b200.experimental.client import Checkpoint, CheckpointOrderBy, CheckpointSortBy, CheckpointState, DownloadMode  # noqa: F401
