from determined_b200.experimental.client import Experiment, ExperimentOrderBy, ExperimentReference, ExperimentSortBy, ExperimentState  # noqa: F401
