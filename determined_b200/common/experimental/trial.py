from determined_b200.experimental.client import Trial, TrialOrderBy, TrialReference, TrialSortBy  # noqa: F401
