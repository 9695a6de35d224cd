from determined_b200.experimental.client import Model, ModelOrderBy, ModelSortBy, ModelVersion  # noqa: F401
