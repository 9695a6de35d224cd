from determined_b200.experimental.client import OrderBy  # noqa: F401
