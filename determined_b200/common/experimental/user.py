from determined_b200.experimental.client import User  # noqa: F401
