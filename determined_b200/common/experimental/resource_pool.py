from determined_b200.experimental.client import ResourcePool  # noqa: F401
