from determined_b200.experimental.client import Project  # noqa: F401
