from determined_b200.experimental.client import Workspace  # noqa: F401
