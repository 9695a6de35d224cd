from determined_b200.experimental.client import Determined  # noqa: F401
