"""Support libraries shared by the harness, CLI and the single-box master (reference L2 layer)."""
from determined_b200.common import util  # noqa: E402,F401
