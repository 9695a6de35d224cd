"""``det`` command line (reference ``harness/determined/cli``)."""
from determined_b200.cli.cli import main  # noqa: F401
