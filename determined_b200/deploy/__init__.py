"""``det deploy``: cluster bring-up.  Only ``local`` exists for the single-box master; the reference's
aws / gcp / gke deployers (``harness/determined/deploy/{aws,gcp,gke}``) drive cloud APIs that have no
meaning for a one-node B200 box and are reported as unsupported by the CLI."""
from determined_b200.deploy import local  # noqa: F401
