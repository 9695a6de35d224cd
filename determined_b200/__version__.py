__version__ = "0.1.0"
# Checkpoint/expconf contracts follow the reference at this version (VERSION file of the reference).
REFERENCE_COMPAT_VERSION = "0.31.1-dev0"
