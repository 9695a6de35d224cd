"""Deploy errors (reference ``harness/determined/deploy/errors.py``)."""


class MasterTimeoutExpired(Exception):
    pass


class PreflightFailure(Exception):
    pass
