"""Wait for a master to answer (reference ``harness/determined/deploy/healthcheck.py``)."""
import time

import requests

from determined_b200.deploy import errors

DEFAULT_TIMEOUT = 100


def wait_for_master(master_url: str, timeout: int = DEFAULT_TIMEOUT) -> None:
    deadline = time.time() + timeout
    while time.time() < deadline:
        try:
            r = requests.get(master_url.rstrip("/") + "/info", timeout=2)
            if r.status_code == 200:
                return
        except requests.RequestException:
            pass
        time.sleep(0.3)
    raise errors.MasterTimeoutExpired(f"master at {master_url} did not come up within {timeout}s")
