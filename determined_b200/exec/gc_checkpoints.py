"""``python -m determined.exec.gc_checkpoints --storage-config JSON --delete UUID[,UUID…]
[--globs …]`` — delete checkpoint files and report what is left.

Parity: reference ``harness/determined/exec/gc_checkpoints.py:22-175``.  The single-box master
runs the same logic in-process (``master.core.Master.run_checkpoint_gc``); this entrypoint exists
for operators and for parity with the reference's GC task."""
import argparse
import json
import logging
import sys
from typing import Dict, List

from determined_b200.common import api, storage

logger = logging.getLogger("determined.exec.gc")


def delete_checkpoints(manager: storage.StorageManager, uuids: List[str], globs: List[str]) -> Dict[str, Dict[str, int]]:
    left: Dict[str, Dict[str, int]] = {}
    for u in uuids:
        try:
            left[u] = manager.delete(u, globs)
        except Exception as e:
            logger.error("failed to delete %s: %s", u, e)
    return left


def main(argv=None) -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--storage-config", required=True, help="JSON or @file")
    p.add_argument("--delete", required=True, help="comma-separated checkpoint uuids")
    p.add_argument("--globs", nargs="*", default=["**/*"])
    p.add_argument("--master-url")
    p.add_argument("--token")
    a = p.parse_args(argv)
    raw = open(a.storage_config[1:]).read() if a.storage_config.startswith("@") else a.storage_config
    mgr = storage.build(json.loads(raw), None)
    left = delete_checkpoints(mgr, [u for u in a.delete.split(",") if u], a.globs)
    if a.master_url:
        sess = api.Session(a.master_url, None, a.token, None)
        sess.patch("/api/v1/checkpoints", json={"checkpoints": [{"uuid": u, "resources": {"resources": {k: str(v) for k, v in r.items()}}} for u, r in left.items()]})
    print(json.dumps(left))
    return 0


if __name__ == "__main__":
    sys.exit(main())
