"""``python -m determined.exec.prep_container [--download_context_directory] [--rendezvous]
[--resources] [--proxy]`` — write the task info files.

Parity: reference ``harness/determined/exec/prep_container.py:28-368``.  The single-box master
extracts the model definition and passes everything through the environment, so this is only
needed by users' custom launch layers: it downloads the context directory, asks the master for
rendezvous info and writes ``$DET_INFO_DIR/{cluster,trial,rendezvous,resources}.json``."""
import argparse
import base64
import io
import logging
import os
import sys
import tarfile

import determined_b200 as det
from determined_b200 import _info
from determined_b200.common import api


def main(argv=None) -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--download_context_directory", action="store_true")
    p.add_argument("--rendezvous", action="store_true")
    p.add_argument("--resources", action="store_true")
    p.add_argument("--proxy", action="store_true")
    p.add_argument("--trial", action="store_true")
    a = p.parse_args(argv)
    info = _info.ClusterInfo._from_env()
    sess = api.Session(info.master_url, None, info.session_token, None)
    os.makedirs(_info._info_dir(), exist_ok=True)
    info._to_file()
    if info._trial_info is not None:
        info._trial_info._to_file()
    if a.download_context_directory:
        r = sess.get(f"/api/v1/tasks/{info.task_id}/context_directory").json()
        blob = base64.b64decode(r.get("b64Tgz") or "")
        dst = os.environ.get("DET_MODEL_DIR", ".")
        if blob:
            with tarfile.open(fileobj=io.BytesIO(blob), mode="r:*") as tf:
                tf.extractall(dst, filter="data")
    if a.resources:
        _info.ResourcesInfo._by_inspection()._to_file()
    if a.rendezvous:
        rid = os.environ.get("DET_RESOURCES_ID", "box")
        r = sess.get(f"/api/v1/allocations/{info.allocation_id}/resources/{rid}/rendezvous").json()["rendezvousInfo"]
        _info.RendezvousInfo(r["addresses"], int(r["rank"]), [int(s) for s in r.get("slots", [])])._to_file()
    if a.proxy:
        sess.post(f"/api/v1/allocations/{info.allocation_id}/proxy_address", json={"proxyAddress": "127.0.0.1"})
    return 0


if __name__ == "__main__":
    sys.exit(main())
