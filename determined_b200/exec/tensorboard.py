"""Tensorboard task: fetch the tfevents of the requested experiments/trials and serve them.

Parity: reference ``harness/determined/exec/tensorboard.py`` (``start_tensorboard``: resolve the
trials, start per-storage fetchers syncing ``cluster_id/tensorboard/experiment/<e>/trial/<t>/`` into
a local directory, exec ``tensorboard --logdir``).  When the ``tensorboard`` package is not
installed (this image) the synced event files are served as a plain directory listing instead, so
the task still schedules, logs and exposes the data.
"""
import argparse
import functools
import http.server
import logging
import os
import pathlib
import shutil
import subprocess
import sys
import tempfile
import threading
import time
from typing import List

import determined_b200 as det
from determined_b200.common import api

logger = logging.getLogger("determined.exec.tensorboard")


def resolve_trials(sess: api.Session, exp_ids: List[int], trial_ids: List[int]) -> List[dict]:
    out = []
    for e in exp_ids:
        out += sess.get(f"/api/v1/experiments/{e}/trials").json()["trials"]
    for t in trial_ids:
        out.append(sess.get(f"/api/v1/trials/{t}").json()["trial"])
    return out


def sync_once(sess: api.Session, trials: List[dict], dest: pathlib.Path) -> int:
    """Copy tfevents of shared_fs-backed trials under ``dest/trial_<id>/``; returns files copied."""
    n = 0
    for t in trials:
        exp = sess.get(f"/api/v1/experiments/{t['experimentId']}").json()["experiment"]
        storage = (exp["config"].get("checkpoint_storage") or {})
        if storage.get("type") != "shared_fs":
            continue
        base = pathlib.Path(storage["host_path"]) / (storage.get("tensorboard_path") or "")
        info = sess.get("/api/v1/master").json()
        src = base / info.get("clusterId", "cluster") / "tensorboard" / "experiment" / str(t["experimentId"]) / "trial" / str(t["id"])
        if not src.exists():
            continue
        for f in src.rglob("*"):
            if f.is_file():
                dst = dest / f"trial_{t['id']}" / f.relative_to(src)
                if not dst.exists() or dst.stat().st_size != f.stat().st_size:
                    dst.parent.mkdir(parents=True, exist_ok=True)
                    shutil.copy2(f, dst)
                    n += 1
    return n


def main(argv: List[str]) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--experiments", default="")
    ap.add_argument("--trials", default="")
    ap.add_argument("--port", type=int, default=0)
    ap.add_argument("--sync-interval", type=float, default=10.0)
    a = ap.parse_args(argv)
    info = det.get_cluster_info()
    master = info.master_url if info is not None else os.environ.get("DET_MASTER", "http://127.0.0.1:8080")
    token = info.session_token if info is not None else os.environ.get("DET_SESSION_TOKEN")
    sess = api.Session(master, None, token, None)
    exp_ids = [int(x) for x in a.experiments.split(",") if x]
    trial_ids = [int(x) for x in a.trials.split(",") if x]
    trials = resolve_trials(sess, exp_ids, trial_ids)
    logdir = pathlib.Path(tempfile.mkdtemp(prefix="detb200-tb-"))
    print(f"tensorboard task: {len(trials)} trial(s), logdir {logdir}", flush=True)
    sync_once(sess, trials, logdir)

    def loop() -> None:
        while True:
            time.sleep(a.sync_interval)
            try:
                sync_once(sess, trials, logdir)
            except Exception as e:     # keep serving what we have
                logger.warning("sync failed: %s", e)

    threading.Thread(target=loop, daemon=True).start()
    try:
        import tensorboard  # noqa: F401

        return subprocess.call([sys.executable, "-m", "tensorboard.main", "--logdir", str(logdir), "--port", str(a.port or 6006), "--bind_all"])
    except ImportError:
        handler = functools.partial(http.server.SimpleHTTPRequestHandler, directory=str(logdir))
        srv = http.server.ThreadingHTTPServer(("127.0.0.1", a.port), handler)
        print(f"tensorboard package not installed; serving event files at http://127.0.0.1:{srv.server_address[1]}/", flush=True)
        srv.serve_forever()
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
