"""``python -m determined.launch.wrap_rank RANK[,ALT…] -- cmd…`` — prefix each output line of the
wrapped command with ``[rank=N]``.  ``RANK`` names environment variables to read the rank from
(first one that is set), e.g. ``RANK,OMPI_COMM_WORLD_RANK``.

Parity: reference ``harness/determined/launch/wrap_rank.py:1-187``.  (The native launcher does
this itself; this module is for the torchrun fallback path and for users' own launch layers.)"""
import argparse
import os
import subprocess
import sys
import threading
from typing import IO, List, Optional


def forward(src: IO[bytes], dst: IO[bytes], prefix: bytes) -> None:
    for line in iter(src.readline, b""):
        if not line.endswith(b"\n"):
            line += b"\n"
        dst.write(prefix + line)
        dst.flush()


def main(argv: Optional[List[str]] = None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    p = argparse.ArgumentParser(prog="python3 -m determined.launch.wrap_rank")
    p.add_argument("--no-redirect-stdio", action="store_true")
    p.add_argument("rank_var_name")
    p.add_argument("cmd", nargs=argparse.REMAINDER)
    a = p.parse_args(argv)
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    if not cmd:
        p.error("a command is required")
    rank = None
    for name in a.rank_var_name.split(","):
        if name in os.environ:
            rank = os.environ[name]
            break
    prefix = f"[rank={rank}] ".encode() if rank is not None else b""
    proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    ts = [threading.Thread(target=forward, args=(proc.stdout, sys.stdout.buffer, prefix), daemon=True),
          threading.Thread(target=forward, args=(proc.stderr, sys.stderr.buffer, prefix), daemon=True)]
    for t in ts:
        t.start()
    from determined_b200 import util

    with util.forward_signals(proc):
        rc = proc.wait()
    for t in ts:
        t.join(timeout=5)
    return rc


if __name__ == "__main__":
    sys.exit(main())
