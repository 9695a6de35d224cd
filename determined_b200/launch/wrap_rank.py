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
    """Copy ``src`` to ``dst`` line by line with ``prefix``; a carriage return ends a line too (progress bars redraw with
    ``\r`` and would otherwise sit in the buffer until the next newline), and every emitted line ends with ``\n``."""
    fd = src.fileno()
    pending = b""
    while True:
        chunk = os.read(fd, 65536)
        if not chunk:
            break
        pending += chunk
        out = []
        while True:
            cut = min((i for i in (pending.find(b"\n"), pending.find(b"\r")) if i >= 0), default=-1)
            if cut < 0:
                break
            was_cr = pending[cut:cut + 1] == b"\r"
            line, pending = pending[:cut], pending[cut + 1:]
            if was_cr and pending[:1] == b"\n":
                pending = pending[1:]           # \r\n is one line end
            out.append(prefix + line + b"\n")
        if out:
            dst.write(b"".join(out))
            dst.flush()
    if pending:
        dst.write(prefix + pending + b"\n")
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
        if name.isdigit():                      # a literal rank instead of a variable name
            rank = name
            break
        if name in os.environ:
            rank = os.environ[name]
            break
    prefix = f"[rank={rank}] ".encode() if rank is not None else b""
    proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    ts = [threading.Thread(target=forward, args=(proc.stdout, sys.stdout.buffer, prefix), daemon=True),
          threading.Thread(target=forward, args=(proc.stderr, sys.stderr.buffer, prefix), daemon=True)]
    for t in ts:
        t.start()
    try:
        from determined_b200 import util
    except ImportError:                         # run as a plain script file (python .../wrap_rank.py): find the package next to it
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
        from determined_b200 import util

    with util.forward_signals(proc):
        rc = proc.wait()
    for t in ts:
        t.join(timeout=5)
    return rc


if __name__ == "__main__":
    sys.exit(main())
