"""``python -m determined.exec.pid_server [--on-fail SIG] [--on-exit SIG|WAIT] [--grace-period N]
ADDR NUM_WORKERS -- cmd…`` (reference ``harness/determined/exec/pid_server.py:1-51``)."""
import argparse
import signal
import sys
from typing import List, Optional

from determined_b200 import ipc


def _sig(name: str) -> Optional[signal.Signals]:
    if name in ("WAIT", "NONE", ""):
        return None
    return getattr(signal, name if name.startswith("SIG") else "SIG" + name)


def main(argv: Optional[List[str]] = None) -> int:
    p = argparse.ArgumentParser()
    p.add_argument("--on-fail", default="SIGTERM")
    p.add_argument("--on-exit", default="WAIT")
    p.add_argument("--grace-period", "-g", type=int, default=3)
    p.add_argument("--signal-children", "-x", action="store_true")
    p.add_argument("addr")
    p.add_argument("num_workers", type=int)
    p.add_argument("cmd", nargs=argparse.REMAINDER)
    a = p.parse_args(argv)
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    addr = int(a.addr) if a.addr.isdigit() else a.addr
    with ipc.PIDServer(addr, a.num_workers) as srv:
        return srv.run_subprocess(cmd, on_fail=_sig(a.on_fail), on_exit=_sig(a.on_exit), grace_period=a.grace_period,
                                  signal_children=a.signal_children)


if __name__ == "__main__":
    sys.exit(main())
