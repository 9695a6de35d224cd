"""``python -m determined.exec.pid_client ADDR -- cmd…`` (reference ``exec/pid_client.py:1-15``)."""
import sys
from typing import List, Optional

from determined_b200 import ipc


def main(argv: Optional[List[str]] = None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    addr_s, rest = argv[0], argv[1:]
    if rest and rest[0] == "--":
        rest = rest[1:]
    addr = int(addr_s) if addr_s.isdigit() else addr_s
    with ipc.PIDClient(addr) as client:
        rc = client.run_subprocess(rest)
        if rc != 0:
            # non-zero: do not send the graceful-exit byte
            client.close(graceful=False)
            client.sock = None
        return rc


if __name__ == "__main__":
    sys.exit(main())
