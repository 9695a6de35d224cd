"""Pack a model-definition directory into the ``v1File`` list that ``CreateExperiment`` takes
(reference ``harness/determined/common/context.py:71-140``): honours ``.detignore``, refuses
contexts over 96 MB, records mode/mtime, directories as type 53 ('5'), files as type 48 ('0')."""
import base64
import os
import pathlib
from typing import Any, Dict, List

from determined_b200 import util

LEGACY_MAX_CONTEXT_SIZE = 96 * 1024 * 1024


def read_v1_context(root: pathlib.Path, limit: int = LEGACY_MAX_CONTEXT_SIZE) -> List[Dict[str, Any]]:
    root = pathlib.Path(root)
    if not root.exists():
        raise ValueError(f"Path '{root}' doesn't exist")
    if root.is_file():
        st = root.stat()
        return [{"path": root.name, "type": 48, "content": base64.b64encode(root.read_bytes()).decode(), "mtime": int(st.st_mtime),
                 "mode": st.st_mode & 0o7777, "uid": 0, "gid": 0}]
    ignore = util.make_shutil_ignore(root)
    items: List[Dict[str, Any]] = []
    total = 0
    for dp, dn, fn in os.walk(root):
        skip = ignore(dp, dn + fn)
        dn[:] = sorted(d for d in dn if d not in skip)
        rel_dir = os.path.relpath(dp, root)
        if rel_dir != ".":
            st = os.stat(dp)
            items.append({"path": rel_dir + "/", "type": 53, "content": "", "mtime": int(st.st_mtime), "mode": st.st_mode & 0o7777, "uid": 0, "gid": 0})
        for f in sorted(fn):
            if f in skip:
                continue
            p = os.path.join(dp, f)
            if not os.path.isfile(p):
                continue
            st = os.stat(p)
            total += st.st_size
            if total > limit:
                raise ValueError(f"The total size of context directory exceeds the {limit // (1024 * 1024)} MB limit; "
                                 "add large files to .detignore or mount them instead")
            with open(p, "rb") as fh:
                items.append({"path": os.path.relpath(p, root), "type": 48, "content": base64.b64encode(fh.read()).decode(),
                              "mtime": int(st.st_mtime), "mode": st.st_mode & 0o7777, "uid": 0, "gid": 0})
    return items
