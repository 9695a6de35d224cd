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


def write_v1_context(files: List[Dict[str, Any]], dest: pathlib.Path) -> int:
    """The inverse of :func:`read_v1_context`: materialise a list of file entries under ``dest``.  Paths that would leave
    ``dest`` are refused.  Returns the number of regular files written."""
    import base64
    import os

    dest = pathlib.Path(dest).resolve()
    n = 0
    for f in files or []:
        rel = str(f.get("path") or "").lstrip("/")
        target = (dest / rel).resolve()
        if not rel or (dest != target and dest not in target.parents):
            raise ValueError(f"context entry {f.get('path')!r} points outside the context directory")
        if f.get("type") in (53, "53") or rel.endswith("/"):
            target.mkdir(parents=True, exist_ok=True)
            continue
        target.parent.mkdir(parents=True, exist_ok=True)
        target.write_bytes(base64.b64decode(f.get("content") or ""))
        os.chmod(target, (int(f.get("mode") or 0o644) & 0o7777) or 0o644)
        n += 1
    return n


def v1_context_to_tarball(files: List[Dict[str, Any]]) -> bytes:
    """The same entries as a gzipped tarball (what a remote agent unpacks)."""
    import base64
    import io
    import tarfile
    import time

    buf = io.BytesIO()
    with tarfile.open(fileobj=buf, mode="w:gz") as tf:
        for f in files or []:
            rel = str(f.get("path") or "").lstrip("/")
            if not rel or ".." in pathlib.PurePosixPath(rel).parts:
                raise ValueError(f"context entry {f.get('path')!r} points outside the context directory")
            ti = tarfile.TarInfo(rel)
            if f.get("type") in (53, "53") or rel.endswith("/"):
                ti.type, ti.mode = tarfile.DIRTYPE, (int(f.get("mode") or 0o755) & 0o7777) or 0o755
                tf.addfile(ti)
                continue
            data = base64.b64decode(f.get("content") or "")
            ti.size, ti.mode, ti.mtime = len(data), (int(f.get("mode") or 0o644) & 0o7777) or 0o644, int(f.get("mtime") or time.time())
            tf.addfile(ti, io.BytesIO(data))
    return buf.getvalue()
