"""Import user code from a directory without polluting (or being polluted by) ``sys.modules``.

Parity: reference ``harness/determined/_import.py:8-107`` (``import_from_path``): used to load a
trial class out of a checkpoint's ``code/`` directory, possibly several versions in one process.
"""
import contextlib
import importlib
import os
import pathlib
import pkgutil
import sys
from typing import Iterator, Set


_active = False


def modules_from_dir(path: str) -> Set[str]:
    """Top-level module names importable from ``path``."""
    names = set()
    for m in pkgutil.iter_modules([path]):
        names.add(m.name)
    return names


@contextlib.contextmanager
def import_from_path(path: os.PathLike) -> Iterator[None]:
    """Within the block, ``import x`` resolves ``x`` from ``path`` first, fresh (no cached module
    of the same name is reused), and whatever was imported from ``path`` is evicted on exit so a
    second checkpoint with different code can be loaded next."""
    global _active
    if _active:
        raise RuntimeError("import_from_path does not support nesting")
    p = str(pathlib.Path(path).resolve())
    local = modules_from_dir(p)
    saved = {}
    _active = True
    write_bytecode, sys.dont_write_bytecode = sys.dont_write_bytecode, True     # a checkpoint directory stays as it was stored
    for name in list(sys.modules):
        if name.split(".")[0] in local:
            saved[name] = sys.modules.pop(name)
    sys.path.insert(0, p)
    importlib.invalidate_caches()
    try:
        yield
    finally:
        _active = False
        sys.dont_write_bytecode = write_bytecode
        with contextlib.suppress(ValueError):
            sys.path.remove(p)
        for name in list(sys.modules):
            if name.split(".")[0] in local:
                mod = sys.modules[name]
                f = getattr(mod, "__file__", None) or ""
                if f.startswith(p) or not f:
                    del sys.modules[name]
        sys.modules.update(saved)
        importlib.invalidate_caches()
