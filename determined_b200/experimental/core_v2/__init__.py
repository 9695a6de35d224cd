"""``core_v2``: detached ("unmanaged") experiments — a plain script reports to the master without having been launched by it
(reference ``harness/determined/experimental/core_v2/__init__.py``).  ``_core_v2`` holds the configuration classes, ``init`` /
``close`` and the module-level singletons, ``_unmanaged`` the experiment / trial rows, ``_core_context_v2`` the context."""
import sys as _sys
import types as _types

from determined_b200.experimental.core_v2 import _core_v2
from determined_b200.experimental.core_v2._core_v2 import (  # noqa: F401
    DefaultConfig, DistributedContext, PreemptMode, TensorboardMode, UnmanagedConfig, close, init, init_context, url_reverse_webui_exp_view)

_SINGLETONS = ("train", "distributed", "preempt", "checkpoint", "searcher", "info", "_context", "_session", "_exp_id")


class _Module(_types.ModuleType):
    """``core_v2.train`` etc. are set by ``init()`` inside ``_core_v2``: read them through."""

    def __getattr__(self, name: str):  # noqa: ANN001
        if name in _SINGLETONS:
            return getattr(_core_v2, name)
        raise AttributeError(f"module {self.__name__!r} has no attribute {name!r}")


_sys.modules[__name__].__class__ = _Module
