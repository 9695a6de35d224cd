"""determined_b200 — a B200-native training harness with Determined's capabilities.

Public names mirror reference ``harness/determined/__init__.py:1-30``.  ``import determined_b200
as det`` is the drop-in for ``import determined as det``; :func:`install_alias` additionally
registers the package under the name ``determined`` for unmodified user code.
"""
from determined_b200.__version__ import __version__
from determined_b200._info import (
    ClusterInfo, RendezvousInfo, ResourcesInfo, TrialInfo, get_cluster_info,
)
from determined_b200.errors import InvalidHP
from determined_b200._experiment_config import ExperimentConfig
from determined_b200._import import import_from_path
from determined_b200._trial_context import EnvContext, LegacyTrial, TrialContext, TrialController
from determined_b200._trial_controller import _DistributedBackend
from determined_b200 import constants, errors, util, gpu, core

LOG_FORMAT = "%(levelname)s: [%(process)s] %(name)s: %(message)s"


# reference module path (below ``determined.``) -> the module here that defines the same public names.  A path whose last
# component is also a function of its parent here (``tensorboard.build``) cannot be mapped: importing a submodule rebinds
# that attribute of the parent to the module.
_REFERENCE_MODULE_PATHS = {
    "common.storage.s3": "common.storage.cloud", "common.storage.gcs": "common.storage.cloud", "common.storage.azure": "common.storage.cloud",
    "common.storage.azure_client": "common.storage.cloud", "common.storage.boto3_credential_manager": "common.storage.cloud",
    "common.storage._util": "common.storage.base",
    "tensorboard.base": "tensorboard", "tensorboard.shared": "tensorboard", "tensorboard.directory": "tensorboard",
    "tensorboard.s3": "tensorboard", "tensorboard.gcs": "tensorboard", "tensorboard.azure": "tensorboard",
    "tensorboard.fetchers.s3": "tensorboard.fetchers.cloud", "tensorboard.fetchers.gcs": "tensorboard.fetchers.cloud",
    "tensorboard.fetchers.azure": "tensorboard.fetchers.cloud", "tensorboard.fetchers.directory": "tensorboard.fetchers.shared",
    "tensorboard.metric_writers.pytorch": "tensorboard.metric_writers", "tensorboard.metric_writers.callback": "tensorboard.metric_writers",
    "searcher._search_method": "searcher", "searcher._search_runner": "searcher", "searcher._remote_search_runner": "searcher",
    "common.constants": "constants", "common.experimental.checkpoint._checkpoint": "common.experimental.checkpoint",
    "common.api._util": "common.api", "common.api.metric": "common.api", "common.api._rbac": "common.api", "common.api.analytics": "common.api",
    "common.v1file_utils": "common.context", "common.detignore": "common.context",
    "_env_context": "_trial_context", "_trial": "_trial_context", "experimental._native": "experimental",
}


def install_alias(name: str = "determined") -> None:
    """Make ``import determined`` (and submodules) resolve to this package, and ``import model_hub`` to ``model_hub`` here."""
    import importlib
    import sys

    if name == "determined":
        install_alias("model_hub")
    real_root = "determined_b200.model_hub" if name == "model_hub" else "determined_b200"

    class _AliasFinder:
        def find_spec(self, fullname, path=None, target=None):  # type: ignore
            if fullname != name and not fullname.startswith(name + "."):
                return None
            real = real_root + fullname[len(name):]
            try:
                mod = importlib.import_module(real)
            except ImportError:
                # a module path of the reference whose contents live in another module here
                moved = _REFERENCE_MODULE_PATHS.get(fullname[len(name) + 1:]) if name == "determined" else None
                if moved is None:
                    return None
                try:
                    mod = importlib.import_module("determined_b200." + moved if moved else "determined_b200")
                except ImportError:
                    return None
            # Do NOT put the module into sys.modules here: importlib's _find_spec would then take the module's own spec (real name,
            # source loader) instead of ours and execute the package a second time under its real name -> duplicate classes.
            # _load_unlocked registers sys.modules[fullname] itself, with the object create_module hands back.
            return importlib.util.spec_from_loader(fullname, loader=_Loader(mod))

    class _Loader:
        def __init__(self, mod):  # type: ignore
            self.mod = mod

        def create_module(self, spec):  # type: ignore
            return self.mod

        def exec_module(self, module):  # type: ignore
            pass

        # runpy (`python -m determined.exec.launch`, runpy.run_module) asks the loader for the code of the module
        def get_code(self, fullname):  # type: ignore
            return self.mod.__spec__.loader.get_code(self.mod.__name__)

        def get_source(self, fullname):  # type: ignore
            return self.mod.__spec__.loader.get_source(self.mod.__name__)

        def is_package(self, fullname):  # type: ignore
            return hasattr(self.mod, "__path__")

    import importlib.util

    _AliasFinder.alias_of = name
    if not any(getattr(f, "alias_of", None) == name for f in sys.meta_path):
        sys.meta_path.insert(0, _AliasFinder())
    if real_root in sys.modules:      # `determined` is this (already imported) package; `model_hub` loads on first import, it pulls in torch
        sys.modules.setdefault(name, sys.modules[real_root])


def __getattr__(attr: str):  # lazy heavy submodules (torch import) – keeps `import determined_b200` light
    if attr in ("pytorch", "launch", "exec", "searcher", "tensorboard", "experimental", "transformers", "cli",
                "master", "models", "ops", "parallel", "utils", "ipc", "keras", "lightning", "layers", "common"):
        import importlib

        return importlib.import_module(f"determined_b200.{attr}")
    raise AttributeError(attr)
