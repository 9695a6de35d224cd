"""Exception hierarchy.

Parity: reference ``harness/determined/errors.py:1-107`` and ``_execution.py:17-24`` (InvalidHP).
The names are part of the public contract (users ``except det.errors.X``); the bodies are ours.
"""
from typing import Any, Dict, Optional


class InternalException(Exception):
    def __init__(self, message: str) -> None:
        super().__init__(f"Internal error: {message}. This is a bug in the harness.")
        self.message = str(self)


class InvalidExperimentException(BaseException):
    """Raised when an experiment (config, trial class, data types) is invalid."""


class InvalidDataTypeException(InvalidExperimentException):
    def __init__(self, typ: type, message: str) -> None:
        super().__init__(f"Invalid data type ({typ.__name__}): {message}.")


class InvalidConfigurationException(InvalidExperimentException):
    def __init__(self, config: Dict[str, Any], message: str) -> None:
        super().__init__(f"Invalid configuration ({config}): {message}.")


class InvalidModelException(InvalidExperimentException):
    pass


class InvalidCheckpointException(Exception):
    def __init__(self, why: Optional[str] = None) -> None:
        super().__init__("Invalid checkpoint." if why is None else f"Invalid checkpoint: {why}")


class CheckpointNotFound(FileNotFoundError):
    """The storage backend holds nothing under the requested checkpoint id (also a ``FileNotFoundError``, so callers that
    predate this class keep working)."""


class StopLoadingImplementation(Exception):
    """Used to interrupt the import of user code (``det.load``)."""


class WorkerError(Exception):
    """A training worker process died without a graceful-exit handshake."""


class SkipWorkloadException(Exception):
    pass


class CheckFailedError(Exception):
    pass


class EnterpriseOnlyError(Exception):
    pass


class InvalidHP(Exception):
    """Raise from trial ``__init__``/``train_batch`` to tell the searcher this hyperparameter
    sample is unusable; the trial is closed gracefully and ASHA replaces it
    (reference ``_execution.py:17-24``, ``core/_context.py:102-106``)."""

    def __init__(self, msg: str = "…") -> None:
        if not isinstance(msg, str):
            raise TypeError("InvalidHP message must be a string")
        super().__init__(msg)


class ShouldExit(Exception):
    """Raised inside the training loop when the trial should checkpoint-and-exit
    (preemption or out of searcher ops).  ``skip_exit_checkpoint`` mirrors the reference's
    ``_pytorch_trial.py:140-160`` semantics."""

    def __init__(self, skip_exit_checkpoint: bool = False):
        super().__init__()
        self.skip_exit_checkpoint = skip_exit_checkpoint


class CheckpointStateException(Exception):
    """A checkpoint is not in a state that allows the operation (deleted, partially deleted, still being written)."""


class NoDirectStorageAccess(Exception):
    """The checkpoint storage cannot be reached from this machine (``DownloadMode.DIRECT``)."""


class ProxiedDownloadFailed(Exception):
    """Downloading a checkpoint through the master failed (``DownloadMode.MASTER``)."""


class MultipleDownloadsFailed(Exception):
    """``DownloadMode.AUTO``: neither the direct nor the proxied download worked; ``args`` holds both errors."""


class WorkerFinishedGracefully(Exception):
    """A worker said it is done before the PID server stopped watching it."""
