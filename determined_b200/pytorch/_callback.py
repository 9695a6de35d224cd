"""``PyTorchCallback`` — user hooks into the trial lifecycle.

Parity: reference ``harness/determined/pytorch/_callback.py:4-153`` (same hook names and call
sites; all default to no-ops).  Callbacks with state participate in checkpoints via
``state_dict``/``load_state_dict``.
"""
import pathlib
from typing import Any, Dict


class PyTorchCallback:
    def on_trial_startup(self, first_batch_idx: int, checkpoint_uuid: Any) -> None:
        """Before training starts (after checkpoint restore).  ``checkpoint_uuid`` is the restored
        checkpoint or None."""

    def on_trial_shutdown(self) -> None:
        """Before exiting, even after an exception (best effort)."""

    def on_training_start(self) -> None:
        pass

    def on_training_epoch_start(self, epoch_idx: int) -> None:
        pass

    def on_training_epoch_end(self, epoch_idx: int) -> None:
        pass

    def on_training_batch_start(self, batch_idx: int) -> None:
        """Before each ``train_batch`` (extension: the reference has no per-batch hooks).  Keep it
        cheap and do not synchronise the device here unless you mean to."""

    def on_training_batch_end(self, batch_idx: int) -> None:
        """After each ``train_batch`` has been *launched* (the device may still be running it)."""

    def on_training_workload_end(self, avg_metrics: Dict[str, Any], batch_metrics: Dict[str, Any]) -> None:
        """After each reporting period, with the (cross-rank averaged) training metrics."""

    def on_validation_start(self) -> None:
        pass

    def on_validation_epoch_start(self) -> None:
        pass

    def on_validation_epoch_end(self, outputs: Any) -> None:
        pass

    def on_validation_end(self, metrics: Dict[str, Any]) -> None:
        pass

    def on_checkpoint_load_start(self, checkpoint: Any) -> None:
        """Before the state dict is applied; may mutate ``checkpoint`` (a dict)."""

    def on_checkpoint_save_start(self, checkpoint: Dict[str, Any]) -> None:
        """Before ``state_dict.pth`` is written; may mutate ``checkpoint``."""

    def on_checkpoint_write_end(self, checkpoint_dir: str) -> None:
        """After files are written locally, before upload: add extra files here."""

    def on_checkpoint_upload_end(self, uuid: str) -> None:
        pass

    def on_checkpoint_end(self, checkpoint_dir: str) -> None:
        """Deprecated alias of ``on_checkpoint_write_end``."""

    def state_dict(self) -> Dict[str, Any]:
        return {}

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        pass
