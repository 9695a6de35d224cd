"""Searcher operations and events (reference ``master/pkg/searcher/operations.go:101-295``).

``Create`` asks the experiment to start a trial with sampled hyperparameters, ``ValidateAfter``
asks a trial to train ``length`` *more* units and then report its searcher metric, ``Close`` ends a
trial, ``Shutdown`` ends the experiment.
"""
import dataclasses
from typing import Any, Dict, Optional

from determined_b200.master import nprand


@dataclasses.dataclass(frozen=True)
class Create:
    request_id: str
    trial_seed: int
    hparams: Dict[str, Any]
    checkpoint: Optional[str] = None          # warm-start checkpoint uuid (PBT-style / custom search)

    def to_json(self) -> Dict[str, Any]:
        return {"type": "create", "request_id": self.request_id, "trial_seed": self.trial_seed,
                "hparams": self.hparams, "checkpoint": self.checkpoint}


@dataclasses.dataclass(frozen=True)
class ValidateAfter:
    request_id: str
    length: int

    def to_json(self) -> Dict[str, Any]:
        return {"type": "validate_after", "request_id": self.request_id, "length": self.length}


@dataclasses.dataclass(frozen=True)
class Close:
    request_id: str

    def to_json(self) -> Dict[str, Any]:
        return {"type": "close", "request_id": self.request_id}


@dataclasses.dataclass(frozen=True)
class Shutdown:
    cancel: bool = False
    failure: bool = False

    def to_json(self) -> Dict[str, Any]:
        return {"type": "shutdown", "cancel": self.cancel, "failure": self.failure}


def new_create(rand: nprand.State, hparams: Dict[str, Any], checkpoint: Optional[str] = None) -> Create:
    """Request id first, then the trial seed — same draw order as the reference's ``NewCreate``."""
    rid = rand.uuid4()
    seed = rand.int64(0, (1 << 31) - 1)
    return Create(rid, seed, hparams, checkpoint)


def op_from_json(d: Dict[str, Any]) -> Any:
    t = d["type"]
    if t == "create":
        return Create(d["request_id"], d["trial_seed"], d["hparams"], d.get("checkpoint"))
    if t == "validate_after":
        return ValidateAfter(d["request_id"], int(d["length"]))
    if t == "close":
        return Close(d["request_id"])
    if t == "shutdown":
        return Shutdown(d.get("cancel", False), d.get("failure", False))
    raise ValueError(f"unknown op type {t}")


class ExitedReason:
    ERRORED = "ERRORED"
    USER_CANCELED = "USER_CANCELED"
    INVALID_HP = "INVALID_HP"
    INIT_INVALID_HP = "INIT_INVALID_HP"
