"""Resolve ``module:Class`` entrypoints (reference ``harness/determined/load.py:11-62``)."""
import importlib
import os
import sys
from typing import Any, Type

from determined_b200 import errors


def trial_class_from_entrypoint(entrypoint_spec: str) -> Type[Any]:
    if ":" not in entrypoint_spec:
        raise errors.InvalidExperimentException(f"entrypoint {entrypoint_spec!r} must look like module:TrialClass")
    module, qual = entrypoint_spec.split(":", 1)
    if os.getcwd() not in sys.path and "" not in sys.path:
        sys.path.insert(0, os.getcwd())
    mod_name = module[:-3] if module.endswith(".py") else module
    mod_name = mod_name.replace("/", ".")
    try:
        mod = importlib.import_module(mod_name)
    except ModuleNotFoundError as e:
        raise errors.InvalidExperimentException(f"cannot import {mod_name!r} for entrypoint {entrypoint_spec!r}: {e}")
    obj: Any = mod
    for part in qual.split("."):
        if not hasattr(obj, part):
            raise errors.InvalidExperimentException(f"{mod_name!r} has no attribute {qual!r}")
        obj = getattr(obj, part)
    if not isinstance(obj, type):
        raise errors.InvalidExperimentException(f"{entrypoint_spec!r} is not a class")
    return obj
