"""Dict-with-accessors view of the (already validated + defaulted) experiment config.

Parity: reference ``harness/determined/_experiment_config.py:4-72``.  Full schema validation
and defaulting lives in :mod:`determined_b200.master.expconf` (the reference does that only in
the Go master, ``master/pkg/schemas/expconf``).
"""
from typing import Any, Dict, List, Optional, Union


class ExperimentConfig(dict):
    def debug_enabled(self) -> bool:
        return bool(self.get("debug", False))

    def scheduling_unit(self) -> int:
        return int(self.get("scheduling_unit", 100))

    def native_parallel_enabled(self) -> bool:
        return bool(self.get("resources", {}).get("native_parallel", False))

    def average_training_metrics_enabled(self) -> bool:
        return bool(self.get("optimizations", {}).get("average_training_metrics", True))

    def slots_per_trial(self) -> int:
        return int(self.get("resources", {}).get("slots_per_trial", 1))

    def experiment_seed(self) -> int:
        return int(self.get("reproducibility", {}).get("experiment_seed", 0))

    def profiling_enabled(self) -> bool:
        return bool(self.get("profiling", {}).get("enabled", False))

    def profiling_interval(self) -> Any:
        p = self.get("profiling", {})
        return int(p.get("begin_on_batch", 0)), p.get("end_after_batch", None)

    def profiling_sync_timings(self) -> bool:
        return bool(self.get("profiling", {}).get("sync_timings", True))

    def get_data_layer_type(self) -> str:
        return str(self.get("data_layer", {}).get("type", "shared_fs"))

    def get_records_per_epoch(self) -> Optional[int]:
        r = self.get("records_per_epoch")
        return int(r) if r else None

    def get_min_validation_period(self) -> Dict[str, int]:
        return dict(self.get("min_validation_period", {}) or {})

    def get_min_checkpoint_period(self) -> Dict[str, int]:
        return dict(self.get("min_checkpoint_period", {}) or {})

    def get_searcher_metric(self) -> str:
        m = self.get("searcher", {}).get("metric")
        if not isinstance(m, str):
            raise KeyError("searcher.metric is required to be a string")
        return m

    def get_smaller_is_better(self) -> bool:
        return bool(self.get("searcher", {}).get("smaller_is_better", True))

    def get_entrypoint(self) -> Union[str, List[str], None]:
        return self.get("entrypoint")

    def get_checkpoint_storage(self) -> Dict[str, Any]:
        return dict(self.get("checkpoint_storage", {}) or {})

    def get_checkpoint_policy(self) -> str:
        return str(self.get("checkpoint_policy", "best"))

    def get_optimizations_config(self) -> Dict[str, Any]:
        from determined_b200 import constants

        out = dict(constants.DEFAULT_OPTIMIZATIONS)
        out.update(self.get("optimizations", {}) or {})
        return out

    def get_perform_initial_validation(self) -> bool:
        return bool(self.get("perform_initial_validation", False))

    def get_max_restarts(self) -> int:
        return int(self.get("max_restarts", 5))
