"""``Searcher`` — wraps a SearchMethod with bookkeeping shared by all methods.

Parity: reference ``master/pkg/searcher/searcher.go:17-295`` (tracks trials requested/created/
closed, per-trial progress, early exits; de-duplicates events; snapshot/restore) and
``simulate.go:51-222`` (offline simulation used by ``det preview-search``).
"""
import json
from typing import Any, Callable, Dict, List, Optional, Tuple

from determined_b200.master import nprand
from determined_b200.master.searcher import methods
from determined_b200.master.searcher.operations import (
    Close, Create, ExitedReason, Shutdown, ValidateAfter, op_from_json,
)


class Searcher:
    def __init__(self, seed: int, method: methods.SearchMethod, hparams: Dict[str, Any]) -> None:
        self.rand = nprand.State(seed)
        self.method = method
        self.hparams = hparams
        self.trials_requested = 0
        self.trials_created: Dict[str, bool] = {}
        self.trials_closed: Dict[str, bool] = {}
        self.exits: Dict[str, bool] = {}
        self.cancels: Dict[str, bool] = {}
        self.failures: Dict[str, bool] = {}
        self.trial_progress: Dict[str, float] = {}
        self.completed_ops: Dict[str, bool] = {}
        self.shutdown = False

    def _ctx(self) -> methods.Context:
        return methods.Context(self.rand, self.hparams)

    def _record(self, ops: List[Any]) -> List[Any]:
        for op in ops:
            if isinstance(op, Create):
                self.trials_requested += 1
            elif isinstance(op, Shutdown):
                self.shutdown = True
        return ops

    def initial_operations(self) -> List[Any]:
        return self._record(self.method.initial_operations(self._ctx()))

    def trial_created(self, request_id: str) -> List[Any]:
        self.trials_created[request_id] = True
        self.trial_progress[request_id] = 0.0
        return self._record(self.method.trial_created(self._ctx(), request_id))

    def trial_is_created(self, request_id: str) -> bool:
        return bool(self.trials_created.get(request_id))

    def trial_is_closed(self, request_id: str) -> bool:
        return bool(self.trials_closed.get(request_id))

    def trial_exited_early(self, request_id: str, reason: str) -> List[Any]:
        if self.exits.get(request_id):
            return []          # already handled (e.g. harness report + process exit both arrive)
        if reason in (ExitedReason.INVALID_HP, ExitedReason.INIT_INVALID_HP):
            self.trial_progress.pop(request_id, None)
        ops = self.method.trial_exited_early(self._ctx(), request_id, reason)
        self.exits[request_id] = True
        if reason == ExitedReason.USER_CANCELED:
            self.cancels[request_id] = True
        elif reason == ExitedReason.ERRORED:
            self.failures[request_id] = True
        out = self._record(ops)
        out += self._maybe_shutdown()
        return out

    def set_trial_progress(self, request_id: str, progress: float) -> None:
        if isinstance(self.method, methods.CustomSearch):
            self.method.trial_progress(request_id, progress)
        self.trial_progress[request_id] = progress

    def validation_completed(self, request_id: str, metric: Any, op: ValidateAfter) -> List[Any]:
        key = f"{request_id}:{op.length}:{len([k for k in self.completed_ops if k.startswith(request_id)])}"
        self.completed_ops[key] = True
        return self._record(self.method.validation_completed(self._ctx(), request_id, metric, op))

    def trial_closed(self, request_id: str) -> List[Any]:
        self.trials_closed[request_id] = True
        ops = self._record(self.method.trial_closed(self._ctx(), request_id))
        ops += self._maybe_shutdown()
        return ops

    def _maybe_shutdown(self) -> List[Any]:
        """All requested trials are closed and the method wants nothing more → end the experiment."""
        if self.shutdown or isinstance(self.method, methods.CustomSearch):
            return []
        if self.trials_requested > 0 and len(self.trials_closed) + sum(
                1 for r in self.exits if r not in self.trials_closed) >= self.trials_requested:
            self.shutdown = True
            n_fail = len(self.failures)
            return [Shutdown(cancel=len(self.failures) == 0 and len(self.cancels) >= self.trials_requested > 0,
                             failure=n_fail >= self.trials_requested > 0)]
        return []

    def progress(self) -> float:
        p = self.method.progress(self.trial_progress, self.trials_closed)
        if p != p or p < 0:
            return 0.0
        return min(p, 1.0)

    def record_external_ops(self, ops: List[Any]) -> List[Any]:
        return self._record(ops)

    def snapshot(self) -> Dict[str, Any]:
        return {"rand": self.rand.snapshot(), "method": self.method.snapshot(),
                "trials_requested": self.trials_requested, "trials_created": self.trials_created,
                "trials_closed": self.trials_closed, "exits": self.exits, "cancels": self.cancels,
                "failures": self.failures, "trial_progress": self.trial_progress,
                "completed_ops": self.completed_ops, "shutdown": self.shutdown}

    def restore(self, snap: Dict[str, Any]) -> None:
        self.rand = nprand.State.restore(snap["rand"])
        self.method.restore(snap["method"])
        for k in ("trials_requested", "trials_created", "trials_closed", "exits", "cancels", "failures",
                  "trial_progress", "completed_ops", "shutdown"):
            setattr(self, k, snap[k])


# ---------------------------------------------------------------------------------------------
# Offline simulation (det preview-search, and the searcher tests)
# ---------------------------------------------------------------------------------------------
def simulate(searcher_cfg: Dict[str, Any], hparams: Dict[str, Any], seed: int = 0,
             metric_fn: Optional[Callable[[Dict[str, Any], int], float]] = None,
             early_exit_fn: Optional[Callable[[Dict[str, Any], int], Optional[str]]] = None,
             max_events: int = 1_000_000) -> Dict[str, Any]:
    """Run a search to completion against a synthetic trial model.

    ``metric_fn(hparams, total_units)`` gives the validation metric (default: a random number);
    ``early_exit_fn`` may return an ExitedReason to kill a trial at that validation instead.
    Returns ``{"trials": {request_id: {"hparams", "lengths": [cumulative validate-after units],
    "closed": bool}}, "results": {total_units_trained: n_trials}}``."""
    method = methods.new_search_method(searcher_cfg)
    s = Searcher(seed, method, hparams)
    sim_rand = nprand.State(seed + 1)
    metric_fn = metric_fn or (lambda hp_, units: sim_rand.unit_interval())
    trials: Dict[str, Dict[str, Any]] = {}
    queues: Dict[str, List[Any]] = {}
    order: List[str] = []

    def absorb(ops: List[Any]) -> None:
        for op in ops:
            if isinstance(op, Create):
                trials[op.request_id] = {"hparams": op.hparams, "lengths": [], "closed": False, "trained": 0,
                                         "seed": op.trial_seed}
                queues[op.request_id] = []
                order.append(op.request_id)
                absorb(s.trial_created(op.request_id))
            elif isinstance(op, (ValidateAfter, Close)):
                queues[op.request_id].append(op)
            elif isinstance(op, Shutdown):
                pass

    absorb(s.initial_operations())
    events = 0
    while events < max_events:
        progressed = False
        for rid in list(order):
            q = queues[rid]
            if not q or trials[rid]["closed"]:
                continue
            op = q.pop(0)
            events += 1
            progressed = True
            if isinstance(op, ValidateAfter):
                t = trials[rid]
                t["trained"] += op.length
                t["lengths"].append(t["trained"])
                reason = early_exit_fn(t["hparams"], t["trained"]) if early_exit_fn else None
                if reason is not None:
                    t["closed"] = True
                    t["exited"] = reason
                    queues[rid].clear()
                    absorb(s.trial_exited_early(rid, reason))
                    absorb(s.trial_closed(rid))      # the experiment closes an exited trial
                else:
                    s.set_trial_progress(rid, float(t["trained"]))
                    absorb(s.validation_completed(rid, metric_fn(t["hparams"], t["trained"]), op))
            else:
                trials[rid]["closed"] = True
                absorb(s.trial_closed(rid))
            break   # one event at a time, oldest trial first: deterministic interleaving
        if not progressed:
            break
    results: Dict[int, int] = {}
    for t in trials.values():
        results[t["trained"]] = results.get(t["trained"], 0) + 1
    return {"trials": trials, "results": results, "searcher": s, "events": events}
