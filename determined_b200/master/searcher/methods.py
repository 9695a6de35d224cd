"""Built-in hyperparameter search methods.

Parity (behavioural; the code is an independent Python design):
  * ``SearchMethod`` interface — reference ``master/pkg/searcher/search_method.go:16-44``
  * single / random — ``random.go:34-138``;  grid — ``grid.go:34-288``
  * asynchronous successive halving (ASHA, promotion flavour) — ``asha.go:56-328``
  * ASHA stopping flavour (``stop_once``) — ``asha_stopping.go:28-260``
  * adaptive_asha = tournament of ASHA brackets — ``adaptive_asha.go:13-143``, ``tournament.go:27-149``
  * custom (event queue driven by a client-side SearchRunner) — ``custom_search.go:27-147``

Every method is a small state machine: events in (``initial_operations``, ``trial_created``,
``validation_completed``, ``trial_closed``, ``trial_exited_early``), operations out.  State is a
plain JSON-able dict (``snapshot``/``restore``) so the master can persist it after every event
batch and survive restarts (reference ``experiment.go:778,980-998``).  ``ValidateAfter.length`` is
*incremental*; the trial state machine accumulates it into the absolute length the harness sees.
"""
import bisect
import math
from typing import Any, Dict, List, Optional, Tuple

from determined_b200.master import nprand
from determined_b200.master.searcher import hparams as hp
from determined_b200.master.searcher.operations import (
    Close, Create, ExitedReason, Shutdown, ValidateAfter, new_create,
)

Op = Any
EXITED_METRIC = 1.7976931348623157e308   # "worst possible" for early-exited trials


class Context:
    def __init__(self, rand: nprand.State, hparams: Dict[str, Any]) -> None:
        self.rand = rand
        self.hparams = hparams


def length_units(max_length: Any) -> Tuple[int, Optional[str]]:
    """``{"batches": N}`` / ``{"records": N}`` / ``{"epochs": N}`` / bare int → (N, unit)."""
    if isinstance(max_length, dict):
        if len(max_length) != 1:
            raise ValueError(f"max_length must have exactly one unit, got {max_length}")
        (unit, n), = max_length.items()
        if unit not in ("batches", "records", "epochs"):
            raise ValueError(f"unknown length unit {unit!r}")
        return int(n), unit
    return int(max_length), None


class SearchMethod:
    name = "base"

    def initial_operations(self, ctx: Context) -> List[Op]:
        raise NotImplementedError

    def trial_created(self, ctx: Context, request_id: str) -> List[Op]:
        return []

    def validation_completed(self, ctx: Context, request_id: str, metric: Any, op: ValidateAfter) -> List[Op]:
        return []

    def trial_closed(self, ctx: Context, request_id: str) -> List[Op]:
        return []

    def trial_exited_early(self, ctx: Context, request_id: str, reason: str) -> List[Op]:
        return []

    def progress(self, trial_progress: Dict[str, float], trials_closed: Dict[str, bool]) -> float:
        return 0.0

    def unit(self) -> Optional[str]:
        return None

    def snapshot(self) -> Dict[str, Any]:
        return {}

    def restore(self, state: Dict[str, Any]) -> None:
        pass


# ------------------------------------------------------------------------------------------------
class RandomSearch(SearchMethod):
    """``max_trials`` independent samples, each trained for ``max_length``; at most
    ``max_concurrent_trials`` at a time (0 = unlimited).  ``single`` is the 1-trial special case."""

    def __init__(self, cfg: Dict[str, Any], single: bool = False) -> None:
        self.name = "single" if single else "random"
        self.max_length, self._unit = length_units(cfg["max_length"])
        self.max_trials = 1 if single else int(cfg["max_trials"])
        self.max_concurrent = 1 if single else int(cfg.get("max_concurrent_trials", 16) or 0)
        self.s = {"created_trials": 0, "pending_trials": 0}

    def _new_trial(self, ctx: Context) -> List[Op]:
        c = new_create(ctx.rand, hp.sample_all(ctx.hparams, ctx.rand))
        self.s["created_trials"] += 1
        self.s["pending_trials"] += 1
        return [c, ValidateAfter(c.request_id, self.max_length), Close(c.request_id)]

    def initial_operations(self, ctx: Context) -> List[Op]:
        n = self.max_trials if self.max_concurrent <= 0 else min(self.max_trials, self.max_concurrent)
        ops: List[Op] = []
        for _ in range(n):
            ops += self._new_trial(ctx)
        return ops

    def trial_closed(self, ctx: Context, request_id: str) -> List[Op]:
        self.s["pending_trials"] -= 1
        if self.s["created_trials"] < self.max_trials:
            return self._new_trial(ctx)
        return []

    def trial_exited_early(self, ctx: Context, request_id: str, reason: str) -> List[Op]:
        self.s["pending_trials"] -= 1
        if self.name == "random" and reason in (ExitedReason.INVALID_HP, ExitedReason.INIT_INVALID_HP):
            self.s["created_trials"] -= 1      # an invalid sample does not count against max_trials
        return []

    def progress(self, trial_progress: Dict[str, float], trials_closed: Dict[str, bool]) -> float:
        done = sum(self.max_length if trials_closed.get(k) else v for k, v in trial_progress.items())
        return done / float(self.max_length * self.max_trials)

    def unit(self) -> Optional[str]:
        return self._unit

    def snapshot(self) -> Dict[str, Any]:
        return dict(self.s)

    def restore(self, state: Dict[str, Any]) -> None:
        self.s.update(state or {})


class GridSearch(SearchMethod):
    name = "grid"

    def __init__(self, cfg: Dict[str, Any]) -> None:
        self.max_length, self._unit = length_units(cfg["max_length"])
        self.max_concurrent = int(cfg.get("max_concurrent_trials", 16) or 0)
        self.s: Dict[str, Any] = {"remaining": [], "pending_trials": 0, "trials": 0}

    def _next(self, ctx: Context) -> List[Op]:
        params = self.s["remaining"].pop()
        c = new_create(ctx.rand, params)
        self.s["pending_trials"] += 1
        return [c, ValidateAfter(c.request_id, self.max_length), Close(c.request_id)]

    def initial_operations(self, ctx: Context) -> List[Op]:
        g = hp.grid(ctx.hparams)
        self.s["trials"] = len(g)
        self.s["remaining"] = list(g)
        n = len(g) if self.max_concurrent <= 0 else min(len(g), self.max_concurrent)
        ops: List[Op] = []
        for _ in range(n):
            ops += self._next(ctx)
        return ops

    def trial_closed(self, ctx: Context, request_id: str) -> List[Op]:
        self.s["pending_trials"] -= 1
        return self._next(ctx) if self.s["remaining"] else []

    def progress(self, trial_progress: Dict[str, float], trials_closed: Dict[str, bool]) -> float:
        done = float(self.max_length * sum(1 for v in trials_closed.values() if v))
        done += sum(v for k, v in trial_progress.items() if not trials_closed.get(k))
        return done / float(self.max_length * max(self.s["trials"], 1))

    def unit(self) -> Optional[str]:
        return self._unit

    def snapshot(self) -> Dict[str, Any]:
        return {"remaining": list(self.s["remaining"]), "pending_trials": self.s["pending_trials"], "trials": self.s["trials"]}

    def restore(self, state: Dict[str, Any]) -> None:
        self.s.update(state or {})


# ------------------------------------------------------------------------------------------------
class _Rung:
    def __init__(self, units_needed: int) -> None:
        self.units_needed = units_needed
        self.metrics: List[List[Any]] = []   # sorted by metric: [metric, request_id, promoted]
        self.outstanding = 0

    def _insert(self, rid: str, metric: float, strict: bool) -> int:
        keys = [m[0] for m in self.metrics]
        i = bisect.bisect_right(keys, metric) if strict else bisect.bisect_left(keys, metric)
        self.metrics.insert(i, [metric, rid, False])
        return i

    def promotions(self, rid: str, metric: float, divisor: float) -> List[str]:
        """Promotion flavour: top 1/divisor of everything seen so far in this rung gets promoted,
        each trial at most once."""
        old_n = int(len(self.metrics) / divisor)
        new_n = int((len(self.metrics) + 1) / divisor)
        i = self._insert(rid, metric, strict=True)
        if i < new_n:
            self.metrics[i][2] = True
            return [rid]
        if new_n != old_n and not self.metrics[old_n][2]:
            self.metrics[old_n][2] = True
            return [self.metrics[old_n][1]]
        return []

    def continue_training(self, rid: str, metric: float, divisor: float) -> bool:
        """Stopping flavour: a trial continues iff it is in the top 1/divisor when it reports."""
        n = max(int((len(self.metrics) + 1) / divisor), 1)
        i = self._insert(rid, metric, strict=False)
        ok = i < n
        self.metrics[i][2] = ok
        return ok

    def drop(self, rid: str) -> None:
        self.metrics = [m for m in self.metrics if m[1] != rid]

    def to_json(self) -> Dict[str, Any]:
        return {"units_needed": self.units_needed, "metrics": [list(m) for m in self.metrics], "outstanding": self.outstanding}

    @classmethod
    def from_json(cls, d: Dict[str, Any]) -> "_Rung":
        r = cls(int(d["units_needed"]))
        r.metrics = [list(m) for m in d["metrics"]]
        r.outstanding = int(d["outstanding"])
        return r


class ASHASearch(SearchMethod):
    """Asynchronous successive halving.  ``num_rungs`` rungs; rung ``i`` trains to
    ``sum_{j<=i} max(max_length / divisor^(num_rungs-j-1), 1)`` units.  ``stop_once=False``:
    promotion ASHA (trials pause at a rung and may be resumed from their checkpoint later);
    ``stop_once=True``: stopping ASHA (a trial either continues immediately or is closed)."""

    name = "async_halving"

    def __init__(self, cfg: Dict[str, Any], smaller_is_better: bool) -> None:
        self.num_rungs = int(cfg["num_rungs"])
        self.max_length, self._unit = length_units(cfg["max_length"])
        self.max_trials = int(cfg["max_trials"])
        self.divisor = float(cfg.get("divisor", 4))
        self.max_concurrent = int(cfg.get("max_concurrent_trials", 16) or 0)
        self.stop_once = bool(cfg.get("stop_once", False))
        self.smaller_is_better = smaller_is_better
        self.rungs: List[_Rung] = []
        need = 0
        for i in range(self.num_rungs):
            need += max(int(self.max_length / math.pow(self.divisor, self.num_rungs - i - 1)), 1)
            self.rungs.append(_Rung(need))
        self.trial_rungs: Dict[str, int] = {}
        self.early_exit: Dict[str, bool] = {}
        self.closed: Dict[str, bool] = {}
        self.trials_completed = 0
        self.invalid_trials = 0
        self.pending_trials = 0

    # -- helpers -----------------------------------------------------------------------------------
    def _new_trial(self, ctx: Context) -> List[Op]:
        c = new_create(ctx.rand, hp.sample_all(ctx.hparams, ctx.rand))
        self.trial_rungs[c.request_id] = 0
        self.pending_trials += 1
        return [c, ValidateAfter(c.request_id, self.rungs[0].units_needed)]

    def initial_operations(self, ctx: Context) -> List[Op]:
        if self.max_concurrent > 0:
            n = min(self.max_concurrent, self.max_trials)
        else:
            n = max(1, min(int(math.pow(self.divisor, self.num_rungs - 1)), self.max_trials))
        ops: List[Op] = []
        for _ in range(n):
            ops += self._new_trial(ctx)
        return ops

    def trial_created(self, ctx: Context, request_id: str) -> List[Op]:
        self.rungs[0].outstanding += 1
        self.trial_rungs[request_id] = 0
        return []

    def trial_closed(self, ctx: Context, request_id: str) -> List[Op]:
        self.trials_completed += 1
        self.closed[request_id] = True
        return []

    def validation_completed(self, ctx: Context, request_id: str, metric: Any, op: ValidateAfter) -> List[Op]:
        self.pending_trials -= 1
        try:
            value = float(metric)
        except (TypeError, ValueError):
            raise ValueError(f"unexpected metric type for ASHA built-in search method: {metric!r}")
        if not self.smaller_is_better:
            value = -value
        return self._promote(ctx, request_id, value)

    def _promote(self, ctx: Context, rid: str, metric: float) -> List[Op]:
        ri = self.trial_rungs[rid]
        rung = self.rungs[ri]
        rung.outstanding -= 1
        added = False
        ops: List[Op] = []
        if ri == self.num_rungs - 1:
            rung.metrics.append([metric, rid, False])
            if not self.early_exit.get(rid):
                ops.append(Close(rid))
                self.closed[rid] = True
        elif self.stop_once:
            nxt = self.rungs[ri + 1]
            keep = rung.continue_training(rid, metric, self.divisor)
            if not self.early_exit.get(rid):
                if keep:
                    self.trial_rungs[rid] = ri + 1
                    nxt.outstanding += 1
                    ops.append(ValidateAfter(rid, max(nxt.units_needed - rung.units_needed, 1)))
                    self.pending_trials += 1
                    added = True
                else:
                    ops.append(Close(rid))
                    self.closed[rid] = True
        else:
            nxt = self.rungs[ri + 1]
            for pid in rung.promotions(rid, metric, self.divisor):
                self.trial_rungs[pid] = ri + 1
                nxt.outstanding += 1
                if not self.early_exit.get(pid):
                    ops.append(ValidateAfter(pid, max(nxt.units_needed - rung.units_needed, 1)))
                    self.pending_trials += 1
                    added = True
                else:
                    # behave as if the promoted-but-dead trial ran and came back with the worst metric
                    return ops + self._promote(ctx, pid, EXITED_METRIC)
        all_trials = len(self.trial_rungs) - self.invalid_trials
        if not added and all_trials < self.max_trials:
            ops += self._new_trial(ctx)
        if not self.stop_once and len(self.rungs[0].metrics) == self.max_trials:
            ops += self._close_out_rungs()
        return ops

    def _close_out_rungs(self) -> List[Op]:
        """Once every trial has reported in rung 0: close the never-promoted trials of each rung
        that has nothing outstanding (bottom-up, stopping at the first busy rung)."""
        ops: List[Op] = []
        for rung in self.rungs:
            if rung.outstanding > 0:
                break
            for metric, rid, promoted in rung.metrics:
                if not promoted and not self.closed.get(rid) and not self.early_exit.get(rid):
                    ops.append(Close(rid))
                    self.closed[rid] = True
        return ops

    def trial_exited_early(self, ctx: Context, request_id: str, reason: str) -> List[Op]:
        self.pending_trials -= 1
        if reason in (ExitedReason.INVALID_HP, ExitedReason.INIT_INVALID_HP):
            self.early_exit[request_id] = True
            self.closed[request_id] = True
            self.invalid_trials += 1
            top = self.trial_rungs.get(request_id, 0)
            self.rungs[top].outstanding -= 1
            for i in range(top + 1):
                self.rungs[i].drop(request_id)
            return [Close(request_id)] + self._new_trial(ctx)    # replace the invalid sample
        self.early_exit[request_id] = True
        self.closed[request_id] = True
        return self._promote(ctx, request_id, EXITED_METRIC)

    def progress(self, trial_progress: Dict[str, float], trials_closed: Dict[str, bool]) -> float:
        seen = len(self.rungs[0].metrics)
        p = seen / (1.2 * self.max_trials)          # 20 % head-room for promotions
        if seen == self.max_trials:
            p = max(p, (self.trials_completed - self.invalid_trials) / float(self.max_trials))
        return p

    def unit(self) -> Optional[str]:
        return self._unit

    def snapshot(self) -> Dict[str, Any]:
        return {"rungs": [r.to_json() for r in self.rungs], "trial_rungs": dict(self.trial_rungs),
                "early_exit_trials": dict(self.early_exit), "closed_trials": dict(self.closed),
                "trials_completed": self.trials_completed, "invalid_trials": self.invalid_trials,
                "pending_trials": self.pending_trials}

    def restore(self, state: Dict[str, Any]) -> None:
        if not state:
            return
        self.rungs = [_Rung.from_json(r) for r in state["rungs"]]
        self.trial_rungs = dict(state["trial_rungs"])
        self.early_exit = dict(state["early_exit_trials"])
        self.closed = dict(state["closed_trials"])
        self.trials_completed = int(state["trials_completed"])
        self.invalid_trials = int(state["invalid_trials"])
        self.pending_trials = int(state["pending_trials"])


# ------------------------------------------------------------------------------------------------
class TournamentSearch(SearchMethod):
    """Runs several sub-searches side by side, routing each event to the sub-search that created
    the trial."""

    def __init__(self, name: str, subs: List[SearchMethod]) -> None:
        self.name = name
        self.subs = subs
        self.trial_table: Dict[str, int] = {}

    def _mark(self, i: int, ops: List[Op]) -> List[Op]:
        for op in ops:
            if isinstance(op, Create):
                self.trial_table[op.request_id] = i
        return ops

    def initial_operations(self, ctx: Context) -> List[Op]:
        out: List[Op] = []
        for i, s in enumerate(self.subs):
            out += self._mark(i, s.initial_operations(ctx))
        return out

    def _route(self, rid: str) -> Tuple[int, SearchMethod]:
        i = self.trial_table[rid]
        return i, self.subs[i]

    def trial_created(self, ctx: Context, request_id: str) -> List[Op]:
        i, s = self._route(request_id)
        return self._mark(i, s.trial_created(ctx, request_id))

    def validation_completed(self, ctx: Context, request_id: str, metric: Any, op: ValidateAfter) -> List[Op]:
        i, s = self._route(request_id)
        return self._mark(i, s.validation_completed(ctx, request_id, metric, op))

    def trial_closed(self, ctx: Context, request_id: str) -> List[Op]:
        i, s = self._route(request_id)
        return self._mark(i, s.trial_closed(ctx, request_id))

    def trial_exited_early(self, ctx: Context, request_id: str, reason: str) -> List[Op]:
        i, s = self._route(request_id)
        return self._mark(i, s.trial_exited_early(ctx, request_id, reason))

    def progress(self, trial_progress: Dict[str, float], trials_closed: Dict[str, bool]) -> float:
        tot = 0.0
        for i, s in enumerate(self.subs):
            tp = {k: v for k, v in trial_progress.items() if self.trial_table.get(k) == i}
            tc = {k: v for k, v in trials_closed.items() if self.trial_table.get(k) == i}
            tot += s.progress(tp, tc)
        return tot / len(self.subs)

    def unit(self) -> Optional[str]:
        return self.subs[0].unit()

    def snapshot(self) -> Dict[str, Any]:
        return {"trial_table": dict(self.trial_table), "sub_search_states": [s.snapshot() for s in self.subs]}

    def restore(self, state: Dict[str, Any]) -> None:
        if not state:
            return
        self.trial_table = dict(state["trial_table"])
        for s, st in zip(self.subs, state["sub_search_states"]):
            s.restore(st)


def bracket_max_trials(max_trials: int, divisor: float, brackets: List[int]) -> List[int]:
    """Split ``max_trials`` so every bracket gets roughly the same total training budget."""
    w = [math.pow(divisor, r - 1) / r for r in brackets]
    tot = sum(w)
    out = [max(int(x / tot * max_trials), 1) for x in w]
    out[0] += max(max_trials - sum(out), 0)
    return out


def bracket_max_concurrent(max_concurrent: int, divisor: float, max_trials: List[int]) -> List[int]:
    n = len(max_trials)
    if max_concurrent == 0:
        return [max(max_trials[-1], int(divisor))] * n
    max_concurrent = max(max_concurrent, n)
    base, rem = divmod(max_concurrent, n)
    return [base + (1 if i < rem else 0) for i in range(n)]


def adaptive_brackets(mode: str, max_rungs: int) -> List[int]:
    if mode == "conservative":
        return list(range(1, max_rungs + 1))
    if mode == "standard":
        return list(range((max_rungs - 1) // 2 + 1, max_rungs + 1))
    if mode == "aggressive":
        return [max_rungs]
    raise ValueError(f"unexpected adaptive mode: {mode}")


def make_adaptive_asha(cfg: Dict[str, Any], smaller_is_better: bool) -> TournamentSearch:
    max_length, _ = length_units(cfg["max_length"])
    divisor = float(cfg.get("divisor", 4))
    max_trials = int(cfg["max_trials"])
    brackets = list(cfg.get("bracket_rungs") or [])
    if not brackets:
        max_rungs = min(int(cfg.get("max_rungs", 5)),
                        int(math.log(max_length) / math.log(divisor)) + 1,
                        int(math.log(max_trials) / math.log(divisor)) + 1)
        brackets = adaptive_brackets(cfg.get("mode", "standard"), max(max_rungs, 1))
    brackets = sorted(brackets, reverse=True)      # most aggressive early stopping first
    trials = bracket_max_trials(max_trials, divisor, brackets)
    conc = bracket_max_concurrent(int(cfg.get("max_concurrent_trials", 16) or 0), divisor, trials)
    subs: List[SearchMethod] = []
    for r, t, c in zip(brackets, trials, conc):
        subs.append(ASHASearch({"num_rungs": r, "max_length": cfg["max_length"], "max_trials": t, "divisor": divisor,
                                "max_concurrent_trials": c, "stop_once": bool(cfg.get("stop_once", False))},
                               smaller_is_better))
    return TournamentSearch("adaptive_asha", subs)


# ------------------------------------------------------------------------------------------------
class CustomSearch(SearchMethod):
    """Search logic lives in a client process (``determined.searcher.SearchRunner``); the master
    only queues events for it and applies the operations it posts back."""

    name = "custom"

    def __init__(self, cfg: Dict[str, Any]) -> None:
        self._unit = cfg.get("unit")
        self.events: List[Dict[str, Any]] = []
        self.next_event_id = 1
        self.progress_value = 0.0

    def _push(self, kind: str, **payload: Any) -> List[Op]:
        self.events.append({"id": self.next_event_id, kind: payload})
        self.next_event_id += 1
        return []

    def initial_operations(self, ctx: Context) -> List[Op]:
        return self._push("initialOperations")

    def trial_created(self, ctx: Context, request_id: str) -> List[Op]:
        return self._push("trialCreated", requestId=request_id)

    def validation_completed(self, ctx: Context, request_id: str, metric: Any, op: ValidateAfter) -> List[Op]:
        return self._push("validationCompleted", requestId=request_id, metric=metric, validateAfterLength=str(op.length))

    def trial_closed(self, ctx: Context, request_id: str) -> List[Op]:
        return self._push("trialClosed", requestId=request_id)

    def trial_exited_early(self, ctx: Context, request_id: str, reason: str) -> List[Op]:
        # the wire enum knows INVALID_HP / USER_REQUESTED_STOP / USER_CANCELED; a crash travels as UNSPECIFIED
        wire = reason if reason in ("INVALID_HP", "USER_REQUESTED_STOP", "USER_CANCELED") else ("INVALID_HP" if reason == "INIT_INVALID_HP" else "UNSPECIFIED")
        return self._push("trialExitedEarly", requestId=request_id, exitedReason=f"EXITED_REASON_{wire}")

    def trial_progress(self, request_id: str, partial_units: float) -> None:
        self._push("trialProgress", requestId=request_id, partialUnits=partial_units)

    def pop_events_through(self, event_id: int) -> None:
        self.events = [e for e in self.events if e["id"] > event_id]

    def progress(self, trial_progress: Dict[str, float], trials_closed: Dict[str, bool]) -> float:
        return self.progress_value

    def unit(self) -> Optional[str]:
        return self._unit

    def snapshot(self) -> Dict[str, Any]:
        return {"events": list(self.events), "next_event_id": self.next_event_id, "progress": self.progress_value}

    def restore(self, state: Dict[str, Any]) -> None:
        if state:
            self.events = list(state["events"])
            self.next_event_id = int(state["next_event_id"])
            self.progress_value = float(state.get("progress", 0.0))


def new_search_method(searcher_cfg: Dict[str, Any]) -> SearchMethod:
    """Factory from the expconf ``searcher`` section (reference ``search_method.go:72-92``)."""
    name = searcher_cfg.get("name")
    sib = bool(searcher_cfg.get("smaller_is_better", True))
    if name == "single":
        return RandomSearch(searcher_cfg, single=True)
    if name == "random":
        return RandomSearch(searcher_cfg)
    if name == "grid":
        return GridSearch(searcher_cfg)
    if name == "async_halving":
        return ASHASearch(searcher_cfg, sib)
    if name == "adaptive_asha":
        return make_adaptive_asha(searcher_cfg, sib)
    if name == "custom":
        return CustomSearch(searcher_cfg)
    raise ValueError(f"unknown searcher name: {name!r}")
