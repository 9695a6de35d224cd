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
import collections
import math
from typing import Any, Dict, List, Optional, Set, Tuple

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
class _Board:
    """The leaderboard of one budget level: every score reported at this budget, best (smallest) first, and which of those
    trials have moved on to the next level.  ``in_flight`` counts trials currently training towards this level."""

    def __init__(self, budget: int) -> None:
        self.budget = budget
        self.scores: List[float] = []         # ascending
        self.ids: List[str] = []              # parallel to scores
        self.advanced: Set[str] = set()
        self.in_flight = 0

    def __len__(self) -> int:
        return len(self.ids)

    def add(self, rid: str, score: float, ahead_of_ties: bool) -> int:
        """File a score; returns its rank.  A newcomer that ties goes behind the incumbents unless ``ahead_of_ties``."""
        rank = (bisect.bisect_left if ahead_of_ties else bisect.bisect_right)(self.scores, score)
        self.scores.insert(rank, score)
        self.ids.insert(rank, rid)
        return rank

    def owed(self, divisor: float) -> List[str]:
        """Trials inside the top ``1/divisor`` of this board that have not moved on yet, best first."""
        return [rid for rid in self.ids[:int(len(self.ids) / divisor)] if rid not in self.advanced]

    def forget(self, rid: str) -> None:
        for i in reversed([i for i, x in enumerate(self.ids) if x == rid]):
            del self.ids[i], self.scores[i]
        self.advanced.discard(rid)

    def to_json(self) -> Dict[str, Any]:
        return {"budget": self.budget, "scores": list(self.scores), "ids": list(self.ids), "advanced": sorted(self.advanced), "in_flight": self.in_flight}

    @classmethod
    def from_json(cls, d: Dict[str, Any]) -> "_Board":
        b = cls(int(d["budget"]))
        b.scores, b.ids, b.advanced, b.in_flight = [float(x) for x in d["scores"]], [str(x) for x in d["ids"]], set(d["advanced"]), int(d["in_flight"])
        return b


class ASHASearch(SearchMethod):
    """Asynchronous successive halving (Li et al., "A System for Massively Parallel Hyperparameter Tuning") over a ladder of
    ``num_rungs`` budget levels; level ``i`` is reached after ``sum_{j<=i} max(max_length / divisor^(num_rungs-1-j), 1)`` units.

    The rule is the paper's, stated on leaderboards: whenever a score is filed at a level, every trial inside the top
    ``1/divisor`` of that level's board that has not moved on yet is *owed* a promotion and gets it now (promotion flavour:
    the others wait, parked on their checkpoint).  ``stop_once`` is the stopping flavour: the reporting trial itself
    continues iff it ranks inside the top ``1/divisor`` at that moment, otherwise it is closed for good.  A trial that died
    is scored with the worst value wherever it would have reported next, so the ladder never waits for it.  Reports are
    settled from a work list, so a dead trial that is owed a promotion is carried upward in the same pass.

    Behaviour parity: reference ``master/pkg/searcher/asha.go`` / ``asha_stopping.go`` (same promotions, same trial counts).
    """

    name = "async_halving"

    def __init__(self, cfg: Dict[str, Any], smaller_is_better: bool) -> None:
        self.max_length, self._unit = length_units(cfg["max_length"])
        self.max_trials = int(cfg["max_trials"])
        self.divisor = float(cfg.get("divisor", 4))
        self.max_concurrent = int(cfg.get("max_concurrent_trials", 16) or 0)
        self.stop_once = bool(cfg.get("stop_once", False))
        self.sign = 1.0 if smaller_is_better else -1.0
        levels = int(cfg["num_rungs"])
        self.boards: List[_Board] = []
        budget = 0
        for i in range(levels):
            budget += max(int(self.max_length / math.pow(self.divisor, levels - 1 - i)), 1)
            self.boards.append(_Board(budget))
        self.level: Dict[str, int] = {}        # trial -> the level it is training towards / parked at
        self.dead: Set[str] = set()            # exited early: never gets another operation
        self.shut: Set[str] = set()            # a Close has been issued (or the trial is dead)
        self.finished = 0                      # trials the master reported closed
        self.rejected = 0                      # invalid hyperparameter samples, replaced by fresh ones
        self.waiting = 0                       # ValidateAfter operations not answered yet

    @property
    def num_rungs(self) -> int:
        return len(self.boards)

    # -- operations ---------------------------------------------------------------------------------
    def _spawn(self, ctx: Context) -> List[Op]:
        c = new_create(ctx.rand, hp.sample_all(ctx.hparams, ctx.rand))
        self.level[c.request_id] = 0
        self.waiting += 1
        return [c, ValidateAfter(c.request_id, self.boards[0].budget)]

    def _climb(self, rid: str, frm: int, ops: List[Op]) -> None:
        self.level[rid] = frm + 1
        self.boards[frm + 1].in_flight += 1
        if rid not in self.dead:
            ops.append(ValidateAfter(rid, max(self.boards[frm + 1].budget - self.boards[frm].budget, 1)))
            self.waiting += 1

    def _close(self, rid: str, ops: List[Op]) -> None:
        if rid not in self.shut:
            self.shut.add(rid)
            ops.append(Close(rid))

    def initial_operations(self, ctx: Context) -> List[Op]:
        if self.max_concurrent > 0:
            n = min(self.max_concurrent, self.max_trials)
        else:       # unlimited: enough starters that one of them can be owed the top level
            n = max(1, min(int(math.pow(self.divisor, self.num_rungs - 1)), self.max_trials))
        return [op for _ in range(n) for op in self._spawn(ctx)]

    def trial_created(self, ctx: Context, request_id: str) -> List[Op]:
        self.boards[0].in_flight += 1
        self.level[request_id] = 0
        return []

    def trial_closed(self, ctx: Context, request_id: str) -> List[Op]:
        self.finished += 1
        self.shut.add(request_id)
        return []

    def validation_completed(self, ctx: Context, request_id: str, metric: Any, op: ValidateAfter) -> List[Op]:
        self.waiting -= 1
        try:
            score = self.sign * float(metric)
        except (TypeError, ValueError):
            raise ValueError(f"unexpected metric type for ASHA built-in search method: {metric!r}")
        return self._settle(ctx, request_id, score)

    def trial_exited_early(self, ctx: Context, request_id: str, reason: str) -> List[Op]:
        self.waiting -= 1
        self.dead.add(request_id)
        self.shut.add(request_id)
        if reason in (ExitedReason.INVALID_HP, ExitedReason.INIT_INVALID_HP):
            # the sample was never a candidate: erase it from the ladder and draw a replacement
            self.rejected += 1
            at = self.level.get(request_id, 0)
            self.boards[at].in_flight -= 1
            for b in self.boards[:at + 1]:
                b.forget(request_id)
            return [Close(request_id)] + self._spawn(ctx)
        return self._settle(ctx, request_id, EXITED_METRIC)

    # -- the rule ------------------------------------------------------------------------------------
    def _settle(self, ctx: Context, rid: str, score: float) -> List[Op]:
        ops: List[Op] = []
        before = self.waiting
        work = collections.deque([(rid, score)])
        top = self.num_rungs - 1
        while work:
            rid, score = work.popleft()
            at = self.level[rid]
            board = self.boards[at]
            board.in_flight -= 1
            if at == top:
                board.add(rid, score, ahead_of_ties=False)
                self._close(rid, ops)
            elif self.stop_once:
                rank = board.add(rid, score, ahead_of_ties=True)
                if rank < max(int(len(board) / self.divisor), 1):
                    board.advanced.add(rid)
                    if rid not in self.dead:
                        self._climb(rid, at, ops)
                else:
                    self._close(rid, ops)
            else:
                board.add(rid, score, ahead_of_ties=False)
                for pid in board.owed(self.divisor):
                    board.advanced.add(pid)
                    self._climb(pid, at, ops)
                    if pid in self.dead:
                        work.append((pid, EXITED_METRIC))
        if self.waiting <= before and len(self.level) - self.rejected < self.max_trials:
            ops += self._spawn(ctx)                   # nobody moved up: the free slot goes to a new sample
        if not self.stop_once and len(self.boards[0]) == self.max_trials:
            self._sweep(ops)
        return ops

    def _sweep(self, ops: List[Op]) -> None:
        """Every sample has reported at the first level: trials parked on a level that nothing can reach any more will
        never be owed anything.  Close them level by level from the bottom, up to the first level still awaited."""
        for board in self.boards:
            if board.in_flight > 0:
                return
            for rid in board.ids:
                if rid not in board.advanced:
                    self._close(rid, ops)

    def progress(self, trial_progress: Dict[str, float], trials_closed: Dict[str, bool]) -> float:
        seen = len(self.boards[0])
        p = seen / (1.2 * self.max_trials)          # 20 % head-room for promotions
        if seen == self.max_trials:
            p = max(p, (self.finished - self.rejected) / float(self.max_trials))
        return p

    def unit(self) -> Optional[str]:
        return self._unit

    def snapshot(self) -> Dict[str, Any]:
        return {"boards": [b.to_json() for b in self.boards], "level": dict(self.level), "dead": sorted(self.dead), "shut": sorted(self.shut),
                "finished": self.finished, "rejected": self.rejected, "waiting": self.waiting}

    def restore(self, state: Dict[str, Any]) -> None:
        if not state:
            return
        if "rungs" in state:        # snapshots written before the leaderboard form: same facts, other names
            state = {"boards": [{"budget": r["units_needed"], "scores": [m[0] for m in r["metrics"]], "ids": [m[1] for m in r["metrics"]],
                                 "advanced": [m[1] for m in r["metrics"] if m[2]], "in_flight": r["outstanding"]} for r in state["rungs"]],
                     "level": state["trial_rungs"], "dead": [k for k, v in state["early_exit_trials"].items() if v],
                     "shut": [k for k, v in state["closed_trials"].items() if v], "finished": state["trials_completed"],
                     "rejected": state["invalid_trials"], "waiting": state["pending_trials"]}
        self.boards = [_Board.from_json(b) for b in state["boards"]]
        self.level = {str(k): int(v) for k, v in state["level"].items()}
        self.dead, self.shut = set(state["dead"]), set(state["shut"])
        self.finished, self.rejected, self.waiting = int(state["finished"]), int(state["rejected"]), int(state["waiting"])


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
