"""Resource pool + schedulers for one 8xB200 box (general over several agents for testing).

Parity (behavioural): reference ``master/internal/rm/agentrm`` — ``resourcePool`` allocate/release/
schedulerTick (``resource_pool.go:126-168,348-482``), **priority** scheduling with preemption and
back-filling (``priority.go:31-198``), **fair share** max-min with weights and per-group
``max_slots`` (``fair_share.go:82-403``), **round robin** (``round_robin.go``), and **fitting**
policies best/worst-fit over single- or multi-agent placements (``fitting.go:72-246``,
``fitting_methods.go:15-77``).  Scheduler config names (``fair_share`` default | ``priority`` |
``round_robin``; ``fitting_policy: best|worst``; ``preemption.enabled``) follow
``master/internal/config/scheduler_config.go:11-122``.

The scheduler is a pure function over (requests, running allocations, agents) returning what to
start and what to preempt — so it is tested with fake agents/tasks exactly like the reference's
``scheduler_test.go``.
"""
import dataclasses
import time
from typing import Any, Callable, Dict, List, Optional, Sequence, Set, Tuple

DEFAULT_PRIORITY = 42     # reference default (lower number = more important)


@dataclasses.dataclass
class Agent:
    id: str
    num_slots: int
    label: str = ""
    enabled: bool = True
    slot_enabled: Optional[List[bool]] = None
    used: Dict[int, str] = dataclasses.field(default_factory=dict)    # slot index -> allocation id
    zero_slot_tasks: Set[str] = dataclasses.field(default_factory=set)
    max_zero_slot_tasks: int = 100

    def free_slots(self) -> List[int]:
        if not self.enabled:
            return []
        en = self.slot_enabled or [True] * self.num_slots
        return [i for i in range(self.num_slots) if en[i] and i not in self.used]

    def num_free(self) -> int:
        return len(self.free_slots())

    def num_used(self) -> int:
        return len(self.used)

    def clone(self) -> "Agent":
        return Agent(self.id, self.num_slots, self.label, self.enabled,
                     list(self.slot_enabled) if self.slot_enabled else None, dict(self.used),
                     set(self.zero_slot_tasks), self.max_zero_slot_tasks)


@dataclasses.dataclass
class AllocateRequest:
    allocation_id: str
    job_id: str
    slots_needed: int
    priority: Optional[int] = None
    weight: float = 1.0
    max_slots: Optional[int] = None           # per-job cap (expconf resources.max_slots)
    preemptible: bool = True
    single_agent: bool = True                 # one box: always true for trials
    agent_label: str = ""
    submit_time: float = dataclasses.field(default_factory=time.time)
    name: str = ""
    queue_position: float = 0.0               # for `det job update --ahead-of`
    backfilled: bool = False

    def prio(self) -> int:
        return DEFAULT_PRIORITY if self.priority is None else self.priority


Assignment = Dict[str, List[int]]            # agent id -> slot indices


# ---------------------------------------------------------------------------------------------
# fitting
# ---------------------------------------------------------------------------------------------
def _score_best(agent: Agent, req: AllocateRequest) -> float:
    """Best fit: prefer the agent that would be left with the fewest free slots (pack tightly,
    keep whole agents free for big jobs)."""
    return float(agent.num_free() - req.slots_needed)


def _score_worst(agent: Agent, req: AllocateRequest) -> float:
    """Worst fit: prefer the emptiest agent (spread load)."""
    return -float(agent.num_free() - req.slots_needed)


FITTING = {"best": _score_best, "worst": _score_worst}


def _pick_slots(free: List[int], n: int) -> List[int]:
    """Prefer an aligned, contiguous run (keeps 2/4-GPU jobs on neighbouring devices so larger
    power-of-two groups stay available); NVSwitch makes any set equally fast, this is only about
    fragmentation."""
    if n <= 0:
        return []
    fs = sorted(free)
    sset = set(fs)
    size = 1
    while size < n:
        size *= 2
    for start in range(0, (max(fs) if fs else 0) + 1, size):
        block = [start + i for i in range(size)]
        if all(b in sset for b in block[:n]) and sum(1 for b in block if b in sset) >= n:
            # tightest aligned block with enough free slots
            return [b for b in block if b in sset][:n]
    return fs[:n]


def find_fit(req: AllocateRequest, agents: Dict[str, Agent], policy: str = "best") -> Optional[Assignment]:
    score = FITTING[policy]
    cands = [a for a in agents.values() if a.enabled and (not req.agent_label or a.label == req.agent_label)]
    if req.slots_needed == 0:
        ok = [a for a in cands if len(a.zero_slot_tasks) < a.max_zero_slot_tasks]
        if not ok:
            return None
        a = min(ok, key=lambda a: (len(a.zero_slot_tasks), a.id))
        return {a.id: []}
    single = [a for a in cands if a.num_free() >= req.slots_needed]
    if single:
        a = min(single, key=lambda a: (score(a, req), a.id))
        return {a.id: _pick_slots(a.free_slots(), req.slots_needed)}
    if req.single_agent:
        return None
    # multi-agent: whole, completely idle agents of identical size only (as the reference requires)
    idle = [a for a in cands if a.num_used() == 0 and a.num_free() == a.num_slots and a.num_slots > 0]
    by_size: Dict[int, List[Agent]] = {}
    for a in idle:
        by_size.setdefault(a.num_slots, []).append(a)
    for size in sorted(by_size, reverse=True):
        if req.slots_needed % size == 0 and len(by_size[size]) * size >= req.slots_needed:
            chosen = sorted(by_size[size], key=lambda a: a.id)[: req.slots_needed // size]
            return {a.id: list(range(size)) for a in chosen}
    return None


def apply_assignment(agents: Dict[str, Agent], alloc_id: str, asg: Assignment) -> None:
    for aid, slots in asg.items():
        if not slots:
            agents[aid].zero_slot_tasks.add(alloc_id)
        for s in slots:
            agents[aid].used[s] = alloc_id


def remove_allocation(agents: Dict[str, Agent], alloc_id: str) -> None:
    for a in agents.values():
        a.zero_slot_tasks.discard(alloc_id)
        for s in [s for s, owner in a.used.items() if owner == alloc_id]:
            del a.used[s]


def _clone_agents(agents: Dict[str, Agent]) -> Dict[str, Agent]:
    return {k: a.clone() for k, a in agents.items()}


# ---------------------------------------------------------------------------------------------
# schedulers
# ---------------------------------------------------------------------------------------------
class Decision:
    def __init__(self) -> None:
        self.to_allocate: List[Tuple[AllocateRequest, Assignment]] = []
        self.to_release: List[str] = []


def _job_usage(running: Dict[str, AllocateRequest]) -> Dict[str, int]:
    use: Dict[str, int] = {}
    for r in running.values():
        use[r.job_id] = use.get(r.job_id, 0) + r.slots_needed
    return use


def _respects_max_slots(req: AllocateRequest, usage: Dict[str, int]) -> bool:
    return req.max_slots is None or usage.get(req.job_id, 0) + req.slots_needed <= req.max_slots


def _order(reqs: Sequence[AllocateRequest]) -> List[AllocateRequest]:
    return sorted(reqs, key=lambda r: (r.queue_position, r.submit_time, r.allocation_id))


def priority_schedule(pending: Sequence[AllocateRequest], running: Dict[str, AllocateRequest],
                      agents: Dict[str, Agent], policy: str = "best", preemption: bool = True,
                      already_releasing: Optional[Set[str]] = None) -> Decision:
    """Strict priority (lower value first), FIFO within a priority.  When a request does not fit:
    lower-priority *preemptible* running tasks are preempted (lowest priority, newest first) if
    that makes it fit; while anything is waiting, lower priorities may only back-fill with
    preemptible tasks."""
    d = Decision()
    local = _clone_agents(agents)
    usage = _job_usage(running)
    # Tasks that were told to stop on an earlier tick still hold their slots: plan as if those
    # slots were free (so the waiting high-priority task reserves them and nothing lower can
    # sneak in); ResourcePool.tick() refuses any placement that touches a slot not yet vacated.
    prior = set(already_releasing or ())
    for rid in prior:
        remove_allocation(local, rid)
    releasing: Set[str] = set()
    backfilling = False
    for prio in sorted({r.prio() for r in pending}):
        unfit: List[AllocateRequest] = []
        for req in _order([r for r in pending if r.prio() == prio]):
            if not _respects_max_slots(req, usage):
                continue
            asg = find_fit(req, local, policy)
            if asg is None:
                unfit.append(req)
                continue
            apply_assignment(local, req.allocation_id, asg)    # reserve: lower priorities cannot take it
            if not backfilling or (preemption and req.preemptible):
                if backfilling:
                    req.backfilled = True
                usage[req.job_id] = usage.get(req.job_id, 0) + req.slots_needed
                d.to_allocate.append((req, asg))
        if unfit:
            backfilling = True
        if not preemption:
            continue
        for req in unfit:
            # would it fit once the already-planned preemptions complete?
            if find_fit(req, local, policy) is not None:
                apply_assignment(local, req.allocation_id, find_fit(req, local, policy) or {})
                continue
            victims = sorted((r for r in running.values()
                              if r.preemptible and r.prio() > prio and r.allocation_id not in releasing
                              and r.allocation_id not in prior),
                             key=lambda r: (-r.prio(), -r.submit_time))
            trial = _clone_agents(local)
            chosen: List[str] = []
            placed = None
            for v in victims:
                remove_allocation(trial, v.allocation_id)
                chosen.append(v.allocation_id)
                placed = find_fit(req, trial, policy)
                if placed is not None:
                    break
            if placed is not None:
                # drop victims that turned out unnecessary (greedy minimisation, newest first kept)
                for vid in list(chosen[:-1]):
                    t2 = _clone_agents(local)
                    for other in chosen:
                        if other != vid:
                            remove_allocation(t2, other)
                    if find_fit(req, t2, policy) is not None:
                        chosen.remove(vid)
                for vid in chosen:
                    remove_allocation(local, vid)
                    releasing.add(vid)
                apply_assignment(local, req.allocation_id, find_fit(req, local, policy) or {})
    d.to_release = sorted(releasing)
    return d


def _fair_shares(demands: Dict[str, int], weights: Dict[str, float], capacity: int) -> Dict[str, int]:
    """Weighted max-min fairness over integer slots (water-filling)."""
    share = {j: 0 for j in demands}
    active = {j for j, dmd in demands.items() if dmd > 0}
    left = capacity
    while left > 0 and active:
        wsum = sum(weights[j] for j in active)
        round_left = left
        progressed = False
        # proportional chunks of what was left at the start of the round; any rounding remainder
        # goes one slot at a time to the most under-served job (by share/weight, then name)
        for j in sorted(active, key=lambda j: (share[j] / weights[j], j)):
            give = min(int(round_left * weights[j] / wsum), demands[j] - share[j], left)
            if give > 0:
                share[j] += give
                left -= give
                progressed = True
        if not progressed:
            j = min(active, key=lambda j: (share[j] / weights[j], j))
            share[j] += 1
            left -= 1
        active = {j for j in active if share[j] < demands[j]}
    return share


def fair_share_schedule(pending: Sequence[AllocateRequest], running: Dict[str, AllocateRequest],
                        agents: Dict[str, Agent], policy: str = "best") -> Decision:
    """Each job (experiment) is offered a weighted max-min share of the slots, capped by its
    demand and ``max_slots``.  Jobs above their share have their newest preemptible tasks released;
    jobs below it get pending tasks started, oldest first."""
    d = Decision()
    local = _clone_agents(agents)
    capacity = sum(len([1 for i in range(a.num_slots) if (a.slot_enabled or [True] * a.num_slots)[i]])
                   for a in agents.values() if a.enabled)
    # zero-slot tasks are always placed
    for req in _order([r for r in pending if r.slots_needed == 0]):
        asg = find_fit(req, local, policy)
        if asg is not None:
            apply_assignment(local, req.allocation_id, asg)
            d.to_allocate.append((req, asg))
    jobs: Dict[str, Dict[str, Any]] = {}
    for r in list(running.values()) + [p for p in pending if 0 < p.slots_needed <= capacity]:
        if r.slots_needed == 0:
            continue
        j = jobs.setdefault(r.job_id, {"demand": 0, "weight": r.weight or 1.0, "max": r.max_slots, "first": r.submit_time})
        j["demand"] += r.slots_needed
        j["first"] = min(j["first"], r.submit_time)
    demands = {j: (min(v["demand"], v["max"]) if v["max"] is not None else v["demand"]) for j, v in jobs.items()}
    weights = {j: float(v["weight"]) for j, v in jobs.items()}
    share = _fair_shares(demands, weights, capacity)
    usage = _job_usage({k: r for k, r in running.items() if r.slots_needed > 0})
    # release from jobs over their share
    for j in sorted(jobs, key=lambda j: jobs[j]["first"]):
        over = usage.get(j, 0) - share.get(j, 0)
        if over <= 0:
            continue
        mine = sorted((r for r in running.values() if r.job_id == j and r.preemptible and r.slots_needed > 0),
                      key=lambda r: -r.submit_time)
        for r in mine:
            if over <= 0:
                break
            # never release if no pending task of an under-served job could use the space
            d.to_release.append(r.allocation_id)
            remove_allocation(local, r.allocation_id)
            usage[j] -= r.slots_needed
            over -= r.slots_needed
    if d.to_release:
        # wait for the preempted tasks to exit before handing out their slots
        under = any(usage.get(j, 0) < share.get(j, 0) for j in jobs)
        if under:
            # still start what fits right now in genuinely free slots
            local = _clone_agents(agents)
            for req, asg in d.to_allocate:
                apply_assignment(local, req.allocation_id, asg)
    for j in sorted(jobs, key=lambda j: (usage.get(j, 0) / weights[j], jobs[j]["first"])):
        for req in _order([p for p in pending if p.job_id == j and p.slots_needed > 0]):
            if usage.get(j, 0) + req.slots_needed > share.get(j, 0):
                continue
            asg = find_fit(req, local, policy)
            if asg is None:
                continue
            apply_assignment(local, req.allocation_id, asg)
            usage[j] = usage.get(j, 0) + req.slots_needed
            d.to_allocate.append((req, asg))
    return d


def round_robin_schedule(pending: Sequence[AllocateRequest], running: Dict[str, AllocateRequest],
                         agents: Dict[str, Agent], policy: str = "best") -> Decision:
    """Cycle over jobs, starting one pending task per job per round while anything fits."""
    d = Decision()
    local = _clone_agents(agents)
    by_job: Dict[str, List[AllocateRequest]] = {}
    for r in _order(pending):
        by_job.setdefault(r.job_id, []).append(r)
    order = sorted(by_job, key=lambda j: by_job[j][0].submit_time)
    usage = _job_usage(running)
    progress = True
    while progress:
        progress = False
        for j in order:
            if not by_job[j]:
                continue
            req = by_job[j][0]
            if not _respects_max_slots(req, usage):
                by_job[j].pop(0)
                continue
            asg = find_fit(req, local, policy)
            if asg is None:
                continue
            by_job[j].pop(0)
            apply_assignment(local, req.allocation_id, asg)
            usage[j] = usage.get(j, 0) + req.slots_needed
            d.to_allocate.append((req, asg))
            progress = True
    return d


class ResourcePool:
    """Holds agents + pending/running requests; ``tick()`` runs the configured scheduler and
    returns the decision after applying starts to the agent state."""

    def __init__(self, name: str = "default", scheduler: str = "fair_share", fitting_policy: str = "best",
                 preemption: bool = True) -> None:
        if scheduler not in ("fair_share", "priority", "round_robin"):
            raise ValueError(f"unknown scheduler type {scheduler!r}")
        if fitting_policy not in FITTING:
            raise ValueError(f"unknown fitting policy {fitting_policy!r}")
        self.name, self.scheduler, self.policy, self.preemption = name, scheduler, fitting_policy, preemption
        self.agents: Dict[str, Agent] = {}
        self.pending: Dict[str, AllocateRequest] = {}
        self.running: Dict[str, AllocateRequest] = {}
        self.assignments: Dict[str, Assignment] = {}
        self.releasing: Set[str] = set()
        self.job_positions: Dict[str, float] = {}      # job id -> queue position set by move_job (0 = submit order)

    def add_agent(self, agent: Agent) -> None:
        self.agents[agent.id] = agent

    def allocate(self, req: AllocateRequest) -> None:
        if req.job_id in self.job_positions:
            req.queue_position = self.job_positions[req.job_id]       # later requests of a moved job keep its place
        self.pending[req.allocation_id] = req

    def release(self, allocation_id: str) -> None:
        """The task exited (or never started): free whatever it holds."""
        self.pending.pop(allocation_id, None)
        self.running.pop(allocation_id, None)
        self.assignments.pop(allocation_id, None)
        self.releasing.discard(allocation_id)
        remove_allocation(self.agents, allocation_id)

    def set_priority(self, job_id: str, priority: int) -> None:
        for r in list(self.pending.values()) + list(self.running.values()):
            if r.job_id == job_id:
                r.priority = priority

    def set_weight(self, job_id: str, weight: float) -> None:
        for r in list(self.pending.values()) + list(self.running.values()):
            if r.job_id == job_id:
                r.weight = weight

    def move_job(self, job_id: str, ahead_of: Optional[str] = None, behind_of: Optional[str] = None) -> None:
        """``det job update JOB --ahead-of OTHER`` / ``--behind-of OTHER`` (reference ``job queue`` position updates): the job takes
        the anchor's priority and a queue position just before / after it; within one priority the queue is ordered by position,
        then submit time."""
        anchor = ahead_of or behind_of
        if anchor is None or anchor == job_id:
            raise ValueError("move_job needs another job as anchor")
        reqs = list(self.pending.values()) + list(self.running.values())
        mine = [r for r in reqs if r.job_id == job_id]
        theirs = [r for r in reqs if r.job_id == anchor]
        if not mine or not theirs:
            raise KeyError(f"job {job_id if not mine else anchor} has no request in pool {self.name}")
        # positions are compared before submit times, so give every job that never moved its submit time as position first
        for r in reqs:
            if r.job_id not in self.job_positions:
                self.job_positions[r.job_id] = min(x.submit_time for x in reqs if x.job_id == r.job_id)
        for r in reqs:
            r.queue_position = self.job_positions[r.job_id]
        ordered = sorted({r.job_id: self.job_positions[r.job_id] for r in reqs if r.job_id != job_id}.items(), key=lambda kv: kv[1])
        idx = [j for j, _ in ordered].index(anchor)
        apos = ordered[idx][1]
        if ahead_of:
            prev = ordered[idx - 1][1] if idx > 0 else apos - 2.0
            pos = (prev + apos) / 2.0
        else:
            nxt = ordered[idx + 1][1] if idx + 1 < len(ordered) else apos + 2.0
            pos = (apos + nxt) / 2.0
        self.job_positions[job_id] = pos
        prio = theirs[0].priority
        for r in mine:
            r.queue_position, r.priority = pos, prio

    def set_max_slots(self, job_id: str, max_slots: Optional[int]) -> None:
        for r in list(self.pending.values()) + list(self.running.values()):
            if r.job_id == job_id:
                r.max_slots = max_slots

    def tick(self) -> Decision:
        pend = list(self.pending.values())
        run = {k: v for k, v in self.running.items() if k not in self.releasing}
        if self.scheduler == "priority":
            d = priority_schedule(pend, run, self.agents, self.policy, self.preemption, set(self.releasing))
        elif self.scheduler == "round_robin":
            d = round_robin_schedule(pend, run, self.agents, self.policy)
        else:
            d = fair_share_schedule(pend, run, self.agents, self.policy)
        started = []
        for req, asg in d.to_allocate:
            # slots of tasks that are being preempted are not free yet
            ok = all(s not in self.agents[aid].used for aid, slots in asg.items() for s in slots)
            if not ok:
                continue
            apply_assignment(self.agents, req.allocation_id, asg)
            self.pending.pop(req.allocation_id, None)
            self.running[req.allocation_id] = req
            self.assignments[req.allocation_id] = asg
            started.append((req, asg))
        d.to_allocate = started
        d.to_release = [a for a in d.to_release if a not in self.releasing]
        self.releasing.update(d.to_release)
        return d

    def job_queue(self) -> List[Dict[str, Any]]:
        rows = []
        for r in _order(list(self.running.values())) + sorted(self.pending.values(), key=lambda r: (r.prio(), r.queue_position, r.submit_time)):
            rows.append({"allocation_id": r.allocation_id, "job_id": r.job_id, "slots": r.slots_needed,
                         "priority": r.prio(), "weight": r.weight,
                         "state": "SCHEDULED" if r.allocation_id in self.running else "QUEUED", "name": r.name})
        return rows
