"""Scheduler tests with fake agents/tasks (test strategy: SURVEY §4, reference
``master/internal/rm/agentrm/{scheduler,priority,fair_share,fitting}_test.go``)."""
import pytest

from determined_b200.master.scheduler import (
    Agent, AllocateRequest, ResourcePool, _fair_shares, find_fit, fair_share_schedule, priority_schedule,
)


def pool(kind, n=8, **kw):
    rp = ResourcePool(scheduler=kind, **kw)
    rp.add_agent(Agent("box", n))
    return rp


def ids(d):
    return sorted(r.allocation_id for r, _ in d.to_allocate)


def test_fit_prefers_aligned_blocks():
    a = {"box": Agent("box", 8)}
    assert find_fit(AllocateRequest("x", "j", 4), a)["box"] == [0, 1, 2, 3]
    a["box"].used = {0: "t"}
    assert find_fit(AllocateRequest("x", "j", 4), a)["box"] == [4, 5, 6, 7]
    assert find_fit(AllocateRequest("x", "j", 2), a)["box"] == [2, 3]
    assert find_fit(AllocateRequest("x", "j", 8), a) is None


def test_best_vs_worst_fit_across_agents():
    agents = {"a": Agent("a", 8, used={0: "t", 1: "t", 2: "t", 3: "t"}), "b": Agent("b", 8)}
    assert list(find_fit(AllocateRequest("x", "j", 2), agents, "best")) == ["a"]
    assert list(find_fit(AllocateRequest("x", "j", 2), agents, "worst")) == ["b"]


def test_multi_agent_fit_needs_whole_idle_agents():
    agents = {"a": Agent("a", 8), "b": Agent("b", 8), "c": Agent("c", 8, used={0: "t"})}
    fit = find_fit(AllocateRequest("x", "j", 16, single_agent=False), agents)
    assert sorted(fit) == ["a", "b"] and all(len(v) == 8 for v in fit.values())
    assert find_fit(AllocateRequest("x", "j", 24, single_agent=False), agents) is None
    assert find_fit(AllocateRequest("x", "j", 16, single_agent=True), agents) is None


def test_zero_slot_tasks_always_fit():
    rp = pool("fair_share")
    rp.allocate(AllocateRequest("nb", "cmd", 0))
    assert ids(rp.tick()) == ["nb"]


def test_priority_fifo_then_preemption_then_resume():
    rp = pool("priority")
    for i in range(3):
        rp.allocate(AllocateRequest(f"lo{i}", "exp-lo", 4, priority=50, submit_time=i))
    assert ids(rp.tick()) == ["lo0", "lo1"]
    rp.allocate(AllocateRequest("hi", "exp-hi", 8, priority=1, submit_time=10))
    d = rp.tick()
    assert d.to_allocate == [] and sorted(d.to_release) == ["lo0", "lo1"]
    assert rp.tick().to_release == []          # already being released: not asked twice
    rp.release("lo0")
    assert rp.tick().to_allocate == []         # 4 free slots are reserved for `hi`, lo2 must not jump in
    rp.release("lo1")
    assert ids(rp.tick()) == ["hi"]
    rp.allocate(AllocateRequest("lo0b", "exp-lo", 4, priority=50, submit_time=20))
    assert rp.tick().to_allocate == []
    rp.release("hi")
    assert ids(rp.tick()) == ["lo0b", "lo2"]


def test_priority_without_preemption_waits():
    rp = pool("priority", preemption=False)
    rp.allocate(AllocateRequest("lo", "a", 8, priority=50))
    rp.tick()
    rp.allocate(AllocateRequest("hi", "b", 4, priority=1))
    d = rp.tick()
    assert d.to_allocate == [] and d.to_release == []


def test_priority_preempts_minimum_victims_newest_first():
    rp = pool("priority")
    for i in range(4):
        rp.allocate(AllocateRequest(f"t{i}", "lo", 2, priority=50, submit_time=i))
    rp.tick()
    rp.allocate(AllocateRequest("hi", "hi", 4, priority=10, submit_time=9))
    d = rp.tick()
    assert len(d.to_release) == 2 and set(d.to_release) <= {"t0", "t1", "t2", "t3"}
    assert "t3" in d.to_release              # newest goes first


def test_non_preemptible_tasks_are_safe():
    rp = pool("priority")
    rp.allocate(AllocateRequest("pinned", "lo", 8, priority=50, preemptible=False))
    rp.tick()
    rp.allocate(AllocateRequest("hi", "hi", 8, priority=1))
    assert rp.tick().to_release == []


def test_fair_share_water_filling():
    assert _fair_shares({"A": 8, "B": 8}, {"A": 1, "B": 1}, 8) == {"A": 4, "B": 4}
    assert _fair_shares({"A": 2, "B": 8}, {"A": 1, "B": 1}, 8) == {"A": 2, "B": 6}
    assert _fair_shares({"A": 8, "B": 8}, {"A": 3, "B": 1}, 8) == {"A": 6, "B": 2}
    s = _fair_shares({"A": 8, "B": 8, "C": 8}, {"A": 1, "B": 1, "C": 1}, 8)
    assert sum(s.values()) == 8 and max(s.values()) - min(s.values()) <= 1


def test_fair_share_rebalances_when_a_second_job_arrives():
    rp = pool("fair_share")
    for i in range(8):
        rp.allocate(AllocateRequest(f"a{i}", "A", 1, submit_time=i))
    assert len(rp.tick().to_allocate) == 8
    for i in range(8):
        rp.allocate(AllocateRequest(f"b{i}", "B", 1, submit_time=100 + i))
    d = rp.tick()
    assert len(d.to_release) == 4 and all(x.startswith("a") for x in d.to_release)
    for x in d.to_release:
        rp.release(x)
    assert ids(rp.tick()) == ["b0", "b1", "b2", "b3"]


def test_fair_share_respects_max_slots():
    rp = pool("fair_share")
    for i in range(8):
        rp.allocate(AllocateRequest(f"a{i}", "A", 1, max_slots=3, submit_time=i))
    assert len(rp.tick().to_allocate) == 3


def test_mixed_slot_sizes_pack_the_box():
    """adaptive_asha packing (BASELINE config #5): trials of 1/2/4 slots share 8 GPUs."""
    rp = pool("fair_share")
    rp.allocate(AllocateRequest("big", "E1", 4, submit_time=0))
    rp.allocate(AllocateRequest("mid", "E2", 2, submit_time=1))
    rp.allocate(AllocateRequest("s1", "E3", 1, submit_time=2))
    rp.allocate(AllocateRequest("s2", "E3", 1, submit_time=3))
    d = rp.tick()
    assert ids(d) == ["big", "mid", "s1", "s2"]
    used = sorted(s for _, asg in d.to_allocate for sl in asg.values() for s in sl)
    assert used == list(range(8))


def test_round_robin_cycles_jobs():
    rp = pool("round_robin", 4)
    for j in "AB":
        for i in range(4):
            rp.allocate(AllocateRequest(f"{j}{i}", j, 1, submit_time=ord(j) * 10 + i))
    assert ids(rp.tick()) == ["A0", "A1", "B0", "B1"]


def test_disabled_slots_are_not_used():
    rp = pool("priority")
    rp.agents["box"].slot_enabled = [True] * 4 + [False] * 4
    rp.allocate(AllocateRequest("x", "j", 8))
    assert rp.tick().to_allocate == []
    rp.allocate(AllocateRequest("y", "j", 4))
    assert ids(rp.tick()) == ["y"]


def test_move_job_ahead_and_behind_reorders_the_queue():
    rp = pool("priority", n=2)
    for i, j in enumerate(["a", "b", "c", "d"]):
        rp.allocate(AllocateRequest(f"r{j}", f"job-{j}", 2, submit_time=100.0 + i))
    d = rp.tick()
    assert ids(d) == ["ra"]                                        # the box is full: b, c, d queue in submit order
    order = lambda: [r["job_id"] for r in rp.job_queue() if r["state"] == "QUEUED"]   # noqa: E731
    assert order() == ["job-b", "job-c", "job-d"]
    rp.move_job("job-d", ahead_of="job-b")
    assert order() == ["job-d", "job-b", "job-c"]
    rp.move_job("job-b", behind_of="job-c")
    assert order() == ["job-d", "job-c", "job-b"]
    rp.release("ra")
    assert ids(rp.tick()) == ["rd"]                                # the moved job really runs first
    rp.allocate(AllocateRequest("rd2", "job-d", 2, submit_time=500.0))   # a later request of a moved job keeps its place
    assert order()[0] == "job-d"
    rp.allocate(AllocateRequest("rlow", "job-low", 2, priority=90, submit_time=50.0))
    rp.move_job("job-low", ahead_of="job-c")                      # crossing a priority lane adopts the anchor's priority
    assert [r for r in rp.job_queue() if r["job_id"] == "job-low"][0]["priority"] == 42
    with pytest.raises(KeyError):
        rp.move_job("job-x", ahead_of="job-c")
    with pytest.raises(ValueError):
        rp.move_job("job-c", ahead_of="job-c")
