"""Master side of the agent protocol: remote agents register their slots, long-poll for actions and
report task output and exits.

Parity: reference ``master/internal/rm/agentrm/agent.go`` (agent actor: connect / reconnect, slot
summary, container state changes), ``master/pkg/aproto`` (``StartContainer`` / ``KillContainer`` /
``ContainerStateChanged`` / ``ContainerLog`` messages, ``master_message.go:44-60``) and
``agent/internal/agent.go`` on the other end.  Differences, on purpose:

* the reference keeps one websocket per agent; here the agent long-polls one REST endpoint
  (``POST /api/v1/agents/{id}/poll``) that carries its events up and the master's actions down —
  the same request/response machinery as every other client of this master, no second transport;
* a "container" is a process group started by the agent (there is no container runtime on the
  box); the container id is the allocation id;
* the master's own box is still driven in-process (``Master.agent_id``): remote agents add slots,
  they do not replace the local ones unless the master runs with ``num_slots=0``.

Agents that stop polling for ``DISCONNECT_AFTER_S`` are marked disconnected, their slots leave the
pool and their tasks are failed (the trial then restarts from its last checkpoint under
``max_restarts``, like the reference after ``agent_reconnect_wait`` expires).  An agent that comes
back re-registers with the list of containers it still runs and is re-attached to them.
"""
import base64
import collections
import logging
import threading
import time
from typing import Any, Deque, Dict, List, Optional

from determined_b200.master import scheduler

logger = logging.getLogger("determined.master.agents")

DISCONNECT_AFTER_S = 30.0
PYTHON_PLACEHOLDER = "{python}"        # argv[0] the agent replaces by its own interpreter


def container_spec(config: Dict[str, Any]) -> Dict[str, Any]:
    """The container-facing part of an experiment / command config (reference ``master/pkg/tasks/task.go`` ToDockerSpec:
    image, mounts, shm size, devices, capabilities, pull policy and registry credentials)."""
    env = config.get("environment") or {}
    res = config.get("resources") or {}
    out = {"image": env.get("image"), "bind_mounts": list(config.get("bind_mounts") or []), "shm_size": res.get("shm_size"),
           "devices": list(res.get("devices") or []), "add_capabilities": list(env.get("add_capabilities") or []),
           "drop_capabilities": list(env.get("drop_capabilities") or []), "force_pull_image": bool(env.get("force_pull_image")),
           "registry_auth": env.get("registry_auth")}
    return {k: v for k, v in out.items() if v not in (None, [], False)}


class RemoteAgent:
    def __init__(self, agent_id: str, info: Dict[str, Any]) -> None:
        self.id = agent_id
        self.slots: List[Dict[str, Any]] = list(info.get("slots") or [])
        self.addresses: List[str] = list(info.get("addresses") or [])
        self.version: str = str(info.get("version") or "")
        self.resource_pool: str = str(info.get("resourcePool") or "default")
        self.registered = time.time()
        self.last_seen = time.time()
        self.connected = True
        self.actions: Deque[Dict[str, Any]] = collections.deque()
        self.cv = threading.Condition()
        self.containers: Dict[str, Any] = {}      # container id -> sink (object with on_logs / on_exit)

    def push(self, action: Dict[str, Any]) -> None:
        with self.cv:
            self.actions.append(action)
            self.cv.notify_all()

    def to_json(self, pool_agent: Optional[scheduler.Agent]) -> Dict[str, Any]:
        slots = {}
        for i, s in enumerate(self.slots):
            used = pool_agent.used.get(i) if pool_agent is not None else None
            en = (pool_agent.slot_enabled or [True] * len(self.slots))[i] if pool_agent is not None else True
            slots[str(i)] = {"id": str(i), "enabled": en, "draining": False,
                             "device": {"id": i, "brand": s.get("brand", ""), "uuid": s.get("uuid", ""), "type": s.get("type", "DEVICE_TYPE_CPU")},
                             "container": {"id": used, "state": "STATE_RUNNING"} if used else None}
        return {"id": self.id, "registeredTime": self.registered, "slots": slots, "resourcePools": [self.resource_pool],
                "enabled": pool_agent.enabled if pool_agent is not None else False, "draining": False, "addresses": self.addresses,
                "version": self.version, "connected": self.connected, "containers": {c: {"id": c} for c in self.containers}}


class AgentRegistry:
    def __init__(self, master: Any) -> None:
        self.master = master
        self.agents: Dict[str, RemoteAgent] = {}
        self.lock = threading.RLock()

    # -- queries -------------------------------------------------------------------------------------
    def is_remote(self, agent_id: Optional[str]) -> bool:
        return agent_id is not None and agent_id in self.agents

    def get(self, agent_id: str) -> Optional[RemoteAgent]:
        return self.agents.get(agent_id)

    # -- agent -> master -----------------------------------------------------------------------------
    def register(self, body: Dict[str, Any]) -> Dict[str, Any]:
        agent_id = str(body.get("id") or "").strip()
        if not agent_id or agent_id == self.master.agent_id:
            raise ValueError("agent id missing or reserved")
        m = self.master
        running = [str(c) for c in body.get("containers") or []]
        with m.lock, self.lock:
            old = self.agents.get(agent_id)
            ag = RemoteAgent(agent_id, body)
            pool = m.pools.get(ag.resource_pool)
            if pool is None:
                raise ValueError(f"unknown resource pool {ag.resource_pool!r} (have: {', '.join(sorted(m.pools))})")
            elsewhere = m.pool_of_agent(agent_id)
            if elsewhere is not None and elsewhere is not pool:
                raise ValueError(f"agent {agent_id} is already registered in pool {elsewhere.name!r}")
            if old is not None:
                # reconnect: keep the tasks the agent still runs, fail the ones it lost
                for cid, sink in old.containers.items():
                    if cid in running:
                        ag.containers[cid] = sink
                    else:
                        self._lost(sink, cid, "agent restarted without this task")
            orphans = [c for c in running if c not in ag.containers]
            self.agents[agent_id] = ag
            pa = pool.agents.get(agent_id)
            if pa is None or pa.num_slots != len(ag.slots):
                pa = scheduler.Agent(agent_id, len(ag.slots))
                pool.add_agent(pa)
            pa.enabled = True
            for cid in orphans:                       # tasks of a previous master life: nobody waits for them
                ag.push({"type": "kill", "container_id": cid, "signal": "KILL"})
        logger.info("agent %s registered: %d slots, %d re-attached tasks", agent_id, len(ag.slots), len(ag.containers))
        m.kick()
        return {"agentId": agent_id, "clusterId": m.cluster_id, "masterVersion": getattr(m, "version", ""), "reattached": sorted(ag.containers),
                "pollTimeoutSeconds": 10}

    def poll(self, agent_id: str, body: Dict[str, Any], wait_s: float = 10.0) -> Dict[str, Any]:
        ag = self.agents.get(agent_id)
        if ag is None or not ag.connected:
            return {"actions": [], "reregister": True}
        ag.last_seen = time.time()
        for ev in body.get("events") or []:
            self._event(ag, ev)
        deadline = time.time() + max(0.0, min(wait_s, 30.0))
        with ag.cv:
            while not ag.actions and time.time() < deadline and ag.connected and not self.master._stop.is_set():
                ag.cv.wait(timeout=min(1.0, max(0.0, deadline - time.time())))
            out = list(ag.actions)
            ag.actions.clear()
        ag.last_seen = time.time()
        return {"actions": out}

    def _event(self, ag: RemoteAgent, ev: Dict[str, Any]) -> None:
        cid = str(ev.get("container_id") or "")
        sink = ag.containers.get(cid)
        typ = ev.get("type")
        if sink is None:
            if typ != "log":
                logger.debug("agent %s: event %s for unknown task %s", ag.id, typ, cid)
            return
        try:
            if typ == "log":
                sink.on_logs([str(x) for x in ev.get("lines") or []])
            elif typ == "exit":
                ag.containers.pop(cid, None)
                sink.on_exit(ev.get("exit_code"))
            elif typ == "start_failed":
                ag.containers.pop(cid, None)
                sink.on_logs([f"agent {ag.id} could not start the task: {ev.get('error')}"])
                sink.on_exit(127)
        except Exception:
            logger.exception("agent %s: handling %s for %s failed", ag.id, typ, cid)

    # -- master -> agent -----------------------------------------------------------------------------
    def start(self, agent_id: str, container_id: str, sink: Any, argv: List[str], env: Dict[str, str], slot_ids: List[int],
              archive: Optional[bytes] = None, workdir: str = "", hpc: Optional[Dict[str, Any]] = None,
              container: Optional[Dict[str, Any]] = None) -> None:
        """Queue a ``start`` for the agent.  ``archive`` (a tarball) is unpacked into ``<run dir>/model`` on the agent, the
        process starts in ``<run dir>/<workdir>``; ``env`` holds only what the master decides (``DET_*``, user variables):
        the agent layers it over its own environment and adds the per-host paths.  ``hpc`` carries the experiment's ``slurm`` /
        ``pbs`` sections for pools backed by a workload manager, ``container`` what an agent with a container runtime needs
        (:func:`container_spec`)."""
        ag = self.agents.get(agent_id)
        if ag is None or not ag.connected:
            raise RuntimeError(f"agent {agent_id} is not connected")
        ag.containers[container_id] = sink
        ag.push({"type": "start", "container_id": container_id, "argv": argv, "env": env, "slot_ids": slot_ids, "workdir": workdir,
                 "archive": base64.b64encode(archive).decode() if archive else None, "hpc": hpc or {}, "container": container or {}})

    def kill(self, agent_id: str, container_id: str, sig: str = "TERM") -> None:
        ag = self.agents.get(agent_id)
        if ag is not None and ag.connected:
            ag.push({"type": "kill", "container_id": container_id, "signal": sig})

    # -- liveness ------------------------------------------------------------------------------------
    def _lost(self, sink: Any, cid: str, why: str) -> None:
        try:
            sink.on_logs([f"task lost: {why}"])
            sink.on_exit(None)
        except Exception:
            logger.exception("failing task %s", cid)

    def reap(self, now: Optional[float] = None) -> None:
        """Called from the scheduler loop: drop agents that stopped polling."""
        now = now if now is not None else time.time()
        for ag in list(self.agents.values()):
            if ag.connected and now - ag.last_seen > DISCONNECT_AFTER_S:
                logger.warning("agent %s has not polled for %.0fs: disconnecting it", ag.id, now - ag.last_seen)
                self.disconnect(ag.id, "agent disconnected")

    def disconnect(self, agent_id: str, why: str = "agent removed") -> None:
        m = self.master
        with m.lock, self.lock:
            ag = self.agents.get(agent_id)
            if ag is None:
                return
            ag.connected = False
            owner = m.pool_of_agent(agent_id)
            pa = owner.agents.get(agent_id) if owner is not None else None
            if pa is not None:
                pa.enabled = False
            lost = list(ag.containers.items())
            ag.containers.clear()
            with ag.cv:
                ag.cv.notify_all()
        for cid, sink in lost:
            self._lost(sink, cid, why)
        m.kick()
