"""ZeRO-style sharded data-parallel engine with a DeepSpeedEngine-compatible surface.

The reference never implements ZeRO: ``DeepSpeedTrial`` users call ``deepspeed.initialize`` and
hand the engine to ``context.wrap_model_engine`` (reference
``pytorch/deepspeed/_deepspeed_context.py:178-209``); DeepSpeed then runs reduce-scatter →
FusedAdam → all-gather over NCCL.  This engine is what plays DeepSpeed's role on B200:

``initialize(model=…, config=…)`` → ``(engine, optimizer, None, lr_scheduler)`` like
``deepspeed.initialize``.  The engine exposes what the trial controller touches:
``train_micro_batch_size_per_gpu()``, ``gradient_accumulation_steps()``, ``train_batch_size()``,
``micro_steps``, ``global_steps``, ``backward(loss)``, ``step()``, ``train()/eval()``,
``save_checkpoint(dir, tag)``, ``load_checkpoint(dir, tag)``, ``zero_optimization_stage()``.

Memory plan (sized for 180 GB HBM3e per GPU): the fp32 working parameters (4 B/param) and the
fp32 gradient accumulation buffer (4 B/param) are replicated in *symmetric* flat buffers; Adam's
moments are **partitioned**: rank r owns slice r of every bucket (8/N B/param).  ``step()`` is ONE
kernel per bucket (``detb200_rs_adam_ag``): read slice r of all peers' gradient buffers over
NVLink, sum*scale, AdamW on the owned moment shard, and push the fresh weights into every peer's
parameter buffer — reduce-scatter, optimizer and all-gather fused, no NCCL, no gradient shard
round trip.  With ``gradient_clipping`` the kernel runs in two phases around a device-side clip
coefficient (global norm needs all shards first).

Stage 3 (``zero_optimization.stage: 3``) additionally **partitions the weights**: every linear
layer the kernels can split (``ops/gemm_comm.partition_linears_``, the role of
``deepspeed.zero.Init``) keeps only its ``out_features / world`` rows as fp32 master, with
matching gradient and moment shards.  The parameter all-gather is fused into the tcgen05 GEMMs
that consume the weights (forward and dgrad: ``detb200_ag_gemm_bf16_ex``) and the wgrad GEMM
reduce-scatters straight into the owner's shard (``detb200_gemm_rs_bf16``), so those weights need
no collective at step time: one local fused AdamW over the shard buffer.  Small / tied /
indivisible parameters (embeddings, norms, biases, an LM head whose vocabulary does not split)
stay replicated on the stage-1/2 path above.  ``stage3_param_persistence_threshold`` is the
size below which a weight is left replicated, as in DeepSpeed.

On CPU (tests) the same API runs on ``torch.distributed`` all-reduce + ``torch.optim.AdamW``.
"""
import json
import logging
import os
import pathlib
from typing import Any, Dict, Iterable, List, Optional, Tuple, Union

import torch
import torch.distributed as dist

from determined_b200.parallel.flat import FlatLayout

logger = logging.getLogger("determined.parallel.zero")


def _use_nvls(world: int) -> bool:
    """Transport of the ZeRO collectives: in-switch reduction + multicast parameter push from four ranks on (the push is
    where it pays most: a rank sends its updated shard once instead of world-1 times)."""
    env = os.environ.get("DETB200_NVLS_MIN_WORLD")
    return world >= (int(env) if env else 4)


def _load_config(config: Union[str, os.PathLike, Dict[str, Any], None]) -> Dict[str, Any]:
    if config is None:
        return {}
    if isinstance(config, (str, os.PathLike)):
        with open(config) as f:
            return json.load(f)
    return dict(config)


def resolve_batch_config(cfg: Dict[str, Any], dp_world: int) -> Tuple[int, int, int]:
    """``train_batch_size = micro_batch * grad_accum * dp_world``: any two determine the third
    (DeepSpeed's rule, reference docstring ``_deepspeed_trial.py:367-374``)."""
    tb, mb, ga = cfg.get("train_batch_size"), cfg.get("train_micro_batch_size_per_gpu"), cfg.get("gradient_accumulation_steps")
    auto = lambda v: v is None or v == "auto"  # noqa: E731
    if not auto(tb) and not auto(mb) and not auto(ga):
        pass
    elif not auto(tb) and not auto(mb):
        ga = tb // (mb * dp_world)
    elif not auto(tb) and not auto(ga):
        mb = tb // (ga * dp_world)
    elif not auto(mb):
        ga = 1 if auto(ga) else ga
        tb = mb * ga * dp_world
    elif not auto(tb):
        ga = 1 if auto(ga) else ga
        mb = tb // (ga * dp_world)
    else:
        raise ValueError("config needs train_batch_size and/or train_micro_batch_size_per_gpu")
    tb, mb, ga = int(tb), int(mb), int(ga)
    if tb != mb * ga * dp_world or min(tb, mb, ga) < 1:
        raise ValueError(f"inconsistent batch config: train_batch_size={tb} != micro_batch={mb} * grad_accum={ga} * dp_world={dp_world}")
    return tb, mb, ga


class ZeroEngine:
    """See module docstring.  Not an ``nn.Module`` subclass on purpose: ``engine.module`` is the
    user's model, ``engine(x)`` forwards to it (under bf16 autocast if configured)."""

    def __init__(self, module: torch.nn.Module, config: Union[str, Dict[str, Any], None] = None,
                 optimizer: Optional[torch.optim.Optimizer] = None, model_parameters: Optional[Iterable] = None,
                 lr_scheduler: Any = None, mpu: Any = None, device: Optional[torch.device] = None) -> None:
        self.config = _load_config(config)
        self.module = module
        self.mpu = mpu
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        self.dp_world = int(mpu.get_data_parallel_world_size()) if mpu is not None and hasattr(mpu, "get_data_parallel_world_size") else self.world
        self._tb, self._mb, self._ga = resolve_batch_config(self.config, self.dp_world)
        zo = self.config.get("zero_optimization") or {}
        self._stage = int(zo.get("stage", 0)) if isinstance(zo, dict) else 0
        self._bucket_elems = int(zo.get("reduce_bucket_size", 5e8)) if isinstance(zo, dict) and zo.get("reduce_bucket_size") not in (None, "auto") else int(5e8)
        self._clip = float(self.config.get("gradient_clipping", 0.0) or 0.0)
        self._bf16 = bool((self.config.get("bf16") or {}).get("enabled", False))
        if (self.config.get("fp16") or {}).get("enabled"):
            logger.warning("fp16 requested; B200 build trains with bf16 autocast (no loss scaling needed)")
            self._bf16 = True
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        self.device = device
        self.module.to(device)
        self._tp: Any = None
        self.partitioned = 0
        if self._stage >= 3 and isinstance(zo, dict) and zo.get("partition_parameters", device.type == "cuda"):
            self._partition_stage3(zo)
            if self.partitioned:
                model_parameters = None        # the caller's list predates the partitioning pass
        self.micro_steps = 0
        self.global_steps = 0
        self.global_samples = 0
        self.skipped_steps = 0
        at = self.config.get("autotuning") or {}
        self._probe: Any = None
        if isinstance(at, dict) and at.get("enabled"):
            from determined_b200.parallel import autotune_probe

            self._probe = autotune_probe.EngineProbe(self, at)
        params = [p for p in (model_parameters if model_parameters is not None else module.parameters()) if p.requires_grad]
        ocfg = self.config.get("optimizer") or {}
        op = dict(ocfg.get("params") or {})
        if optimizer is not None:
            g0 = optimizer.param_groups[0]
            self._hp = {"lr": float(g0.get("lr", 1e-3)), "betas": tuple(g0.get("betas", (0.9, 0.999))), "eps": float(g0.get("eps", 1e-8)),
                        "weight_decay": float(g0.get("weight_decay", 0.0))}
            if not self.partitioned:
                params = [p for g in optimizer.param_groups for p in g["params"] if p.requires_grad]
        else:
            typ = str(ocfg.get("type", "AdamW")).lower()
            if typ not in ("adam", "adamw", "fusedadam"):
                raise ValueError(f"optimizer type {ocfg.get('type')!r} is not supported by the ZeRO engine (Adam/AdamW only)")
            self._hp = {"lr": float(op.get("lr", 1e-3)), "betas": tuple(op.get("betas", (0.9, 0.999))), "eps": float(op.get("eps", 1e-8)),
                        "weight_decay": float(op.get("weight_decay", 0.01 if typ == "adamw" else 0.0))}
        self._params = params
        self._native = device.type == "cuda"
        self._sharded = [p for p in params if getattr(p, "_zero3_shard", False)]
        self._replicated = [p for p in params if not getattr(p, "_zero3_shard", False)]
        self.param_groups = [{"params": params, **{k: v for k, v in self._hp.items()}}]   # optimizer-like for LR schedulers
        if self._native:
            self._setup_native()
        else:
            self._setup_fallback()
        self.optimizer = self
        self.lr_scheduler = lr_scheduler
        self._build_scheduler()
        self.training = True

    # ------------------------------------------------------------------ set-up
    def _partition_stage3(self, zo: Dict[str, Any]) -> None:
        from determined_b200.ops import gemm_comm

        if self.world > 1:
            with torch.no_grad():          # shards must be cut from identical weights
                for t in list(self.module.parameters()) + list(self.module.buffers()):
                    dist.broadcast(t.data, src=0)
        thresh = zo.get("stage3_param_persistence_threshold", 1 << 20)
        thresh = (1 << 20) if thresh in (None, "auto") else int(float(thresh))
        if self.device.type == "cuda":
            from determined_b200.ops import _lib, comm

            _lib.load()
            with torch.cuda.device(self.device):
                self.group = comm.PeerGroup.from_torch_distributed(self.device)
                self._tp = gemm_comm.TensorParallelGemm(self.group)
        else:
            self._tp = gemm_comm.HostTensorParallel()      # CPU tests: same layers over gloo collectives
        self._tp.external_bias_reduce = True       # biases are replicated parameters: the bucketed path reduces them
        self.partitioned = gemm_comm.partition_linears_(self.module, self._tp, min_params=thresh)
        if self.partitioned and self.device.type == "cuda":
            torch.cuda.empty_cache()
        logger.info("ZeRO-3: partitioned %d weight matrices over %d ranks", self.partitioned, self.world)

    def _setup_native(self) -> None:
        """Buffers of the native path (see the module docstring for the memory plan per stage)."""
        from determined_b200.ops import _lib, comm, optim

        _lib.load()
        world = self.world
        align = 4 * world
        self._bufs: Dict[str, Any] = {}
        zo = self.config.get("zero_optimization") or {}
        zo = zo if isinstance(zo, dict) else {}
        # stage >= 2 partitions the gradients: no full-size gradient buffer, a ring of landing slots instead
        self._grad_partitioned = self._stage >= 2 and world > 1 and bool(zo.get("partition_gradients", True))
        self._ring_slots = int(zo.get("landing_slots", 2))
        self._max_ctas = zo.get("comm_ctas")

        with torch.cuda.device(self.device):
            if getattr(self, "group", None) is None:
                self.group = comm.PeerGroup.from_torch_distributed(self.device)

            def symm(name: str):
                def alloc(n: int) -> torch.Tensor:
                    b = self.group.alloc(4 * n)
                    self._bufs[name] = b
                    return b.tensor(torch.float32, n)
                return alloc

            self.layout = FlatLayout([{"params": self._replicated}], self.device, align=align, segment_align=align,
                                     grad_alloc=symm("grad"), param_alloc=symm("param"), bind_grads=not self._grad_partitioned)
            self.shard_layout: Optional[FlatLayout] = None
            if self._sharded:
                self.shard_layout = FlatLayout([{"params": self._sharded}], self.device, align=4, segment_align=4)
                self.shard_m = torch.zeros(self.shard_layout.total, dtype=torch.float32, device=self.device)
                self.shard_v = torch.zeros(self.shard_layout.total, dtype=torch.float32, device=self.device)
                self._norm_scratch = optim.NormScratch(self.device) if self._clip > 0 else None
            total = self.layout.total
            # buckets in reverse parameter order (bucket 0 = last layers = first gradients), cut at parameter boundaries
            cap_elems = max(min(self._bucket_elems, 1 << 26), align)
            self._zbuckets = self.layout.make_buckets(4 * cap_elems, align)
            self.buckets: List[Tuple[int, int]] = [(b.offset, b.numel) for b in self._zbuckets]
            self._soff: List[int] = []
            soff = 0
            for _, n in self.buckets:
                self._soff.append(soff)
                soff += n // world
            shard_total = soff
            self.peer_params = self.group.map(self._bufs["param"])
            self.mc_param = self.group.multicast(self._bufs["param"]) if world > 1 else None
            if self._grad_partitioned:
                self._slot_elems = max(n for _, n in self.buckets)
                self._bufs["land"] = self.group.alloc(4 * self._slot_elems * self._ring_slots)
                self.landing = self._bufs["land"].tensor(torch.float32, self._slot_elems * self._ring_slots)
                self.peer_grads = self.group.map(self._bufs["land"])
                self.mc_grad = self.group.multicast(self._bufs["land"])
            else:
                self.peer_grads = self.group.map(self._bufs["grad"])
                self.mc_grad = self.group.multicast(self._bufs["grad"]) if world > 1 else None
            if not _use_nvls(world):
                self.mc_grad = self.mc_param = None
            self.m = torch.zeros(shard_total, dtype=torch.float32, device=self.device)
            self.v = torch.zeros(shard_total, dtype=torch.float32, device=self.device)
            # the reduced gradient shard: its own buffer when gradients are partitioned; with replicated gradients the
            # owner's slice of the gradient buffer itself (only the owner ever reads that slice) takes the result in place
            self.gshard = torch.zeros(shard_total, dtype=torch.float32, device=self.device) if self._grad_partitioned else None
            self.lr_dev = torch.full((), self._hp["lr"], dtype=torch.float32, device=self.device)
            self.step_dev = torch.zeros((), dtype=torch.int32, device=self.device)
            self.coef = torch.ones(1, dtype=torch.float32, device=self.device)
            self.norm = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._lr_host = self._hp["lr"]
            _, hi = torch.cuda.Stream.priority_range()
            self.comm_stream = torch.cuda.Stream(device=self.device, priority=hi)
            if self._sharded and world > 1 and self._clip > 0:
                # every rank's weight-shard gradient norm reaches every rank through one tiny in-kernel all-reduce
                self._bufs["nx"] = self.group.alloc(4 * 4 * world)
                self._nx = self._bufs["nx"].tensor(torch.float32, 4 * world)
                self._nx_peers = self.group.map(self._bufs["nx"])
        if world > 1:
            # identical starting weights everywhere (DeepSpeed broadcasts from rank 0 too)
            comm.broadcast(self.group, self.peer_params, 0, self.layout.total, root=0)
            torch.cuda.synchronize()
        self.kernel_launches = 0
        # --- overlap of the reduce-scatter with backward: gradient hooks count bucket readiness
        seen: Dict[int, int] = {}
        for mod in self.module.modules():
            for p in mod._parameters.values():
                if p is not None:
                    seen[id(p)] = seen.get(id(p), 0) + 1
        tied = any(seen.get(id(p), 0) > 1 for p in self._replicated)
        self._overlap = world > 1 and bool(zo.get("overlap_comm", True)) and (not tied or self._grad_partitioned)
        if tied and self._overlap:
            raise ValueError("ZeRO stage >= 2 on the native path needs untied parameters (a tied weight receives its gradient "
                             "in two pieces far apart in backward); untie them or set zero_optimization.partition_gradients=false")
        self._bucket_of: Dict[int, int] = {}
        for bi, b in enumerate(self._zbuckets):
            for sl in b.owned:
                self._bucket_of[id(sl.param)] = bi
        self._slot_of_param = {id(sl.param): sl for sl in self.layout.slots}
        self._hooks: List[Any] = []
        if self._overlap or self._grad_partitioned:
            for sl in self.layout.slots:
                self._hooks.append(sl.param.register_post_accumulate_grad_hook(self._grad_hook))
        self._arm_micro_step()
        self._first_kernel_of_step = True
        self._comm_this_pass = False
        self._boundary_pass = False
        self._micro_in_window = 0
        self._ticked = False
        logger.info("ZeRO stage %d native path: %d replicated-path params in %d buckets (%s gradients, %s, %s), %d partitioned weights",
                    self._stage, len(self._replicated), len(self.buckets),
                    "partitioned: landing ring of %d x %.0f MB" % (self._ring_slots, 4 * self._slot_elems / 2**20) if self._grad_partitioned else "replicated",
                    "reduce-scatter overlapped with backward" if self._overlap else "reduce-scatter at step time",
                    "NVLS" if self.mc_param else "peer-to-peer", len(self._sharded))

    # ------------------------------------------------------------------ gradient hooks (native path)
    def _arm_micro_step(self) -> None:
        nb = len(self.buckets)
        self._fired: List[set] = [set() for _ in range(nb)]
        self._launched = [False] * nb
        self._next_bucket = 0
        self._deferred: List[List[Any]] = [[] for _ in range(nb)]
        self._slot_owner: List[Optional[int]] = [None] * max(1, getattr(self, "_ring_slots", 1))
        if not hasattr(self, "_slot_free"):
            self._slot_free: List[Optional[torch.cuda.Event]] = [None] * max(1, getattr(self, "_ring_slots", 1))

    def _landing_view(self, bi: int, sl: Any) -> torch.Tensor:
        off, _ = self.buckets[bi]
        base = (bi % self._ring_slots) * self._slot_elems + (sl.offset - off)
        return self.landing[base: base + sl.numel]

    def _claim_slot(self, bi: int) -> bool:
        """Make bucket ``bi`` the owner of its landing slot if the slot is free (the compute stream then waits for the
        reduce-scatter that last read it)."""
        slot = bi % self._ring_slots
        if self._slot_owner[slot] == bi:
            return True
        if self._slot_owner[slot] is not None:
            return False
        self._slot_owner[slot] = bi
        ev = self._slot_free[slot]
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)
        return True

    def _land(self, bi: int, p: torch.nn.Parameter) -> None:
        sl = self._slot_of_param[id(p)]
        self._landing_view(bi, sl).copy_(p.grad.reshape(-1))
        p.grad = None
        self._fired[bi].add(id(p))

    def _grad_hook(self, p: torch.nn.Parameter) -> None:
        bi = self._bucket_of.get(id(p))
        if bi is None:
            return
        if self._grad_partitioned:
            if self._launched[bi]:
                raise RuntimeError("a parameter received a second gradient after its bucket was reduce-scattered in this "
                                   "micro-batch; ZeRO stage >= 2 needs one gradient per parameter per backward pass")
            if self._claim_slot(bi):
                self._land(bi, p)
            else:
                self._deferred[bi].append(p)       # its slot still holds an earlier bucket: keep .grad until it is free
                return
        else:
            if not self._comm_this_pass:
                return
            self._fired[bi].add(id(p))
        if self._overlap:
            self._launch_ready()

    def _bucket_complete(self, bi: int) -> bool:
        return len(self._fired[bi]) == len(self._zbuckets[bi].owned)

    def _launch_ready(self) -> None:
        """Launch, strictly in bucket order (every rank issues the same sequence of collectives), whatever is complete."""
        progressed = True
        while progressed and self._next_bucket < len(self.buckets):
            progressed = False
            bi = self._next_bucket
            if self._grad_partitioned and self._deferred[bi] and self._claim_slot(bi):
                for p in self._deferred[bi]:
                    self._land(bi, p)
                self._deferred[bi] = []
            if self._bucket_complete(bi) and (not self._grad_partitioned or self._slot_owner[bi % self._ring_slots] == bi):
                self._launch_bucket(bi)
                self._next_bucket += 1
                progressed = True

    def _rs_mode_in_backward(self) -> int:
        if self._grad_partitioned:
            return 1 if self._boundary_pass else 3
        return 1 if self._clip > 0 else 0

    def _launch_bucket(self, bi: int) -> None:
        """Reduce-scatter of bucket ``bi`` on the communication stream (mode 1/3: into the owner's gradient shard; mode 0:
        fused with AdamW and the parameter push)."""
        from determined_b200.ops import _lib

        lib = _lib.load()
        cur = torch.cuda.current_stream(self.device)
        mode = self._rs_mode_in_backward()
        if mode == 0 and self._first_kernel_of_step:
            self._sync_lr()                 # (a fill on the compute stream: before the event the other stream waits for)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.comm_stream.wait_event(ev)
        if mode == 0 and self._first_kernel_of_step:
            with torch.cuda.stream(self.comm_stream):
                _lib.check(lib.detb200_optim_tick(self.step_dev.data_ptr(), None, _lib.stream_ptr(self.comm_stream)), "tick")
            self._ticked = True
        off, n = self.buckets[bi]
        goff = (bi % self._ring_slots) * self._slot_elems if self._grad_partitioned else off
        accumulate = 1 if (self._grad_partitioned and self._micro_in_window > 0) else 0
        first = 1 if (self._first_kernel_of_step and mode != 3) else 0
        self._rs_call(bi, mode, goff, off, n, accumulate, first, None, self.comm_stream)
        if mode != 3:
            self._first_kernel_of_step = False
        self._launched[bi] = True
        if self._grad_partitioned:
            slot = bi % self._ring_slots
            done = torch.cuda.Event()
            done.record(self.comm_stream)
            self._slot_free[slot] = done
            self._slot_owner[slot] = None

    def _rs_call(self, bi: int, mode: int, goff: int, poff: int, n: int, accumulate: int, first: int, grad_scale_ptr: Any,
                 stream: torch.cuda.Stream) -> None:
        from determined_b200.ops import _lib

        lib = _lib.load()
        world, rank, hp = self.world, self.rank, self._hp
        shard = n // world
        so = self._soff[bi]
        _lib.check(lib.detb200_rs_adam_ag_ex(
            self.peer_grads, self.group.ctrl_ptrs, self.peer_params, self.mc_grad, self.mc_param, 0, rank, world, goff, poff, n,
            self.layout.flat_param.data_ptr() + 4 * (poff + rank * shard),
            self.m.data_ptr() + 4 * so, self.v.data_ptr() + 4 * so,
            (self.gshard.data_ptr() + 4 * so) if self.gshard is not None else (self.layout.flat_grad.data_ptr() + 4 * (poff + rank * shard)),
            1.0 / world, self.lr_dev.data_ptr(), self.step_dev.data_ptr(), grad_scale_ptr, None,
            float(hp["betas"][0]), float(hp["betas"][1]), float(hp["eps"]), float(hp["weight_decay"]),
            mode, accumulate, first, 7 if world > 1 else 2, int(self._max_ctas or 0), 120.0, _lib.stream_ptr(stream)), "rs_adam_ag")
        self.kernel_launches += 1

    def _setup_fallback(self) -> None:
        self.shard_layout = None
        self._torch_opt = torch.optim.AdamW(self._params, lr=self._hp["lr"], betas=self._hp["betas"], eps=self._hp["eps"],
                                            weight_decay=self._hp["weight_decay"])
        self.param_groups = self._torch_opt.param_groups
        if self.world > 1:
            with torch.no_grad():
                for p in self.module.parameters():
                    if not getattr(p, "_zero3_shard", False):      # weight shards differ per rank by construction
                        dist.broadcast(p.data, src=0)

    def _build_scheduler(self) -> None:
        sc = self.config.get("scheduler")
        if self.lr_scheduler is not None or not sc:
            return
        typ, sp = sc.get("type"), sc.get("params") or {}
        if typ == "WarmupLR":
            lo, hi, n = float(sp.get("warmup_min_lr", 0)), float(sp.get("warmup_max_lr", self._hp["lr"])), int(sp.get("warmup_num_steps", 1000))
            self._sched_fn = lambda s: lo + (hi - lo) * min(1.0, s / max(n, 1))
        elif typ == "WarmupDecayLR":
            lo, hi, n, tot = (float(sp.get("warmup_min_lr", 0)), float(sp.get("warmup_max_lr", self._hp["lr"])),
                              int(sp.get("warmup_num_steps", 1000)), int(sp.get("total_num_steps", 10000)))
            self._sched_fn = lambda s: (lo + (hi - lo) * s / max(n, 1)) if s < n else hi * max(0.0, (tot - s) / max(tot - n, 1))
        else:
            logger.warning("scheduler type %s not built in; pass lr_scheduler= instead", typ)
            return
        self.lr_scheduler = self

    # LR-scheduler-like surface when the config asked for a built-in schedule
    def get_last_lr(self) -> List[float]:
        return [float(self.param_groups[0]["lr"])]

    # ------------------------------------------------------------------ DeepSpeedEngine surface
    def train_micro_batch_size_per_gpu(self) -> int:
        return self._mb

    def gradient_accumulation_steps(self) -> int:
        return self._ga

    def train_batch_size(self) -> int:
        return self._tb

    def zero_optimization_stage(self) -> int:
        return self._stage

    def is_gradient_accumulation_boundary(self) -> bool:
        return (self.micro_steps + 1) % self._ga == 0

    def train(self, mode: bool = True) -> "ZeroEngine":
        self.training = mode
        self.module.train(mode)
        return self

    def eval(self) -> "ZeroEngine":
        return self.train(False)

    def to(self, device: Any) -> "ZeroEngine":
        return self      # placement is fixed at construction (symmetric buffers are per device)

    def parameters(self) -> Iterable[torch.nn.Parameter]:
        return self.module.parameters()

    def named_parameters(self) -> Any:
        return self.module.named_parameters()

    def state_dict(self) -> Dict[str, Any]:
        return self.module.state_dict()

    def load_state_dict(self, sd: Dict[str, Any], strict: bool = True) -> Any:
        out = self.module.load_state_dict(dict(sd), strict=strict)
        self._rebind()
        return out

    def _rebind(self) -> None:
        if self._native:
            self.layout.rebind()
            if self.shard_layout is not None:
                self.shard_layout.rebind()

    def consolidated_state_dict(self) -> Optional[Dict[str, Any]]:
        """State dict of the *unpartitioned* model on the CPU (rank 0; other ranks get None).  With stage-3 partitioning this
        all-gathers one weight at a time — every rank must call it (DeepSpeed: ``stage3_gather_16bit_weights_on_model_save``)."""
        if not self.partitioned:
            return {k: v.detach().cpu() for k, v in self.module.state_dict().items()} if self.rank == 0 else None
        from determined_b200.ops.gemm_comm import Zero3Linear

        sd: Dict[str, Any] = {k: v.detach().cpu() for k, v in self.module.state_dict().items()} if self.rank == 0 else {}
        for name, m in self.module.named_modules():
            if isinstance(m, Zero3Linear):
                full = m.full_weight()
                if self.rank == 0:
                    prefix = name + "." if name else ""
                    sd.pop(prefix + "weight_shard", None)
                    sd[prefix + "weight"] = full.cpu()
                del full
        return sd if self.rank == 0 else None

    def __call__(self, *args: Any, **kwargs: Any) -> Any:
        return self.forward(*args, **kwargs)

    def forward(self, *args: Any, **kwargs: Any) -> Any:
        if self._probe is not None:
            return self._probe.forward(self._forward, args, kwargs)
        return self._forward(*args, **kwargs)

    def _forward(self, *args: Any, **kwargs: Any) -> Any:
        if self._bf16 and self.device.type == "cuda":
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return self.module(*args, **kwargs)
        return self.module(*args, **kwargs)

    def backward(self, loss: torch.Tensor, **kw: Any) -> torch.Tensor:
        if self._probe is not None:
            self._probe.backward_begin()
            try:
                return self._backward(loss)
            finally:
                self._probe.backward_end()
        return self._backward(loss)

    def _backward(self, loss: torch.Tensor) -> torch.Tensor:
        scaled = loss / self._ga if self._ga > 1 else loss
        if self._native:
            self._boundary_pass = self.is_gradient_accumulation_boundary()
            self._micro_in_window = self.micro_steps % self._ga
            # replicated gradients: only the last micro-batch of the window communicates (the others accumulate locally);
            # partitioned gradients: every micro-batch is reduce-scattered into the owner's shard
            self._comm_this_pass = self._boundary_pass
            if self._grad_partitioned:
                self._arm_micro_step()
        scaled.backward()
        if self._native and self._grad_partitioned:
            self._drain()
        return loss

    def _drain(self) -> None:
        """Reduce-scatter whatever the hooks could not launch: buckets whose parameters got no gradient in this pass
        (their landing region is zero-filled) and buckets that were waiting for a slot."""
        nb = len(self.buckets)
        while self._next_bucket < nb:
            bi = self._next_bucket
            if self._grad_partitioned:
                if not self._claim_slot(bi):
                    raise RuntimeError("landing slot still owned by an earlier bucket while draining")     # cannot happen in order
                for p in self._deferred[bi]:
                    self._land(bi, p)
                self._deferred[bi] = []
                for sl in self._zbuckets[bi].owned:
                    if id(sl.param) not in self._fired[bi]:
                        if sl.param.grad is not None:
                            self._land(bi, sl.param)
                        else:
                            self._landing_view(bi, sl).zero_()
                            self._fired[bi].add(id(sl.param))
            self._launch_bucket(bi)
            self._next_bucket += 1

    # ------------------------------------------------------------------ step
    def _sync_lr(self) -> None:
        lr = float(self.param_groups[0]["lr"])
        if self._native and lr != self._lr_host:
            self.lr_dev.fill_(lr)
            self._lr_host = lr

    def step(self) -> None:
        """Counts a micro step; on accumulation boundaries runs the fused reduce-scatter + AdamW +
        all-gather, steps the LR schedule and clears gradients."""
        boundary = self.is_gradient_accumulation_boundary()
        self.micro_steps += 1
        self.global_samples += self._mb * self.dp_world
        if not boundary:
            return
        if self._native:
            self._native_step()
        else:
            self._fallback_step()
        self.global_steps += 1
        if self.lr_scheduler is self and hasattr(self, "_sched_fn"):
            self.param_groups[0]["lr"] = float(self._sched_fn(self.global_steps))
        elif self.lr_scheduler is not None and self.lr_scheduler is not self:
            self.lr_scheduler.step()
        if self._probe is not None:
            self._probe.step_done()          # autotuning probe: may write its result file and exit the process

    def _fallback_step(self) -> None:
        if self.world > 1:
            grads = [p.grad for p in self._replicated if p.grad is not None]
            if grads:
                flat = torch._utils._flatten_dense_tensors(grads)
                dist.all_reduce(flat)
                flat.div_(self.world)
                for g, r in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
                    g.copy_(r)
        if self._clip > 0:
            sq = sum((p.grad.float().pow(2).sum() for p in self._replicated if p.grad is not None), torch.zeros(()))
            mine = sum((p.grad.float().pow(2).sum() for p in self._sharded if p.grad is not None), torch.zeros(()))
            if self.world > 1 and self._sharded:
                dist.all_reduce(mine)
            coef = min(1.0, self._clip / (float((sq + mine).sqrt()) + 1e-6))
            if coef < 1.0:
                for p in self._params:
                    if p.grad is not None:
                        p.grad.mul_(coef)
        self._torch_opt.step()
        self._torch_opt.zero_grad(set_to_none=False)

    def _native_step(self) -> None:
        from determined_b200.ops import _lib, optim

        lib = _lib.load()
        self._sync_lr()
        world = self.world
        hp = self._hp
        cur = torch.cuda.current_stream(self.device)
        # whatever the hooks did not launch during backward
        if not self._grad_partitioned:
            if not getattr(self, "_boundary_pass", False):
                # step() without a backward through the engine (gradients were filled some other way)
                self._first_kernel_of_step = True
            while self._next_bucket < len(self.buckets):
                self._launch_bucket(self._next_bucket)
                self._next_bucket += 1
        st = self.comm_stream
        st.wait_stream(cur)
        sp = _lib.stream_ptr(st)
        with torch.cuda.stream(st):
            if not getattr(self, "_ticked", False):
                _lib.check(lib.detb200_optim_tick(self.step_dev.data_ptr(), None, sp), "tick")
            self._ticked = False

            def shard_update(grad_scale_ptr: Any) -> None:
                # partitioned weights: their gradient shards are already reduced (wgrad GEMM epilogue), so the update is local
                lay = self.shard_layout
                _lib.check(lib.detb200_adam_flat(lay.flat_param.data_ptr(), lay.flat_grad.data_ptr(), self.shard_m.data_ptr(),
                                                 self.shard_v.data_ptr(), None, lay.total, self.lr_dev.data_ptr(), self.step_dev.data_ptr(),
                                                 grad_scale_ptr, None, float(hp["betas"][0]), float(hp["betas"][1]), float(hp["eps"]),
                                                 float(hp["weight_decay"]), 1, 1.0, 1, 0, sp), "adam_flat")
                self.kernel_launches += 1

            second_phase = self._clip > 0 or self._grad_partitioned
            gs_ptr = None
            if self._clip > 0:
                terms = self.group.stats_tensor()[0, :world]
                if self.shard_layout is not None:
                    # global norm = replicated part (per-rank slices, exchanged by the kernel) + every rank's weight shards
                    mine = optim.sumsq(self.shard_layout.flat_grad, self._norm_scratch)
                    if world > 1:
                        from determined_b200.ops import comm

                        self._nx.zero_()
                        self._nx[self.rank:self.rank + 1].copy_(mine)
                        saved_first = self.group._first_of_step
                        # statistics of this helper collective must not disturb the step's table: read the table first
                        terms = terms.clone()
                        comm.allreduce_fused(self.group, self._nx_peers, 0, 4 * world, 1.0, stream=st, timeout_s=120.0)
                        self.group._first_of_step = saved_first
                        terms = torch.cat([terms, self._nx[:world]])
                    else:
                        terms = torch.cat([terms, mine])
                _lib.check(lib.detb200_clip_coef(terms.data_ptr(), terms.numel(), self._clip, 1.0, self.coef.data_ptr(), self.norm.data_ptr(), sp), "clip")
                gs_ptr = self.coef.data_ptr()
            if second_phase:
                for bi, (off, n) in enumerate(self.buckets):
                    self._rs_call(bi, 2, off, off, n, 0, 0, gs_ptr, st)
            if self.shard_layout is not None:
                shard_update(gs_ptr)
            if not self._grad_partitioned:
                # every peer finished reading our gradients (end barrier of the last bucket): clear them
                self.layout.flat_grad.zero_()
        cur.wait_stream(st)
        self._arm_micro_step()
        self._first_kernel_of_step = True
        self._boundary_pass = False

    def get_global_grad_norm(self) -> Optional[float]:
        if self._native and self._clip > 0:
            return float(self.norm.item())
        return None

    def get_lr(self) -> List[float]:
        return [float(self.param_groups[0]["lr"])]

    def zero_grad(self, set_to_none: bool = False) -> None:
        if self._native:
            if self.layout.flat_grad is not None:
                self.layout.flat_grad.zero_()
        else:
            self._torch_opt.zero_grad(set_to_none=False)

    # ------------------------------------------------------------------ checkpoints (DeepSpeed directory layout)
    def save_checkpoint(self, save_dir: Union[str, os.PathLike], tag: Optional[str] = None, client_state: Optional[Dict[str, Any]] = None,
                        save_latest: bool = True) -> bool:
        tag = tag or f"global_step{self.global_steps}"
        d = pathlib.Path(save_dir) / str(tag)
        d.mkdir(parents=True, exist_ok=True)
        module_sd = self.consolidated_state_dict() if self.partitioned else self.module.state_dict()
        if self.rank == 0:
            torch.save({"module": module_sd, "global_steps": self.global_steps, "micro_steps": self.micro_steps,
                        "global_samples": self.global_samples, "dp_world_size": self.world, "lr": float(self.param_groups[0]["lr"]),
                        "client_state": client_state or {}, "ds_config": self.config}, d / "mp_rank_00_model_states.pt")
            if save_latest:
                (pathlib.Path(save_dir) / "latest").write_text(str(tag))
        if self._native:
            opt = {"m": self.m.cpu(), "v": self.v.cpu(), "step": int(self.step_dev.item()), "buckets": self.buckets, "world": self.world,
                   # where every parameter sits in the flat buffer (the padding depends on the world size): what a run with
                   # another world size needs to find its own slices in these shards
                   "slots": [(sl.offset, sl.numel) for sl in self.layout.slots],
                   "shard_slots": [(sl.offset, sl.numel) for sl in self.shard_layout.slots] if self.shard_layout is not None else []}
            if self.shard_layout is not None:
                opt.update(shard_m=self.shard_m.cpu(), shard_v=self.shard_v.cpu(), shard_total=self.shard_layout.total)
        else:
            opt = {"torch_opt": self._torch_opt.state_dict()}
        torch.save(opt, d / f"zero_pp_rank_{self.rank}_mp_rank_00_optim_states.pt")
        if self.world > 1:
            dist.barrier()
        return True

    def load_checkpoint(self, load_dir: Union[str, os.PathLike], tag: Optional[str] = None, load_module_strict: bool = True,
                        load_optimizer_states: bool = True, load_lr_scheduler_states: bool = True) -> Tuple[Optional[str], Dict[str, Any]]:
        load_dir = pathlib.Path(load_dir)
        if tag is None:
            latest = load_dir / "latest"
            if not latest.exists():
                return None, {}
            tag = latest.read_text().strip()
        d = load_dir / str(tag)
        ms = torch.load(d / "mp_rank_00_model_states.pt", map_location="cpu", weights_only=False)
        self.module.load_state_dict(dict(ms["module"]), strict=load_module_strict)
        self._rebind()
        self.global_steps, self.micro_steps = int(ms["global_steps"]), int(ms["micro_steps"])
        self.global_samples = int(ms.get("global_samples", 0))
        self.param_groups[0]["lr"] = float(ms.get("lr", self.param_groups[0]["lr"]))
        of = d / f"zero_pp_rank_{self.rank}_mp_rank_00_optim_states.pt"
        if load_optimizer_states and of.exists():
            opt = torch.load(of, map_location="cpu", weights_only=False)
            if self._native and "m" in opt:
                same = opt.get("world") == self.world and [tuple(b) for b in opt.get("buckets", [])] == [tuple(b) for b in self.buckets]
                if same:
                    self.m.copy_(opt["m"])
                    self.v.copy_(opt["v"])
                    if self.shard_layout is not None and "shard_m" in opt:
                        if int(opt.get("shard_total", -1)) != self.shard_layout.total:
                            raise RuntimeError("weight-shard optimizer state was saved for a different partitioning")
                        self.shard_m.copy_(opt["shard_m"])
                        self.shard_v.copy_(opt["shard_v"])
                else:
                    # saved by another world size (or bucket plan): rebuild the full moments from every rank's file and cut
                    # this rank's slices out of them
                    old_world = int(opt["world"])
                    files = [torch.load(d / f"zero_pp_rank_{r}_mp_rank_00_optim_states.pt", map_location="cpu", weights_only=False)
                             for r in range(old_world)]
                    new_slots = [(sl.offset, sl.numel) for sl in self.layout.slots]
                    for name, dst in (("m", self.m), ("v", self.v)):
                        dst.copy_(reshard_flat_state([f[name] for f in files], files[0]["buckets"], files[0]["slots"], old_world,
                                                     self.buckets, new_slots, self.layout.total, self.rank, self.world))
                    if self.shard_layout is not None and "shard_m" in files[0]:
                        ns = [(sl.offset, sl.numel) for sl in self.shard_layout.slots]
                        for name, dst in (("shard_m", self.shard_m), ("shard_v", self.shard_v)):
                            dst.copy_(reshard_row_shards([f[name] for f in files], files[0]["shard_slots"], ns, self.shard_layout.total,
                                                         self.rank, self.world))
                self.step_dev.fill_(int(opt["step"]))
            elif not self._native and "torch_opt" in opt:
                self._torch_opt.load_state_dict(opt["torch_opt"])
        return str(d), ms.get("client_state", {})


def reshard_flat_state(shards: List[torch.Tensor], old_buckets: Any, old_slots: Any, old_world: int, new_buckets: Any, new_slots: Any,
                       new_total: int, rank: int, world: int) -> torch.Tensor:
    """Optimizer state of the replicated-path parameters, saved as per-rank bucket shards by ``old_world`` ranks, re-cut for
    rank ``rank`` of ``world``.  Within a bucket ``(off, n)`` rank r holds elements ``[off + r n/W, off + (r+1) n/W)`` of the flat
    buffer; parameters sit at different flat offsets under another world size (alignment is 4 x world), so the state goes
    shards -> old flat buffer -> per parameter -> new flat buffer -> new shards."""
    old_total = max(off + n for off, n in old_buckets)
    full_old = torch.zeros(old_total, dtype=shards[0].dtype)
    cursor = [0] * old_world
    for off, n in old_buckets:
        per = n // old_world
        for r in range(old_world):
            full_old[off + r * per: off + (r + 1) * per] = shards[r][cursor[r]: cursor[r] + per]
            cursor[r] += per
    if len(old_slots) != len(new_slots) or any(a[1] != b[1] for a, b in zip(old_slots, new_slots)):
        raise RuntimeError("optimizer shards belong to a different set of parameters")
    full_new = torch.zeros(new_total, dtype=shards[0].dtype)
    for (oo, n), (no, _) in zip(old_slots, new_slots):
        full_new[no: no + n] = full_old[oo: oo + n]
    out = []
    for off, n in new_buckets:
        per = n // world
        out.append(full_new[off + rank * per: off + (rank + 1) * per])
    return torch.cat(out) if out else full_new[:0]


def reshard_row_shards(shards: List[torch.Tensor], old_slots: Any, new_slots: Any, new_total: int, rank: int, world: int) -> torch.Tensor:
    """State of row-partitioned weights (stage 3): rank r of W holds rows ``[r rows/W, (r+1) rows/W)`` of every weight,
    flattened row-major, so the full state of a weight is the concatenation of the old ranks' slots."""
    out = torch.zeros(new_total, dtype=shards[0].dtype)
    if len(old_slots) != len(new_slots):
        raise RuntimeError("weight-shard optimizer state belongs to a different set of partitioned weights")
    for (oo, on), (no, nn_) in zip(old_slots, new_slots):
        full = torch.cat([s[oo: oo + on] for s in shards])
        if full.numel() != nn_ * world:
            raise RuntimeError("a partitioned weight changed size between save and load")
        out[no: no + nn_] = full[rank * nn_: (rank + 1) * nn_]
    return out


def initialize(args: Any = None, model: Optional[torch.nn.Module] = None, optimizer: Optional[torch.optim.Optimizer] = None,
               model_parameters: Optional[Iterable] = None, training_data: Any = None, lr_scheduler: Any = None, mpu: Any = None,
               dist_init_required: Optional[bool] = None, collate_fn: Any = None, config: Union[str, Dict[str, Any], None] = None,
               config_params: Union[str, Dict[str, Any], None] = None) -> Tuple[ZeroEngine, Any, Any, Any]:
    """Drop-in for ``deepspeed.initialize``: returns ``(engine, optimizer, training_dataloader,
    lr_scheduler)``."""
    if model is None:
        raise ValueError("initialize() requires model=")
    cfg = config if config is not None else config_params
    if cfg is None and args is not None and getattr(args, "deepspeed_config", None):
        cfg = args.deepspeed_config
    engine = ZeroEngine(model, cfg, optimizer=optimizer, model_parameters=model_parameters, lr_scheduler=lr_scheduler, mpu=mpu)
    loader = None
    if training_data is not None:
        sampler = None
        if engine.dp_world > 1:
            sampler = torch.utils.data.distributed.DistributedSampler(training_data, num_replicas=engine.dp_world, rank=engine.rank % engine.dp_world)
        loader = torch.utils.data.DataLoader(training_data, batch_size=engine.train_micro_batch_size_per_gpu(), sampler=sampler,
                                             shuffle=sampler is None, collate_fn=collate_fn)
    return engine, engine.optimizer, loader, engine.lr_scheduler


def init_distributed(dist_backend: Optional[str] = None, auto_mpi_discovery: bool = False, **kw: Any) -> None:
    """Drop-in for ``deepspeed.init_distributed``."""
    if dist.is_available() and not dist.is_initialized() and int(os.environ.get("WORLD_SIZE", "1")) > 1:
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count())
        dist.init_process_group(backend=dist_backend or ("nccl" if torch.cuda.is_available() else "gloo"))
