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
        from determined_b200.ops import _lib, comm, optim

        _lib.load()
        world = self.world
        align = 4 * world
        self._bufs: Dict[str, Any] = {}

        def symm(name: str):
            def alloc(n: int) -> torch.Tensor:
                b = comm.SymmetricBuffer(4 * n, self.device)
                self._bufs[name] = b
                return b.tensor(torch.float32, n)
            return alloc

        with torch.cuda.device(self.device):
            self.layout = FlatLayout([{"params": self._replicated}], self.device, align=align, segment_align=align,
                                     grad_alloc=symm("grad"), param_alloc=symm("param"))
            if getattr(self, "group", None) is None:
                self.group = comm.PeerGroup.from_torch_distributed(self.device)
            self.shard_layout: Optional[FlatLayout] = None
            if self._sharded:
                self.shard_layout = FlatLayout([{"params": self._sharded}], self.device, align=4, segment_align=4)
                self.shard_m = torch.zeros(self.shard_layout.total, dtype=torch.float32, device=self.device)
                self.shard_v = torch.zeros(self.shard_layout.total, dtype=torch.float32, device=self.device)
                self._norm_scratch = optim.NormScratch(self.device) if self._clip > 0 else None
            self.peer_grads = self.group.map(self._bufs["grad"])
            self.peer_params = self.group.map(self._bufs["param"])
            total = self.layout.total
            cap = max(min(self._bucket_elems, total) // align * align, align)
            self.buckets: List[Tuple[int, int]] = []
            off = 0
            while off < total:
                n = min(cap, total - off)
                self.buckets.append((off, n))
                off += n
            shard_total = sum(n // world for _, n in self.buckets)
            self.m = torch.zeros(shard_total, dtype=torch.float32, device=self.device)
            self.v = torch.zeros(shard_total, dtype=torch.float32, device=self.device)
            self.gshard = torch.zeros(shard_total, dtype=torch.float32, device=self.device) if self._clip > 0 else None
            self.lr_dev = torch.full((), self._hp["lr"], dtype=torch.float32, device=self.device)
            self.step_dev = torch.zeros((), dtype=torch.int32, device=self.device)
            self.coef = torch.ones(1, dtype=torch.float32, device=self.device)
            self.norm = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._lr_host = self._hp["lr"]
            self.comm_stream = torch.cuda.Stream(device=self.device)
        if world > 1:
            # identical starting weights everywhere (DeepSpeed broadcasts from rank 0 too)
            comm.broadcast(self.group, self.peer_params, 0, self.layout.total, root=0)
            torch.cuda.synchronize()
        self.kernel_launches = 0

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
        if self._bf16 and self.device.type == "cuda":
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return self.module(*args, **kwargs)
        return self.module(*args, **kwargs)

    def backward(self, loss: torch.Tensor, **kw: Any) -> torch.Tensor:
        scaled = loss / self._ga if self._ga > 1 else loss
        scaled.backward()
        return loss

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
        from determined_b200.ops import _lib, comm, optim

        lib = _lib.load()
        self._sync_lr()
        cur = torch.cuda.current_stream(self.device)
        self.comm_stream.wait_stream(cur)
        st = self.comm_stream
        sp = _lib.stream_ptr(st)
        world, rank = self.world, self.rank
        hp = self._hp
        scale = 1.0 / world
        with torch.cuda.stream(st):
            _lib.check(lib.detb200_optim_tick(self.step_dev.data_ptr(), None, sp), "tick")
            self.group.begin_step()

            def run(mode: int, grad_scale_ptr: Any) -> None:
                soff = 0
                for bi, (off, n) in enumerate(self.buckets):
                    shard = n // world
                    first = 1 if (bi == 0 and mode != 2) else 0
                    _lib.check(lib.detb200_rs_adam_ag(
                        self.peer_grads, self.group.ctrl_ptrs, self.peer_params, 0, rank, world, off, n,
                        self.layout.flat_param.data_ptr() + 4 * (off + rank * shard),
                        self.m.data_ptr() + 4 * soff, self.v.data_ptr() + 4 * soff,
                        (self.gshard.data_ptr() + 4 * soff) if self.gshard is not None else None,
                        scale, self.lr_dev.data_ptr(), self.step_dev.data_ptr(), grad_scale_ptr, None,
                        float(hp["betas"][0]), float(hp["betas"][1]), float(hp["eps"]), float(hp["weight_decay"]),
                        mode, first, 7 if world > 1 else 2, 120.0, sp), "rs_adam_ag")
                    self.kernel_launches += 1
                    soff += shard

            def shard_update(grad_scale_ptr: Any) -> None:
                # partitioned weights: their gradient shards are already reduced (wgrad GEMM epilogue), so the update is local
                lay = self.shard_layout
                _lib.check(lib.detb200_adam_flat(lay.flat_param.data_ptr(), lay.flat_grad.data_ptr(), self.shard_m.data_ptr(),
                                                 self.shard_v.data_ptr(), None, lay.total, self.lr_dev.data_ptr(), self.step_dev.data_ptr(),
                                                 grad_scale_ptr, None, float(hp["betas"][0]), float(hp["betas"][1]), float(hp["eps"]),
                                                 float(hp["weight_decay"]), 1, 1.0, 1, 0, sp), "adam_flat")
                self.kernel_launches += 1

            if self._clip > 0:
                run(1, None)
                terms = self.group.stats_tensor()[0, :world]
                if self.shard_layout is not None:
                    # global norm = replicated part (per-rank slices, exchanged by the kernel) + every rank's weight shards
                    mine = optim.sumsq(self.shard_layout.flat_grad, self._norm_scratch)
                    if world > 1:
                        dist.all_reduce(mine)
                    terms = torch.cat([terms, mine])
                _lib.check(lib.detb200_clip_coef(terms.data_ptr(), terms.numel(), self._clip, 1.0, self.coef.data_ptr(), self.norm.data_ptr(), sp), "clip")
                run(2, self.coef.data_ptr())
                if self.shard_layout is not None:
                    shard_update(self.coef.data_ptr())
            else:
                run(0, None)
                if self.shard_layout is not None:
                    shard_update(None)
            # every peer finished reading our gradients (end barrier of the last bucket): clear them
            self.layout.flat_grad.zero_()
        cur.wait_stream(st)

    def get_global_grad_norm(self) -> Optional[float]:
        if self._native and self._clip > 0:
            return float(self.norm.item())
        return None

    def get_lr(self) -> List[float]:
        return [float(self.param_groups[0]["lr"])]

    def zero_grad(self, set_to_none: bool = False) -> None:
        if self._native:
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
            opt = {"m": self.m.cpu(), "v": self.v.cpu(), "step": int(self.step_dev.item()), "buckets": self.buckets, "world": self.world}
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
                if opt.get("world") != self.world or [tuple(b) for b in opt.get("buckets", [])] != [tuple(b) for b in self.buckets]:
                    raise RuntimeError("optimizer shards were saved with a different world size / bucket plan; "
                                       "re-sharding is not supported yet")
                self.m.copy_(opt["m"])
                self.v.copy_(opt["v"])
                if self.shard_layout is not None and "shard_m" in opt:
                    if int(opt.get("shard_total", -1)) != self.shard_layout.total:
                        raise RuntimeError("weight-shard optimizer state was saved for a different partitioning")
                    self.shard_m.copy_(opt["shard_m"])
                    self.shard_v.copy_(opt["shard_v"])
                self.step_dev.fill_(int(opt["step"]))
            elif not self._native and "torch_opt" in opt:
                self._torch_opt.load_state_dict(opt["torch_opt"])
        return str(d), ms.get("client_state", {})


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
