"""Data-parallel gradient engines behind ``context.backward`` / ``context.step_optimizer``.

The reference delegates gradient exchange to ``torch.nn.parallel.DistributedDataParallel`` or
``hvd.DistributedOptimizer`` (``_pytorch_context.py:296-299,357-368``) and then averages over the
aggregation window one parameter at a time (``:805-812``).  Here the harness owns the engine:

``FusedGradEngine`` (CUDA + native library)
    * every optimizer's parameters/gradients live in flat buffers (:mod:`.flat`); the gradient
      buffer is symmetric memory mapped by all ranks;
    * post-accumulate-grad hooks count bucket readiness during ``loss.backward()`` and launch the
      fused P2P all-reduce kernel for each finished bucket on a communication stream, overlapping
      the rest of backward (buckets are launched strictly in index order on every rank);
    * the all-reduce epilogue applies 1/(world*agg) and accumulates the squared norm + non-finite
      count that clipping needs — no separate div_/unscale_/norm passes;
    * ``step()`` of a fused optimizer consumes the clip coefficient from device memory and zeroes
      the gradients in the same pass.
    Accumulation batches (``aggregation_frequency`` > 1) simply do not launch the collectives:
    autograd keeps adding into the flat buffer (the ``no_sync()`` analogue, ``:789-799``).

``GenericGradEngine`` (CPU / gloo, or fused path disabled)
    flat all-reduce through ``torch.distributed`` after backward; stock optimizer step.
"""
import logging
import os
from typing import Any, Callable, Dict, Iterator, List, Optional, Tuple

import torch

from determined_b200 import errors
from determined_b200.parallel.flat import Bucket, FlatLayout

logger = logging.getLogger("determined.parallel")


class FusedClipGradNorm:
    """Pass as ``clip_grads`` to ``context.step_optimizer`` for global L2-norm clipping.  On the
    fused engine the norm comes for free from the all-reduce epilogue and the coefficient is
    applied inside the optimizer kernel; on the generic engine it falls back to
    ``torch.nn.utils.clip_grad_norm_``."""

    def __init__(self, max_norm: float) -> None:
        self.max_norm = float(max_norm)
        self.last_norm: Optional[torch.Tensor] = None

    def __call__(self, params: Any) -> None:
        self.last_norm = torch.nn.utils.clip_grad_norm_(list(params), self.max_norm)


def _prefer_nvls(world: int) -> bool:
    """In-switch reduction against two-shot peer-to-peer.  With two ranks the peer-to-peer kernel moves the same bytes
    per link as the switch would and its loads have half the latency: measured 0.095 ms (p2p) vs 0.143 ms (NVLS) for 45 MB
    on 2 x B200 (profiles/r2_allreduce_sweep_n2.json).  From four ranks on the switch halves the wire volume."""
    env = os.environ.get("DETB200_NVLS_MIN_WORLD")
    return world >= (int(env) if env else 4)


class _OptRecord:
    def __init__(self) -> None:
        self.optimizer: Any = None
        self.layout: Optional[FlatLayout] = None
        self.buckets: List[Bucket] = []
        self.bucket_of: Dict[int, Bucket] = {}
        self.group: Any = None            # ops.comm.PeerGroup
        self.peer_bufs: Any = None
        self.grad_buf: Any = None
        self.next_bucket = 0
        self.fused_step = False
        self.average_aggregated_gradients = True
        self.scratch: Any = None
        self.hooks: List[Any] = []
        self.mc_ptr: Optional[int] = None      # multicast mapping of the gradient buffer (NVLS) or None
        self.backward_passes = 1               # communicate on the last of this many backward passes (hvd semantics)
        self.passes_seen = 0
        self.stats_stale = False               # a bucket was reduced twice this step: the running norm counts it twice
        self.wire_bf16 = False                 # 16-bit wire format (fp16_compression / gradient_compression)
        self.stage_buf: Any = None
        self.stage: Optional[torch.Tensor] = None
        self.peer_stage: Any = None
        self.mc_stage: Optional[int] = None
        self.inv_scale: Optional[torch.Tensor] = None   # 1/loss_scale of the wrapped GradScaler, device
        self.found_inf: Optional[torch.Tensor] = None   # int32[1], device
        self.unscaled_in_flight = False        # this step's buckets were unscaled + inf-checked by the all-reduce
        self.overlapped = True                 # bucketed exchange during backward (False: one all-reduce at step time)
        self.sink: Any = None                  # parallel.sink.GradSink: deferred multi-tensor gradient accumulation
        self.shadow: Any = None                # parallel.sink.LowpShadow: bf16 mirror of the flat parameters (conv weights)
        self.transport_nvls = False


class GradEngineBase:
    fused = False

    def __init__(self, dist_ctx: Any, device: torch.device, aggregation_frequency: int) -> None:
        self.dist = dist_ctx
        self.device = device
        self.agg = max(1, aggregation_frequency)
        self.records: Dict[int, _OptRecord] = {}
        self._communicating = True

    def add_optimizer(self, optimizer: torch.optim.Optimizer, average_aggregated_gradients: bool,
                      backward_passes_per_step: int = 1, fp16_compression: bool = False,
                      models: Optional[List[torch.nn.Module]] = None) -> torch.optim.Optimizer:
        raise NotImplementedError

    def backward_begin(self, communicate: bool, scaler: Any = None) -> None:
        self._communicating = communicate

    def backward_end(self) -> None:
        pass

    def step(self, optimizer: torch.optim.Optimizer, clip_grads: Optional[Callable], auto_zero_grads: bool,
             scaler: Any, unscale_for_clip: bool) -> None:
        raise NotImplementedError

    def sync_params_from_chief(self, models: List[torch.nn.Module]) -> None:
        """Make every rank start from rank 0's weights and buffers (reference
        ``_pytorch_trial.py:602-606`` / DDP constructor broadcast)."""
        if self.dist.size < 2:
            return
        import torch.distributed as dist

        if not (dist.is_available() and dist.is_initialized()):
            return
        with torch.no_grad():
            tensors = [t for m in models for t in ([p.data for p in m.parameters()] + [b.data for b in m.buffers()])]
            self._broadcast_tensors(tensors)

    def _broadcast_tensors(self, tensors: List[torch.Tensor]) -> None:
        """One collective per dtype instead of one per tensor (ResNet-18: 122 tensors)."""
        import torch.distributed as dist

        by_dtype: Dict[Any, List[torch.Tensor]] = {}
        for t in tensors:
            by_dtype.setdefault((t.dtype, t.device), []).append(t)
        for group in by_dtype.values():
            flat = torch.cat([t.reshape(-1) for t in group])
            dist.broadcast(flat, src=0)
            off = 0
            for t in group:
                n = t.numel()
                t.copy_(flat[off:off + n].view_as(t))
                off += n

    def close(self) -> None:
        pass


class GenericGradEngine(GradEngineBase):
    def add_optimizer(self, optimizer: torch.optim.Optimizer, average_aggregated_gradients: bool,
                      backward_passes_per_step: int = 1, fp16_compression: bool = False,
                      models: Optional[List[torch.nn.Module]] = None) -> torch.optim.Optimizer:
        rec = _OptRecord()
        rec.optimizer = optimizer
        rec.average_aggregated_gradients = average_aggregated_gradients
        self.records[id(optimizer)] = rec
        return optimizer

    def _allreduce(self, optimizer: torch.optim.Optimizer) -> None:
        import torch.distributed as dist

        grads = [p.grad for g in optimizer.param_groups for p in g["params"] if p.grad is not None]
        # ranks may disagree on which grads are None (unused params): give them zeros
        missing = [p for g in optimizer.param_groups for p in g["params"] if p.requires_grad and p.grad is None]
        for p in missing:
            p.grad = torch.zeros_like(p)
            grads.append(p.grad)
        if not grads:
            return
        by_dtype: Dict[Any, List[torch.Tensor]] = {}
        for gt in grads:
            by_dtype.setdefault((gt.dtype, gt.device), []).append(gt)
        for tensors in by_dtype.values():
            flat = torch._utils._flatten_dense_tensors(tensors)
            dist.all_reduce(flat)
            flat.div_(self.dist.size)
            for t, r in zip(tensors, torch._utils._unflatten_dense_tensors(flat, tensors)):
                t.copy_(r)

    def step(self, optimizer: torch.optim.Optimizer, clip_grads: Optional[Callable], auto_zero_grads: bool,
             scaler: Any, unscale_for_clip: bool) -> None:
        rec = self.records[id(optimizer)]
        if self.dist.size > 1:
            self._allreduce(optimizer)
        params = [p for g in optimizer.param_groups for p in g["params"]]
        if rec.average_aggregated_gradients and self.agg > 1:
            inv = 1.0 / self.agg
            for p in params:
                if p.grad is not None:
                    p.grad.mul_(inv)
        if clip_grads is not None:
            if scaler is not None and unscale_for_clip:
                scaler.unscale_(optimizer)
            clip_grads(params)
        if scaler is not None:
            scaler.step(optimizer)
        else:
            optimizer.step()
        if auto_zero_grads:
            optimizer.zero_grad()


class FusedGradEngine(GradEngineBase):
    fused = True

    @property
    def kernel_launches(self) -> int:
        """Native kernels launched by this process so far (every entry point of the library counts itself)."""
        from determined_b200.ops import _lib as _l

        return _l.LAUNCHES[0]

    def __init__(self, dist_ctx: Any, device: torch.device, aggregation_frequency: int,
                 bucket_cap_mb: float = 25.0, timeout_s: float = 120.0, tail_bucket_mb: float = 1.0,
                 overlap: Optional[bool] = None, overlap_min_mb: float = 128.0) -> None:
        """``overlap``: True = bucketed all-reduces launched from gradient hooks on a side stream while backward runs;
        False = ONE all-reduce over the whole gradient buffer at step time, on the compute stream, with the whole GPU;
        None = decide per optimizer: overlap when it has at least ``overlap_min_mb`` of gradients.  Measured on
        2 x B200 with the ResNet-18 trial replayed as a CUDA graph (44.7 MB of gradients, 2.69 ms/step on one GPU):
        overlapped 2.88 ms/step, everything on the compute stream 2.82 — the collective is ~0.1 ms at NVLink speed, a
        side branch of the graph that waits for a peer costs more than it hides, and backward is a train of 10-50 us
        kernels that any resident spinning CTA disturbs.  For models whose exchange takes milliseconds the overlap wins."""
        super().__init__(dist_ctx, device, aggregation_frequency)
        env = os.environ.get("DETB200_GRAD_OVERLAP", "")
        if env in ("0", "1"):
            overlap = env == "1"
        self.overlap = overlap
        self.overlap_min_bytes = int(overlap_min_mb * (1 << 20))
        from determined_b200.ops import _lib

        _lib.load()
        self.bucket_cap = int(bucket_cap_mb * (1 << 20))
        self.tail_cap = int(tail_bucket_mb * (1 << 20))
        self.timeout_s = timeout_s
        # highest priority: a bucket's CTAs should get the first SM slots that free up, not queue behind backward
        self.comm_stream = None
        # diagnosis knob (timing experiments only): "inline" keeps the buckets but runs them on the compute stream
        self._debug_comm = os.environ.get("DETB200_DEBUG_COMM", "")
        if dist_ctx.size > 1:
            _, hi = torch.cuda.Stream.priority_range()
            self.comm_stream = torch.cuda.Stream(device=device, priority=hi)

    # ------------------------------------------------------------------ set-up
    def add_optimizer(self, optimizer: torch.optim.Optimizer, average_aggregated_gradients: bool,
                      backward_passes_per_step: int = 1, fp16_compression: bool = False,
                      models: Optional[List[torch.nn.Module]] = None) -> torch.optim.Optimizer:
        from determined_b200.ops import comm, optim
        from determined_b200.ops import conv as conv_ops
        from determined_b200.parallel import sink as sink_mod

        world = self.dist.size
        rec = _OptRecord()
        rec.optimizer = optimizer
        rec.average_aggregated_gradients = average_aggregated_gradients
        rec.backward_passes = max(1, int(backward_passes_per_step))
        rec.wire_bf16 = bool(fp16_compression) and world > 1
        align = (8 if rec.wire_bf16 else 4) * world
        grad_alloc = None
        with torch.cuda.device(self.device):
            if world > 1:
                rec.group = comm.PeerGroup.from_torch_distributed(self.device)

                def grad_alloc(n: int) -> torch.Tensor:  # noqa: E306
                    rec.grad_buf = rec.group.alloc(4 * n)
                    return rec.grad_buf.tensor(torch.float32, n)
            rec.layout = FlatLayout(optimizer.param_groups, self.device, align=align, segment_align=align,
                                    grad_alloc=grad_alloc)
            rec.fused_step = optim.fuse_optimizer_(optimizer, rec.layout)
            rec.scratch = optim.NormScratch(self.device)
            # native modules hand their parameter gradients to the sink; converted convolutions read bf16 weights from the shadow
            rec.sink = sink_mod.GradSink(rec.layout)
            for sl in rec.layout.slots:
                sl.param._detb200_sink = rec.sink            # type: ignore[attr-defined]
            if models and os.environ.get("DETB200_LOWP_SHADOW", "1") != "0":
                owned = {id(sl.param) for sl in rec.layout.slots}
                if any(isinstance(mm, conv_ops.FusedConv2d) and id(mm.weight) in owned for md in models for mm in md.modules()):
                    rec.shadow = sink_mod.LowpShadow(rec.layout)
                    conv_ops.attach_shadow_(models, rec.shadow)
            rec.found_inf = torch.zeros(1, dtype=torch.int32, device=self.device)
            rec.found_inf_f = torch.zeros(1, dtype=torch.float32, device=self.device)
            rec.inv_scale = torch.ones(1, dtype=torch.float32, device=self.device)
            if world > 1:
                rec.peer_bufs = rec.group.map(rec.grad_buf)
                rec.mc_ptr = rec.group.multicast(rec.grad_buf)
                if rec.wire_bf16:
                    rec.stage_buf = rec.group.alloc(2 * rec.layout.total)
                    rec.stage = rec.stage_buf.tensor(torch.bfloat16, rec.layout.total)
                    rec.peer_stage = rec.group.map(rec.stage_buf)
                    rec.mc_stage = rec.group.multicast(rec.stage_buf)
                rec.overlapped = self.overlap if self.overlap is not None else (4 * rec.layout.total >= self.overlap_min_bytes)
                if rec.overlapped:
                    rec.buckets = rec.layout.make_buckets(self.bucket_cap, align, tail_cap_bytes=self.tail_cap)
                    for b in rec.buckets:
                        for s in b.owned:
                            rec.bucket_of[id(s.param)] = b
                    for s in rec.layout.slots:
                        rec.hooks.append(s.param.register_post_accumulate_grad_hook(self._make_hook(rec)))
                else:
                    # one bucket = the whole buffer, reduced at step time with every SM
                    rec.buckets = [Bucket(0, 0, rec.layout.total, list(rec.layout.slots), list(rec.layout.slots))]
                rec.transport_nvls = bool(rec.mc_ptr) and _prefer_nvls(world)
                self._arm(rec)
                logger.info("fused gradient engine: %d params, %s, %s%s", len(rec.layout.slots),
                            f"{len(rec.buckets)} buckets overlapped with backward" if rec.overlapped else "one all-reduce at step time",
                            "NVLS in-switch reduction" if rec.transport_nvls else "peer-to-peer two-shot",
                            ", bf16 wire format" if rec.wire_bf16 else "")
        if not rec.fused_step:
            logger.info("optimizer %s is not fusable; using its own step() on flat gradient views",
                        type(optimizer).__name__)
        self.records[id(optimizer)] = rec
        return optimizer

    def sync_params_from_chief(self, models: List[torch.nn.Module]) -> None:
        """Rank 0's weights reach every rank through the in-kernel broadcast (K8): the flat parameter buffer is staged in
        the (still empty) symmetric gradient buffer, pulled from rank 0 over NVLink by one kernel, and copied back.
        Buffers and parameters outside any optimizer go through one coalesced NCCL broadcast."""
        if self.dist.size < 2:
            return
        from determined_b200.ops import comm

        covered = set()
        with torch.no_grad():
            for rec in self.records.values():
                if rec.group is None or rec.layout is None:
                    continue
                lay = rec.layout
                lay.flat_grad.copy_(lay.flat_param)
                comm.broadcast(rec.group, rec.peer_bufs, 0, lay.total, root=0, timeout_s=self.timeout_s)
                lay.flat_param.copy_(lay.flat_grad)
                lay.flat_grad.zero_()
                if rec.shadow is not None:
                    rec.shadow.invalidate()
                covered.update(id(s.param) for s in lay.slots)
            rest = [t for m in models for t in ([p.data for p in m.parameters() if id(p) not in covered] + [b.data for b in m.buffers()])]
            if rest:
                self._broadcast_tensors(rest)
        torch.cuda.synchronize(self.device)

    # ------------------------------------------------------------------ backward-time overlap
    def _make_hook(self, rec: _OptRecord) -> Callable[[torch.Tensor], None]:
        def hook(p: torch.Tensor) -> None:
            if not self._communicating or rec.passes_seen < rec.backward_passes:
                return
            b = rec.bucket_of.get(id(p))
            if b is None:
                return
            if b.launched:
                raise errors.InvalidExperimentException(
                    "a parameter received a second gradient inside one backward pass after its bucket was already "
                    "all-reduced (a weight used by two branches whose gradients arrive far apart); run the backward "
                    "passes through context.backward(), which re-reduces buckets touched by a later pass, or turn the "
                    "fused engine off with context.experimental.use_fused_engine(False)")
            b.pending -= 1
            if b.pending == 0:
                self._launch_ready(rec)
        return hook

    def _arm(self, rec: _OptRecord) -> None:
        """Reset bucket bookkeeping for the next optimizer step."""
        for b in rec.buckets:
            b.pending = len(b.owned)
            b.launched = False
        rec.next_bucket = 0
        rec.passes_seen = 0
        rec.stats_stale = False
        rec.unscaled_in_flight = False
        if rec.group is not None:
            rec.group.begin_step()

    def _scale_of(self, rec: _OptRecord) -> float:
        scale = 1.0 / self.dist.size
        if rec.average_aggregated_gradients:
            scale /= self.agg
        return scale

    def _launch_bucket(self, rec: _OptRecord, b: Bucket) -> None:
        from determined_b200.ops import comm

        cur = torch.cuda.current_stream(self.device)
        if rec.overlapped and self._debug_comm != "inline":
            stream = self.comm_stream
            ev = torch.cuda.Event()
            ev.record(cur)
            stream.wait_event(ev)
            ctas = None                      # default budget: the rest of the GPU is busy with backward
        else:
            stream = cur
            # nothing else is running: the peer-to-peer kernel wants every SM (its loads are latency-bound), the in-switch
            # reduction is fastest from 8-16 CTAs (8 x B200, 45 MB: 0.127 ms with 16 CTAs, 0.154 ms with 148;
            # profiles/r2_allreduce_sweep_n8.json)
            ctas = 16 if rec.transport_nvls else 148
        scale = self._scale_of(rec)
        ils = rec.inv_scale if rec.unscaled_in_flight else None
        finf = rec.found_inf if rec.unscaled_in_flight else None
        if rec.wire_bf16:
            grad = rec.layout.flat_grad[b.offset:b.offset + b.numel]
            stage = rec.stage[b.offset:b.offset + b.numel]
            comm.narrow_bf16(grad, stage, stream=stream)
            comm.allreduce_bf16(rec.group, rec.peer_stage, b.offset, b.numel, scale, inv_loss_scale=ils, found_inf=finf,
                                stream=stream, timeout_s=self.timeout_s, mc_ptr=rec.mc_stage if rec.transport_nvls else None,
                                max_ctas=ctas)
            comm.widen_bf16(stage, grad, stream=stream)
        else:
            comm.allreduce_fused(rec.group, rec.peer_bufs, b.offset, b.numel, scale, inv_loss_scale=ils, found_inf=finf,
                                 stream=stream, timeout_s=self.timeout_s, mc_ptr=rec.mc_ptr if rec.transport_nvls else None,
                                 max_ctas=ctas)
        b.launched = True

    def _launch_ready(self, rec: _OptRecord) -> None:
        while rec.next_bucket < len(rec.buckets):
            b = rec.buckets[rec.next_bucket]
            if b.pending > 0 or b.launched:
                if b.launched:
                    rec.next_bucket += 1
                    continue
                return
            self._launch_bucket(rec, b)
            rec.next_bucket += 1

    def _rearm_for_another_pass(self, rec: _OptRecord, scaler_scale: Optional[torch.Tensor]) -> None:
        """A second communicating backward pass in one optimizer step (two losses flowing into the same parameters, the
        GAN pattern; torch DDP reduces every backward).  Buckets already reduced hold A = c * sum_r(g_r) on every rank with
        c = scale [/ loss_scale]; multiplying them by 1/(world*c) makes the next reduction of (that + new local gradient)
        come out as A + c * sum_r(g2_r): the sum stays linear.  All buckets are then re-armed."""
        cur = torch.cuda.current_stream(self.device)
        if rec.overlapped:
            cur.wait_stream(self.comm_stream)
        factor = 1.0 / (self.dist.size * self._scale_of(rec))
        for b in rec.buckets:
            if b.launched:
                region = rec.layout.flat_grad[b.offset:b.offset + b.numel]
                if rec.unscaled_in_flight and scaler_scale is not None:
                    region.mul_(scaler_scale.to(torch.float32) * factor)
                elif factor != 1.0:
                    region.mul_(factor)
                rec.stats_stale = True
            b.launched = False
            b.pending = len(b.owned)
        rec.next_bucket = 0

    def backward_begin(self, communicate: bool, scaler: Any = None) -> None:
        for rec in self.records.values():
            if rec.sink is not None:
                rec.sink.active = True
        self._communicating = communicate and self.dist.size > 1
        if not self._communicating:
            return
        amp = scaler is not None and scaler.is_enabled() and getattr(scaler, "_scale", None) is not None
        for rec in self.records.values():
            if not rec.buckets:
                continue
            rec.passes_seen += 1
            if rec.passes_seen < rec.backward_passes:
                continue
            if amp and rec.fused_step and not rec.unscaled_in_flight and not any(b.launched for b in rec.buckets):
                # K4: the all-reduce epilogue unscales and checks for inf/nan; the optimizer reads the flag
                torch.reciprocal(scaler._scale.reshape(1).to(torch.float32), out=rec.inv_scale)
                rec.found_inf.zero_()
                rec.unscaled_in_flight = True
            if any(b.launched for b in rec.buckets):
                self._rearm_for_another_pass(rec, scaler._scale if amp else None)

    def backward_end(self) -> None:
        """One multi-tensor kernel adds the gradients that native modules handed to the sinks during this pass into the flat
        buffers; their parameters then count as ready for the bucketed exchange."""
        for rec in self.records.values():
            if rec.sink is None:
                continue
            touched = rec.sink.flush()
            rec.sink.active = False
            if touched and rec.overlapped and rec.hooks and self._communicating and rec.passes_seen >= rec.backward_passes:
                hook = self._make_hook(rec)
                for p in touched:
                    b = rec.bucket_of.get(id(p))
                    if b is not None and not b.launched:
                        hook(p)

    def invalidate_shadows(self) -> None:
        """Parameters may have changed behind the engine's back (or a CUDA graph is about to be captured and must contain
        its own refresh): the next module that reads a bf16 weight refreshes the mirror first."""
        for rec in self.records.values():
            if rec.shadow is not None:
                rec.shadow.invalidate()

    def shadows_stale(self) -> bool:
        return any(rec.shadow is not None and rec.shadow.stale() for rec in self.records.values())

    def _flush(self, rec: _OptRecord) -> None:
        """Launch whatever has not been launched yet (buckets holding parameters that received no
        gradient this batch — the ``find_unused_parameters`` case — or everything, if the user
        stepped without a communicating backward), then make the compute stream wait."""
        if not rec.buckets:
            return
        for b in rec.buckets:
            if not b.launched:
                self._launch_bucket(rec, b)
        rec.next_bucket = len(rec.buckets)
        if rec.overlapped:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)

    # ------------------------------------------------------------------ optimizer step
    def step(self, optimizer: torch.optim.Optimizer, clip_grads: Optional[Callable], auto_zero_grads: bool,
             scaler: Any, unscale_for_clip: bool) -> None:
        from determined_b200.ops import optim

        rec = self.records[id(optimizer)]
        lay = rec.layout
        assert lay is not None
        world = self.dist.size
        if world > 1:
            self._flush(rec)
            static_scale = 1.0      # averaging already applied by the all-reduce epilogue
        else:
            static_scale = (1.0 / self.agg) if (rec.average_aggregated_gradients and self.agg > 1) else 1.0

        params = [p for g in optimizer.param_groups for p in g["params"]]
        amp = scaler is not None and scaler.is_enabled()
        if scaler is not None and not amp:
            scaler = None           # a disabled GradScaler is a pass-through
        # Fused AMP (K4): no scaler.unscale_ / scaler.step.  Multi-GPU: the all-reduce already unscaled the gradients and
        # raised found_inf.  One GPU: a sum-of-squares pass finds non-finite values (and gives the norm), and the
        # optimizer kernel multiplies by 1/loss_scale.
        fused_amp = amp and rec.fused_step and (rec.unscaled_in_flight or world == 1) \
            and (clip_grads is None or isinstance(clip_grads, FusedClipGradNorm)) and getattr(scaler, "_scale", None) is not None
        use_fused_step = rec.fused_step and (scaler is None or fused_amp)
        grad_scale = None
        skip = None
        if fused_amp:
            skip = rec.found_inf
            if world == 1:
                torch.reciprocal(scaler._scale.reshape(1).to(torch.float32), out=rec.inv_scale)
                rec.scratch.nonfinite.zero_()
                optim.sumsq(lay.flat_grad, rec.scratch, check_finite=True)
                skip = rec.scratch.nonfinite
                grad_scale = rec.inv_scale
        if isinstance(clip_grads, FusedClipGradNorm) and use_fused_step:
            if world > 1 and not rec.stats_stale:
                terms = rec.group.stats_tensor()[0, :world]
                grad_scale = optim.clip_coef(terms, clip_grads.max_norm, rec.scratch)
            else:
                if not (fused_amp and world == 1):
                    optim.sumsq(lay.flat_grad, rec.scratch)
                # the norm is of the *averaged*, unscaled gradient: fold the pending factors in
                if static_scale != 1.0:
                    rec.scratch.sumsq.mul_(static_scale * static_scale)
                if fused_amp and world == 1:
                    rec.scratch.sumsq.mul_(rec.inv_scale * rec.inv_scale)
                grad_scale = optim.clip_coef(rec.scratch.sumsq, clip_grads.max_norm, rec.scratch)
                if fused_amp and world == 1:
                    grad_scale = grad_scale.mul_(rec.inv_scale)
            clip_grads.last_norm = rec.scratch.norm
        elif clip_grads is not None:
            if static_scale != 1.0:
                lay.flat_grad.mul_(static_scale)
                static_scale = 1.0
            if scaler is not None and unscale_for_clip and not rec.unscaled_in_flight:
                scaler.unscale_(optimizer)
            clip_grads(params)

        if use_fused_step:
            optimizer.step(grad_scale=grad_scale, skip=skip, static_scale=static_scale, zero_grad=auto_zero_grads)
            if fused_amp:
                self._record_inf_check(scaler, optimizer, skip)
        else:
            if static_scale != 1.0:
                lay.flat_grad.mul_(static_scale)
            if scaler is not None and rec.unscaled_in_flight:
                # gradients are already unscaled and checked: tell the scaler so, then let it decide about the step
                self._record_inf_check(scaler, optimizer, rec.found_inf, stepped=False)
                scaler.step(optimizer)
            elif scaler is not None:
                scaler.step(optimizer)
            else:
                optimizer.step()
            if auto_zero_grads:
                lay.flat_grad.zero_()
        if rec.shadow is not None:
            rec.shadow.invalidate()          # the parameters just changed: the next forward refreshes the bf16 mirror
        if world > 1:
            self._arm(rec)

    def _record_inf_check(self, scaler: Any, optimizer: torch.optim.Optimizer, flag: torch.Tensor, stepped: bool = True) -> None:
        """Give ``scaler.update()`` what ``unscale_`` + ``step`` would have left behind: the found-inf tensor of this
        optimizer and its stage (reference flow ``_pytorch_context.py:886-897``; torch ``GradScaler.update`` reads
        ``_per_optimizer_states[id(opt)]["found_inf_per_device"]``)."""
        from torch.amp.grad_scaler import OptState

        rec = self.records[id(optimizer)]
        rec.found_inf_f.copy_(flag)
        st = scaler._per_optimizer_states[id(optimizer)]
        st["found_inf_per_device"] = {self.device: rec.found_inf_f}
        st["stage"] = OptState.STEPPED if stepped else OptState.UNSCALED

    def close(self) -> None:
        for rec in self.records.values():
            for h in rec.hooks:
                h.remove()
            if rec.group is not None:
                rec.group.close()


def make_engine(dist_ctx: Any, device: torch.device, aggregation_frequency: int, fused: Optional[bool],
                bucket_cap_mb: float = 25.0, overlap: Optional[bool] = None) -> GradEngineBase:
    """Pick the engine.  ``fused=None`` means auto: fused when on CUDA (the native library is then
    mandatory — a GPU box without it is an error, not a silent fallback)."""
    want = fused if fused is not None else (device.type == "cuda")
    if want:
        if device.type != "cuda":
            raise errors.InvalidExperimentException("the fused engine requires a CUDA device")
        return FusedGradEngine(dist_ctx, device, aggregation_frequency, bucket_cap_mb, overlap=overlap)
    return GenericGradEngine(dist_ctx, device, aggregation_frequency)


# ---------------------------------------------------------------------------------------------------------------------
# torch DDP communication hook: gradients of a stock ``DistributedDataParallel`` model through the fused all-reduce
# ---------------------------------------------------------------------------------------------------------------------
class FusedAllReduceHook:
    """``ddp_model.register_comm_hook(state, fused_allreduce_hook)`` state.

    HuggingFace ``Trainer`` (BASELINE config #3; reference ``examples/hf_trainer_api/hf_language_modeling/run_clm.py:711-735``
    with ``harness/determined/transformers/_hf_callback.py:14-47``) wraps the model in torch DDP itself, so the harness
    cannot hand the optimizer to :class:`FusedGradEngine`.  DDP does let a hook replace its NCCL all-reduce per bucket.
    DDP owns the bucket tensors (ordinary torch allocations the peers cannot map), so every bucket is staged: copied into
    its range of ONE symmetric buffer as it becomes ready (a local copy on a side stream, overlapping backward), reduced
    there by ``detb200_allreduce_fused`` (in-switch reduction from four ranks on, two-shot peer-to-peer below; averaging
    folded in), and copied back.

    ``overlap=True`` reduces each bucket as soon as it is staged, on the side stream with a small CTA budget;
    ``overlap=False`` (default) stages during backward and runs ONE all-reduce over all buckets with the whole GPU when the
    last bucket arrives.  The bucket layout is learned during the first backward pass, which runs bucket by bucket."""

    def __init__(self, device: Optional[torch.device] = None, timeout_s: float = 120.0, overlap: Optional[bool] = None,
                 overlap_ctas: int = 8) -> None:
        from determined_b200.ops import _lib, comm

        _lib.load()
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.group = comm.PeerGroup.from_torch_distributed(self.device)
        self.world = self.group.world
        self.timeout_s = timeout_s
        env = os.environ.get("DETB200_HOOK_OVERLAP", "")
        self.overlap = (env == "1") if env in ("0", "1") else bool(overlap)
        self.overlap_ctas = int(os.environ.get("DETB200_HOOK_CTAS", overlap_ctas))
        _, hi = torch.cuda.Stream.priority_range()
        self.stream = torch.cuda.Stream(device=self.device, priority=hi)
        self.use_nvls = _prefer_nvls(self.world)
        self.launches = 0
        self.layout: Dict[int, Tuple[int, int]] = {}      # bucket index -> (offset, numel) in the staging buffer
        self.buf: Any = None
        self.stage: Optional[torch.Tensor] = None
        self.peers: Any = None
        self.mc: Optional[int] = None
        self._seen: List[Tuple[int, int]] = []
        self._pending: List[Tuple[int, torch.Tensor]] = []
        self._tmp: List[Any] = []
        self._learn_passes = 0
        self.total = 0

    # -- learning passes: DDP rebuilds its buckets once, before the second backward pass, so the layout is only stable from
    #    then on; until it is known each bucket goes through a throw-away symmetric slot, one collective per bucket
    def _learning_bucket(self, bucket: Any, grad: torch.Tensor) -> None:
        from determined_b200.ops import comm

        n = grad.numel()
        align = 4 * self.world
        padded = (n + align - 1) // align * align
        buf = self.group.alloc(4 * padded)
        t = buf.tensor(torch.float32, padded)
        peers = self.group.map(buf)
        self._tmp.append(buf)
        t[:n].copy_(grad)
        self.group.begin_step()
        comm.allreduce_fused(self.group, peers, 0, padded, 1.0 / self.world, timeout_s=self.timeout_s)
        grad.copy_(t[:n])
        self.launches += 1
        self._seen.append((bucket.index(), n))
        if bucket.is_last():
            torch.cuda.synchronize(self.device)
            for tb in self._tmp:
                tb.free()
            self._tmp = []
            self._learn_passes += 1
            if self._learn_passes >= 2:
                # every rank saw the same buckets in the same order: build the single staging buffer
                off = 0
                self.layout = {}
                for idx, numel in sorted(self._seen):
                    self.layout[idx] = (off, numel)
                    off += (numel + align - 1) // align * align
                self.buf = self.group.alloc(4 * off)
                self.stage = self.buf.tensor(torch.float32, off)
                self.peers = self.group.map(self.buf)
                self.mc = self.group.multicast(self.buf) if self.use_nvls else None
                self.total = off
            self._seen = []

    def _invalidate(self, mid_pass: bool = False) -> None:
        """The bucket layout changed under us (DDP rebuilt its buckets again): finish what is staged, bucket by bucket,
        and learn the new layout."""
        from determined_b200.ops import comm

        cur = torch.cuda.current_stream(self.device)
        cur.wait_stream(self.stream)
        if not self.overlap:
            align = 4 * self.world
            for idx, g in self._pending:
                o, m = self.layout[idx]
                self.group.begin_step()
                comm.allreduce_fused(self.group, self.peers, o, (m + align - 1) // align * align, 1.0 / self.world, stream=cur,
                                     timeout_s=self.timeout_s, mc_ptr=self.mc)
                g.copy_(self.stage[o:o + m])
        self._pending = []
        torch.cuda.synchronize(self.device)
        if self.buf is not None:
            self.buf.free()
        self.buf = self.stage = self.peers = self.mc = None
        self.layout = {}
        self._learn_passes = 0 if mid_pass else 1       # a layout is only built from a pass seen from its first bucket

    def __call__(self, bucket: Any) -> "torch.futures.Future[torch.Tensor]":
        from determined_b200.ops import comm

        grad = bucket.buffer()
        if grad.dtype != torch.float32:
            raise TypeError(f"fused all-reduce hook handles fp32 gradient buckets, got {grad.dtype}")
        fut: "torch.futures.Future[torch.Tensor]" = torch.futures.Future()
        if self.stage is not None and (bucket.index() not in self.layout or self.layout[bucket.index()][1] != grad.numel()):
            self._invalidate(mid_pass=bool(self._pending))
        if self.stage is None:
            self._learning_bucket(bucket, grad)
            fut.set_result(grad)
            return fut
        off, n = self.layout[bucket.index()]
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.stream.wait_event(ev)
        align = 4 * self.world
        padded = (n + align - 1) // align * align
        with torch.cuda.stream(self.stream):
            self.stage[off:off + n].copy_(grad, non_blocking=True)
            if self.overlap:
                if not self._pending:
                    self.group.begin_step()
                comm.allreduce_fused(self.group, self.peers, off, padded, 1.0 / self.world, stream=self.stream,
                                     timeout_s=self.timeout_s, mc_ptr=self.mc, max_ctas=self.overlap_ctas)
                grad.copy_(self.stage[off:off + n], non_blocking=True)
                self.launches += 1
        self._pending.append((bucket.index(), grad))
        if bucket.is_last():
            cur.wait_stream(self.stream)
            if not self.overlap:
                # everything is staged: one collective over the whole buffer with every SM, then scatter the result back
                self.group.begin_step()
                comm.allreduce_fused(self.group, self.peers, 0, self.total, 1.0 / self.world, stream=cur,
                                     timeout_s=self.timeout_s, mc_ptr=self.mc, max_ctas=16 if self.mc else 148)
                self.launches += 1
                for idx, g in self._pending:
                    o, m = self.layout[idx]
                    g.copy_(self.stage[o:o + m], non_blocking=True)
            self._pending = []
        fut.set_result(grad)
        return fut

    def close(self) -> None:
        self.group.close()


def fused_allreduce_hook(state, bucket):  # type: ignore[no-untyped-def]
    """(state: FusedAllReduceHook, bucket: dist.GradBucket) -> Future[Tensor].  Left unannotated on purpose: DDP compares
    the annotations of a hook by identity with its own classes when the hook is registered."""
    return state(bucket)


def install_comm_hook(ddp_model: torch.nn.Module, device: Optional[torch.device] = None) -> FusedAllReduceHook:
    """Register the fused all-reduce on a ``DistributedDataParallel`` module; returns the hook state."""
    state = FusedAllReduceHook(device)
    ddp_model.register_comm_hook(state, fused_allreduce_hook)
    return state


def install_hf_comm_hook(trainer: Any) -> Any:
    """Route a HuggingFace ``Trainer``'s gradient exchange through the fused all-reduce.

    ``Trainer`` only wraps its model in DDP inside ``train()``, so this adds a ``TrainerCallback`` whose
    ``on_train_begin`` registers the hook on the wrapped model (``trainer.model_wrapped``).  Returns the callback; its
    ``state`` attribute is the :class:`FusedAllReduceHook` once training has begun.  No-op on one process."""
    try:
        import transformers

        base = transformers.TrainerCallback
    except Exception:  # pragma: no cover
        base = object

    class _InstallHook(base):  # type: ignore[misc, valid-type]
        state: Optional[FusedAllReduceHook] = None

        def on_train_begin(self, args: Any, state: Any, control: Any, **kw: Any) -> None:
            wrapped = getattr(trainer, "model_wrapped", None)
            if isinstance(wrapped, torch.nn.parallel.DistributedDataParallel) and _InstallHook.state is None:
                _InstallHook.state = install_comm_hook(wrapped)
                logger.info("HF Trainer: gradient buckets go through the fused all-reduce (%s)",
                            "NVLS" if _InstallHook.state.use_nvls else "peer-to-peer")

    cb = _InstallHook()
    trainer.add_callback(cb)
    return cb
