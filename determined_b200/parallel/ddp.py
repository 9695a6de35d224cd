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
from typing import Any, Callable, Dict, Iterator, List, Optional

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


class GradEngineBase:
    fused = False

    def __init__(self, dist_ctx: Any, device: torch.device, aggregation_frequency: int) -> None:
        self.dist = dist_ctx
        self.device = device
        self.agg = max(1, aggregation_frequency)
        self.records: Dict[int, _OptRecord] = {}
        self._communicating = True

    def add_optimizer(self, optimizer: torch.optim.Optimizer, average_aggregated_gradients: bool) -> torch.optim.Optimizer:
        raise NotImplementedError

    def backward_begin(self, communicate: bool) -> None:
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
            for m in models:
                tensors = [p.data for p in m.parameters()] + [b.data for b in m.buffers()]
                for t in tensors:
                    dist.broadcast(t, src=0)

    def close(self) -> None:
        pass


class GenericGradEngine(GradEngineBase):
    def add_optimizer(self, optimizer: torch.optim.Optimizer, average_aggregated_gradients: bool) -> torch.optim.Optimizer:
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
                 bucket_cap_mb: float = 25.0, timeout_s: float = 120.0) -> None:
        super().__init__(dist_ctx, device, aggregation_frequency)
        from determined_b200.ops import _lib

        _lib.load()
        self.bucket_cap = int(bucket_cap_mb * (1 << 20))
        self.timeout_s = timeout_s
        self.comm_stream = torch.cuda.Stream(device=device) if dist_ctx.size > 1 else None

    # ------------------------------------------------------------------ set-up
    def add_optimizer(self, optimizer: torch.optim.Optimizer, average_aggregated_gradients: bool) -> torch.optim.Optimizer:
        from determined_b200.ops import comm, optim

        world = self.dist.size
        rec = _OptRecord()
        rec.optimizer = optimizer
        rec.average_aggregated_gradients = average_aggregated_gradients
        align = 4 * world
        grad_alloc = None
        if world > 1:
            def grad_alloc(n: int) -> torch.Tensor:  # noqa: E306
                rec.grad_buf = comm.SymmetricBuffer(4 * n, self.device)
                return rec.grad_buf.tensor(torch.float32, n)
        with torch.cuda.device(self.device):
            rec.layout = FlatLayout(optimizer.param_groups, self.device, align=align, segment_align=align,
                                    grad_alloc=grad_alloc)
            rec.fused_step = optim.fuse_optimizer_(optimizer, rec.layout)
            rec.scratch = optim.NormScratch(self.device)
            if world > 1:
                rec.group = comm.PeerGroup.from_torch_distributed(self.device)
                rec.peer_bufs = rec.group.map(rec.grad_buf)
                rec.buckets = rec.layout.make_buckets(self.bucket_cap, align)
                for b in rec.buckets:
                    for s in b.owned:
                        rec.bucket_of[id(s.param)] = b
                for s in rec.layout.slots:
                    rec.hooks.append(s.param.register_post_accumulate_grad_hook(self._make_hook(rec)))
                self._arm(rec)
        if not rec.fused_step:
            logger.info("optimizer %s is not fusable; using its own step() on flat gradient views",
                        type(optimizer).__name__)
        self.records[id(optimizer)] = rec
        return optimizer

    # ------------------------------------------------------------------ backward-time overlap
    def _make_hook(self, rec: _OptRecord) -> Callable[[torch.Tensor], None]:
        def hook(p: torch.Tensor) -> None:
            if not self._communicating:
                return
            b = rec.bucket_of.get(id(p))
            if b is None:
                return
            if b.launched:
                raise errors.InvalidExperimentException(
                    "a parameter received a second gradient in the same batch after its bucket was already "
                    "all-reduced; with distributed training each parameter's gradient may be computed once per "
                    "batch (use optimizations.aggregation_frequency for accumulation)")
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
        if rec.group is not None:
            rec.group.begin_step()

    def _launch_bucket(self, rec: _OptRecord, b: Bucket) -> None:
        from determined_b200.ops import comm

        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.comm_stream.wait_event(ev)
        scale = 1.0 / self.dist.size
        if rec.average_aggregated_gradients:
            scale /= self.agg
        comm.allreduce_fused(rec.group, rec.peer_bufs, b.offset, b.numel, scale, stream=self.comm_stream,
                             timeout_s=self.timeout_s)
        b.launched = True

    def _launch_ready(self, rec: _OptRecord) -> None:
        while rec.next_bucket < len(rec.buckets):
            b = rec.buckets[rec.next_bucket]
            if b.pending > 0:
                return
            self._launch_bucket(rec, b)
            rec.next_bucket += 1

    def backward_begin(self, communicate: bool) -> None:
        self._communicating = communicate and self.dist.size > 1

    def _flush(self, rec: _OptRecord) -> None:
        """Launch whatever has not been launched yet (buckets holding parameters that received no
        gradient this batch — the ``find_unused_parameters`` case — or everything, if the user
        stepped without a communicating backward), then make the compute stream wait."""
        if not rec.buckets:
            return
        while rec.next_bucket < len(rec.buckets):
            b = rec.buckets[rec.next_bucket]
            if not b.launched:
                self._launch_bucket(rec, b)
            rec.next_bucket += 1
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
        grad_scale = None
        use_fused_step = rec.fused_step and scaler is None
        if isinstance(clip_grads, FusedClipGradNorm) and use_fused_step:
            if world > 1:
                terms = rec.group.stats_tensor()[0, :world]
                grad_scale = optim.clip_coef(terms, clip_grads.max_norm, rec.scratch)
            else:
                optim.sumsq(lay.flat_grad, rec.scratch)
                # the norm is of the *averaged* gradient: fold static_scale^2 in through extra handling
                if static_scale != 1.0:
                    rec.scratch.sumsq.mul_(static_scale * static_scale)
                grad_scale = optim.clip_coef(rec.scratch.sumsq, clip_grads.max_norm, rec.scratch)
            clip_grads.last_norm = rec.scratch.norm
        elif clip_grads is not None:
            if static_scale != 1.0:
                lay.flat_grad.mul_(static_scale)
                static_scale = 1.0
            if scaler is not None and unscale_for_clip:
                scaler.unscale_(optimizer)
            clip_grads(params)

        if use_fused_step:
            optimizer.step(grad_scale=grad_scale, static_scale=static_scale, zero_grad=auto_zero_grads)
        else:
            if static_scale != 1.0:
                lay.flat_grad.mul_(static_scale)
            if scaler is not None:
                scaler.step(optimizer)
            else:
                optimizer.step()
            if auto_zero_grads:
                lay.flat_grad.zero_()
        if world > 1:
            self._arm(rec)

    def close(self) -> None:
        for rec in self.records.values():
            for h in rec.hooks:
                h.remove()
            if rec.group is not None:
                rec.group.close()


def make_engine(dist_ctx: Any, device: torch.device, aggregation_frequency: int, fused: Optional[bool],
                bucket_cap_mb: float = 25.0) -> GradEngineBase:
    """Pick the engine.  ``fused=None`` means auto: fused when on CUDA (the native library is then
    mandatory — a GPU box without it is an error, not a silent fallback)."""
    want = fused if fused is not None else (device.type == "cuda")
    if want:
        if device.type != "cuda":
            raise errors.InvalidExperimentException("the fused engine requires a CUDA device")
        return FusedGradEngine(dist_ctx, device, aggregation_frequency, bucket_cap_mb)
    return GenericGradEngine(dist_ctx, device, aggregation_frequency)
