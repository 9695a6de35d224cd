"""Context parallelism: attention over a sequence that is split across the ranks of a group (SURVEY §5.7 — the reference
has no mechanism for this; its only hook is ``ModelParallelUnit``, which here carries ``context_parallel_rank / world_size``).

Rank ``r`` of ``W`` holds tokens ``[r*S/W, (r+1)*S/W)`` of every sequence: its Q, K and V shards.  ``ring_attention`` keeps Q
in place and passes the K/V shards around the ring, one hop per step, while the block that has already arrived is being
worked on (the next hop's ``isend`` / ``irecv`` are posted before the block's math starts).  Each block yields a partial
output and the log-sum-exp of its scores; partials are merged with the usual rescaling, so the result equals attention
over the whole sequence.  With ``causal=True`` blocks that lie entirely in the future are skipped, the rank's own block is
masked on the diagonal.  The backward pass runs the ring again: every K/V block travels with its gradient accumulator, each
rank adds its contribution (probabilities recomputed from the saved global log-sum-exp — no second softmax pass), and after
``W`` hops every accumulator is back at its owner.

The block math is plain torch (fp32 scores, matmuls on the device's GEMM); on a B200 the forward block uses the tcgen05
attention kernel when shapes allow (``use_native=True``; it returns the log-sum-exp the merge needs).  A variant that pulls
K/V tiles from peer memory inside the attention kernel is not built.

    mpu = ModelParallelUnit(dp_rank, dp_size, True, True, context_parallel_rank=cp.rank, context_parallel_world_size=cp.size)
    out = ring_attention(q, k, v, group=cp_group, causal=True)          # q, k, v: [B, H, S/W, D] shards
"""
import math
from typing import Any, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_sequence(x: torch.Tensor, rank: int, world: int, dim: int = 1) -> torch.Tensor:
    """This rank's contiguous slice of the sequence dimension (what a context-parallel data loader hands out)."""
    n = x.shape[dim]
    if n % world:
        raise ValueError(f"sequence length {n} is not divisible by the context-parallel size {world}")
    return x.narrow(dim, rank * (n // world), n // world)


def _block_forward(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, scale: float, diagonal: bool, use_native: bool) -> Tuple[torch.Tensor, torch.Tensor]:
    """Attention of ``q`` against one K/V block: (output normalised within the block, log-sum-exp of the scores)."""
    if use_native and q.is_cuda and q.dtype == torch.bfloat16 and q.shape[-1] in (64, 128):
        from determined_b200.ops import attention as native

        o, lse = native.attention_forward(q, k, v, diagonal, scale, "bhsd", need_lse=True)
        return o.float(), lse
    if k.shape[1] != q.shape[1]:
        rep = q.shape[1] // k.shape[1]
        k, v = k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1)
    s = torch.matmul(q.float(), k.float().transpose(-1, -2)) * scale
    if diagonal:
        mask = torch.ones(s.shape[-2], s.shape[-1], dtype=torch.bool, device=s.device).tril()
        s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    return torch.matmul(torch.exp(s - lse.unsqueeze(-1)), v.float()), lse


def _merge(o: Optional[torch.Tensor], lse: Optional[torch.Tensor], o_b: torch.Tensor, lse_b: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    if o is None or lse is None:
        return o_b, lse_b
    new = torch.logaddexp(lse, lse_b)
    return o * torch.exp(lse - new).unsqueeze(-1) + o_b * torch.exp(lse_b - new).unsqueeze(-1), new


class _Ring:
    """Neighbours of this rank in ``group`` and one hop of a list of tensors (send to the next rank, receive from the previous)."""

    def __init__(self, group: Any) -> None:
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        ranks = dist.get_process_group_ranks(group) if group is not None else list(range(self.world))
        self.next, self.prev = ranks[(self.rank + 1) % self.world], ranks[(self.rank - 1) % self.world]

    def start(self, tensors: List[torch.Tensor]) -> Tuple[List[torch.Tensor], List[Any]]:
        recv = [torch.empty(t.shape, dtype=t.dtype, device=t.device) for t in tensors]        # contiguous whatever t's strides are
        ops = []
        for t, r in zip(tensors, recv):
            ops.append(dist.P2POp(dist.isend, t.contiguous(), self.next, self.group))
            ops.append(dist.P2POp(dist.irecv, r, self.prev, self.group))
        return recv, dist.batch_isend_irecv(ops)

    @staticmethod
    def finish(reqs: List[Any]) -> None:
        for r in reqs:
            r.wait()


class _RingAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, group: Any, causal: bool, scale: float, use_native: bool) -> torch.Tensor:
        ring = _Ring(group)
        o: Optional[torch.Tensor] = None
        lse: Optional[torch.Tensor] = None
        kb, vb = k, v
        for step in range(ring.world):
            src = (ring.rank - step) % ring.world                 # whose shard kb / vb is
            nxt, reqs = ring.start([kb, vb]) if step + 1 < ring.world else ([], [])
            if not causal or src <= ring.rank:
                o, lse = _merge(o, lse, *_block_forward(q, kb, vb, scale, causal and src == ring.rank, use_native))
            ring.finish(reqs)
            if nxt:
                kb, vb = nxt
        assert o is not None and lse is not None
        out = o.to(q.dtype)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.cfg = (group, causal, scale)
        return out

    @staticmethod
    def backward(ctx: Any, do: torch.Tensor) -> Tuple[Any, ...]:
        q, k, v, out, lse = ctx.saved_tensors
        group, causal, scale = ctx.cfg
        ring = _Ring(group)
        rep = q.shape[1] // k.shape[1]
        qf, dof = q.float(), do.float()
        delta = (dof * out.float()).sum(-1)                        # rowsum(dO * O): the softmax Jacobian's diagonal term
        dq = torch.zeros_like(qf)
        kb, vb = k, v
        dkb, dvb = torch.zeros_like(k, dtype=torch.float32), torch.zeros_like(v, dtype=torch.float32)
        for step in range(ring.world):
            src = (ring.rank - step) % ring.world
            if not causal or src <= ring.rank:
                kf, vf = kb.float(), vb.float()
                if rep > 1:
                    kf, vf = kf.repeat_interleave(rep, 1), vf.repeat_interleave(rep, 1)
                s = torch.matmul(qf, kf.transpose(-1, -2)) * scale
                if causal and src == ring.rank:
                    mask = torch.ones(s.shape[-2], s.shape[-1], dtype=torch.bool, device=s.device).tril()
                    s = s.masked_fill(~mask, float("-inf"))
                p = torch.exp(s - lse.unsqueeze(-1))               # probabilities against the GLOBAL normaliser
                dv_c = torch.matmul(p.transpose(-1, -2), dof)
                ds = p * (torch.matmul(dof, vf.transpose(-1, -2)) - delta.unsqueeze(-1)) * scale
                dq += torch.matmul(ds, kf)
                dk_c = torch.matmul(ds.transpose(-1, -2), qf)
                if rep > 1:                                        # grouped-query attention: fold the query heads of a group back
                    B, H, S, D = dk_c.shape
                    dk_c, dv_c = dk_c.view(B, H // rep, rep, S, D).sum(2), dv_c.view(B, H // rep, rep, S, D).sum(2)
                dkb, dvb = dkb + dk_c, dvb + dv_c
            # the block and its gradient accumulator move on together; after `world` hops they are home
            (kb, vb, dkb, dvb), reqs = ring.start([kb, vb, dkb, dvb])
            ring.finish(reqs)
        return dq.to(q.dtype), dkb.to(k.dtype), dvb.to(v.dtype), None, None, None, None


def ring_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, group: Any = None, causal: bool = False, scale: Optional[float] = None,
                   use_native: bool = True) -> torch.Tensor:
    """Attention over the full sequence from per-rank shards ``[B, H, S_local, D]`` (K/V may have fewer heads: GQA).
    ``group`` is the context-parallel process group (default: the world).  Without an initialised process group, or with a
    group of one, this is ordinary attention on the local tensors."""
    scale = float(scale) if scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        o, _ = _block_forward(q, k, v, scale, causal, use_native)
        return o.to(q.dtype)
    return _RingAttention.apply(q, k, v, group, causal, scale, use_native)


class ContextParallelAttention(torch.nn.Module):
    """Drop-in core-attention module for a transformer block whose activations are sharded along the sequence."""

    def __init__(self, group: Any = None, causal: bool = True, scale: Optional[float] = None) -> None:
        super().__init__()
        self.group, self.causal, self.scale = group, causal, scale

    def forward(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
        return ring_attention(q, k, v, self.group, self.causal, self.scale)


def position_ids(local_len: int, rank: int, world: int, device: Any = None) -> torch.Tensor:
    """Global positions of this rank's tokens (rotary / learned position embeddings must see these, not 0..local_len)."""
    return torch.arange(rank * local_len, (rank + 1) * local_len, device=device)
