"""Flat parameter / gradient storage.

The fused engine keeps, per optimizer, ONE contiguous fp32 buffer for parameters and one for
gradients; every ``param.data`` and ``param.grad`` is a view into them.  That is what lets the
gradient all-reduce run over a handful of large buckets with no pack/unpack copies (autograd
accumulates straight into the communication buffer — the reference's DDP does this with
``gradient_as_bucket_view``, Horovod with a fusion-buffer memcpy, ``horovod.py:115-120``) and lets
the optimizer update a whole param group in a single launch (reference steps one tensor at a
time, ``_pytorch_context.py:899-907``).

The gradient buffer is "symmetric": allocated through ``ops.comm.SymmetricBuffer`` so peers can
map it over NVLink when world_size > 1.
"""
import dataclasses
import math
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch


@dataclasses.dataclass
class Slot:
    param: torch.nn.Parameter
    offset: int      # element offset in the flat buffer
    numel: int
    group: int       # param-group index


@dataclasses.dataclass
class Segment:
    """Contiguous range of one param group inside the flat buffers."""
    group: int
    offset: int
    numel: int       # padded


def _dense_view(flat: torch.Tensor, offset: int, like: torch.Tensor) -> torch.Tensor:
    """A view of ``flat[offset: offset+numel]`` with the same sizes/strides as ``like`` (which must
    be dense & non-overlapping, e.g. contiguous or channels_last)."""
    return flat.as_strided(like.size(), like.stride(), storage_offset=flat.storage_offset() + offset)


def _is_dense(t: torch.Tensor) -> bool:
    if t.is_contiguous():
        return True
    if t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last):
        return True
    if t.dim() == 5 and t.is_contiguous(memory_format=torch.channels_last_3d):
        return True
    return False


class FlatLayout:
    """Lay the parameters of ``param_groups`` out in flat buffers.

    ``align`` is the element alignment of every parameter (4 floats = 16 bytes for vector access);
    ``segment_align`` pads each group's segment (4*world so every bucket splits evenly into
    per-rank float4 slices)."""

    def __init__(self, param_groups: Sequence[Dict[str, Any]], device: torch.device, align: int = 4,
                 segment_align: int = 4, grad_alloc: Optional[Any] = None, param_alloc: Optional[Any] = None,
                 bind_grads: bool = True) -> None:
        self.device = device
        self.slots: List[Slot] = []
        self.segments: List[Segment] = []
        off = 0
        for gi, g in enumerate(param_groups):
            seg_start = off
            for p in g["params"]:
                if not p.requires_grad:
                    continue
                if p.dtype != torch.float32:
                    raise TypeError(f"fused engine keeps fp32 master parameters; got {p.dtype}")
                if not _is_dense(p.data):
                    p.data = p.data.contiguous()
                n = p.numel()
                self.slots.append(Slot(p, off, n, gi))
                off += (n + align - 1) // align * align
            off = seg_start + (off - seg_start + segment_align - 1) // segment_align * segment_align
            self.segments.append(Segment(gi, seg_start, off - seg_start))
        self.total = max(off, segment_align)
        self.generation = 0          # bumped whenever parameters are re-homed (rebind): mirrors of the buffer are then stale
        alloc = lambda n: torch.zeros(n, dtype=torch.float32, device=device)  # noqa: E731
        self.flat_param = (param_alloc or alloc)(self.total)
        # bind_grads=False: no full-size gradient buffer at all (ZeRO-2/3: gradients live in a small landing ring and in
        # the owner's shard); .grad stays None between backward passes
        self.flat_grad = (grad_alloc or alloc)(self.total) if bind_grads else None
        self._by_param = {id(s.param): s for s in self.slots}
        with torch.no_grad():
            for s in self.slots:
                view = _dense_view(self.flat_param, s.offset, s.param.data)
                view.copy_(s.param.data)
                s.param.data = view
                if not bind_grads:
                    s.param.grad = None
                    continue
                gview = _dense_view(self.flat_grad, s.offset, s.param.data)
                if s.param.grad is not None:
                    gview.copy_(s.param.grad)
                s.param.grad = gview

    def slot_of(self, p: torch.nn.Parameter) -> Optional[Slot]:
        return self._by_param.get(id(p))

    def state_view(self, flat_state: torch.Tensor, s: Slot) -> torch.Tensor:
        return _dense_view(flat_state, s.offset, s.param.data)

    def rebind(self) -> None:
        """Re-point ``.data``/``.grad`` at the flat buffers (after something replaced them, e.g.
        ``model.load_state_dict(assign=True)`` or ``zero_grad(set_to_none=True)``)."""
        self.generation += 1
        with torch.no_grad():
            for s in self.slots:
                want = _dense_view(self.flat_param, s.offset, s.param.data)
                if s.param.data.data_ptr() != want.data_ptr():
                    want.copy_(s.param.data)
                    s.param.data = want
                if self.flat_grad is None:
                    continue
                gwant = _dense_view(self.flat_grad, s.offset, s.param.data)
                if s.param.grad is None:
                    s.param.grad = gwant
                elif s.param.grad.data_ptr() != gwant.data_ptr():
                    gwant.copy_(s.param.grad)
                    s.param.grad = gwant

    def make_buckets(self, cap_bytes: int, multiple: int, tail_cap_bytes: int = 0) -> List["Bucket"]:
        """Partition [0,total) into buckets of about ``cap_bytes``, cut at parameter boundaries,
        in *reverse* parameter order (gradients become ready roughly last-layer-first).  Bucket
        boundaries are multiples of ``multiple`` elements.

        ``tail_cap_bytes`` > 0 additionally splits off the *last-ready* bucket (the first parameters of the model) at
        about that size: it can only be reduced once backward is over, so it is the part of the exchange nothing can
        hide, and it should be small (torch DDP makes its first-ready bucket 1 MB for the mirror-image reason)."""
        cap = max(cap_bytes // 4, multiple)
        bounds = [self.total]
        cur_hi = self.total
        for s in reversed(self.slots):
            lo = s.offset // multiple * multiple
            if cur_hi - lo >= cap and lo < cur_hi and lo > 0:
                bounds.append(lo)
                cur_hi = lo
        if tail_cap_bytes > 0 and self.slots:
            tail = max(tail_cap_bytes // 4, multiple)
            for s in self.slots:
                lo = s.offset // multiple * multiple
                if lo >= tail:
                    if lo < self.total:
                        bounds.append(lo)
                    break
        bounds.append(0)
        bounds = sorted(set(bounds))
        buckets = []
        for i in range(len(bounds) - 1, 0, -1):
            lo, hi = bounds[i - 1], bounds[i]
            members = [s for s in self.slots if s.offset < hi and s.offset + s.numel > lo]
            # a parameter straddling a boundary counts for the bucket holding its first element
            owned = [s for s in self.slots if lo <= s.offset < hi]
            buckets.append(Bucket(len(buckets), lo, hi - lo, owned, members))
        return buckets


@dataclasses.dataclass
class Bucket:
    index: int
    offset: int
    numel: int
    owned: List[Slot]      # params whose readiness gates this bucket
    members: List[Slot]    # params overlapping the range
    pending: int = 0
    launched: bool = False
