"""Deferred gradient accumulation and bf16 weight shadows for the fused engine.

What the reference's ``loss.backward()`` costs per parameter on a GPU (``_pytorch_context.py:786-803`` under autocast): a cast
of the fp32 weight to bf16 in forward, a cast of the bf16 weight gradient back to fp32 in backward, and one AccumulateGrad
``add`` kernel.  On the ResNet-18 trial those are ~250 launches of a few microseconds each per optimizer step, 13 % of the
step's device time (``profiles/r1_launches_resnet_step_v3_fused_blocks.txt``).  With the flat buffers of the fused engine
they collapse into two launches:

* ``GradSink``: native modules hand their parameter gradients (bf16 or fp32, any number) to the sink of the optimizer that
  owns the parameter instead of returning them to autograd; when backward ends ONE multi-tensor kernel adds all of them
  into the flat fp32 gradient buffer (``detb200_accum_multi``).  Only between ``backward_begin`` and ``backward_end`` of the
  engine (i.e. inside ``context.backward``); a bare ``loss.backward()`` gets ordinary autograd gradients.
* ``LowpShadow``: a flat bf16 copy of the flat fp32 parameters; modules read their weight from it.  The engine marks it
  invalid whenever parameters may have changed (every optimizer step, parameter broadcast, re-homing, and before a CUDA
  graph is captured); the first module that needs it afterwards refreshes the whole mirror with one cast kernel.  Writes
  through torch ops outside those points bump the parameter's version counter, which is checked per use as well.
"""
import ctypes
from typing import Any, Dict, List, Optional, Tuple

import torch

from determined_b200.ops import _lib


def _same_layout(a: torch.Tensor, b: torch.Tensor) -> bool:
    """Same element order in memory (strides of size-1 dimensions are arbitrary and do not count)."""
    return all(sa == sb for n, sa, sb in zip(a.shape, a.stride(), b.stride()) if n > 1)


class GradSink:
    """One per optimizer record of the fused engine."""

    def __init__(self, layout: Any) -> None:
        self.layout = layout
        self.active = False
        self._pending: List[Tuple[torch.Tensor, torch.Tensor]] = []      # (destination view in the flat buffer, gradient)
        self._seen: List[torch.nn.Parameter] = []

    def accepts(self, p: torch.nn.Parameter, grad: torch.Tensor) -> bool:
        """Gradients the kernel can add in place: dense, same memory layout as the parameter, bf16 or fp32."""
        if not self.active or grad.dtype not in (torch.bfloat16, torch.float32) or p.grad is None:
            return False
        return grad.shape == p.shape and grad.is_cuda and _same_layout(grad, p.grad)

    def add(self, p: torch.nn.Parameter, grad: torch.Tensor) -> None:
        self._pending.append((p.grad, grad))
        self._seen.append(p)

    def flush(self) -> List[torch.nn.Parameter]:
        """Add everything handed in since the last flush into the flat gradient buffer; returns the parameters touched."""
        if not self._pending:
            return []
        lib = _lib.load()
        n = len(self._pending)
        srcs = (ctypes.c_void_p * n)(*[g.data_ptr() for _, g in self._pending])
        dsts = (ctypes.c_void_p * n)(*[d.data_ptr() for d, _ in self._pending])
        nums = (ctypes.c_int64 * n)(*[g.numel() for _, g in self._pending])
        kinds = bytes(1 if g.dtype == torch.bfloat16 else 0 for _, g in self._pending)
        _lib.check(lib.detb200_accum_multi(srcs, dsts, nums, kinds, n, _lib.stream_ptr()), "accum_multi")
        _lib.count_launches((n + 47) // 48 - 1)
        seen, self._seen, self._pending = self._seen, [], []
        return seen


def sink_of(p: Optional[torch.nn.Parameter]) -> Optional[GradSink]:
    return getattr(p, "_detb200_sink", None) if p is not None else None


class LowpShadow:
    """bf16 mirror of one flat fp32 parameter buffer."""

    def __init__(self, layout: Any) -> None:
        self.layout = layout
        n = (layout.total + 7) // 8 * 8
        self.flat = torch.zeros(n, dtype=torch.bfloat16, device=layout.device)
        self._versions: Dict[int, int] = {}
        self._generation = -1
        self.valid = False
        self.refresh_all()

    def view_for(self, p: torch.nn.Parameter) -> Optional[torch.Tensor]:
        s = self.layout.slot_of(p)
        if s is None:
            return None
        return self.flat.as_strided(p.size(), p.stride(), storage_offset=s.offset)

    def refresh_all(self) -> None:
        """One cast kernel over the whole buffer (after the fused optimizer wrote the parameters)."""
        lib = _lib.load()
        lay = self.layout
        n = lay.total // 8 * 8
        if n:
            _lib.check(lib.detb200_cast_bf16(lay.flat_param.data_ptr(), self.flat.data_ptr(), n, _lib.stream_ptr()), "cast_bf16")
        if lay.total > n:
            self.flat[n:lay.total].copy_(lay.flat_param[n:lay.total])
        for s in lay.slots:
            self._versions[id(s.param)] = s.param._version
        self._generation = lay.generation
        self.valid = True

    def invalidate(self) -> None:
        self.valid = False

    def stale(self) -> bool:
        """Cheap host-side check (no GPU work): did anything but the fused optimizer write a parameter since the last
        refresh?  Used before replaying a captured step, whose own checks ran at capture time only."""
        if not self.valid or self._generation != self.layout.generation:
            return True
        return any(self._versions.get(id(s.param)) != s.param._version for s in self.layout.slots)

    def fresh(self, p: torch.nn.Parameter, view: torch.Tensor) -> torch.Tensor:
        """``view`` after making sure it still mirrors ``p`` (an in-place write by torch bumps ``p._version``)."""
        if not self.valid or self._generation != self.layout.generation:
            self.refresh_all()
        elif self._versions.get(id(p)) != p._version:
            with torch.no_grad():
                view.copy_(p)
            self._versions[id(p)] = p._version
        return view
