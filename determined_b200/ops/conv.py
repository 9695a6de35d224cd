"""Convolution module of the fused engine's module pass (``context.wrap_model``).

The convolution itself stays on cuDNN (not on a communication-bound path).  What this module removes is the
per-layer bookkeeping around it under bf16 autocast, all of which exists only because the parameter is fp32: the weight
cast in forward (the weight comes from the engine's bf16 shadow of the flat parameter buffer instead), the cast of the bf16
weight gradient to fp32 and the AccumulateGrad add in backward (the bf16 gradient goes to the optimizer's gradient sink,
which adds every parameter's gradient into the flat fp32 buffer in one launch).  Outside these conditions (no engine, no
autocast, bias, a bare ``loss.backward()``) it behaves exactly like ``nn.Conv2d``.

Reference call sites replaced: ``autocast_forward_pass`` ``_pytorch_context.py:236-283`` + ``loss.backward()`` ``:786-803``.
"""
from typing import Any, Optional, Tuple

import torch
import torch.nn as nn

from determined_b200.parallel import sink as _sink


def _bf16_autocast_on() -> bool:
    try:
        return bool(torch.is_autocast_enabled("cuda")) and torch.get_autocast_dtype("cuda") == torch.bfloat16
    except TypeError:          # older signatures
        return bool(torch.is_autocast_enabled()) and torch.get_autocast_gpu_dtype() == torch.bfloat16


class _ConvLowpFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, w_master: torch.Tensor, w_lowp: torch.Tensor, mod: "FusedConv2d") -> torch.Tensor:
        xb = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
        y = torch.ops.aten.convolution(xb, w_lowp, None, mod.stride, mod.padding, mod.dilation, False, (0, 0), mod.groups)
        ctx.save_for_backward(xb, w_lowp)
        ctx.mod = mod
        ctx.x_dtype = x.dtype
        return y

    @staticmethod
    def backward(ctx: Any, gy: torch.Tensor) -> Tuple[Any, ...]:
        xb, w = ctx.saved_tensors
        mod = ctx.mod
        if gy.dtype != torch.bfloat16:
            gy = gy.to(torch.bfloat16)
        gx, gw, _ = torch.ops.aten.convolution_backward(gy, xb, w, None, mod.stride, mod.padding, mod.dilation, False, (0, 0),
                                                        mod.groups, (ctx.needs_input_grad[0], True, False))
        p = mod.weight
        s = _sink.sink_of(p)
        if s is not None and s.accepts(p, gw):
            s.add(p, gw)            # bf16, added into the flat fp32 gradient buffer with everybody else's when backward ends
            gw_out = None
        else:
            gw_out = gw.to(p.dtype)
        if gx is not None and gx.dtype != ctx.x_dtype:
            gx = gx.to(ctx.x_dtype)
        return gx, gw_out, None, None


class FusedConv2d(nn.Conv2d):
    """``nn.Conv2d`` re-classed in place by :func:`convert_conv_` (same parameters, same state-dict keys)."""

    _detb200_shadow: Optional[Any] = None          # parallel.sink.LowpShadow of the optimizer that owns the weight
    _detb200_lowp_weight: Optional[torch.Tensor] = None

    def _lowp_path(self, x: torch.Tensor) -> bool:
        return (self._detb200_lowp_weight is not None and self.bias is None and self.padding_mode == "zeros"
                and not isinstance(self.padding, str) and x.is_cuda and x.dim() == 4
                and self.weight.dtype == torch.float32 and _bf16_autocast_on())

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._lowp_path(x):
            w = self._detb200_shadow.fresh(self.weight, self._detb200_lowp_weight)
            if torch.is_grad_enabled() and (self.weight.requires_grad or x.requires_grad):
                return _ConvLowpFn.apply(x, self.weight, w, self)
            xb = x if x.dtype == torch.bfloat16 else x.to(torch.bfloat16)
            return torch.ops.aten.convolution(xb, w, None, self.stride, self.padding, self.dilation, False, (0, 0), self.groups)
        return super().forward(x)


def convert_conv_(model: nn.Module) -> int:
    """Re-class every plain ``nn.Conv2d`` of ``model`` (exact type: subclasses may override forward)."""
    n = 0
    for m in model.modules():
        if type(m) is nn.Conv2d:
            m.__class__ = FusedConv2d
            n += 1
    return n


def attach_shadow_(models: Any, shadow: Any) -> int:
    """Give every converted convolution whose weight lives in ``shadow``'s layout its bf16 view."""
    n = 0
    for model in models:
        for m in model.modules():
            if isinstance(m, FusedConv2d) and m._detb200_lowp_weight is None:
                v = shadow.view_for(m.weight)
                if v is not None:
                    m._detb200_shadow = shadow
                    m._detb200_lowp_weight = v
                    n += 1
    return n
