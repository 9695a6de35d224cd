"""Fused channels-last BatchNorm2d (native: ``csrc/batchnorm.cu``).

``convert_batchnorm_(model)`` re-classes every eligible ``nn.BatchNorm2d`` in place into
:class:`FusedBatchNorm2d` (same parameters/buffers/state-dict keys, so checkpoints remain
interchangeable with the reference's).  The harness calls it from ``context.wrap_model`` when the
fused engine is on.  Inputs that are not channels-last CUDA tensors of a supported width fall back
to ``F.batch_norm`` for that call.
"""
from typing import Any, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from determined_b200.ops import _lib
from determined_b200.parallel import sink as _sink

_WS = {}


def _workspace(device: torch.device, C: int) -> torch.Tensor:
    # one workspace per (device, stream): kernels of different streams must not share the ticket
    key = (device.index, torch.cuda.current_stream(device).cuda_stream, C)
    ws = _WS.get(key)
    if ws is None:
        lib = _lib.load()
        ws = torch.zeros(int(lib.detb200_bn_workspace_bytes(C)), dtype=torch.uint8, device=device)
        _WS[key] = ws
    return ws


def _eligible(x: torch.Tensor) -> bool:
    if not (x.is_cuda and x.dim() == 4 and x.dtype in (torch.float32, torch.bfloat16)):
        return False
    C = x.shape[1]
    V = 8 if x.dtype == torch.bfloat16 else 4
    if C % V or (C // V) > 256 or 256 % (C // V):
        return False
    return x.is_contiguous(memory_format=torch.channels_last) and x.numel() > 0


class _BNTrainFn(torch.autograd.Function):
    """y = relu?(bn(x) + residual?) with batch statistics; one stats pass + one apply pass forward, one reduce pass +
    one apply pass backward (the ReLU gate is recomputed from the saved output, the residual gradient is the gated dy)."""

    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, residual: Optional[torch.Tensor], weight: Optional[torch.Tensor], bias: Optional[torch.Tensor],
                running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
                nbt: Optional[torch.Tensor], momentum: float, eps: float, relu: bool) -> torch.Tensor:
        lib = _lib.load()
        N, C, H, W = x.shape
        M = N * H * W
        y = torch.empty_like(x, memory_format=torch.channels_last)
        stats = torch.empty(4, C, dtype=torch.float32, device=x.device)   # mean, invstd, scale, shift
        w = weight.float() if weight is not None and weight.dtype != torch.float32 else weight
        b = bias.float() if bias is not None and bias.dtype != torch.float32 else bias
        _lib.check(lib.detb200_bn_fwd_train_fused(
            x.data_ptr(), _lib.ptr(residual), y.data_ptr(), M, C, _lib.ptr(w), _lib.ptr(b), _lib.ptr(running_mean), _lib.ptr(running_var),
            _lib.ptr(nbt), stats[0].data_ptr(), stats[1].data_ptr(), stats[2].data_ptr(), stats[3].data_ptr(),
            float(momentum), float(eps), int(relu), _workspace(x.device, C).data_ptr(), _lib.dtype_code(x.dtype),
            _lib.stream_ptr()), "bn_fwd_train")
        ctx.save_for_backward(x, w, stats, y if relu else None)
        ctx.has_w, ctx.has_b, ctx.has_res = weight is not None, bias is not None, residual is not None
        ctx.wdtype = weight.dtype if weight is not None else None
        ctx.bdtype = bias.dtype if bias is not None else None
        ctx.wparam, ctx.bparam = weight, bias
        ctx.mark_non_differentiable(*(t for t in (running_mean, running_var, nbt) if t is not None))
        return y

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        lib = _lib.load()
        x, w, stats, y = ctx.saved_tensors
        N, C, H, W = x.shape
        M = N * H * W
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        if dy.dtype != x.dtype:
            dy = dy.to(x.dtype)
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = None
        if ctx.has_res:
            # without a ReLU the residual gradient is dy itself; with one it is the gated dy the kernel writes
            dres = torch.empty_like(x, memory_format=torch.channels_last) if y is not None else dy
        grads = torch.empty(5, C, dtype=torch.float32, device=x.device)   # dgamma, dbeta, coef[3]
        _lib.check(lib.detb200_bn_bwd_fused(dy.data_ptr(), x.data_ptr(), _lib.ptr(y), dx.data_ptr(), _lib.ptr(dres) if y is not None else None, M, C,
                                            _lib.ptr(w), stats[0].data_ptr(), stats[1].data_ptr(), grads[0].data_ptr(),
                                            grads[1].data_ptr(), grads[2].data_ptr(), _workspace(x.device, C).data_ptr(),
                                            _lib.dtype_code(x.dtype), _lib.stream_ptr()), "bn_bwd")
        dw = grads[0].to(ctx.wdtype) if ctx.has_w else None
        db = grads[1].to(ctx.bdtype) if ctx.has_b else None
        # inside context.backward the engine adds all parameter gradients into its flat buffer in one launch
        # (parallel/sink.py) instead of one AccumulateGrad add per tensor
        if dw is not None:
            s = _sink.sink_of(ctx.wparam)
            if s is not None and s.accepts(ctx.wparam, dw):
                s.add(ctx.wparam, dw)
                dw = None
        if db is not None:
            s = _sink.sink_of(ctx.bparam)
            if s is not None and s.accepts(ctx.bparam, db):
                s.add(ctx.bparam, db)
                db = None
        return dx, dres, dw, db, None, None, None, None, None, None


def batch_norm_train(x: torch.Tensor, weight: Optional[torch.Tensor], bias: Optional[torch.Tensor],
                     running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
                     num_batches_tracked: Optional[torch.Tensor], momentum: float, eps: float,
                     residual: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
    return _BNTrainFn.apply(x, residual, weight, bias, running_mean, running_var, num_batches_tracked, momentum, eps, relu)


class FusedBatchNorm2d(nn.BatchNorm2d):
    """Drop-in ``nn.BatchNorm2d`` whose training forward/backward run the native kernels."""

    def _native_ok(self, x: torch.Tensor) -> bool:
        return bool(self.training and self.momentum is not None and self.track_running_stats and _eligible(x)
                    and self.running_mean is not None and self.running_mean.dtype == torch.float32)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._native_ok(x):
            return batch_norm_train(x, self.weight, self.bias, self.running_mean, self.running_var,
                                    self.num_batches_tracked, self.momentum, self.eps)
        return super().forward(x)

    def forward_fused(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
        """``relu?(bn(x) + residual?)`` in the BatchNorm kernels' own passes."""
        if self._native_ok(x) and (residual is None or (residual.shape == x.shape and residual.dtype == x.dtype
                                                        and residual.is_contiguous(memory_format=torch.channels_last))):
            return batch_norm_train(x, self.weight, self.bias, self.running_mean, self.running_var, self.num_batches_tracked,
                                    self.momentum, self.eps, residual, relu)
        y = self.forward(x)
        if residual is not None:
            y = y + residual
        return F.relu(y) if relu else y


def _is_basic_block(m: nn.Module) -> bool:
    names = {n for n, _ in m.named_children()}
    if not {"conv1", "bn1", "conv2", "bn2"} <= names or not names <= {"conv1", "bn1", "conv2", "bn2", "shortcut", "downsample", "relu"}:
        return False
    if not (isinstance(m.conv1, nn.Conv2d) and isinstance(m.conv2, nn.Conv2d) and isinstance(m.bn1, FusedBatchNorm2d) and isinstance(m.bn2, FusedBatchNorm2d)):
        return False
    return hasattr(m, "shortcut") or hasattr(m, "downsample")


def _basic_block_forward(self: Any, x: torch.Tensor) -> torch.Tensor:
    out = self.bn1.forward_fused(self.conv1(x), None, True)
    if hasattr(self, "shortcut"):
        sc = self.shortcut(x)
    else:
        sc = x if self.downsample is None else self.downsample(x)
    return self.bn2.forward_fused(self.conv2(out), sc, True)


def fuse_residual_blocks_(model: nn.Module) -> int:
    """Give ResNet-style basic blocks (conv-bn-relu-conv-bn-(+shortcut)-relu; our CIFAR ResNet and torchvision's
    ``BasicBlock``) a forward that folds ReLU and the residual add into the BatchNorm kernels.  A block is only patched
    after its original forward and the fused composition agree on a random input (so look-alike blocks that compute
    something else are left alone).  Returns the number of patched blocks."""
    import types

    n = 0
    for m in model.modules():
        if not _is_basic_block(m) or getattr(m, "_detb200_fused_block", False):
            continue
        was_training = m.training
        try:
            m.eval()
            w = m.conv1.weight
            with torch.no_grad():
                x = torch.randn(2, m.conv1.in_channels, 8, 8, device=w.device, dtype=w.dtype)
                ref = m(x)
                got = _basic_block_forward(m, x)
            same = ref.shape == got.shape and torch.allclose(ref, got, rtol=1e-4, atol=1e-5)
        except Exception:
            same = False
        finally:
            m.train(was_training)
        if same:
            m.forward = types.MethodType(_basic_block_forward, m)
            m._detb200_fused_block = True
            n += 1
    return n


def convert_batchnorm_(model: nn.Module) -> int:
    """Re-class eligible BatchNorm2d modules in place; returns how many were converted."""
    n = 0
    for m in model.modules():
        if type(m) is nn.BatchNorm2d and m.track_running_stats and m.momentum is not None:
            m.__class__ = FusedBatchNorm2d
            n += 1
    return n
