"""bf16 GEMM / Linear / MLP over the tcgen05 kernel (``csrc/gemm_sm100.cu``).

``gemm`` is the raw entry point; ``linear`` and ``mlp`` are autograd functions whose backward GEMMs
read the saved operands in place (MN-major descriptors, no transposes) and which fuse bias,
GELU/ReLU and their derivatives into the GEMM epilogues.  On CUDA the native kernel is the only
path; CPU tensors take the equivalent torch math so models stay testable without a GPU.

Reference: the model's ``nn.Linear`` layers run through ATen/cuBLAS under the reference's DDP /
DeepSpeed wrappers (``harness/determined/pytorch/_pytorch_context.py:296-299``); it has no GEMM
code of its own.
"""
from typing import Any, Optional, Tuple

import torch
import torch.nn.functional as F

from determined_b200.ops import _lib, fused

EPILOGUES = {"none": 0, "gelu": 1, "gelu_erf": 1, "gelu_tanh": 2, "relu": 3, "dgelu": 4, "dgelu_erf": 4, "dgelu_tanh": 5, "drelu": 6}
_BWD_OF = {1: 4, 2: 5, 3: 6}


def _check_operand(t: torch.Tensor, name: str) -> None:
    if t.dtype != torch.bfloat16:
        raise TypeError(f"gemm: {name} must be bfloat16, got {t.dtype}")
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"gemm: {name} must be 2-D with a contiguous last dimension, got shape {tuple(t.shape)} strides {t.stride()}")
    if t.stride(0) % 8 != 0 or t.data_ptr() % 16 != 0:
        raise ValueError(f"gemm: {name} needs a 16-byte aligned base and a row stride that is a multiple of 8 elements "
                         f"(stride {t.stride(0)}, ptr {t.data_ptr():#x})")


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False, bias: Optional[torch.Tensor] = None,
         out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16, epilogue: str = "none",
         aux: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """``out[M,N] (+)= op(a) @ op(b)^T``.

    ``a`` is ``[M,K]`` (or ``[K,M]`` when ``a_mn_major``); ``b`` is ``[N,K]`` (or ``[K,N]`` when
    ``b_mn_major``).  ``epilogue`` in :data:`EPILOGUES`; forward activations write the
    pre-activation to ``aux`` when given, derivative epilogues read it from ``aux``.
    """
    _check_operand(a, "a")
    _check_operand(b, "b")
    M, K = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn_major else (b.shape[0], b.shape[1])
    if K != Kb:
        raise ValueError(f"gemm: reduction sizes differ ({K} vs {Kb})")
    epi = EPILOGUES[epilogue]
    if out is None:
        if accumulate:
            raise ValueError("gemm: accumulate=True needs `out`")
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    if out.shape != (M, N) or out.stride(1) != 1 or out.dtype not in (torch.bfloat16, torch.float32):
        raise ValueError(f"gemm: bad output {tuple(out.shape)} {out.dtype} strides {out.stride()} for M={M} N={N}")
    if aux is not None and (aux.shape != (M, N) or aux.dtype != torch.bfloat16 or aux.stride(1) != 1):
        raise ValueError("gemm: aux must be a bf16 [M,N] tensor with a contiguous last dimension")
    if bias is not None and (bias.numel() != N or not bias.is_contiguous() or bias.dtype not in (torch.bfloat16, torch.float32)):
        raise ValueError("gemm: bias must be a contiguous bf16/fp32 vector of length N")
    if not a.is_cuda:
        return _gemm_cpu(a, b, a_mn_major, b_mn_major, bias, out, epi, aux, accumulate)
    lib = _lib.load()
    _lib.check(lib.detb200_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), _lib.ptr(bias), _lib.ptr(aux), M, N, K,
                                     a.stride(0), b.stride(0), out.stride(0), aux.stride(0) if aux is not None else 0,
                                     int(a_mn_major), int(b_mn_major), int(out.dtype == torch.float32),
                                     int(bias is not None and bias.dtype == torch.float32), epi, int(accumulate), _lib.stream_ptr()),
               "gemm_bf16")
    return out


def _act(x: torch.Tensor, epi: int) -> torch.Tensor:
    return F.gelu(x) if epi == 1 else F.gelu(x, approximate="tanh") if epi == 2 else F.relu(x)


def _dact(pre: torch.Tensor, epi: int) -> torch.Tensor:
    """Closed-form derivative of the activation selected by a *derivative* epilogue code (4/5/6)."""
    x = pre.detach().float()
    if epi == 4:
        return 0.5 * (1.0 + torch.erf(x * 0.7071067811865476)) + x * torch.exp(-0.5 * x * x) * 0.3989422804014327
    if epi == 5:
        u = 0.7978845608028654 * (x + 0.044715 * x ** 3)
        t = torch.tanh(u)
        return 0.5 * (1.0 + t) + 0.5 * x * (1.0 - t * t) * 0.7978845608028654 * (1.0 + 3 * 0.044715 * x * x)
    return (x > 0).float()


def _gemm_cpu(a: torch.Tensor, b: torch.Tensor, a_mn: bool, b_mn: bool, bias: Optional[torch.Tensor], out: torch.Tensor, epi: int,
              aux: Optional[torch.Tensor], accumulate: bool) -> torch.Tensor:
    A = (a.t() if a_mn else a).float()
    B = (b if b_mn else b.t()).float()
    r = A @ B
    if bias is not None:
        r = r + bias.float()
    if 1 <= epi <= 3:
        if aux is not None:
            aux.copy_(r)
        r = _act(r, epi)
    elif epi >= 4:
        assert aux is not None
        r = r * _dact(aux, epi)
    if accumulate:
        r = r + out.float()
    out.copy_(r)
    return out


def _as_bf16_2d(x: torch.Tensor) -> torch.Tensor:
    x2 = x.reshape(-1, x.shape[-1])
    if x2.dtype != torch.bfloat16:
        x2 = x2.to(torch.bfloat16)
    if x2.stride(1) != 1 or x2.stride(0) % 8 != 0 or x2.data_ptr() % 16 != 0:
        x2 = x2.contiguous()
    return x2


class _LinearFn(torch.autograd.Function):
    """y = act(x W^T + b).  Saves x, W (bf16) and, for an activation, the pre-activation."""

    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], epi: int) -> torch.Tensor:
        x2 = _as_bf16_2d(x)
        w = _as_bf16_2d(weight)
        pre = torch.empty((x2.shape[0], w.shape[0]), dtype=torch.bfloat16, device=x.device) if epi and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]) else None
        y = gemm(x2, w, bias=bias.contiguous() if bias is not None else None, epilogue={0: "none", 1: "gelu", 2: "gelu_tanh", 3: "relu"}[epi], aux=pre)
        ctx.save_for_backward(x2, w, pre)
        ctx.epi, ctx.x_shape, ctx.w_dtype = epi, x.shape, weight.dtype
        ctx.b_dtype = bias.dtype if bias is not None else None
        return y.reshape(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        x2, w, pre = ctx.saved_tensors
        dy2 = _as_bf16_2d(dy)
        if ctx.epi:
            # dpre = dy * act'(pre): fused into a tiny "GEMM with identity" is not worth it; one elementwise pass
            dy2 = (dy2.float() * _dact(pre, ctx.epi + 3)).to(torch.bfloat16)
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w, b_mn_major=True).reshape(ctx.x_shape)                     # [T,N] x [N,K](k-rows) -> [T,K]
        if ctx.needs_input_grad[1]:
            dw = gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32 if ctx.w_dtype == torch.float32 else torch.bfloat16)
            if dw.dtype != ctx.w_dtype:
                dw = dw.to(ctx.w_dtype)
        if ctx.b_dtype is not None and ctx.needs_input_grad[2]:
            db = fused.colsum(dy2).to(ctx.b_dtype)
        return dx, dw, db, None


def linear(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None, activation: str = "none") -> torch.Tensor:
    return _LinearFn.apply(x, weight, bias, EPILOGUES[activation])


class _MLPFn(torch.autograd.Function):
    """y = act(x W1^T + b1) W2^T + b2 with every elementwise step living in a GEMM epilogue:
    forward act + saved pre-activation in GEMM1; backward act' in the dH GEMM."""

    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, w1: torch.Tensor, b1: Optional[torch.Tensor], w2: torch.Tensor, b2: Optional[torch.Tensor], epi: int) -> torch.Tensor:
        x2, w1b, w2b = _as_bf16_2d(x), _as_bf16_2d(w1), _as_bf16_2d(w2)
        T = x2.shape[0]
        pre = torch.empty((T, w1b.shape[0]), dtype=torch.bfloat16, device=x.device)
        h = gemm(x2, w1b, bias=b1, epilogue={1: "gelu", 2: "gelu_tanh", 3: "relu"}[epi], aux=pre)
        y = gemm(h, w2b, bias=b2)
        ctx.save_for_backward(x2, w1b, w2b, pre, h)
        ctx.epi, ctx.x_shape = epi, x.shape
        ctx.dtypes = (w1.dtype, b1.dtype if b1 is not None else None, w2.dtype, b2.dtype if b2 is not None else None)
        return y.reshape(*x.shape[:-1], w2b.shape[0])

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        x2, w1b, w2b, pre, h = ctx.saved_tensors
        d1, db1t, d2, db2t = ctx.dtypes
        dy2 = _as_bf16_2d(dy)
        f32 = torch.float32

        def wgrad(a: torch.Tensor, b: torch.Tensor, dt: torch.dtype) -> torch.Tensor:
            g = gemm(a, b, a_mn_major=True, b_mn_major=True, out_dtype=f32 if dt == f32 else torch.bfloat16)
            return g if g.dtype == dt else g.to(dt)

        dw2 = wgrad(dy2, h, d2) if ctx.needs_input_grad[3] else None
        db2 = fused.colsum(dy2).to(db2t) if db2t is not None and ctx.needs_input_grad[4] else None
        dpre = gemm(dy2, w2b, b_mn_major=True, epilogue={1: "dgelu", 2: "dgelu_tanh", 3: "drelu"}[ctx.epi], aux=pre)
        dw1 = wgrad(dpre, x2, d1) if ctx.needs_input_grad[1] else None
        db1 = fused.colsum(dpre).to(db1t) if db1t is not None and ctx.needs_input_grad[2] else None
        dx = gemm(dpre, w1b, b_mn_major=True).reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        return dx, dw1, db1, dw2, db2, None


def mlp(x: torch.Tensor, w1: torch.Tensor, b1: Optional[torch.Tensor], w2: torch.Tensor, b2: Optional[torch.Tensor], activation: str = "gelu") -> torch.Tensor:
    return _MLPFn.apply(x, w1, b1, w2, b2, EPILOGUES[activation])


class FusedLinear(torch.nn.Linear):
    """``nn.Linear`` whose forward/backward GEMMs run on the tcgen05 kernel (bf16 compute, fp32
    accumulation; fp32 master weights are cast once per call).  State-dict compatible."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, activation: str = "none", **kw: Any) -> None:
        super().__init__(in_features, out_features, bias=bias, **kw)
        self.activation = activation

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return linear(x, self.weight, self.bias, self.activation)


class FusedMLP(torch.nn.Module):
    def __init__(self, hidden: int, inner: int, activation: str = "gelu", bias: bool = True, **kw: Any) -> None:
        super().__init__()
        self.fc1 = torch.nn.Linear(hidden, inner, bias=bias, **kw)
        self.fc2 = torch.nn.Linear(inner, hidden, bias=bias, **kw)
        self.activation = activation

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return mlp(x, self.fc1.weight, self.fc1.bias, self.fc2.weight, self.fc2.bias, self.activation)
