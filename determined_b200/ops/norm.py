"""LayerNorm / RMSNorm as autograd functions over the native kernels (``csrc/norm.cu``).

On CUDA the native kernel is the only path (no silent fallback); on CPU the functions compute the
same math with torch ops so models remain testable without a GPU.
"""
from typing import Any, Optional, Tuple

import torch

from determined_b200.ops import _lib


class _NormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, gamma: torch.Tensor, beta: Optional[torch.Tensor], eps: float, rms: bool) -> torch.Tensor:
        lib = _lib.load()
        cols = x.shape[-1]
        x2 = x.reshape(-1, cols)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        gamma_c = gamma.to(x.dtype).contiguous()
        beta_c = beta.to(x.dtype).contiguous() if beta is not None else None
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        _lib.check(lib.detb200_norm_fwd(x2.data_ptr(), gamma_c.data_ptr(), _lib.ptr(beta_c), y.data_ptr(),
                                        mean.data_ptr(), rstd.data_ptr(), rows, cols, float(eps),
                                        _lib.dtype_code(x.dtype), int(rms), _lib.stream_ptr()), "norm_fwd")
        ctx.save_for_backward(x2, gamma_c, mean, rstd)
        ctx.rms, ctx.has_beta, ctx.shape = rms, beta is not None, x.shape
        ctx.gamma_dtype = gamma.dtype
        ctx.beta_dtype = beta.dtype if beta is not None else None
        return y.reshape(x.shape)

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        lib = _lib.load()
        x2, gamma_c, mean, rstd = ctx.saved_tensors
        rows, cols = x2.shape
        dy2 = dy.reshape(rows, cols)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = torch.empty_like(x2)
        dgamma = torch.empty_like(gamma_c)
        dbeta = torch.empty_like(gamma_c) if ctx.has_beta else None
        ws = torch.empty(int(lib.detb200_norm_bwd_workspace_elems(rows, cols, int(ctx.rms))),
                         dtype=torch.float32, device=x2.device)
        _lib.check(lib.detb200_norm_bwd(dy2.data_ptr(), x2.data_ptr(), gamma_c.data_ptr(), mean.data_ptr(),
                                        rstd.data_ptr(), dx.data_ptr(), dgamma.data_ptr(), _lib.ptr(dbeta),
                                        ws.data_ptr(), rows, cols, _lib.dtype_code(x2.dtype), int(ctx.rms),
                                        _lib.stream_ptr()), "norm_bwd")
        return (dx.reshape(ctx.shape), dgamma.to(ctx.gamma_dtype),
                dbeta.to(ctx.beta_dtype) if dbeta is not None else None, None, None)


def _supported(x: torch.Tensor) -> bool:
    c = x.shape[-1]
    return x.is_cuda and x.dtype in (torch.float32, torch.bfloat16) and c % 2 == 0 and 2 <= c <= 8192


def layer_norm(x: torch.Tensor, gamma: torch.Tensor, beta: Optional[torch.Tensor], eps: float = 1e-5) -> torch.Tensor:
    if x.is_cuda:
        if not _supported(x):
            raise _lib.NativeError(f"layer_norm: unsupported input {tuple(x.shape)} {x.dtype}")
        return _NormFn.apply(x, gamma, beta, eps, False)
    return torch.nn.functional.layer_norm(x, (x.shape[-1],), gamma, beta, eps)


def rms_norm(x: torch.Tensor, gamma: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    if x.is_cuda:
        if not _supported(x):
            raise _lib.NativeError(f"rms_norm: unsupported input {tuple(x.shape)} {x.dtype}")
        return _NormFn.apply(x, gamma, None, eps, True)
    xf = x.float()
    return (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * gamma


class LayerNorm(torch.nn.Module):
    def __init__(self, hidden: int, eps: float = 1e-5, bias: bool = True) -> None:
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(hidden))
        self.bias = torch.nn.Parameter(torch.zeros(hidden)) if bias else None
        self.eps = eps

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return layer_norm(x, self.weight, self.bias, self.eps)


class RMSNorm(torch.nn.Module):
    def __init__(self, hidden: int, eps: float = 1e-6) -> None:
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(hidden))
        self.eps = eps

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return rms_norm(x, self.weight, self.eps)
