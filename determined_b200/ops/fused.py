"""Thin wrappers over the small fused kernels of ``csrc/elementwise.cu`` (CPU tensors take torch math)."""
from typing import Optional

import torch

from determined_b200.ops import _lib


def colsum(x: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """fp32 column sums of a bf16 ``[M, N]`` matrix (bias gradients)."""
    M, N = x.shape
    if not x.is_cuda or x.dtype != torch.bfloat16 or N % 4 or x.stride(1) != 1 or x.stride(0) % 4 or x.data_ptr() % 8:
        r = x.float().sum(0)
        if out is None:
            return r
        return out.add_(r) if accumulate else out.copy_(r)
    if out is None:
        out = torch.empty(N, dtype=torch.float32, device=x.device)
        accumulate = False
    _lib.check(_lib.load().detb200_colsum_bf16(x.data_ptr(), out.data_ptr(), M, N, x.stride(0), int(accumulate), _lib.stream_ptr()), "colsum_bf16")
    return out


def rotary_qkv_(qkv: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, rot: int, conj: bool = False) -> torch.Tensor:
    """Rotate Q and K of a packed ``[B, S, 3, H, D]`` projection in place.  ``cos``/``sin``: fp32 ``[S, rot/2]``."""
    B, S, three, H, D = qkv.shape
    assert three == 3
    if qkv.is_cuda:
        _lib.check(_lib.load().detb200_rotary_qkv(qkv.data_ptr(), cos.data_ptr(), sin.data_ptr(), B, S, H, D, rot, qkv.stride(0), qkv.stride(1),
                                                  qkv.stride(2), qkv.stride(3), int(conj), _lib.stream_ptr()), "rotary_qkv")
        return qkv
    half = rot // 2
    c = cos[None, :, None, None, :].to(qkv.dtype)
    s = (-sin if conj else sin)[None, :, None, None, :].to(qkv.dtype)
    x = qkv[:, :, :2]
    x1, x2 = x[..., :half].clone(), x[..., half:rot].clone()
    x[..., :half] = x1 * c - x2 * s
    x[..., half:rot] = x2 * c + x1 * s
    return qkv


def add(a: torch.Tensor, b: torch.Tensor, c: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``a + b (+ c)`` in one pass (bf16, contiguous, numel % 8 == 0 on CUDA; otherwise torch)."""
    ok = a.is_cuda and a.dtype == torch.bfloat16 and a.is_contiguous() and b.is_contiguous() and b.dtype == a.dtype and a.numel() % 8 == 0 \
        and (c is None or (c.is_contiguous() and c.dtype == a.dtype)) and not (a.requires_grad or b.requires_grad or (c is not None and c.requires_grad))
    if not ok:
        return a + b if c is None else a + b + c
    y = torch.empty_like(a)
    _lib.check(_lib.load().detb200_add_bf16(a.data_ptr(), b.data_ptr(), _lib.ptr(c), y.data_ptr(), a.numel(), _lib.stream_ptr()), "add_bf16")
    return y
