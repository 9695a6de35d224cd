"""Flash attention over the tcgen05 forward kernel (``csrc/attention.cu``).

``flash_attention(q, k, v, causal=..., scale=..., layout=...)`` takes ``[B, S, H, D]`` (``layout=
"bshd"``, the shape a fused QKV projection produces — any strides with D contiguous, so slices of
a packed QKV tensor are consumed in place) or ``[B, H, S, D]`` (``"bhsd"``).  Grouped-query
attention is expressed by K/V having fewer heads.  The forward (and its log-sum-exp) is native;
the packed-QKV path (`attention_qkvpacked`, used by the GPT models) has a native tcgen05 backward too;
the generic `flash_attention` backward recomputes through PyTorch's fused SDPA (a library call).

Reference: GPT-NeoX / GPT-2 example attention (``examples/deepspeed/gpt_neox``; Megatron fused
softmax + bmm), K13 in SURVEY.md.
"""
import math
from typing import Any, Optional, Tuple

import torch
import torch.nn.functional as F

from determined_b200.ops import _lib, fused


def _strides_bshd(t: torch.Tensor, layout: str) -> Tuple[int, int, int]:
    if layout == "bshd":
        return t.stride(0), t.stride(1), t.stride(2)
    return t.stride(0), t.stride(2), t.stride(1)


def _dims(t: torch.Tensor, layout: str) -> Tuple[int, int, int, int]:
    """(B, S, H, D)"""
    if layout == "bshd":
        return t.shape[0], t.shape[1], t.shape[2], t.shape[3]
    return t.shape[0], t.shape[2], t.shape[1], t.shape[3]


def _ok(t: torch.Tensor) -> bool:
    return t.stride(-1) == 1 and all(s % 8 == 0 for s in t.stride()[:-1]) and t.data_ptr() % 16 == 0


def attention_forward(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool, scale: float, layout: str,
                      need_lse: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    lib = _lib.load()
    q, k, v = (t if _ok(t) else t.contiguous() for t in (q, k, v))
    B, Sq, H, D = _dims(q, layout)
    _, Skv, Hkv, _ = _dims(k, layout)
    o = torch.empty((B, Sq, H, D), dtype=torch.bfloat16, device=q.device)      # bshd storage
    lse = torch.empty((B, H, Sq), dtype=torch.float32, device=q.device) if need_lse else None
    i64x3 = _lib.c_int64 * 3
    _lib.check(lib.detb200_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), _lib.ptr(lse), B, H, Hkv, Sq, Skv, D,
                                    i64x3(*_strides_bshd(q, layout)), i64x3(*_strides_bshd(k, layout)), i64x3(*_strides_bshd(v, layout)),
                                    i64x3(o.stride(0), o.stride(1), o.stride(2)), float(scale), int(causal), _lib.stream_ptr()), "attn_fwd")
    return (o if layout == "bshd" else o.permute(0, 2, 1, 3)), lse


def _sdpa(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool, scale: float, layout: str) -> torch.Tensor:
    if layout == "bshd":
        q, k, v = (t.transpose(1, 2) for t in (q, k, v))
    if k.shape[1] != q.shape[1]:
        rep = q.shape[1] // k.shape[1]
        k, v = k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1)
    if causal and q.shape[2] != k.shape[2]:
        Sq, Sk = q.shape[2], k.shape[2]
        mask = torch.ones(Sq, Sk, dtype=torch.bool, device=q.device).tril(Sk - Sq)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, scale=scale)
    else:
        o = F.scaled_dot_product_attention(q, k, v, is_causal=causal, scale=scale)
    return o.transpose(1, 2) if layout == "bshd" else o


class _FlashAttnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool, scale: float, layout: str) -> torch.Tensor:
        o, _ = attention_forward(q, k, v, causal, scale, layout, need_lse=False)
        ctx.save_for_backward(q, k, v)
        ctx.cfg = (causal, scale, layout)
        return o

    @staticmethod
    def backward(ctx: Any, do: torch.Tensor) -> Tuple[Any, ...]:
        q, k, v = ctx.saved_tensors
        causal, scale, layout = ctx.cfg
        with torch.enable_grad():
            qq, kk, vv = (t.detach().requires_grad_(True) for t in (q, k, v))
            o = _sdpa(qq, kk, vv, causal, scale, layout)
        dq, dk, dv = torch.autograd.grad(o, (qq, kk, vv), do)
        return dq, dk, dv, None, None, None


def flash_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool = False, scale: Optional[float] = None,
                    layout: str = "bhsd") -> torch.Tensor:
    if layout not in ("bhsd", "bshd"):
        raise ValueError(f"layout must be 'bhsd' or 'bshd', got {layout!r}")
    D = q.shape[-1]
    scale = float(scale) if scale is not None else 1.0 / math.sqrt(D)
    if not q.is_cuda:
        return _sdpa(q, k, v, causal, scale, layout)
    if q.dtype != torch.bfloat16 or D not in (64, 128):
        _warn_fallback(D, q.dtype)
        return _sdpa(q, k, v, causal, scale, layout)
    return _FlashAttnFn.apply(q, k, v, causal, scale, layout)


class _QKVPackedFn(torch.autograd.Function):
    """Attention over a packed ``[B, S, 3, H, D]`` projection: rotary is applied to Q/K *in place* (the
    projection output is private to this op), the forward kernel reads the three strided slices, and the
    backward produces one packed ``dqkv`` (no per-slice zero-fill + scatter through autograd)."""

    @staticmethod
    def forward(ctx: Any, qkv: torch.Tensor, cos: Optional[torch.Tensor], sin: Optional[torch.Tensor], rot: int, causal: bool,
                scale: float) -> torch.Tensor:
        qkv = qkv.data              # same storage, detached from the producer's version counter: we own this buffer
        if rot:
            fused.rotary_qkv_(qkv, cos, sin, rot)
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
        if qkv.is_cuda:
            o, lse = attention_forward(q, k, v, causal, scale, "bshd", need_lse=True)
        else:
            o, lse = _sdpa(q, k, v, causal, scale, "bshd").contiguous(), None
        ctx.save_for_backward(qkv, o, lse, cos, sin)
        ctx.cfg = (rot, causal, scale)
        return o

    @staticmethod
    def backward(ctx: Any, do: torch.Tensor) -> Tuple[Any, ...]:
        qkv, o, lse, cos, sin = ctx.saved_tensors
        rot, causal, scale = ctx.cfg
        dqkv = torch.empty_like(qkv)
        if qkv.is_cuda and qkv.shape[-1] in (64, 128) and _native_bwd_available():
            attention_backward_packed(qkv, o, lse, do, dqkv, causal, scale)
        else:
            with torch.enable_grad():
                qq, kk, vv = (qkv[:, :, i].detach().requires_grad_(True) for i in range(3))
                oo = _sdpa(qq, kk, vv, causal, scale, "bshd")
            dq, dk, dv = torch.autograd.grad(oo, (qq, kk, vv), do)
            dqkv[:, :, 0].copy_(dq)
            dqkv[:, :, 1].copy_(dk)
            dqkv[:, :, 2].copy_(dv)
        if rot:
            fused.rotary_qkv_(dqkv, cos, sin, rot, conj=True)
        return dqkv, None, None, None, None, None


def _native_bwd_available() -> bool:
    return hasattr(_lib.load(), "detb200_attn_bwd")


def attention_backward_packed(qkv: torch.Tensor, o: torch.Tensor, lse: torch.Tensor, do: torch.Tensor, dqkv: torch.Tensor, causal: bool,
                              scale: float) -> None:
    lib = _lib.load()
    B, S, _, H, D = qkv.shape
    do = do if (_ok(do) and do.stride(-1) == 1) else do.contiguous()
    i64x3 = _lib.c_int64 * 3
    dq_acc = torch.zeros((B, S, H, D), dtype=torch.float32, device=qkv.device)
    delta = torch.empty((2, B, H, (S + 63) // 64 * 64), dtype=torch.float32, device=qkv.device)   # row-dot(O, dO) and lse * log2(e), padded rows

    def st(t: torch.Tensor) -> Any:
        return i64x3(t.stride(0), t.stride(1), t.stride(2))

    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    dq, dk, dv = dqkv[:, :, 0], dqkv[:, :, 1], dqkv[:, :, 2]
    _lib.check(lib.detb200_attn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(), delta.data_ptr(),
                                    dq_acc.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), B, H, S, D, st(q), st(k), st(v), st(o), st(do),
                                    st(dq), st(dk), st(dv), float(scale), int(causal), _lib.stream_ptr()), "attn_bwd")


def attention_qkvpacked(qkv: torch.Tensor, rope: Optional[Tuple[torch.Tensor, torch.Tensor]] = None, rot: int = 0, causal: bool = True,
                        scale: Optional[float] = None) -> torch.Tensor:
    """``qkv``: ``[B, S, 3, H, D]`` (consumed: rotary is applied in place).  ``rope``: fp32 ``(cos, sin)`` of shape
    ``[S, rot/2]``.  Returns ``[B, S, H, D]``."""
    D = qkv.shape[-1]
    scale = float(scale) if scale is not None else 1.0 / math.sqrt(D)
    cos, sin = (rope if (rope is not None and rot) else (None, None))
    if qkv.is_cuda and (qkv.dtype != torch.bfloat16 or D not in (64, 128)):
        # the tcgen05 kernels are built for head dims 64 and 128 (TMEM column budget); anything else (GPT-NeoX 2.7B: 80)
        # goes through torch's attention with a differentiable rotary, so the model still trains, at library speed
        _warn_fallback(D, qkv.dtype)
        return _qkvpacked_torch(qkv, cos, sin, rot if cos is not None else 0, causal, scale)
    return _QKVPackedFn.apply(qkv, cos, sin, rot if cos is not None else 0, causal, scale)


_warned = set()


def _warn_fallback(D: int, dtype: torch.dtype) -> None:
    if (D, dtype) not in _warned:
        _warned.add((D, dtype))
        import logging

        logging.getLogger("determined.ops").warning(
            "attention: head dim %d / %s is outside the native kernels (bf16, 64 or 128): using torch scaled_dot_product_attention", D, dtype)


def _rope_torch(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, rot: int) -> torch.Tensor:
    """Out-of-place, differentiable form of ``fused.rotary_qkv_`` for one of Q / K: ``x`` is ``[B, S, H, D]``."""
    half = rot // 2
    c = cos[None, :, None, :].to(x.dtype)
    s_ = sin[None, :, None, :].to(x.dtype)
    x1, x2 = x[..., :half], x[..., half:rot]
    return torch.cat([x1 * c - x2 * s_, x2 * c + x1 * s_, x[..., rot:]], dim=-1)


def _qkvpacked_torch(qkv: torch.Tensor, cos: Optional[torch.Tensor], sin: Optional[torch.Tensor], rot: int, causal: bool,
                     scale: float) -> torch.Tensor:
    q, k, v = qkv.unbind(2)
    if rot:
        q, k = _rope_torch(q, cos, sin, rot), _rope_torch(k, cos, sin, rot)
    return _sdpa(q, k, v, causal, scale, "bshd").contiguous()
