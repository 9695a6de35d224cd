"""Fused softmax cross-entropy (native: ``csrc/xent.cu``): mean over non-ignored targets.

The backward writes ``dlogits`` in place over the saved logits when ``inplace_backward`` is set
(the logits tensor must not be used afterwards), so the [tokens, vocab] activation is stored once.
"""
from typing import Any, Tuple

import torch

from determined_b200.ops import _lib


class _XentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, logits: torch.Tensor, target: torch.Tensor, ignore_index: int, inplace_backward: bool) -> torch.Tensor:
        lib = _lib.load()
        vocab = logits.shape[-1]
        l2 = logits.reshape(-1, vocab)
        if not l2.is_contiguous():
            l2 = l2.contiguous()
        rows = l2.shape[0]
        tgt = target.reshape(-1).to(torch.int64).contiguous()
        loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
        lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
        _lib.check(lib.detb200_xent_fwd(l2.data_ptr(), tgt.data_ptr(), loss.data_ptr(), lse.data_ptr(), rows, vocab,
                                        int(ignore_index), _lib.dtype_code(l2.dtype), _lib.stream_ptr()), "xent_fwd")
        nvalid = (tgt != ignore_index).sum().clamp_min(1).to(torch.float32)
        ctx.save_for_backward(l2, tgt, lse, nvalid)
        ctx.ignore_index, ctx.inplace, ctx.shape = ignore_index, inplace_backward, logits.shape
        return loss.sum() / nvalid

    @staticmethod
    def backward(ctx: Any, dloss: torch.Tensor) -> Tuple[Any, ...]:
        lib = _lib.load()
        l2, tgt, lse, nvalid = ctx.saved_tensors
        rows, vocab = l2.shape
        scale = (dloss.to(torch.float32) / nvalid).reshape(1).contiguous()
        out = l2 if ctx.inplace else torch.empty_like(l2)
        _lib.check(lib.detb200_xent_bwd(l2.data_ptr(), tgt.data_ptr(), lse.data_ptr(), scale.data_ptr(), out.data_ptr(),
                                        rows, vocab, int(ctx.ignore_index), _lib.dtype_code(l2.dtype),
                                        _lib.stream_ptr()), "xent_bwd")
        return out.reshape(ctx.shape), None, None, None


def cross_entropy(logits: torch.Tensor, target: torch.Tensor, ignore_index: int = -100,
                  inplace_backward: bool = False) -> torch.Tensor:
    if logits.is_cuda:
        if logits.dtype not in (torch.float32, torch.bfloat16):
            raise _lib.NativeError(f"cross_entropy: unsupported dtype {logits.dtype}")
        return _XentFn.apply(logits, target, ignore_index, inplace_backward)
    return torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), target.reshape(-1),
                                             ignore_index=ignore_index)
