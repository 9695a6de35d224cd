"""GEMMs fused with their collective over NVLink peer memory (``csrc/gemm_sm100.cu``).

``TensorParallelGemm`` owns the symmetric buffers and counters of one tensor-parallel group:

``all_gather_gemm``   sequence-sharded activations ``x_r [T/W, K]`` -> ``act(x_full @ W_r^T + b)``.
                      One kernel: a few CTAs push this rank's shard into every peer's gather buffer
                      while the rest run the tcgen05 GEMM, own shard first, consuming remote shards
                      as their arrival counters tick.
``gemm_reduce_scatter``  ``x [T, K/W] @ W_r[N, K/W]^T`` summed over ranks and scattered by rows:
                      the epilogue TMA-stores bf16 partial tiles straight into the owner's receive
                      slots; each rank then sums its shard in fp32 inside the same kernel.

Reference: tensor/sequence-parallel layers of the GPT-NeoX example run through DeepSpeed/Megatron
``mpu`` (``examples/deepspeed/gpt_neox``, wired by ``harness/determined/pytorch/deepspeed/_mpu.py``)
as separate NCCL all-gather / reduce-scatter calls around cuBLAS GEMMs.
"""
import ctypes
from ctypes import c_void_p
from typing import Any, Optional, Tuple

import torch

from determined_b200.ops import _lib, comm, fused
from determined_b200.ops import gemm as G

PUSH_CTAS = 16         # pusher CTAs of the gather-fused kernels (measured best at world 2 and 8; each drives ~40 GB/s of NVLink)


class TensorParallelGemm:
    def __init__(self, group: comm.PeerGroup) -> None:
        self.group = group
        self.rank, self.world = group.rank, group.world
        lib = _lib.load()
        self.ctrl = comm.SymmetricBuffer(int(lib.detb200_gemm_ctrl_bytes()), group.device)
        self.ctrl_ptrs = group.map(self.ctrl)
        self.epoch = 0                 # number of fused calls issued so far (identical on every rank)
        self._ready = 0                # cumulative a_ready arrivals per shard
        self._rs = 0                   # cumulative rs_count arrivals per source
        self._gather: dict = {}
        self._recv: dict = {}

    # -- buffers -----------------------------------------------------------------------------------
    def gather_buffer(self, T: int, K: int) -> Tuple[torch.Tensor, torch.Tensor, Any]:
        """Symmetric ``[T, K]`` bf16 buffer; returns (full, this rank's row shard view, peer ptrs).
        Write the local activations into the shard view, then call :meth:`all_gather_gemm`."""
        key = (T, K)
        if key not in self._gather:
            buf = comm.SymmetricBuffer(T * K * 2, self.group.device)
            full = buf.tensor(torch.bfloat16, T * K).view(T, K)
            self._gather[key] = (buf, full, self.group.map(buf))
        buf, full, ptrs = self._gather[key]
        rows = T // self.world
        return full, full[self.rank * rows:(self.rank + 1) * rows], ptrs

    def _recv_slots(self, T: int, N: int) -> Tuple[torch.Tensor, Any]:
        key = (T, N)
        if key not in self._recv:
            rows = T // self.world
            buf = comm.SymmetricBuffer(self.world * rows * N * 2, self.group.device)
            self._recv[key] = (buf, buf.tensor(torch.bfloat16, self.world * rows * N).view(self.world, rows, N), self.group.map(buf))
        _, t, ptrs = self._recv[key]
        return t, ptrs

    # -- fused ops ---------------------------------------------------------------------------------
    def all_gather_gemm(self, T: int, weight: torch.Tensor, *, b_mn_major: bool = False, bias: Optional[torch.Tensor] = None,
                        epilogue: str = "none", aux: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                        out_dtype: torch.dtype = torch.bfloat16, push_ctas: int = PUSH_CTAS) -> torch.Tensor:
        """``weight`` is ``[N, K]`` (or ``[K, N]`` with ``b_mn_major``).  The local shard must already
        be in ``gather_buffer(T, K)[1]`` (stream-ordered before this call)."""
        G._check_operand(weight, "weight")
        N, K = (weight.shape[1], weight.shape[0]) if b_mn_major else (weight.shape[0], weight.shape[1])
        full, _, ptrs = self.gather_buffer(T, K)
        if out is None:
            out = torch.empty((T, N), dtype=out_dtype, device=full.device)
        lib = _lib.load()
        self.epoch += 1
        npush = (push_ctas or 16) if self.world > 1 else 0
        self._ready += npush
        _lib.check(lib.detb200_ag_gemm_bf16(ptrs, self.ctrl_ptrs, self.rank, self.world, weight.data_ptr(), out.data_ptr(), _lib.ptr(bias),
                                            _lib.ptr(aux), T, N, K, full.stride(0), weight.stride(0), out.stride(0),
                                            aux.stride(0) if aux is not None else 0, int(b_mn_major), int(out.dtype == torch.float32),
                                            int(bias is not None and bias.dtype == torch.float32), G.EPILOGUES[epilogue], self.epoch,
                                            self._ready, npush, _lib.stream_ptr()), "ag_gemm_bf16")
        return out

    def gather_weight_gemm(self, x: torch.Tensor, rows_total: int, cols: int, *, transposed: bool = False, bias: Optional[torch.Tensor] = None,
                           out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16, push_ctas: int = PUSH_CTAS) -> torch.Tensor:
        """ZeRO-3 style: the *weight* ``W [rows_total, cols]`` is row-sharded over the ranks inside ``gather_buffer(rows_total,
        cols)`` (this rank's rows already written).  ``transposed=False``: ``x [T, cols] @ W^T -> [T, rows_total]`` (forward; W's
        shards arrive as column blocks of the output).  ``transposed=True``: ``x [T, rows_total] @ W -> [T, cols]`` (input
        gradient; W's shards are ranges of the reduction, which starts at the own shard)."""
        G._check_operand(x, "x")
        full, _, ptrs = self.gather_buffer(rows_total, cols)
        T = x.shape[0]
        N, K = (cols, rows_total) if transposed else (rows_total, cols)
        if x.shape[1] != K:
            raise ValueError(f"gather_weight_gemm: x has {x.shape[1]} columns, expected {K}")
        if out is None:
            out = torch.empty((T, N), dtype=out_dtype, device=x.device)
        lib = _lib.load()
        self.epoch += 1
        npush = (push_ctas or 16) if self.world > 1 else 0
        self._ready += npush
        _lib.check(lib.detb200_ag_gemm_bf16_ex(ptrs, self.ctrl_ptrs, self.rank, self.world, x.data_ptr(), out.data_ptr(), _lib.ptr(bias), None, T, N, K,
                                               full.stride(0), x.stride(0), out.stride(0), 0, 1, int(transposed), 0, int(out.dtype == torch.float32),
                                               int(bias is not None and bias.dtype == torch.float32), 0, self.epoch, self._ready, npush,
                                               _lib.stream_ptr()), "ag_gemm_bf16")
        return out

    def gemm_reduce_scatter(self, a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False,
                            bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                            out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
        """Returns this rank's ``[T/W, N]`` row shard of ``sum_r a_r @ b_r^T`` (+ bias)."""
        G._check_operand(a, "a")
        G._check_operand(b, "b")
        T, K = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
        N = b.shape[1] if b_mn_major else b.shape[0]
        rows = T // self.world
        _, ptrs = self._recv_slots(T, N)
        if out is None:
            out = torch.empty((rows, N), dtype=out_dtype, device=a.device)
        lib = _lib.load()
        self.epoch += 1
        base = self._rs
        self._rs += int(lib.detb200_gemm_rs_arrivals(T, N, self.world))
        _lib.check(lib.detb200_gemm_rs_bf16(a.data_ptr(), b.data_ptr(), ptrs, self.ctrl_ptrs, self.rank, self.world, out.data_ptr(),
                                            _lib.ptr(bias), T, N, K, a.stride(0), b.stride(0), out.stride(0), int(a_mn_major),
                                            int(b_mn_major), int(out.dtype == torch.float32),
                                            int(bias is not None and bias.dtype == torch.float32), self.epoch, base & 0xFFFFFFFF,
                                            _lib.stream_ptr()), "gemm_rs_bf16")
        return out


class _ColumnParallelFn(torch.autograd.Function):
    """y_full[T, N/W] = act(all_gather(x)[T, K] @ W_r^T + b_r);  dx_shard = reduce_scatter(dy @ W_r)."""

    @staticmethod
    def forward(ctx: Any, tp: TensorParallelGemm, x_shard: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
        rows, K = x_shard.shape
        T = rows * tp.world
        full, mine, _ = tp.gather_buffer(T, K)
        mine.copy_(x_shard)
        w = G._as_bf16_2d(weight)
        y = tp.all_gather_gemm(T, w, bias=bias)
        ctx.tp, ctx.T = tp, T
        ctx.save_for_backward(full.clone(), w)     # the gather buffer is reused by the next layer
        ctx.w_dtype, ctx.b_dtype = weight.dtype, bias.dtype if bias is not None else None
        return y

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        x_full, w = ctx.saved_tensors
        dy2 = G._as_bf16_2d(dy)
        dx = ctx.tp.gemm_reduce_scatter(dy2, w, b_mn_major=True)                     # [T, N/W] x [N/W, K] -> rows of [T, K]
        dw = G.gemm(dy2, x_full, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32 if ctx.w_dtype == torch.float32 else torch.bfloat16)
        db = fused.colsum(dy2).to(ctx.b_dtype) if ctx.b_dtype is not None else None
        return None, dx, dw.to(ctx.w_dtype), db


class _RowParallelFn(torch.autograd.Function):
    """y_shard[T/W, N] = reduce_scatter(x[T, K/W] @ W_r^T) + b;  dx = all_gather(dy) @ W_r."""

    @staticmethod
    def forward(ctx: Any, tp: TensorParallelGemm, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
        x2, w = G._as_bf16_2d(x), G._as_bf16_2d(weight)
        y = tp.gemm_reduce_scatter(x2, w, bias=bias)
        ctx.tp = tp
        ctx.save_for_backward(x2, w)
        ctx.w_dtype, ctx.b_dtype = weight.dtype, bias.dtype if bias is not None else None
        return y

    @staticmethod
    def backward(ctx: Any, dy_shard: torch.Tensor) -> Tuple[Any, ...]:
        x2, w = ctx.saved_tensors
        tp = ctx.tp
        rows, N = dy_shard.shape
        T = rows * tp.world
        full, mine, _ = tp.gather_buffer(T, N)
        mine.copy_(dy_shard)
        dx = tp.all_gather_gemm(T, w, b_mn_major=True)                                # [T, N] x [N, K/W]
        dw = G.gemm(full, x2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32 if ctx.w_dtype == torch.float32 else torch.bfloat16)
        db = fused.colsum(G._as_bf16_2d(dy_shard)).to(ctx.b_dtype) if ctx.b_dtype is not None else None
        return None, dx, dw.to(ctx.w_dtype), db


def column_parallel_linear(tp: TensorParallelGemm, x_shard: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    return _ColumnParallelFn.apply(tp, x_shard, weight, bias)


def row_parallel_linear(tp: TensorParallelGemm, x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    return _RowParallelFn.apply(tp, x, weight, bias)


class _Zero3LinearFn(torch.autograd.Function):
    """Linear layer whose weight is partitioned by output rows across data-parallel ranks (ZeRO stage 3):

        forward   y  = x W^T       weight shards gathered *inside* the GEMM (they arrive as column blocks of y)
        backward  dx = dy W        gathered again inside the dgrad GEMM (as ranges of its reduction)
                  dW_shard = reduce_scatter_r(dy_r^T x_r)   the wgrad GEMM ships partial tiles to the owning rank

    so the full weight never exists outside the transient gather buffer shared by all layers of that shape, and
    neither collective runs as a separate kernel.  Gradients are averaged over the ranks (DDP convention)."""

    @staticmethod
    def forward(ctx: Any, tp: TensorParallelGemm, x: torch.Tensor, w_shard: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
        x2 = G._as_bf16_2d(x)
        rows, K = w_shard.shape
        N = rows * tp.world
        _, mine, _ = tp.gather_buffer(N, K)
        mine.copy_(w_shard)                                    # fp32 master shard -> bf16 slot of the gather buffer
        y = tp.gather_weight_gemm(x2, N, K, bias=bias)
        ctx.tp, ctx.shape, ctx.x_shape = tp, (N, K), x.shape
        ctx.save_for_backward(x2, w_shard)
        ctx.w_dtype, ctx.b_dtype = w_shard.dtype, bias.dtype if bias is not None else None
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        x2, w_shard = ctx.saved_tensors
        tp = ctx.tp
        N, K = ctx.shape
        dy2 = G._as_bf16_2d(dy)
        _, mine, _ = tp.gather_buffer(N, K)
        mine.copy_(w_shard)                                    # another layer of the same shape may have used the buffer since
        dx = tp.gather_weight_gemm(dy2, N, K, transposed=True).reshape(ctx.x_shape)
        dw = tp.gemm_reduce_scatter(dy2, x2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32)     # rows of dW owned by this rank
        dw = dw.mul_(1.0 / tp.world).to(ctx.w_dtype)
        db = None
        if ctx.b_dtype is not None:
            db = fused.colsum(dy2)
            if tp.world > 1:
                import torch.distributed as dist

                dist.all_reduce(db)
                db.mul_(1.0 / tp.world)
            db = db.to(ctx.b_dtype)
        return None, dx, dw, db


def zero3_linear(tp: TensorParallelGemm, x: torch.Tensor, w_shard: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    return _Zero3LinearFn.apply(tp, x, w_shard, bias)


class Zero3Linear(torch.nn.Module):
    """``nn.Linear`` holding only this rank's ``out_features / world`` rows of the weight (fp32 master)."""

    def __init__(self, tp: TensorParallelGemm, in_features: int, out_features: int, bias: bool = True) -> None:
        super().__init__()
        if out_features % tp.world or (out_features // tp.world) % 256:
            raise ValueError("Zero3Linear: out_features / world must be a multiple of 256")
        self.tp, self.in_features, self.out_features = tp, in_features, out_features
        self.weight_shard = torch.nn.Parameter(torch.empty(out_features // tp.world, in_features))
        self.bias = torch.nn.Parameter(torch.zeros(out_features)) if bias else None
        torch.nn.init.normal_(self.weight_shard, std=0.02)

    @torch.no_grad()
    def load_full_weight(self, w: torch.Tensor) -> None:
        rows = self.out_features // self.tp.world
        self.weight_shard.copy_(w[self.tp.rank * rows:(self.tp.rank + 1) * rows])

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return zero3_linear(self.tp, x, self.weight_shard, self.bias)
