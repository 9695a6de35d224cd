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
        self.external_bias_reduce = False   # an engine that all-reduces replicated gradients itself sets this
        self._recv: dict = {}

    # -- buffers -----------------------------------------------------------------------------------
    def _quiesce(self) -> None:
        """Mapping a new symmetric buffer exchanges IPC handles through a ``torch.distributed`` collective, i.e. an NCCL kernel on
        another stream.  The fused kernels hold every SM while they wait for their peers, so an NCCL kernel that needs an SM can
        close a cycle across ranks (A's NCCL waits for B's, B's cannot start until B's GEMM gets A's tiles, A's tile producers
        queue behind A's NCCL kernel).  Drain the device first; steady-state steps never come here."""
        if self.world > 1:
            torch.cuda.synchronize(self.group.device)

    def prepare(self, rows_total: int, cols: int) -> None:
        """Allocate and map everything a ``[rows_total, cols]`` partitioned weight needs (gather buffer, wgrad receive slots) now,
        at model-construction time, instead of inside the first forward / backward."""
        self.gather_buffer(rows_total, cols)
        self._recv_slots(rows_total, cols)

    def gather_buffer(self, T: int, K: int) -> Tuple[torch.Tensor, torch.Tensor, Any]:
        """Symmetric ``[T, K]`` bf16 buffer; returns (full, this rank's row shard view, peer ptrs).
        Write the local activations into the shard view, then call :meth:`all_gather_gemm`."""
        key = (T, K)
        if key not in self._gather:
            self._quiesce()
            buf = comm.SymmetricBuffer(T * K * 2, self.group.device)
            full = buf.tensor(torch.bfloat16, T * K).view(T, K)
            self._gather[key] = (buf, full, self.group.map(buf))
        buf, full, ptrs = self._gather[key]
        rows = T // self.world
        return full, full[self.rank * rows:(self.rank + 1) * rows], ptrs

    def _recv_slots(self, T: int, N: int) -> Tuple[torch.Tensor, Any]:
        key = (T, N)
        if key not in self._recv:
            self._quiesce()
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
                           epilogue: str = "none", aux: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                           out_dtype: torch.dtype = torch.bfloat16, push_ctas: int = PUSH_CTAS) -> torch.Tensor:
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
        _lib.check(lib.detb200_ag_gemm_bf16_ex(ptrs, self.ctrl_ptrs, self.rank, self.world, x.data_ptr(), out.data_ptr(), _lib.ptr(bias), _lib.ptr(aux),
                                               T, N, K, full.stride(0), x.stride(0), out.stride(0), aux.stride(0) if aux is not None else 0, 1,
                                               int(transposed), 0, int(out.dtype == torch.float32),
                                               int(bias is not None and bias.dtype == torch.float32), G.EPILOGUES[epilogue], self.epoch, self._ready,
                                               npush, _lib.stream_ptr()), "ag_gemm_bf16")
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


class HostTensorParallel:
    """``torch.distributed`` twin of :class:`TensorParallelGemm` with the same methods — separate collectives around
    plain matmuls.  It exists so the layers and engines built on the fused kernels (ZeRO-3 partitioning, checkpoint
    consolidation) run under ``gloo`` in the CPU test-suite; it is never selected on a GPU."""

    def __init__(self, group: Any = None) -> None:
        import torch.distributed as dist

        self.pg = group
        on = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if on else 0
        self.world = dist.get_world_size(group) if on else 1
        self.external_bias_reduce = False
        self._gather: dict = {}

    def prepare(self, rows_total: int, cols: int) -> None:
        self.gather_buffer(rows_total, cols)

    def gather_buffer(self, T: int, K: int) -> Tuple[torch.Tensor, torch.Tensor, Any]:
        if (T, K) not in self._gather:
            self._gather[(T, K)] = torch.zeros(T, K, dtype=torch.bfloat16)
        full = self._gather[(T, K)]
        rows = T // self.world
        return full, full[self.rank * rows:(self.rank + 1) * rows], None

    def _all_gather(self, T: int, K: int) -> torch.Tensor:
        import torch.distributed as dist

        full, mine, _ = self.gather_buffer(T, K)
        if self.world > 1:
            src = mine.contiguous().view(torch.int32)            # gloo has no 16-bit types
            parts = [torch.empty_like(src) for _ in range(self.world)]
            dist.all_gather(parts, src, group=self.pg)
            full.copy_(torch.cat(parts).view(torch.bfloat16))
        return full

    def gather_weight_gemm(self, x: torch.Tensor, rows_total: int, cols: int, *, transposed: bool = False, bias: Optional[torch.Tensor] = None,
                           epilogue: str = "none", aux: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                           out_dtype: torch.dtype = torch.bfloat16, push_ctas: int = 0) -> torch.Tensor:
        w = self._all_gather(rows_total, cols)
        if out is None:
            out = torch.empty((x.shape[0], cols if transposed else rows_total), dtype=out_dtype)
        return G._gemm_cpu(x, w, False, transposed, bias, out, G.EPILOGUES[epilogue], aux, False)

    def gemm_reduce_scatter(self, a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False,
                            bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                            out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
        import torch.distributed as dist

        T = a.shape[1] if a_mn_major else a.shape[0]
        N = b.shape[1] if b_mn_major else b.shape[0]
        full = G._gemm_cpu(a, b, a_mn_major, b_mn_major, None, torch.empty((T, N), dtype=torch.float32), 0, None, False)
        if self.world > 1:
            dist.all_reduce(full, group=self.pg)
        rows = T // self.world
        r = full[self.rank * rows:(self.rank + 1) * rows]
        if bias is not None:
            r = r + bias.float()
        if out is None:
            return r.to(out_dtype)
        return out.copy_(r)


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

        forward   y  = act(x W^T + b)   weight shards gathered *inside* the GEMM (they arrive as column blocks of y)
        backward  dx = dy W             gathered again inside the dgrad GEMM (as ranges of its reduction)
                  dW_shard = reduce_scatter_r(dy_r^T x_r)   the wgrad GEMM ships partial tiles to the owning rank

    so the full weight never exists outside the transient gather buffer shared by all layers of that shape, and
    neither collective runs as a separate kernel.  Gradients are averaged over the ranks (DDP convention).  The bias
    is replicated; its gradient is all-reduced here unless an engine does that (``tp.external_bias_reduce``)."""

    @staticmethod
    def forward(ctx: Any, tp: TensorParallelGemm, x: torch.Tensor, w_shard: torch.Tensor, bias: Optional[torch.Tensor], epi: int) -> torch.Tensor:
        x2 = G._as_bf16_2d(x)
        rows, K = w_shard.shape
        N = rows * tp.world
        _, mine, _ = tp.gather_buffer(N, K)
        mine.copy_(w_shard)                                    # fp32 master shard -> bf16 slot of the gather buffer
        pre = torch.empty((x2.shape[0], N), dtype=torch.bfloat16, device=x.device) if epi else None
        y = tp.gather_weight_gemm(x2, N, K, bias=bias, epilogue=_FWD_EPI[epi], aux=pre)
        ctx.tp, ctx.shape, ctx.x_shape, ctx.epi = tp, (N, K), x.shape, epi
        ctx.save_for_backward(x2, w_shard, pre)
        ctx.w_dtype, ctx.b_dtype = w_shard.dtype, bias.dtype if bias is not None else None
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        x2, w_shard, pre = ctx.saved_tensors
        tp = ctx.tp
        N, K = ctx.shape
        dy2 = G._as_bf16_2d(dy)
        if ctx.epi:
            dy2 = (dy2.float() * G._dact(pre, ctx.epi + 3)).to(torch.bfloat16)
        _, mine, _ = tp.gather_buffer(N, K)
        mine.copy_(w_shard)                                    # another layer of the same shape may have used the buffer since
        dx = tp.gather_weight_gemm(dy2, N, K, transposed=True).reshape(ctx.x_shape) if ctx.needs_input_grad[1] else None
        dw = _shard_wgrad(tp, dy2, x2, ctx.w_dtype)
        db = _bias_grad(tp, dy2, ctx.b_dtype)
        return None, dx, dw, db, None


_FWD_EPI = {0: "none", 1: "gelu", 2: "gelu_tanh", 3: "relu"}
_BWD_EPI = {1: "dgelu", 2: "dgelu_tanh", 3: "drelu"}


def _shard_wgrad(tp: TensorParallelGemm, dy2: torch.Tensor, x2: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """Rows of ``mean_r(dy_r^T x_r)`` owned by this rank (the reduce-scatter happens in the GEMM epilogue)."""
    dw = tp.gemm_reduce_scatter(dy2, x2, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32)
    if tp.world > 1:
        dw.mul_(1.0 / tp.world)
    return dw if dw.dtype == dtype else dw.to(dtype)


def _bias_grad(tp: TensorParallelGemm, dy2: torch.Tensor, dtype: Optional[torch.dtype]) -> Optional[torch.Tensor]:
    if dtype is None:
        return None
    db = fused.colsum(dy2)
    if tp.world > 1 and not tp.external_bias_reduce:
        import torch.distributed as dist

        dist.all_reduce(db)
        db.mul_(1.0 / tp.world)
    return db.to(dtype)


class _Zero3MLPFn(torch.autograd.Function):
    """``act(x W1^T + b1) W2^T + b2`` with both weights row-partitioned: four gather-fused GEMMs and two reduce-scatter-fused
    wgrads; the activation and its derivative live in GEMM epilogues like :class:`determined_b200.ops.gemm._MLPFn`."""

    @staticmethod
    def forward(ctx: Any, tp: TensorParallelGemm, x: torch.Tensor, w1: torch.Tensor, b1: Optional[torch.Tensor], w2: torch.Tensor,
                b2: Optional[torch.Tensor], epi: int) -> torch.Tensor:
        x2 = G._as_bf16_2d(x)
        inner, hidden = w1.shape[0] * tp.world, w1.shape[1]
        out = w2.shape[0] * tp.world
        T = x2.shape[0]
        tp.gather_buffer(inner, hidden)[1].copy_(w1)
        pre = torch.empty((T, inner), dtype=torch.bfloat16, device=x.device)
        h = tp.gather_weight_gemm(x2, inner, hidden, bias=b1, epilogue=_FWD_EPI[epi], aux=pre)
        tp.gather_buffer(out, inner)[1].copy_(w2)
        y = tp.gather_weight_gemm(h, out, inner, bias=b2)
        ctx.tp, ctx.dims, ctx.epi, ctx.x_shape = tp, (inner, hidden, out), epi, x.shape
        ctx.save_for_backward(x2, w1, w2, pre, h)
        ctx.dtypes = (w1.dtype, b1.dtype if b1 is not None else None, w2.dtype, b2.dtype if b2 is not None else None)
        return y.reshape(*x.shape[:-1], out)

    @staticmethod
    def backward(ctx: Any, dy: torch.Tensor) -> Tuple[Any, ...]:
        x2, w1, w2, pre, h = ctx.saved_tensors
        tp = ctx.tp
        inner, hidden, out = ctx.dims
        d1, db1t, d2, db2t = ctx.dtypes
        dy2 = G._as_bf16_2d(dy)
        dw2 = _shard_wgrad(tp, dy2, h, d2)
        db2 = _bias_grad(tp, dy2, db2t)
        tp.gather_buffer(out, inner)[1].copy_(w2)
        dpre = tp.gather_weight_gemm(dy2, out, inner, transposed=True, epilogue=_BWD_EPI[ctx.epi], aux=pre)
        dw1 = _shard_wgrad(tp, dpre, x2, d1)
        db1 = _bias_grad(tp, dpre, db1t)
        dx = None
        if ctx.needs_input_grad[1]:
            tp.gather_buffer(inner, hidden)[1].copy_(w1)
            dx = tp.gather_weight_gemm(dpre, inner, hidden, transposed=True).reshape(ctx.x_shape)
        return None, dx, dw1, db1, dw2, db2, None


def zero3_linear(tp: TensorParallelGemm, x: torch.Tensor, w_shard: torch.Tensor, bias: Optional[torch.Tensor] = None,
                 activation: str = "none") -> torch.Tensor:
    return _Zero3LinearFn.apply(tp, x, w_shard, bias, G.EPILOGUES[activation])


def zero3_mlp(tp: TensorParallelGemm, x: torch.Tensor, w1: torch.Tensor, b1: Optional[torch.Tensor], w2: torch.Tensor,
              b2: Optional[torch.Tensor], activation: str = "gelu") -> torch.Tensor:
    return _Zero3MLPFn.apply(tp, x, w1, b1, w2, b2, G.EPILOGUES[activation])


def shard_eligible(out_features: int, in_features: int, world: int) -> bool:
    """Row-partitioning constraints of the fused kernels: 256-row blocks per rank in the forward, 128-row blocks in the
    reduce-scattered wgrad, 64-wide reduction blocks, at most 8 ranks."""
    return 1 <= world <= 8 and out_features % world == 0 and (out_features // world) % 256 == 0 and in_features % 64 == 0


class Zero3Linear(torch.nn.Module):
    """``nn.Linear`` holding only this rank's ``out_features / world`` rows of the weight (fp32 master).

    State dicts: ``weight_shard`` is what ``state_dict()`` carries; loading accepts either that or a full ``weight``
    (sliced on the way in), so checkpoints of the unpartitioned model load unchanged.  :func:`consolidated_state_dict`
    goes the other way."""

    def __init__(self, tp: TensorParallelGemm, in_features: int, out_features: int, bias: bool = True, activation: str = "none",
                 device: Any = None) -> None:
        super().__init__()
        if not shard_eligible(out_features, in_features, tp.world):
            raise ValueError("Zero3Linear: out_features / world must be a multiple of 256 and in_features of 64")
        self.tp, self.in_features, self.out_features, self.activation = tp, in_features, out_features, activation
        self.weight_shard = torch.nn.Parameter(torch.empty(out_features // tp.world, in_features, device=device))
        self.weight_shard._zero3_shard = True          # type: ignore[attr-defined]
        tp.prepare(out_features, in_features)
        self.bias = torch.nn.Parameter(torch.zeros(out_features, device=device)) if bias else None
        torch.nn.init.normal_(self.weight_shard, std=0.02)

    @property
    def rows(self) -> slice:
        n = self.out_features // self.tp.world
        return slice(self.tp.rank * n, (self.tp.rank + 1) * n)

    @torch.no_grad()
    def load_full_weight(self, w: torch.Tensor) -> None:
        self.weight_shard.copy_(w[self.rows])

    @torch.no_grad()
    def full_weight(self) -> torch.Tensor:
        """All-gathers the shards (a collective: call on every rank)."""
        if self.tp.world == 1:
            return self.weight_shard.detach().clone()
        import torch.distributed as dist

        full = torch.empty(self.out_features, self.in_features, dtype=self.weight_shard.dtype, device=self.weight_shard.device)
        dist.all_gather_into_tensor(full, self.weight_shard.detach().contiguous())
        return full

    def _load_from_state_dict(self, state_dict: Any, prefix: str, *args: Any, **kw: Any) -> None:
        full = state_dict.pop(prefix + "weight", None)
        if full is not None and prefix + "weight_shard" not in state_dict:
            state_dict[prefix + "weight_shard"] = full[self.rows]
        super()._load_from_state_dict(state_dict, prefix, *args, **kw)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return zero3_linear(self.tp, x, self.weight_shard, self.bias, self.activation)

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features} (rows {self.rows.start}:{self.rows.stop} here), bias={self.bias is not None}"


class Zero3MLP(torch.nn.Module):
    """Partitioned twin of :class:`determined_b200.ops.gemm.FusedMLP` (same parameter names under ``fc1`` / ``fc2``)."""

    def __init__(self, tp: TensorParallelGemm, hidden: int, inner: int, activation: str = "gelu", bias: bool = True, device: Any = None,
                 fc1: Optional[Zero3Linear] = None, fc2: Optional[Zero3Linear] = None) -> None:
        super().__init__()
        self.tp, self.activation = tp, activation
        self.fc1 = fc1 if fc1 is not None else Zero3Linear(tp, hidden, inner, bias=bias, device=device)
        self.fc2 = fc2 if fc2 is not None else Zero3Linear(tp, inner, hidden, bias=bias, device=device)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return zero3_mlp(self.tp, x, self.fc1.weight_shard, self.fc1.bias, self.fc2.weight_shard, self.fc2.bias, self.activation)


def _from_linear(tp: TensorParallelGemm, lin: torch.nn.Linear, activation: str = "none") -> Zero3Linear:
    z = Zero3Linear(tp, lin.in_features, lin.out_features, bias=lin.bias is not None, activation=activation, device=lin.weight.device)
    with torch.no_grad():
        z.load_full_weight(lin.weight)
        if lin.bias is not None:
            z.bias.copy_(lin.bias)
    return z


def partition_linears_(module: torch.nn.Module, tp: TensorParallelGemm, min_params: int = 1 << 20) -> int:
    """ZeRO stage 3 partitioning pass (the role of ``deepspeed.zero.Init``): replaces every eligible ``FusedLinear`` /
    ``nn.Linear`` / ``FusedMLP`` below ``module`` by its row-partitioned twin, keeping this rank's rows of the current
    weights.  Tied or small (``< min_params`` elements) weights and shapes the kernels cannot split stay replicated.
    Returns the number of weights partitioned.  Every rank must hold identical weights beforehand."""
    # a tied weight shows up once in parameters(); detect ties through the owning modules
    owners: dict = {}
    for m in module.modules():
        w = getattr(m, "weight", None)
        if isinstance(w, torch.nn.Parameter):
            owners[id(w)] = owners.get(id(w), 0) + 1
    n = 0

    def ok(lin: torch.nn.Linear) -> bool:
        return (owners.get(id(lin.weight), 0) == 1 and lin.weight.numel() >= min_params
                and shard_eligible(lin.out_features, lin.in_features, tp.world))

    def walk(parent: torch.nn.Module) -> None:
        nonlocal n
        for name, child in list(parent.named_children()):
            if isinstance(child, G.FusedMLP) and ok(child.fc1) and ok(child.fc2):
                setattr(parent, name, Zero3MLP(tp, child.fc1.in_features, child.fc1.out_features, child.activation,
                                               fc1=_from_linear(tp, child.fc1), fc2=_from_linear(tp, child.fc2)))
                n += 2
            elif isinstance(child, torch.nn.Linear) and not isinstance(parent, G.FusedMLP) and ok(child):
                setattr(parent, name, _from_linear(tp, child, getattr(child, "activation", "none")))
                n += 1
            else:
                walk(child)

    walk(module)
    return n


def consolidated_state_dict(module: torch.nn.Module) -> dict:
    """``module.state_dict()`` with every partitioned weight gathered back under its original ``weight`` key, i.e. the state
    dict of the unpartitioned model (a collective: call on every rank; tensors stay on the device)."""
    sd = module.state_dict()
    for name, m in module.named_modules():
        if isinstance(m, Zero3Linear):
            prefix = name + "." if name else ""
            sd.pop(prefix + "weight_shard", None)
            sd[prefix + "weight"] = m.full_weight()
    return sd
