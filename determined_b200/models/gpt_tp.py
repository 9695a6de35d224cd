"""Tensor + sequence parallel GPT blocks over the fused GEMM+collective kernels.

The residual stream is sharded by token rows (``[T/W, E]`` per rank, T = batch * seq flattened batch-major, so
shard boundaries fall on whole sequences when ``batch % W == 0``); attention heads and MLP columns are
sharded across the W ranks (Megatron-style).  Per block:

    ln1 (local rows) -> all-gather->GEMM  QKV   [T, 3E/W]     one kernel (``detb200_ag_gemm_bf16``)
    flash attention over the local heads of every sequence
    GEMM->reduce-scatter  out-proj               [T/W, E]      one kernel (``detb200_gemm_rs_bf16``)
    ln2 -> all-gather->GEMM fc1 (+GELU epilogue) [T, 4E/W] -> GEMM->reduce-scatter fc2 [T/W, E]

so a block issues four collectives, each hidden inside the GEMM that needs it.  Backward mirrors this
(the gradient of an all-gather->GEMM is a GEMM->reduce-scatter and vice versa).

Reference: the GPT-NeoX example's Megatron ``mpu`` layers (``examples/deepspeed/gpt_neox``,
``model_parallel_size`` in ``zero1.yaml:16``) with NCCL all-gather / reduce-scatter around cuBLAS.
"""
from typing import Any, List, Optional, Tuple

import torch
import torch.nn as nn

from determined_b200.models import gpt
from determined_b200.ops import attention as attn_ops
from determined_b200.ops import gemm_comm, norm as norm_ops, xent as xent_ops


class _ScaleGrad(torch.autograd.Function):
    @staticmethod
    def forward(ctx: Any, x: torch.Tensor, s: float) -> torch.Tensor:
        ctx.s = s
        return x.view_as(x)

    @staticmethod
    def backward(ctx: Any, g: torch.Tensor) -> Tuple[Any, None]:
        return g * ctx.s, None


class TPBlock(nn.Module):
    def __init__(self, cfg: gpt.GPTConfig, tp: gemm_comm.TensorParallelGemm) -> None:
        super().__init__()
        W = tp.world
        if cfg.n_head % W or (cfg.mlp_ratio * cfg.n_embd) % W:
            raise ValueError(f"heads ({cfg.n_head}) and MLP width must be divisible by the tensor-parallel size {W}")
        self.cfg, self.tp = cfg, tp
        E, self.h_local = cfg.n_embd, cfg.n_head // W
        F4 = cfg.mlp_ratio * E // W
        self.ln1 = norm_ops.LayerNorm(E, eps=cfg.layer_norm_eps)
        self.ln2 = norm_ops.LayerNorm(E, eps=cfg.layer_norm_eps)
        # sharded parameters: rows of this rank's heads / MLP columns
        self.qkv_w = nn.Parameter(torch.empty(3 * self.h_local * cfg.head_dim, E))
        self.qkv_b = nn.Parameter(torch.zeros(3 * self.h_local * cfg.head_dim))
        self.proj_w = nn.Parameter(torch.empty(E, self.h_local * cfg.head_dim))
        self.proj_b = nn.Parameter(torch.zeros(E))            # replicated
        self.fc1_w = nn.Parameter(torch.empty(F4, E))
        self.fc1_b = nn.Parameter(torch.zeros(F4))
        self.fc2_w = nn.Parameter(torch.empty(E, F4))
        self.fc2_b = nn.Parameter(torch.zeros(E))             # replicated
        self.rot = int(cfg.head_dim * cfg.rotary_pct) // 2 * 2
        self.replicated = [self.ln1.weight, self.ln1.bias, self.ln2.weight, self.ln2.bias, self.proj_b, self.fc2_b]

    @torch.no_grad()
    def load_from_full(self, blk: gpt.Block) -> None:
        """Take this rank's shard of an unsharded block's parameters."""
        r, W, E, D, hl = self.tp.rank, self.tp.world, self.cfg.n_embd, self.cfg.head_dim, self.h_local
        qkv_w = blk.qkv.weight.view(3, self.cfg.n_head, D, E)[:, r * hl:(r + 1) * hl].reshape(-1, E)
        qkv_b = blk.qkv.bias.view(3, self.cfg.n_head, D)[:, r * hl:(r + 1) * hl].reshape(-1)
        self.qkv_w.copy_(qkv_w)
        self.qkv_b.copy_(qkv_b)
        self.proj_w.copy_(blk.proj.weight[:, r * hl * D:(r + 1) * hl * D])
        self.proj_b.copy_(blk.proj.bias)
        F4 = self.fc1_w.shape[0]
        self.fc1_w.copy_(blk.mlp.fc1.weight[r * F4:(r + 1) * F4])
        self.fc1_b.copy_(blk.mlp.fc1.bias[r * F4:(r + 1) * F4])
        self.fc2_w.copy_(blk.mlp.fc2.weight[:, r * F4:(r + 1) * F4])
        self.fc2_b.copy_(blk.mlp.fc2.bias)
        for mine, full in ((self.ln1, blk.ln1), (self.ln2, blk.ln2)):
            mine.weight.copy_(full.weight)
            mine.bias.copy_(full.bias)

    def attend(self, x_shard: torch.Tensor, B: int, S: int, rope: Any) -> torch.Tensor:
        D = self.cfg.head_dim
        qkv = gemm_comm.column_parallel_linear(self.tp, x_shard, self.qkv_w, self.qkv_b)             # [T, 3*hl*D]
        o = attn_ops.attention_qkvpacked(qkv.view(B, S, 3, self.h_local, D), (rope[2], rope[3]) if rope is not None else None, self.rot, causal=True)
        # the bias of a row-parallel layer is added once, after the reduction (inside the kernel's reduce phase)
        return gemm_comm.row_parallel_linear(self.tp, o.reshape(B * S, self.h_local * D), self.proj_w, self.proj_b)

    def mlp(self, x_shard: torch.Tensor) -> torch.Tensor:
        h = gemm_comm.column_parallel_linear(self.tp, x_shard, self.fc1_w, self.fc1_b)
        h = torch.nn.functional.gelu(h, approximate="tanh" if self.cfg.activation == "gelu_tanh" else "none")
        return gemm_comm.row_parallel_linear(self.tp, h, self.fc2_w, self.fc2_b)

    def forward(self, x: torch.Tensor, B: int, S: int, rope: Any) -> torch.Tensor:
        if self.cfg.parallel_residual:
            return x + self.attend(self.ln1(x), B, S, rope) + self.mlp(self.ln2(x))
        x = x + self.attend(self.ln1(x), B, S, rope)
        return x + self.mlp(self.ln2(x))


class TPGPT(nn.Module):
    """Sequence-sharded GPT: embeddings, final norm and LM head are replicated and run on this rank's token rows."""

    def __init__(self, cfg: gpt.GPTConfig, tp: gemm_comm.TensorParallelGemm) -> None:
        super().__init__()
        self.cfg, self.tp = cfg, tp
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.block_size, cfg.n_embd) if cfg.rotary_pct == 0.0 else None
        self.blocks = nn.ModuleList(TPBlock(cfg, tp) for _ in range(cfg.n_layer))
        self.ln_f = norm_ops.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps)
        from determined_b200.ops import gemm as gemm_ops

        self.lm_head = gemm_ops.FusedLinear(cfg.n_embd, cfg.vocab_size, bias=False)
        if cfg.tie_embeddings:
            self.lm_head.weight = self.wte.weight
        self._rope: dict = {}

    @torch.no_grad()
    def load_from_full(self, full: gpt.GPT) -> None:
        self.wte.weight.copy_(full.wte.weight)
        if self.wpe is not None:
            self.wpe.weight.copy_(full.wpe.weight)
        for mine, theirs in zip(self.blocks, full.blocks):
            mine.load_from_full(theirs)
        self.ln_f.weight.copy_(full.ln_f.weight)
        self.ln_f.bias.copy_(full.ln_f.bias)
        if not self.cfg.tie_embeddings:
            self.lm_head.weight.copy_(full.lm_head.weight)

    def replicated_parameters(self) -> List[nn.Parameter]:
        out = [self.wte.weight, self.ln_f.weight, self.ln_f.bias]
        if self.wpe is not None:
            out.append(self.wpe.weight)
        if not self.cfg.tie_embeddings:
            out.append(self.lm_head.weight)
        for b in self.blocks:
            out += b.replicated
        return out

    def sync_replicated_grads(self) -> None:
        """Replicated parameters see only this rank's tokens: sum their gradients over the TP group (one flat all-reduce)."""
        import torch.distributed as dist

        ps = [p for p in self.replicated_parameters() if p.grad is not None]
        if self.tp.world == 1 or not ps:
            return
        flat = torch.cat([p.grad.reshape(-1).float() for p in ps])
        dist.all_reduce(flat)
        off = 0
        for p in ps:
            n = p.numel()
            p.grad.copy_(flat[off:off + n].view_as(p.grad))
            off += n

    def forward(self, idx: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        """``idx``/``targets``: the *full* ``[B, S]`` batch (identical on every rank).  Returns this rank's share of the
        mean loss; summing over ranks (or calling backward on each) gives the unsharded model's loss / gradients."""
        B, S = idx.shape
        W, r = self.tp.world, self.tp.rank
        T = B * S
        rows = T // W
        flat_idx = idx.reshape(T)[r * rows:(r + 1) * rows]
        x = self.wte(flat_idx)
        if self.wpe is not None:
            pos = (torch.arange(T, device=idx.device) % S)[r * rows:(r + 1) * rows]
            x = x + self.wpe(pos)
        x = x.to(torch.bfloat16)
        rope = None
        rot = int(self.cfg.head_dim * self.cfg.rotary_pct) // 2 * 2
        if rot:
            key = (S, str(idx.device))
            if key not in self._rope:
                self._rope[key] = gpt._rotary_tables(S, rot, self.cfg.rotary_base, idx.device)
            rope = self._rope[key]
        for blk in self.blocks:
            x = blk(x, B, S, rope)
        x = self.ln_f(x)
        logits = self.lm_head(x)
        tgt = targets.reshape(T)[r * rows:(r + 1) * rows]
        local = xent_ops.cross_entropy(logits, tgt, inplace_backward=True)      # mean over this rank's rows
        return local / W
