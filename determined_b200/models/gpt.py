"""GPT-2 / GPT-NeoX decoder family built from the native ops.

Every hot op is ours: tcgen05 GEMMs (QKV / out-proj / fused GELU MLP / LM head), tcgen05 flash
attention consuming the packed QKV projection in place, vector LayerNorm, fused softmax
cross-entropy.  What stays in PyTorch: embedding gathers, rotary rotation and residual adds.

Reference: the GPT-2 (HF ``run_clm``) and GPT-NeoX (Megatron/DeepSpeed) examples —
``examples/hf_trainer_api/hf_language_modeling``, ``examples/deepspeed/gpt_neox`` — whose models
live in external libraries; SURVEY.md §6 configs #3 / #4.
"""
import dataclasses
import math
from typing import Any, Dict, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from determined_b200.ops import attention as attn_ops
from determined_b200.ops import gemm as gemm_ops
from determined_b200.ops import norm as norm_ops
from determined_b200.ops import xent as xent_ops


@dataclasses.dataclass
class GPTConfig:
    vocab_size: int = 50304
    n_layer: int = 12
    n_head: int = 12
    n_embd: int = 768
    block_size: int = 1024
    mlp_ratio: int = 4
    rotary_pct: float = 0.0            # 0 -> learned absolute positions (GPT-2); >0 -> rotary on that share of head dim (NeoX)
    rotary_base: float = 10000.0
    parallel_residual: bool = False    # NeoX: x + attn(ln1(x)) + mlp(ln2(x))
    layer_norm_eps: float = 1e-5
    tie_embeddings: bool = True
    activation: str = "gelu"           # "gelu" (erf) | "gelu_tanh" | "relu"
    bias: bool = True
    init_std: float = 0.02

    @property
    def head_dim(self) -> int:
        return self.n_embd // self.n_head


PRESETS: Dict[str, Dict[str, Any]] = {
    "gpt2": dict(n_layer=12, n_head=12, n_embd=768, activation="gelu_tanh"),
    "gpt2-medium": dict(n_layer=24, n_head=16, n_embd=1024, activation="gelu_tanh"),
    "gpt2-large": dict(n_layer=36, n_head=20, n_embd=1280, activation="gelu_tanh"),
    "gpt2-xl": dict(n_layer=48, n_head=25, n_embd=1600, activation="gelu_tanh"),
    # NeoX-style (rotary, parallel residual, untied head); sizes follow the public NeoX configs
    "neox-125m": dict(n_layer=12, n_head=12, n_embd=768, rotary_pct=0.25, parallel_residual=True, tie_embeddings=False, block_size=2048),
    "neox-1.3b": dict(n_layer=24, n_head=16, n_embd=2048, rotary_pct=0.25, parallel_residual=True, tie_embeddings=False, block_size=2048),
    "neox-2.7b": dict(n_layer=32, n_head=32, n_embd=2560, rotary_pct=0.25, parallel_residual=True, tie_embeddings=False, block_size=2048),
    "neox-6.7b": dict(n_layer=32, n_head=32, n_embd=4096, rotary_pct=0.25, parallel_residual=True, tie_embeddings=False, block_size=2048),
    "tiny": dict(n_layer=2, n_head=2, n_embd=128, block_size=128, vocab_size=512),
}


def config_for(name: str, **overrides: Any) -> GPTConfig:
    kw = dict(PRESETS[name])
    kw.update(overrides)
    return GPTConfig(**kw)


def _rotary_tables(seq: int, dim: int, base: float, device: torch.device) -> Tuple[torch.Tensor, ...]:
    inv = 1.0 / (base ** (torch.arange(0, dim, 2, device=device, dtype=torch.float32) / dim))
    t = torch.arange(seq, device=device, dtype=torch.float32)
    f = torch.outer(t, inv)
    emb = torch.cat((f, f), dim=-1)
    return emb.cos(), emb.sin(), f.cos().contiguous(), f.sin().contiguous()       # full tables + the [S, rot/2] halves


def apply_rotary(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, rot: int) -> torch.Tensor:
    """x: [B, S, H, D]; rotate the first ``rot`` channels (NeoX "rotate half" convention)."""
    xr, xp = x[..., :rot], x[..., rot:]
    half = rot // 2
    x1, x2 = xr[..., :half], xr[..., half:]
    rotated = torch.cat((-x2, x1), dim=-1)
    c, s = cos[None, :, None, :].to(x.dtype), sin[None, :, None, :].to(x.dtype)
    return torch.cat((xr * c + rotated * s, xp), dim=-1)


class _TorchMLP(nn.Module):
    """Library-only twin of FusedMLP (same parameter names) for the comparison arm of benchmarks."""

    def __init__(self, hidden: int, inner: int, activation: str, bias: bool) -> None:
        super().__init__()
        self.fc1 = nn.Linear(hidden, inner, bias=bias)
        self.fc2 = nn.Linear(inner, hidden, bias=bias)
        self.activation = activation

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        h = self.fc1(x)
        h = F.gelu(h) if self.activation == "gelu" else F.gelu(h, approximate="tanh") if self.activation == "gelu_tanh" else F.relu(h)
        return self.fc2(h)


class Block(nn.Module):
    def __init__(self, cfg: GPTConfig, native: bool = True) -> None:
        super().__init__()
        self.cfg, self.native = cfg, native
        E = cfg.n_embd
        if native:
            self.ln1 = norm_ops.LayerNorm(E, eps=cfg.layer_norm_eps)
            self.ln2 = norm_ops.LayerNorm(E, eps=cfg.layer_norm_eps)
            self.qkv = gemm_ops.FusedLinear(E, 3 * E, bias=cfg.bias)
            self.proj = gemm_ops.FusedLinear(E, E, bias=cfg.bias)
            self.mlp = gemm_ops.FusedMLP(E, cfg.mlp_ratio * E, activation=cfg.activation, bias=cfg.bias)
        else:
            self.ln1 = nn.LayerNorm(E, eps=cfg.layer_norm_eps)
            self.ln2 = nn.LayerNorm(E, eps=cfg.layer_norm_eps)
            self.qkv = nn.Linear(E, 3 * E, bias=cfg.bias)
            self.proj = nn.Linear(E, E, bias=cfg.bias)
            self.mlp = _TorchMLP(E, cfg.mlp_ratio * E, cfg.activation, cfg.bias)
        self.rot = int(cfg.head_dim * cfg.rotary_pct) // 2 * 2

    def attend(self, x: torch.Tensor, rope: Optional[Tuple[torch.Tensor, ...]]) -> torch.Tensor:
        B, S, E = x.shape
        H, D = self.cfg.n_head, self.cfg.head_dim
        qkv = self.qkv(x).view(B, S, 3, H, D)
        if getattr(self, "cp_group", None) is not None:
            # the sequence is sharded over the context-parallel group: K/V shards travel the ring (parallel/context_parallel.py);
            # `rope` already holds the tables of this rank's global positions
            from determined_b200.parallel import context_parallel as cp

            q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
            if self.rot and rope is not None:
                q, k = apply_rotary(q, rope[0], rope[1], self.rot), apply_rotary(k, rope[0], rope[1], self.rot)
            o = cp.ring_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), self.cp_group, causal=True).transpose(1, 2)
            return self.proj(o.reshape(B, S, E))
        if self.native:
            # rotary in place on the packed projection, attention reads the strided slices, one packed dqkv comes back
            o = attn_ops.attention_qkvpacked(qkv, (rope[2], rope[3]) if rope is not None else None, self.rot, causal=True)
            return self.proj(o.reshape(B, S, E))
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
        if self.rot and rope is not None:
            q, k = apply_rotary(q, rope[0], rope[1], self.rot), apply_rotary(k, rope[0], rope[1], self.rot)
        if self.native:
            o = attn_ops.flash_attention(q, k, v, causal=True, layout="bshd")
        else:
            o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True).transpose(1, 2)
        return self.proj(o.reshape(B, S, E))

    def forward(self, x: torch.Tensor, rope: Optional[Tuple[torch.Tensor, ...]]) -> torch.Tensor:
        if self.cfg.parallel_residual:
            return x + self.attend(self.ln1(x), rope) + self.mlp(self.ln2(x))
        x = x + self.attend(self.ln1(x), rope)
        return x + self.mlp(self.ln2(x))


class GPT(nn.Module):
    """``native=False`` builds the same network from stock torch modules (cuBLAS / cuDNN-SDPA / ATen),
    to be run under ``torch.autocast`` — the comparison arm; state dicts are interchangeable."""

    def __init__(self, cfg: GPTConfig, native: bool = True) -> None:
        super().__init__()
        self.cfg, self.native = cfg, native
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.block_size, cfg.n_embd) if cfg.rotary_pct == 0.0 else None
        self.blocks = nn.ModuleList(Block(cfg, native) for _ in range(cfg.n_layer))
        if native:
            self.ln_f = norm_ops.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps)
            self.lm_head = gemm_ops.FusedLinear(cfg.n_embd, cfg.vocab_size, bias=False)
        else:
            self.ln_f = nn.LayerNorm(cfg.n_embd, eps=cfg.layer_norm_eps)
            self.lm_head = nn.Linear(cfg.n_embd, cfg.vocab_size, bias=False)
        if cfg.tie_embeddings:
            self.lm_head.weight = self.wte.weight
        self._rope: Dict[Tuple[int, str], Tuple[torch.Tensor, ...]] = {}
        self.apply(self._init)
        for name, p in self.named_parameters():
            if name.endswith("proj.weight") or name.endswith("fc2.weight"):
                nn.init.normal_(p, mean=0.0, std=cfg.init_std / math.sqrt(2 * cfg.n_layer))

    def _init(self, m: nn.Module) -> None:
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, mean=0.0, std=self.cfg.init_std)
            if m.bias is not None:
                nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, mean=0.0, std=self.cfg.init_std)

    def num_params(self, non_embedding: bool = True) -> int:
        n = sum(p.numel() for p in self.parameters())
        if non_embedding and self.wpe is not None:
            n -= self.wpe.weight.numel()
        return n

    def flops_per_token(self, seq: int) -> float:
        """Forward+backward model FLOPs per token (6N + attention), the convention of the NeoX/PaLM MFU."""
        c = self.cfg
        n = sum(p.numel() for name, p in self.named_parameters() if "wte" not in name and "wpe" not in name and "lm_head" not in name)
        n += c.vocab_size * c.n_embd            # LM head matmul (tied or not)
        return 6.0 * n + 12.0 * c.n_layer * c.n_embd * seq * 0.5

    cp_group: Any = None

    def set_context_parallel(self, group: Any) -> None:
        """Shard the sequence over ``group``: ``forward`` then takes this rank's contiguous slice of the tokens (and targets),
        attention runs as a ring over the group, positions are the global ones.  The loss returned is the mean over the local
        tokens; with equal shards the mean over the group's ranks is the loss of the whole sequence, and parameter gradients
        are averaged over the group like over data-parallel replicas.  ``None`` turns it off."""
        self.cp_group = group
        for blk in self.blocks:
            blk.cp_group = group

    def _cp_offset(self, S: int) -> Tuple[int, int]:
        """(first global position of the local shard, global sequence length)."""
        if self.cp_group is None:
            return 0, S
        import torch.distributed as dist

        return dist.get_rank(self.cp_group) * S, dist.get_world_size(self.cp_group) * S

    def rope(self, S: int, device: torch.device) -> Optional[Tuple[torch.Tensor, ...]]:
        rot = int(self.cfg.head_dim * self.cfg.rotary_pct) // 2 * 2
        if rot == 0:
            return None
        key = (S, str(device))
        if key not in self._rope:
            self._rope[key] = _rotary_tables(S, rot, self.cfg.rotary_base, device)
        return self._rope[key]

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
        """Returns (logits, loss).  With ``targets`` the logits are consumed by the fused
        cross-entropy and not returned (saves a [B*S, V] round trip)."""
        B, S = idx.shape
        first, total = self._cp_offset(S)
        if total > self.cfg.block_size:
            raise ValueError(f"sequence length {total} exceeds block size {self.cfg.block_size}")
        x = self.wte(idx)
        if self.wpe is not None:
            x = x + self.wpe(torch.arange(first, first + S, device=idx.device))[None]
        x = x.to(torch.bfloat16) if (idx.is_cuda and self.native) else x
        rope = self.rope(total, idx.device)
        if rope is not None and total != S:
            rope = tuple(t[first:first + S] for t in rope)
        for blk in self.blocks:
            x = blk(x, rope)
        x = self.ln_f(x)
        logits = self.lm_head(x)
        if targets is None:
            return logits, None
        if self.native:
            loss = xent_ops.cross_entropy(logits.view(B * S, -1), targets.reshape(B * S), inplace_backward=True)
        else:
            loss = F.cross_entropy(logits.view(B * S, -1).float(), targets.reshape(B * S))
        return None, loss


def build(name: str = "gpt2", native: bool = True, **overrides: Any) -> GPT:
    return GPT(config_for(name, **overrides), native=native)
