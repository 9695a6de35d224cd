"""Pipeline-parallel engine (1F1B schedule) with the DeepSpeed ``PipelineEngine`` surface.

What the reference needs from a pipeline engine (``pytorch/deepspeed/_deepspeed_context.py:187-193``,
``_deepspeed_trial.py:364-499``, ``_mpu.py:38-50``): ``train_batch(data_iter)`` / ``eval_batch(data_iter)``
consuming ``gradient_accumulation_steps`` micro batches per call, ``mpu`` with pipe/data ranks (data
loaders only on the first and last stage, metrics from the last), micro-batch / GAS getters and
per-rank checkpoints.  The reference gets all of that from DeepSpeed; this is a native equivalent:

* layers are split into contiguous stages (balanced by parameter count), rank = pipe * dp + data;
* one-forward-one-backward schedule: stage s runs ``min(P-s-1, M)`` warm-up forwards, then alternates,
  then drains; activations and activation-gradients travel with ``torch.distributed`` p2p (NCCL over
  NVLink on GPUs, gloo on CPU); sends are non-blocking so neighbouring stages never deadlock;
* gradients are averaged over micro batches and all-reduced across the stage's data-parallel group;
* the mean loss is broadcast from the last stage so every rank reports the same metric.
"""
import logging
import os
import pathlib
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Tuple, Union

import torch
import torch.distributed as dist

logger = logging.getLogger("determined.parallel")

_DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.int64, torch.int32]


class PipeGrid:
    """Topology helper with the accessor names DeepSpeed's ``PipelineParallelGrid`` exposes."""

    def __init__(self, rank: int, world: int, num_stages: int) -> None:
        if world % num_stages != 0:
            raise ValueError(f"world size {world} is not divisible by the number of pipeline stages {num_stages}")
        self.rank, self.world, self.pp, self.dp = rank, world, num_stages, world // num_stages
        self.stage_id, self.data_id = rank // self.dp, rank % self.dp

    def get_pipe_parallel_rank(self) -> int:
        return self.stage_id

    def get_pipe_parallel_world_size(self) -> int:
        return self.pp

    def get_data_parallel_rank(self) -> int:
        return self.data_id

    def get_data_parallel_world_size(self) -> int:
        return self.dp

    def get_model_parallel_rank(self) -> int:
        return 0

    def get_model_parallel_world_size(self) -> int:
        return 1

    def stage_to_global(self, stage_id: int) -> int:
        return stage_id * self.dp + self.data_id


def partition_balanced(weights: Sequence[int], parts: int) -> List[int]:
    """Boundaries ``b[0]=0 <= ... <= b[parts]=n`` of contiguous chunks minimising the heaviest chunk."""
    n = len(weights)
    if parts > n:
        raise ValueError(f"cannot split {n} layers into {parts} stages")
    prefix = [0]
    for w in weights:
        prefix.append(prefix[-1] + max(int(w), 1))

    def feasible(cap: int) -> Optional[List[int]]:
        bounds, start = [0], 0
        for _ in range(parts):
            end = start
            while end < n and prefix[end + 1] - prefix[start] <= cap:
                end += 1
            if end == start:
                return None
            # leave at least one layer for each remaining stage
            remaining = parts - len(bounds)
            end = min(end, n - remaining)
            if end <= start:
                return None
            bounds.append(end)
            start = end
        return bounds if bounds[-1] == n else None

    lo, hi = max(prefix[i + 1] - prefix[i] for i in range(n)), prefix[-1]
    best = feasible(hi)
    while lo < hi:
        mid = (lo + hi) // 2
        b = feasible(mid)
        if b is not None:
            best, hi = b, mid
        else:
            lo = mid + 1
    assert best is not None
    return best


class PipelineEngine:
    is_pipe_parallel = True

    def __init__(self, layers: Sequence[torch.nn.Module], num_stages: int, loss_fn: Callable[[torch.Tensor, Any], torch.Tensor],
                 config: Dict[str, Any], optimizer_factory: Optional[Callable[[Any], torch.optim.Optimizer]] = None,
                 device: Optional[torch.device] = None, partition: str = "parameters") -> None:
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.grid = self.mpu = PipeGrid(self.rank, self.world, num_stages)
        self.num_stages, self.stage_id = num_stages, self.grid.stage_id
        self.device = device or torch.device("cpu")
        self._micro_bs = int(config.get("train_micro_batch_size_per_gpu", 1))
        self._gas = int(config.get("gradient_accumulation_steps", 1))
        self.config = config
        weights = [sum(p.numel() for p in l.parameters()) for l in layers] if partition == "parameters" else [1] * len(layers)
        self.bounds = partition_balanced(weights, num_stages)
        lo, hi = self.bounds[self.stage_id], self.bounds[self.stage_id + 1]
        self.layer_range = (lo, hi)
        self.module = torch.nn.Sequential(*layers[lo:hi]).to(self.device)
        self.loss_fn = loss_fn
        params = [p for p in self.module.parameters() if p.requires_grad]
        if optimizer_factory is not None:
            self.optimizer = optimizer_factory(params)
        else:
            oc = (config.get("optimizer") or {"type": "AdamW", "params": {}})
            kind, kw = oc.get("type", "AdamW").lower(), dict(oc.get("params") or {})
            if "betas" in kw:
                kw["betas"] = tuple(kw["betas"])
            cls = {"adamw": torch.optim.AdamW, "adam": torch.optim.Adam, "sgd": torch.optim.SGD}[kind]
            self.optimizer = cls(params, **kw) if params else None
        self.clip = float(config.get("gradient_clipping", 0.0) or 0.0)
        self.global_steps = 0
        self.micro_steps = 0
        # groups: data-parallel peers of this stage; every rank must create every group
        self.dp_group = None
        if self.world > 1 and self.grid.dp > 1:
            for s in range(num_stages):
                g = dist.new_group([s * self.grid.dp + d for d in range(self.grid.dp)])
                if s == self.stage_id:
                    self.dp_group = g
        self._recv_meta: Dict[str, Tuple[torch.Size, torch.dtype]] = {}
        self._sent_meta: set = set()
        self._pending: List[Any] = []

    # -- DeepSpeed-compatible getters -----------------------------------------------------------------
    def train_micro_batch_size_per_gpu(self) -> int:
        return self._micro_bs

    def gradient_accumulation_steps(self) -> int:
        return self._gas

    def train_batch_size(self) -> int:
        return self._micro_bs * self._gas * self.grid.dp

    def is_first_stage(self) -> bool:
        return self.stage_id == 0

    def is_last_stage(self) -> bool:
        return self.stage_id == self.num_stages - 1

    def is_gradient_accumulation_boundary(self) -> bool:
        return True

    def to(self, device: Any) -> "PipelineEngine":
        self.device = torch.device(device)
        self.module.to(self.device)
        return self

    def train(self, mode: bool = True) -> "PipelineEngine":
        self.module.train(mode)
        return self

    def eval(self) -> "PipelineEngine":
        return self.train(False)

    def parameters(self) -> Any:
        return self.module.parameters()

    # -- p2p -------------------------------------------------------------------------------------------
    def _send(self, t: torch.Tensor, dst_stage: int, tag: str) -> None:
        dst = self.grid.stage_to_global(dst_stage)
        t = t.contiguous()
        if tag not in self._sent_meta:
            meta = torch.zeros(10, dtype=torch.int64, device=self.device)
            meta[0], meta[1] = t.dim(), _DTYPES.index(t.dtype)
            for i, s in enumerate(t.shape):
                meta[2 + i] = s
            dist.send(meta, dst)
            self._sent_meta.add(tag)
        self._pending.append((dist.isend(t, dst), t))       # keep the tensor alive until the send completes

    def _recv(self, src_stage: int, tag: str) -> torch.Tensor:
        src = self.grid.stage_to_global(src_stage)
        if tag not in self._recv_meta:
            meta = torch.zeros(10, dtype=torch.int64, device=self.device)
            dist.recv(meta, src)
            m = meta.tolist()
            self._recv_meta[tag] = (torch.Size(m[2:2 + m[0]]), _DTYPES[m[1]])
        shape, dtype = self._recv_meta[tag]
        buf = torch.empty(shape, dtype=dtype, device=self.device)
        dist.recv(buf, src)
        return buf

    def _drain_sends(self) -> None:
        for h, _ in self._pending:
            h.wait()
        self._pending.clear()

    # -- schedule -------------------------------------------------------------------------------------
    def _forward_micro(self, data_iter: Optional[Iterator[Any]], train: bool, store: List[Any]) -> Optional[torch.Tensor]:
        labels = None
        if self.is_first_stage() or self.is_last_stage():
            batch = next(data_iter)  # type: ignore[arg-type]
            inputs, labels = (batch[0], batch[1]) if isinstance(batch, (tuple, list)) else (batch, None)
            inputs = inputs.to(self.device)
            labels = labels.to(self.device) if isinstance(labels, torch.Tensor) else labels
        if self.is_first_stage():
            x = inputs
        else:
            x = self._recv(self.stage_id - 1, "act")
            if train and x.is_floating_point():
                x.requires_grad_(True)
        y = self.module(x)
        loss = None
        if self.is_last_stage():
            loss = self.loss_fn(y, labels)
            out = loss
        else:
            self._send(y.detach(), self.stage_id + 1, "act")
            out = y
        if train:
            store.append((x, out))
        return loss.detach() if loss is not None else None

    def _backward_micro(self, store: List[Any]) -> None:
        x, out = store.pop(0)
        if self.is_last_stage():
            (out / self._gas).backward()
        else:
            g = self._recv(self.stage_id + 1, "grad")
            out.backward(g)
        if not self.is_first_stage() and x.grad is not None:
            self._send(x.grad, self.stage_id - 1, "grad")

    def train_batch(self, data_iter: Optional[Iterator[Any]] = None) -> torch.Tensor:
        """One optimizer step over ``gradient_accumulation_steps`` micro batches; returns the mean loss."""
        self.module.train()
        M, P, s = self._gas, self.num_stages, self.stage_id
        store: List[Any] = []
        losses: List[torch.Tensor] = []
        warm = min(P - s - 1, M)
        for _ in range(warm):
            l = self._forward_micro(data_iter, True, store)
            if l is not None:
                losses.append(l)
        for _ in range(M - warm):
            l = self._forward_micro(data_iter, True, store)
            if l is not None:
                losses.append(l)
            self._backward_micro(store)
        for _ in range(warm):
            self._backward_micro(store)
        self._drain_sends()
        self.micro_steps += M
        params = [p for p in self.module.parameters() if p.grad is not None]
        if self.dp_group is not None:
            for p in params:
                dist.all_reduce(p.grad, group=self.dp_group)
                p.grad.div_(self.grid.dp)
        if self.clip > 0:
            # global norm spans all stages: combine the per-stage squared norms
            sq = torch.zeros(1, device=self.device)
            for p in params:
                sq += p.grad.float().pow(2).sum()
            if self.world > 1:
                dist.all_reduce(sq)
                sq /= self.grid.dp
            coef = min(1.0, self.clip / (float(sq.sqrt()) + 1e-6))
            if coef < 1.0:
                for p in params:
                    p.grad.mul_(coef)
        if self.optimizer is not None:
            self.optimizer.step()
            self.optimizer.zero_grad(set_to_none=True)
        self.global_steps += 1
        return self._share_loss(losses)

    def eval_batch(self, data_iter: Optional[Iterator[Any]] = None, num_micro_batches: Optional[int] = None) -> torch.Tensor:
        self.module.eval()
        losses: List[torch.Tensor] = []
        with torch.no_grad():
            for _ in range(num_micro_batches or self._gas):
                l = self._forward_micro(data_iter, False, [])
                if l is not None:
                    losses.append(l)
        self._drain_sends()
        return self._share_loss(losses)

    def _share_loss(self, losses: List[torch.Tensor]) -> torch.Tensor:
        mean = torch.stack(losses).float().mean().reshape(1) if losses else torch.zeros(1, device=self.device)
        if self.world > 1:
            if self.dp_group is not None and self.is_last_stage():
                dist.all_reduce(mean, group=self.dp_group)
                mean /= self.grid.dp
            dist.broadcast(mean, src=(self.num_stages - 1) * self.grid.dp)     # first data-parallel replica of the last stage
        return mean[0]

    # -- checkpoints (per-stage layer files, DeepSpeed naming) -----------------------------------------------
    def save_checkpoint(self, save_dir: Union[str, os.PathLike], tag: Optional[str] = None, client_state: Optional[Dict[str, Any]] = None) -> None:
        tag = tag or f"global_step{self.global_steps}"
        d = pathlib.Path(save_dir) / tag
        d.mkdir(parents=True, exist_ok=True)
        if self.grid.data_id == 0:
            for i, layer in enumerate(self.module):
                torch.save(layer.state_dict(), d / f"layer_{self.layer_range[0] + i:02d}-model_states.pt")
            torch.save({"optimizer": self.optimizer.state_dict() if self.optimizer is not None else None, "global_steps": self.global_steps,
                        "micro_steps": self.micro_steps, "client_state": client_state or {}}, d / f"mp_rank_{self.stage_id:02d}_model_states.pt")
        if self.rank == 0:
            (pathlib.Path(save_dir) / "latest").write_text(tag)
        if self.world > 1:
            dist.barrier()          # the checkpoint is complete (all stages + `latest`) when any rank returns

    def load_checkpoint(self, load_dir: Union[str, os.PathLike], tag: Optional[str] = None, **_: Any) -> Tuple[Optional[str], Dict[str, Any]]:
        base = pathlib.Path(load_dir)
        if tag is None:
            if not (base / "latest").exists():
                return None, {}
            tag = (base / "latest").read_text().strip()
        d = base / tag
        for i, layer in enumerate(self.module):
            layer.load_state_dict(torch.load(d / f"layer_{self.layer_range[0] + i:02d}-model_states.pt", map_location=self.device))
        st = torch.load(d / f"mp_rank_{self.stage_id:02d}_model_states.pt", map_location=self.device, weights_only=False)
        if self.optimizer is not None and st.get("optimizer") is not None:
            self.optimizer.load_state_dict(st["optimizer"])
        self.global_steps, self.micro_steps = int(st["global_steps"]), int(st["micro_steps"])
        return str(d), st.get("client_state", {})
