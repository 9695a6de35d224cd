"""BASELINE config #4 through the public API: GPT-NeoX ``DeepSpeedTrial`` with ZeRO stage 3, bf16, on N GPUs
(``bench.py --config neox_zero3``).

Ours: ``determined_b200.pytorch.deepspeed.init`` + ``Trainer.fit`` around a ``DeepSpeedTrial`` whose engine is the native
ZeRO engine (weights of the linear layers partitioned, gathered inside the tcgen05 GEMMs that use them, weight gradients
reduce-scattered by the wgrad GEMM's epilogue; the rest reduce-scattered from gradient hooks during backward and updated by
the fused reduce-scatter + AdamW + multicast parameter push kernel).
Reference arm: DeepSpeed cannot be installed offline, so the reference's ``DeepSpeedTrial`` (which imports ``deepspeed``)
cannot run here.  What does run is the unmodified reference harness's own ``PyTorchTrial`` path (``pytorch.init`` +
``Trainer.fit``, torch DDP over NCCL) around the same network built from stock torch modules, with
``torch.optim.AdamW(fused=True)``, bf16 autocast and gradient clipping.  The JSON line says so.

A step is one ``train_batch`` (one micro batch of ``batch`` sequences of ``seq`` tokens per GPU, gradient accumulation 1,
optimizer step included).  Every step's tokens are copied from pinned host memory and every step's loss is copied back
into pinned host memory, inside the timed region; the region is bracketed by barrier + synchronize, timed with CUDA
events, max over ranks.  Token batches come from a pool of 16 distinct batches; the step's working set (weights,
activations: tens of GB) is far larger than the 126 MB L2.
"""
import os
import sys
import tempfile
from typing import Any, Dict, Iterator

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

POOL = 16


def ds_config(batch: int, stage: int) -> Dict[str, Any]:
    return {"train_micro_batch_size_per_gpu": batch, "gradient_accumulation_steps": 1, "gradient_clipping": 1.0,
            "bf16": {"enabled": True},
            "optimizer": {"type": "AdamW", "params": {"lr": 1e-4, "betas": [0.9, 0.95], "eps": 1e-8, "weight_decay": 0.1}},
            "zero_optimization": {"stage": stage, "reduce_bucket_size": 5e8, "stage3_param_persistence_threshold": 1 << 20}}


class TokenPool:
    def __init__(self, vocab: int, batch: int, seq: int, seed: int) -> None:
        import torch

        g = torch.Generator().manual_seed(seed)
        self.batches = [torch.randint(0, vocab, (batch, seq + 1), generator=g).pin_memory() for _ in range(POOL)]

    def __len__(self) -> int:
        return len(self.batches)

    def __iter__(self) -> Iterator[Any]:
        while True:
            for b in self.batches:
                yield b


def run_ours(args, bracket, model_name: str, batch: int, seq: int, vocab: int, stage: int):
    import torch

    from determined_b200 import pytorch
    from determined_b200.models import gpt
    from determined_b200.ops import _lib
    from determined_b200.pytorch import deepspeed as det_ds

    info: Dict[str, Any] = {}
    loss_ring = torch.zeros(4096, dtype=torch.float32).pin_memory()

    class NeoXTrial(det_ds.DeepSpeedTrial):
        def __init__(self, context: Any) -> None:
            self.context = context
            context.disable_dataset_reproducibility_checks()          # the loader is a pinned in-memory token pool
            torch.manual_seed(0)
            model = gpt.build(model_name, block_size=seq, vocab_size=vocab)
            info["flops_per_token"] = model.flops_per_token(seq)
            info["params"] = sum(p.numel() for p in model.parameters())
            engine, _, _, _ = det_ds.initialize(model=model, config=ds_config(batch, stage))
            self.engine = context.wrap_model_engine(engine)
            info["engine"] = engine

        def build_training_data_loader(self) -> Any:
            return TokenPool(vocab, batch, seq, seed=1234 + self.context.distributed.rank)

        def build_validation_data_loader(self) -> Any:
            return _Len(TokenPool(vocab, batch, seq, seed=99), 2)

        def train_batch(self, it: Iterator[Any], epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
            if batch_idx == bracket.first:
                info["launch0"] = _lib.LAUNCHES[0]
                torch.cuda.reset_peak_memory_stats()
                bracket.begin()
            tokens = self.context.to_device(next(it))                  # H2D of this step's inputs from pinned memory
            _, loss = self.engine(tokens[:, :-1], tokens[:, 1:])
            self.engine.backward(loss)
            self.engine.step()
            loss_ring[batch_idx % 4096: batch_idx % 4096 + 1].copy_(loss.detach().reshape(1).float(), non_blocking=True)   # D2H of the result
            if batch_idx == bracket.last:
                bracket.end()
                info["launch1"] = _lib.LAUNCHES[0]
                info["peak_mem_gb"] = torch.cuda.max_memory_allocated() / 2**30
            return {"lm_loss": loss}

        def evaluate_batch(self, it: Iterator[Any], batch_idx: int) -> Dict[str, Any]:
            tokens = self.context.to_device(next(it))
            _, loss = self.engine(tokens[:, :-1], tokens[:, 1:])
            return {"lm_loss": loss}

        def save(self, context: Any, path: Any) -> None:
            # the controller's exit checkpoint comes after the timed region; a benchmark run is never resumed, so it keeps
            # only a marker instead of several GB of weights and optimizer shards
            (path / "benchmark-run-not-resumable").write_text("")

    total = args.warmup + args.steps
    ckdir = tempfile.mkdtemp(prefix="bench-ckpt-")
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp(prefix="bench-cwd-"))
    try:
        with det_ds.init(hparams={}, exp_conf={"checkpoint_storage": {"type": "shared_fs", "host_path": ckdir}}) as ctx:
            trial = NeoXTrial(ctx)
            det_ds.Trainer(trial, ctx).fit(max_length=pytorch.Batch(total), reporting_period=pytorch.Batch(total),
                                           checkpoint_policy="none")
            ms = bracket.max_over_ranks_ms()
            eng = info["engine"]
            info["losses"] = loss_ring[:total].tolist()
            info["partitioned_weights"] = eng.partitioned
            info["transport"] = "nvls" if getattr(eng, "mc_param", None) else "p2p"
            info["overlap"] = bool(getattr(eng, "_overlap", False))
            info["grad_partitioned"] = bool(getattr(eng, "_grad_partitioned", False))
            info["zero_kernel_launches"] = info.get("launch1", 0) - info.get("launch0", 0)
    finally:
        os.chdir(cwd)
    info.pop("engine", None)
    return ms, info


def run_reference(args, bracket, model_name: str, batch: int, seq: int, vocab: int):
    """Unmodified reference harness (baseline/_ref): PyTorchTrial + torch DDP/NCCL + AdamW(fused) on the stock-torch twin of
    the network.  Returns (ms, info) or (None, reason)."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "determined")):
        return None, "reference not installed under baseline/_ref (see DESIGN.md)"
    sys.path.insert(0, os.path.join(ROOT, "baseline", "stubs"))
    sys.path.insert(0, ref)
    import torch
    import torch.distributed as dist

    rank, local, world = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        os.environ["USE_TORCH_DISTRIBUTED"] = "1"
    try:
        from determined import core, pytorch
    except Exception as e:  # pragma: no cover
        return None, f"reference import failed: {type(e).__name__}: {e}"
    from determined_b200.models import gpt          # only the stock-torch definition of the network (native=False)

    distributed = None
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group(backend="nccl")
        distributed = core.DistributedContext.from_torch_distributed()
    info: Dict[str, Any] = {}
    loss_ring = torch.zeros(4096, dtype=torch.float32).pin_memory()

    class NeoXRefTrial(pytorch.PyTorchTrial):  # type: ignore[misc]
        def __init__(self, context: Any) -> None:
            self.context = context
            context.experimental.disable_dataset_reproducibility_checks()
            torch.manual_seed(0)
            self.model = context.wrap_model(gpt.build(model_name, native=False, block_size=seq, vocab_size=vocab))
            self.opt = context.wrap_optimizer(torch.optim.AdamW(self.model.parameters(), lr=1e-4, weight_decay=0.1, betas=(0.9, 0.95),
                                                                fused=True))

        def build_training_data_loader(self) -> Any:
            return _Len(TokenPool(vocab, batch, seq, seed=1234 + self.context.distributed.rank))

        def build_validation_data_loader(self) -> Any:
            src = TokenPool(vocab, batch, seq, seed=99)
            return torch.utils.data.DataLoader(src.batches[:2], batch_size=None, shuffle=False)

        def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, Any]:
            batch_ = batch
            if batch_idx == bracket.first:
                torch.cuda.reset_peak_memory_stats()
                bracket.begin()
            elif batch_idx == bracket.last + 1:
                bracket.end()
                info["peak_mem_gb"] = torch.cuda.max_memory_allocated() / 2**30
            tokens = batch_
            with torch.autocast("cuda", dtype=torch.bfloat16):
                _, loss = self.model(tokens[:, :-1], tokens[:, 1:])
            self.context.backward(loss)
            self.context.step_optimizer(self.opt, clip_grads=lambda ps: torch.nn.utils.clip_grad_norm_(ps, 1.0))
            return {"lm_loss": loss}

        def evaluate_batch(self, batch: Any, batch_idx: int) -> Dict[str, Any]:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                _, loss = self.model(batch[:, :-1], batch[:, 1:])
            return {"lm_loss": loss}

    total = args.warmup + args.steps + 1
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp(prefix="bench-ref-cwd-"))
    try:
        with pytorch.init(hparams={"global_batch_size": batch * world}, distributed=distributed,
                          enable_tensorboard_logging=False) as ctx:
            trial = NeoXRefTrial(ctx)
            pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(total), reporting_period=pytorch.Batch(total),
                                            checkpoint_policy="none")
        ms = bracket.max_over_ranks_ms()
    finally:
        os.chdir(cwd)
    del loss_ring
    return ms, info


class _Len:
    """A loader-like view of the pool: ``n`` batches per pass (None: the endless training stream)."""

    def __init__(self, src: TokenPool, n: Any = None) -> None:
        self.src, self.n = src, n

    def __len__(self) -> int:
        return self.n if self.n is not None else len(self.src)

    def __iter__(self) -> Iterator:
        return iter(self.src) if self.n is None else iter(self.src.batches[: self.n])
