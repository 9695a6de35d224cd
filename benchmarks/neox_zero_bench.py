"""GPT-NeoX data-parallel training step over N GPUs: native model + ZeRO engine (reduce-scatter + AdamW + parameter push in
one kernel per bucket) vs the same network on stock torch modules under DDP + AdamW(fused=True).  BASELINE config #4 shape
(synthetic tokens, random init).  Device-timed, max over ranks.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 benchmarks/neox_zero_bench.py --model neox-1.3b
"""
import argparse
import json
import os
import pathlib
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from determined_b200.models import gpt  # noqa: E402
from determined_b200.parallel import zero  # noqa: E402


def timed(step, steps, warmup, dev):
    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(steps):
        step(i)
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="neox-1.3b")
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--stage", type=int, default=1)
    ap.add_argument("--vocab", type=int, default=50304, help="51200 also lets the LM head split over 8 ranks under stage 3")
    ap.add_argument("--json")
    a = ap.parse_args()
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl")
    out = {"model": a.model, "world": world, "micro_batch": a.batch, "seq": a.seq, "zero_stage": a.stage}
    pool = [torch.randint(0, a.vocab, (a.batch, a.seq + 1), device=dev) for _ in range(4)]
    # ---- native: our model + ZeRO engine
    torch.manual_seed(0)
    model = gpt.build(a.model, block_size=a.seq, vocab_size=a.vocab)
    flops_tok = model.flops_per_token(a.seq)            # before stage 3 cuts the weights into shards
    cfg = {"train_micro_batch_size_per_gpu": a.batch, "gradient_accumulation_steps": 1, "bf16": {"enabled": True}, "gradient_clipping": 1.0,
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-4, "weight_decay": 0.1, "betas": [0.9, 0.95]}}, "zero_optimization": {"stage": a.stage}}
    engine, _, _, _ = zero.initialize(model=model, config=cfg)

    def native_step(i):
        idx = pool[i % 4]
        _, loss = engine(idx[:, :-1], idx[:, 1:])
        engine.backward(loss)
        engine.step()

    torch.cuda.reset_peak_memory_stats()
    ms = timed(native_step, a.steps, a.warmup, dev)
    tok = world * a.batch * a.seq / ms * 1e3
    out["native"] = {"ms_per_step": ms, "tokens_per_s": tok, "model_tflops_per_gpu": tok / world * flops_tok / 1e12,
                     "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30, "partitioned_weights": engine.partitioned,
                     "optimizer_kernel_launches_total": engine.kernel_launches}
    del engine, model, native_step
    torch.cuda.empty_cache()
    # ---- library arm: stock torch modules, DDP, fused AdamW, bf16 autocast
    torch.manual_seed(0)
    ref = gpt.build(a.model, native=False, block_size=a.seq, vocab_size=a.vocab).to(dev)
    arm = "torch_ddp"
    ddp = None
    if a.stage >= 3:
        # the library equivalent of ZeRO-3: FSDP full sharding per transformer block, bf16 compute, NCCL all-gather / reduce-scatter
        try:
            from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
            from torch.distributed.fsdp import MixedPrecision
            from torch.distributed.fsdp.wrap import ModuleWrapPolicy

            ddp = FSDP(ref, auto_wrap_policy=ModuleWrapPolicy({gpt.Block}), device_id=local, use_orig_params=True,
                       mixed_precision=MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.float32))
            arm = "torch_fsdp"
        except Exception as e:  # noqa: BLE001
            out["fsdp_error"] = repr(e)[:300]
            ddp = None
    if ddp is None:
        ddp = torch.nn.parallel.DistributedDataParallel(ref, device_ids=[local], gradient_as_bucket_view=True)
    opt = torch.optim.AdamW(ddp.parameters(), lr=1e-4, weight_decay=0.1, betas=(0.9, 0.95), fused=True)

    def torch_step(i):
        idx = pool[i % 4]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            _, loss = ddp(idx[:, :-1], idx[:, 1:])
        loss.backward()
        if arm == "torch_fsdp":
            ddp.clip_grad_norm_(1.0)
        else:
            torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)

    torch.cuda.reset_peak_memory_stats()
    ms2 = timed(torch_step, a.steps, a.warmup, dev)
    tok2 = world * a.batch * a.seq / ms2 * 1e3
    out[arm] = {"ms_per_step": ms2, "tokens_per_s": tok2, "peak_mem_gb": torch.cuda.max_memory_allocated() / 2**30}
    out["speedup"] = ms2 / ms
    if rank == 0:
        print("NEOX_ZERO_BENCH", json.dumps(out), flush=True)
        if a.json:
            pathlib.Path(a.json).write_text(json.dumps(out, indent=1))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
