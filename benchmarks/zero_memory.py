"""Model-state memory of the ZeRO engine per stage against the model DSAT uses to pick stages (reference
``harness/determined/pytorch/dsat/_dsat_search_method.py:383-430``; bf16 counts as "not fp16" there: 4 B params, 4 B
gradients, 8 B Adam moments per parameter; stage 1 divides the moments by the world size, stage 2 also the gradients,
stage 3 everything).  Measured as the growth of *device* memory in use (``cudaMemGetInfo``, so symmetric buffers that do
not come from torch's allocator are included) from before the engine is built to after two training steps, minus the
activation / workspace memory that remains cached (reported separately from torch's allocator statistics).

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 benchmarks/zero_memory.py --json out.json
"""
import argparse
import gc
import json
import os
import pathlib
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from determined_b200.models import gpt  # noqa: E402
from determined_b200.parallel import zero  # noqa: E402


def used_bytes() -> int:
    free, total = torch.cuda.mem_get_info()
    return total - free


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=12)
    ap.add_argument("--embd", type=int, default=2048)
    ap.add_argument("--vocab", type=int, default=2048)
    ap.add_argument("--seq", type=int, default=256)
    ap.add_argument("--json")
    a = ap.parse_args()
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl")
    dist.barrier()
    rows = []
    for stage in (1, 2, 3):
        gc.collect()
        torch.cuda.empty_cache()
        torch.cuda.synchronize()
        dist.barrier()
        base = used_bytes()
        torch.manual_seed(0)
        model = gpt.build("tiny", native=True, n_embd=a.embd, n_head=a.embd // 128, n_layer=a.layers, vocab_size=a.vocab, block_size=a.seq,
                          rotary_pct=0.25, parallel_residual=True, tie_embeddings=False)
        P = sum(p.numel() for p in model.parameters())
        cfg = {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1, "gradient_clipping": 1.0, "bf16": {"enabled": True},
               "optimizer": {"type": "AdamW", "params": {"lr": 1e-4}},
               "zero_optimization": {"stage": stage, "reduce_bucket_size": 3.2e7, "stage3_param_persistence_threshold": 1 << 16}}
        eng, _, _, _ = zero.initialize(model=model, config=cfg)
        g = torch.Generator().manual_seed(rank)
        for _ in range(2):
            idx = torch.randint(0, a.vocab, (2, a.seq), generator=g).to(dev)
            _, loss = eng(idx, idx)
            eng.backward(loss)
            eng.step()
        torch.cuda.synchronize()
        del loss, idx
        gc.collect()
        torch.cuda.empty_cache()            # drops cached activation blocks; parameters / states / symmetric buffers stay
        torch.cuda.synchronize()
        state = used_bytes() - base
        model_pred = {1: 4 * P + 4 * P + 8 * P // world, 2: 4 * P + 12 * P // world, 3: 16 * P // world}[stage]
        rows.append({"stage": stage, "params": P, "partitioned_weights": eng.partitioned, "measured_state_bytes": state,
                     "dsat_model_bytes": model_pred, "measured_over_model": state / model_pred,
                     "bytes_per_param": state / P})
        for h in getattr(eng, "_hooks", []):
            h.remove()
        del eng, model
    if rank == 0:
        out = {"world": world, "rows": rows,
               "note": "measured = device memory in use after training minus before building the engine, caches emptied; the "
                       "model counts 4 B params + 4 B grads + 8 B moments per parameter divided as the stage says. Ours adds a "
                       "4 B/W gradient shard (clipping needs the global norm before the update), the landing ring of two buckets "
                       "in stages 2-3, bf16 gather buffers of the partitioned layers in stage 3, and keeps embeddings / norms / biases "
                       "replicated in stage 3"}
        print("ZERO_MEMORY", json.dumps(out), flush=True)
        if a.json:
            pathlib.Path(a.json).parent.mkdir(parents=True, exist_ok=True)
            pathlib.Path(a.json).write_text(json.dumps(out, indent=1))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
