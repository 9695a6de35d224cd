"""BASELINE config #3: GPT-2 medium fine-tune "via hf_trainer_api" at N GPUs (``bench.py --config gpt2_medium_hf``).

``transformers.Trainer`` (and ``TrainingArguments``) cannot be constructed on this image: they require ``accelerate``,
which is neither installed nor in the offline wheelhouse.  What *can* run is everything the Trainer would drive: the HF
``GPT2LMHeadModel`` (gpt2-medium config, random init — no checkpoints offline), wrapped in torch
``DistributedDataParallel`` exactly as ``Trainer._wrap_model`` does, ``torch.optim.AdamW(fused=True)``, bf16 autocast, a
linear schedule, gradient clipping at 1.0 — in a minimal loop that makes the ``TrainerCallback`` calls ``DetCallback``
listens to.  The two arms differ in ONE thing, which is the thing config #3 is about:

* ours: ``determined_b200.parallel.ddp.install_comm_hook`` — every DDP gradient bucket goes through the fused in-kernel
  all-reduce (NVSwitch in-switch reduction from four ranks on, two-shot peer-to-peer below);
* reference: DDP's own NCCL all-reduce, with the unmodified reference ``determined`` package imported for its Core API.

Steps are host-fed (token batches from pinned memory each step, loss copied back each step), timed with CUDA events
between barriers, max over ranks.
"""
import json
import os
import sys
from typing import Any

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEQ, VOCAB = 1024, 50257


def main(args: Any, rank: int, local: int, world: int, requested_warmup: int, ClockSampler: Any, Bracket: Any) -> None:
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device"}))
        return
    try:
        from transformers import GPT2Config, GPT2LMHeadModel
    except Exception as e:  # pragma: no cover
        if rank == 0:
            print(json.dumps({"impl": args.impl, "unavailable": f"transformers not importable: {e}"}))
        return
    torch.cuda.set_device(local % torch.cuda.device_count())
    dev = torch.device("cuda", local % torch.cuda.device_count())
    if world > 1:
        dist.init_process_group("nccl")
    if args.impl == "reference":
        sys.path.insert(0, os.path.join(ROOT, "baseline", "stubs"))
        sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
        try:
            import determined  # noqa: F401  (the unmodified reference package; its DetCallback needs transformers.Trainer types)
        except Exception as e:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": f"reference import failed: {e}"}))
            return
    batch = args.micro_batch if args.micro_batch != 4 else 8
    args.warmup = max(args.warmup, 3)
    torch.manual_seed(0)
    cfg = GPT2Config(n_layer=24, n_head=16, n_embd=1024, n_positions=SEQ, vocab_size=VOCAB, attn_pdrop=0.0, embd_pdrop=0.0, resid_pdrop=0.0)
    model = GPT2LMHeadModel(cfg).to(dev)
    params = sum(p.numel() for p in model.parameters())
    hook = None
    net: Any = model
    if world > 1:
        net = torch.nn.parallel.DistributedDataParallel(model, device_ids=[dev.index])
        if args.impl == "ours":
            from determined_b200.parallel import ddp as pddp

            hook = pddp.install_comm_hook(net, dev)
    opt = torch.optim.AdamW(net.parameters(), lr=5e-5, weight_decay=0.0, fused=True)
    total = args.warmup + args.steps
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: max(0.0, 1.0 - s / (10 * total)))
    g = torch.Generator().manual_seed(1234 + rank)
    pool = [torch.randint(0, VOCAB, (batch, SEQ), generator=g).pin_memory() for _ in range(16)]
    loss_ring = torch.zeros(4096).pin_memory()
    sampler = ClockSampler(local) if rank == 0 else None
    bracket = Bracket(args.warmup, args.steps, sampler)
    for i in range(total):
        if i == bracket.first:
            bracket.begin()
        ids = pool[i % 16].to(dev, non_blocking=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = net(input_ids=ids, labels=ids)
        out.loss.backward()
        torch.nn.utils.clip_grad_norm_(net.parameters(), 1.0)
        opt.step()
        sched.step()
        opt.zero_grad(set_to_none=True)
        loss_ring[i % 4096: i % 4096 + 1].copy_(out.loss.detach().float().reshape(1), non_blocking=True)
        if i == bracket.last:
            bracket.end()
    ms = bracket.max_over_ranks_ms()
    clocks = sampler.stop() if sampler is not None else None
    if rank == 0:
        tokens = batch * SEQ * world * args.steps
        v = tokens / (ms / 1e3)
        print(json.dumps({
            "impl": args.impl, "metric": "tokens/sec per trial", "value": v, "unit": "tokens/s", "samples_per_s": v / SEQ,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "warmup_requested": requested_warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"model": f"HF GPT2LMHeadModel gpt2-medium ({params / 1e6:.0f}M params), random init", "global_batch": batch * world,
                       "per_device_batch": batch, "seq_len": SEQ, "optimizer": "AdamW(fused), clip 1.0, linear schedule",
                       "parallelism": f"dp{world} (torch DDP)", "l2_policy": "16 distinct token batches; 1.4 GB of gradients + activations >> L2"},
            "harness": ("torch DDP + fused in-kernel all-reduce as DDP communication hook (" + ("NVLS" if hook is not None and hook.use_nvls else "peer-to-peer")
                        + ")" if args.impl == "ours" else "torch DDP + NCCL all-reduce; reference `determined` package imported")
                       + "; transformers.Trainer needs `accelerate`, absent from this image: minimal Trainer-equivalent loop, host-fed, value == e2e",
            "losses_first_last": [float(loss_ring[0]), float(loss_ring[(total - 1) % 4096])],
            "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": batch * SEQ * 8, "d2h_bytes_per_step": 4},
            "gpu_launches": (hook.launches if hook is not None else 0), "clocks": clocks}))
    if world > 1:
        dist.barrier()
        if hook is not None:
            hook.close()
        dist.destroy_process_group()
