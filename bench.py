#!/usr/bin/env python3
"""Headline benchmark: samples/sec per trial for BASELINE config #2 — CIFAR-10 ResNet-18
PyTorchTrial, bf16 autocast, ``aggregation_frequency=2``, fused gradient all-reduce — at
``slots_per_trial`` = N GPUs of one node (weak scaling: fixed per-slot batch).

    python bench.py --gpus 1 --steps 200 --warmup 20
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 200 --warmup 20
    python bench.py --impl reference ...      # the UNMODIFIED reference harness, same trial

A "step" is one ``train_batch`` call (one micro-batch per slot); with aggregation_frequency=2
every second step all-reduces and runs the optimizer, so K is rounded up to an even number.
``value`` is measured with the input pool resident on the device; ``e2e`` repeats the run through
the same public API (``pytorch.init`` + ``Trainer.fit``) with every step's inputs copied from pinned
host memory and every step's loss copied back to the host.  Both are timed on the device with CUDA
events bracketed by barrier + synchronize; the max over ranks is reported.  Data is synthetic and
weights are random-init (no network on the box).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PER_SLOT_BATCH = 256
AGG_FREQ = 2
POOL = 48            # 48 x 256 x 3x32x32 fp32 = 151 MB of distinct inputs (> 126 MB L2)
BASELINE_PUBLISHED = None   # the reference publishes no throughput number


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int) -> None:
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self) -> None:
        assert self.proc is not None and self.proc.stdout is not None
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [c.strip() for c in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def _dist_env():
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, local, world


class Bracket:
    """Barrier + synchronize on both sides of exactly K steps; device time by CUDA events."""

    def __init__(self, warmup: int, steps: int, sampler) -> None:
        import torch

        self.torch = torch
        self.first, self.last = warmup, warmup + steps - 1
        self.ev0 = torch.cuda.Event(enable_timing=True)
        self.ev1 = torch.cuda.Event(enable_timing=True)
        self.wall0 = self.wall1 = None
        self.sampler = sampler
        self.ms = None

    def _barrier(self) -> None:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            dist.barrier()

    def begin(self) -> None:
        self._barrier()
        self.torch.cuda.synchronize()
        if self.sampler is not None:
            self.sampler.start()
        self.wall0 = time.perf_counter()
        self.ev0.record()

    def end(self) -> None:
        self.ev1.record()
        self.torch.cuda.synchronize()
        self.wall1 = time.perf_counter()
        self._barrier()
        self.ms = self.ev0.elapsed_time(self.ev1)

    def max_over_ranks_ms(self) -> float:
        import torch.distributed as dist

        t = self.torch.tensor([self.ms], device="cuda", dtype=self.torch.float64)
        if dist.is_available() and dist.is_initialized():
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())


def allreduce_self_check(world: int) -> str:
    """At N > 1, before anything is timed: the in-kernel all-reduce (whichever transport the engine will pick on this box:
    NVSwitch in-switch reduction, else two-shot peer-to-peer) against ``torch.distributed.all_reduce`` on rank-seeded data,
    and the propagation of a non-finite value from one rank to every rank's found_inf flag.  Raises on mismatch."""
    import torch
    import torch.distributed as dist

    from determined_b200.ops import comm

    dev = torch.device("cuda", torch.cuda.current_device())
    group = comm.PeerGroup.from_torch_distributed(dev)
    n = 4 * world * 300_000
    buf = group.alloc(4 * n)
    peers = group.map(buf)
    mc = group.multicast(buf)
    t = buf.tensor(torch.float32, n)
    transports = [("nvls", mc)] if mc else []
    transports.append(("p2p", None))
    for name, m in transports:
        torch.manual_seed(1000 + group.rank)
        t.copy_(torch.randn(n, device=dev))
        ref = t.clone()
        dist.all_reduce(ref)
        ref.mul_(1.0 / world)
        group.begin_step()
        group.found_inf.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_fused(group, peers, 0, n, 1.0 / world, found_inf=group.found_inf, timeout_s=30, mc_ptr=m)
        torch.cuda.synchronize()
        err = float((t - ref).abs().max())
        if not err <= 1e-5 or int(group.found_inf.item()) != 0:
            raise RuntimeError(f"allreduce self-check failed ({name}): max abs error {err}, found_inf {int(group.found_inf.item())}")
        t.zero_()
        if group.rank == world - 1:
            t[3] = float("inf")
        group.begin_step()
        torch.cuda.synchronize()
        dist.barrier()
        comm.allreduce_fused(group, peers, 0, n, 1.0, found_inf=group.found_inf, timeout_s=30, mc_ptr=m)
        torch.cuda.synchronize()
        if int(group.found_inf.item()) != 1:
            raise RuntimeError(f"allreduce self-check failed ({name}): a non-finite value on one rank was not reported on rank {group.rank}")
    dist.barrier()
    group.close()
    return "ok (" + "+".join(n for n, _ in transports) + ", max abs err <= 1e-5 vs NCCL, found_inf propagates)"


def run_ours(args, on_device: bool, sampler):
    import torch

    from benchmarks import cifar_trial
    from determined_b200 import pytorch

    torch.backends.cudnn.benchmark = True
    bracket = Bracket(args.warmup, args.steps, sampler)
    counters = {}

    class BenchCallback(pytorch.PyTorchCallback):
        def on_training_batch_start(self, batch_idx: int) -> None:
            if batch_idx == 0 and args.world > 1 and not getattr(args, "_checked", None):
                args._checked = allreduce_self_check(args.world)      # fails the run before anything is timed
            if batch_idx == bracket.first:
                eng = counters["ctx"]._get_engine()
                counters["launch0"] = getattr(eng, "kernel_launches", 0)
                bracket.begin()

        def on_training_batch_end(self, batch_idx: int) -> None:
            if batch_idx == bracket.last:
                bracket.end()
                eng = counters["ctx"]._get_engine()
                counters["launch1"] = getattr(eng, "kernel_launches", 0)

    timing = {"callback": BenchCallback()}
    Trial = cifar_trial.make_trial(pytorch, per_slot_batch=PER_SLOT_BATCH, pool=POOL, on_device=on_device,
                                   use_b200_extras=not args.no_graphs, timing=timing)
    total = args.warmup + args.steps
    ckdir = tempfile.mkdtemp(prefix="bench-ckpt-")
    exp_conf = {"checkpoint_storage": {"type": "shared_fs", "host_path": ckdir},
                "optimizations": {"aggregation_frequency": AGG_FREQ, "average_aggregated_gradients": True}}
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp(prefix="bench-cwd-"))   # checkpoints copy the cwd as code/: keep it empty
    try:
        with pytorch.init(hparams={"global_batch_size": PER_SLOT_BATCH * args.world}, exp_conf=exp_conf,
                          aggregation_frequency=AGG_FREQ, enable_tensorboard_logging=False) as ctx:
            counters["ctx"] = ctx
            trial = Trial(ctx)
            pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(total), reporting_period=pytorch.Batch(total),
                                            checkpoint_policy="none")
            ms = bracket.max_over_ranks_ms()
            launches = counters.get("launch1", 0) - counters.get("launch0", 0)
            fused = type(trial.opt).__name__
            eng = ctx._get_engine()
            rec = next(iter(eng.records.values()))
            info = {"optimizer": fused, "buckets": len(rec.buckets),
                    "transport": (("nvls" if rec.transport_nvls else "p2p") + (", overlapped with backward" if rec.overlapped
                                  else ", one launch at step time")) if args.world > 1 else "none"}
            info["allreduce_check"] = getattr(args, "_checked", None)
    finally:
        os.chdir(cwd)
    return ms, launches, info


def run_reference(args, sampler):
    """Same trial through the unmodified reference harness (baseline/_ref) and its stock DDP path."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "determined")):
        return None, "reference not installed under baseline/_ref (see DESIGN.md)"
    sys.path.insert(0, os.path.join(ROOT, "baseline", "stubs"))
    sys.path.insert(0, ref)
    import torch
    import torch.distributed as dist

    rank, local, world = _dist_env()
    os.environ.setdefault("USE_TORCH_DISTRIBUTED", "1" if world > 1 else "")
    if not os.environ["USE_TORCH_DISTRIBUTED"]:
        os.environ.pop("USE_TORCH_DISTRIBUTED")
    try:
        import determined as det  # noqa: F401
        from determined import core, pytorch
    except Exception as e:  # pragma: no cover
        return None, f"reference import failed: {type(e).__name__}: {e}"
    from benchmarks import cifar_trial

    torch.backends.cudnn.benchmark = True
    distributed = None
    if world > 1:
        torch.cuda.set_device(local)
        dist.init_process_group(backend="nccl")
        distributed = core.DistributedContext.from_torch_distributed()
    bracket = Bracket(args.warmup, args.steps, sampler)

    def hook(batch_idx: int) -> None:
        # called at the top of train_batch: batch W starts the clock, batch W+K stops it, so the
        # bracket holds exactly steps W .. W+K-1 including their per-step D2H of the loss
        if batch_idx == bracket.first:
            bracket.begin()
        elif batch_idx == bracket.last + 1:
            bracket.end()

    timing = {"hook": hook}
    Trial = cifar_trial.make_trial(pytorch, per_slot_batch=PER_SLOT_BATCH, pool=POOL, on_device=False,
                                   use_b200_extras=False, timing=timing)
    total = args.warmup + args.steps + 1
    cwd = os.getcwd()
    os.chdir(tempfile.mkdtemp(prefix="bench-ref-cwd-"))
    try:
        with pytorch.init(hparams={"global_batch_size": PER_SLOT_BATCH * world}, distributed=distributed,
                          aggregation_frequency=AGG_FREQ, enable_tensorboard_logging=False) as ctx:
            trial = Trial(ctx)
            pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(total), reporting_period=pytorch.Batch(total),
                                            checkpoint_policy="none")
        ms = bracket.max_over_ranks_ms()
    finally:
        os.chdir(cwd)
    return ms, None


def main_neox(args, rank: int, local: int, world: int, requested_warmup: int) -> None:
    """BASELINE config #4: GPT-NeoX DeepSpeedTrial, ZeRO-3, bf16 (see benchmarks/bench_neox.py)."""
    import torch

    from benchmarks import bench_neox

    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device"}))
        return
    torch.cuda.set_device(local % torch.cuda.device_count())
    args.warmup = max(args.warmup, 3)
    model = args.model or "neox-1.3b"
    batch, seq = args.micro_batch, args.seq
    # the LM head only splits into 256-row shards per rank when the vocabulary is a multiple of 256*world: pad it the
    # way Megatron / gpt-neox pad theirs (make_vocab_size_divisible_by)
    vocab = 50304 if world == 1 else (50304 + 256 * world - 1) // (256 * world) * (256 * world)
    sampler = ClockSampler(local) if rank == 0 else None
    bracket = Bracket(args.warmup, args.steps, sampler)
    tokens = batch * seq * world * args.steps
    config = {"model": f"gpt-{model} (GPT-NeoX architecture: rotary, parallel residual, untied embeddings), vocab {vocab}",
              "global_batch": batch * world, "micro_batch_per_gpu": batch, "seq_len": seq, "gradient_accumulation": 1,
              "optimizer": "AdamW, gradient clipping 1.0", "parallelism": f"dp{world} ZeRO stage {args.zero_stage}",
              "l2_policy": "16 distinct token batches; weights + activations of a step (tens of GB) >> 126 MB L2"}
    if args.impl == "reference":
        ms, info = bench_neox.run_reference(args, bracket, model, batch, seq, vocab)
        if ms is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": info}))
            return
        clocks = sampler.stop() if sampler is not None else None
        if rank == 0:
            v = tokens / (ms / 1e3)
            print(json.dumps({"impl": "reference", "metric": "tokens/sec per trial", "value": v, "unit": "tokens/s",
                              "samples_per_s": v / seq, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                              "warmup_requested": requested_warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                              "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": config,
                              "harness": "reference determined 0.31.1-dev0 (baseline/_ref, unmodified) PyTorchTrial + torch DDP/NCCL + "
                                         "AdamW(fused=True) on the stock-torch twin of the network: DeepSpeed is not installable "
                                         "offline, so the reference's DeepSpeedTrial (ZeRO-3) itself cannot run on this box",
                              "peak_mem_gb": info.get("peak_mem_gb"),
                              "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": batch * (seq + 1) * 8, "d2h_bytes_per_step": 4},
                              "clocks": clocks}))
        return
    ms, info = bench_neox.run_ours(args, bracket, model, batch, seq, vocab, args.zero_stage)
    clocks = sampler.stop() if sampler is not None else None
    if rank == 0:
        v = tokens / (ms / 1e3)
        print(json.dumps({"impl": "ours", "metric": "tokens/sec per trial", "value": v, "unit": "tokens/s", "samples_per_s": v / seq,
                          "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "warmup_requested": requested_warmup,
                          "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "bf16", "data": "synthetic", "config": config,
                          "harness": f"determined_b200 DeepSpeedTrial + native ZeRO engine (stage {args.zero_stage}: {info['partitioned_weights']} "
                                     f"weight matrices partitioned, gradients {'partitioned' if info['grad_partitioned'] else 'replicated'}, "
                                     f"reduce-scatter {'overlapped with backward' if info['overlap'] else 'at step time'}, transport "
                                     f"{info['transport']}); host-fed through pytorch.deepspeed.init + Trainer.fit: value == e2e",
                          "model_tflops_per_gpu": v / world * info["flops_per_token"] / 1e12, "params": info["params"],
                          "peak_mem_gb": info.get("peak_mem_gb"), "losses_first_last": [info["losses"][0], info["losses"][-1]],
                          "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": batch * (seq + 1) * 8, "d2h_bytes_per_step": 4},
                          "gpu_launches": info.get("zero_kernel_launches"), "clocks": clocks}))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graphs", action="store_true", help="ours: disable CUDA-graph replay of train_batch")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--config", default="cifar_resnet", choices=["cifar_resnet", "neox_zero3", "gpt2_medium_hf", "asha_pack"],
                    help="BASELINE.json config: cifar_resnet = #2 (default, the headline), neox_zero3 = #4, gpt2_medium_hf = #3, "
                         "asha_pack = #5")
    ap.add_argument("--model", default=None, help="neox_zero3: neox-1.3b (default) | neox-2.7b | ...")
    ap.add_argument("--micro-batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=2048)
    ap.add_argument("--zero-stage", type=int, default=3)
    args = ap.parse_args()
    rank, local, world = _dist_env()
    args.world = world
    if args.gpus != world and rank == 0 and world == 1 and args.gpus > 1:
        # convenience: re-launch ourselves under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"),
               os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    requested_warmup = args.warmup
    if args.config == "neox_zero3":
        return main_neox(args, rank, local, world, requested_warmup)
    if args.config == "gpt2_medium_hf":
        from benchmarks import bench_hf

        return bench_hf.main(args, rank, local, world, requested_warmup, ClockSampler, Bracket)
    if args.config == "asha_pack":
        from benchmarks import bench_asha

        return bench_asha.main(args)
    args.warmup = max(args.warmup, 4)
    args.warmup += args.warmup % AGG_FREQ          # start the bracket on an aggregation boundary
    args.steps += args.steps % AGG_FREQ

    import torch

    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device"}))
        return
    torch.cuda.set_device(local % torch.cuda.device_count())
    sampler = ClockSampler(local) if rank == 0 else None
    samples = PER_SLOT_BATCH * world * args.steps
    config = {"model": "resnet18-cifar10 (11.2M params), PyTorchTrial, SGD nesterov", "global_batch": PER_SLOT_BATCH * world,
              "per_slot_batch": PER_SLOT_BATCH, "image": "3x32x32", "aggregation_frequency": AGG_FREQ,
              "parallelism": f"dp{world}", "l2_policy": f"input pool of {POOL} distinct batches (151 MB) > 126 MB L2"}

    if args.impl == "reference":
        ms, why = run_reference(args, sampler)
        if ms is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": why}))
            return
        clocks = sampler.stop() if sampler is not None else None
        if rank == 0:
            v = samples / (ms / 1e3)
            print(json.dumps({"impl": "reference", "metric": "samples/sec per trial", "value": v, "unit": "samples/s",
                              "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                              "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                              "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                              "warmup_requested": requested_warmup, "config": config,
                              "harness": "reference determined 0.31.1-dev0 (baseline/_ref, unmodified), torch DDP/NCCL, eager; "
                                         "host-fed: value == e2e (the reference has no device-resident input mode)",
                              "e2e": {"value": v, "unit": "samples/s",
                                      "h2d_bytes_per_step": PER_SLOT_BATCH * (3 * 32 * 32 * 4 + 8), "d2h_bytes_per_step": 4},
                              "clocks": clocks}))
        return

    ms, launches, info = run_ours(args, on_device=True, sampler=sampler)
    clocks = sampler.stop() if sampler is not None else None
    e2e = None
    if not args.skip_e2e:
        ms2, _, _ = run_ours(args, on_device=False, sampler=None)
        e2e = {"value": samples / (ms2 / 1e3), "unit": "samples/s", "ms_per_step": ms2 / args.steps,
               "h2d_bytes_per_step": PER_SLOT_BATCH * (3 * 32 * 32 * 4 + 8), "d2h_bytes_per_step": 4}
    if rank == 0:
        v = samples / (ms / 1e3)
        print(json.dumps({"impl": "ours", "metric": "samples/sec per trial", "value": v, "unit": "samples/s",
                          "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
                          "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": (v / BASELINE_PUBLISHED) if BASELINE_PUBLISHED else None,
                          "dtype": "bf16", "data": "synthetic",
                          "warmup_requested": requested_warmup, "config": config,
                          "harness": f"determined_b200 fused engine ({info['optimizer']}, {info['buckets']} buckets, "
                                     f"all-reduce transport {info['transport']}), "
                                     f"{'eager' if args.no_graphs else 'CUDA-graph replay'}; value = device-resident input "
                                     "pool, e2e = host-fed through pytorch.init + Trainer.fit (compare e2e with the reference)",
                          "allreduce_check": info.get("allreduce_check"),
                          "e2e": e2e, "gpu_launches": launches, "clocks": clocks}))


if __name__ == "__main__":
    main()
