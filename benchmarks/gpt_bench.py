"""GPT training-step throughput on one GPU: native stack vs the same network on stock torch modules
(autocast bf16, cuBLAS, fused SDPA, torch AdamW(fused=True)).  Synthetic tokens, random init.

    python benchmarks/gpt_bench.py [--model gpt2] [--batch 16] [--seq 1024] [--steps 10] [--json out.json]
"""
import argparse
import json
import pathlib
import sys

import torch

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
from determined_b200.models import gpt  # noqa: E402
from determined_b200.ops import optim as fused_optim  # noqa: E402


def run(model_name, native, batch, seq, steps, warmup):
    torch.manual_seed(0)
    m = gpt.build(model_name, native=native, block_size=max(seq, 1024)).cuda()
    opt = torch.optim.AdamW(m.parameters(), lr=1e-4, weight_decay=0.1, fused=not native)
    if native:
        from determined_b200.parallel.flat import FlatLayout

        layout = FlatLayout(opt.param_groups, torch.device("cuda"))     # params/grads re-homed into flat fp32 buffers
        assert fused_optim.fuse_optimizer_(opt, layout)
    pool = [torch.randint(0, m.cfg.vocab_size, (batch, seq + 1), device="cuda") for _ in range(4)]

    def step(i):
        idx = pool[i % len(pool)]
        if native:
            _, loss = m(idx[:, :-1], idx[:, 1:])
        else:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                _, loss = m(idx[:, :-1], idx[:, 1:])
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=False)
        return loss

    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(steps):
        loss = step(i)
    e.record()
    torch.cuda.synchronize()
    ms = s.elapsed_time(e) / steps
    tok = batch * seq / ms * 1e3
    return {"native": native, "ms_per_step": ms, "tokens_per_s": tok, "model_tflops": tok * m.flops_per_token(seq) / 1e12, "loss": float(loss.detach()),
            "params": m.num_params()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="gpt2")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--json")
    ap.add_argument("--only", choices=["native", "torch"])
    a = ap.parse_args()
    out = {"model": a.model, "batch": a.batch, "seq": a.seq, "steps": a.steps}
    for native in (True, False):
        if a.only and (a.only == "native") != native:
            continue
        r = run(a.model, native, a.batch, a.seq, a.steps, a.warmup)
        out["native" if native else "torch"] = r
        print(json.dumps(r), flush=True)
        torch.cuda.empty_cache()
    if "torch" in out and "native" in out:
        out["speedup"] = out["torch"]["ms_per_step"] / out["native"]["ms_per_step"]
        print(json.dumps({"speedup_native_over_torch": out["speedup"]}))
    if a.json:
        pathlib.Path(a.json).write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
