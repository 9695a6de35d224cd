"""Micro-benchmarks of the native kernels against their roofline (MEASURED_PEAKS.json).

Timing hygiene (B200_PROFILING.md): >= 3 warm-ups, L2 flushed between timed iterations by writing a
256 MB buffer, CUDA events on the launching stream, synchronize on both sides.
    python benchmarks/kernel_micro.py [bn|norm|xent|optim|all] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


_flush = None


def timeit(fn, iters=20, warmup=3):
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        _flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def bench_bn(rows):
    from determined_b200.ops import _lib

    lib = _lib.load()
    hbm, _, src = peaks()
    for (N, C, H, W) in [(256, 64, 32, 32), (256, 128, 16, 16), (256, 256, 8, 8), (256, 512, 4, 4)]:
        x = torch.randn(N, C, H, W, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        dy = torch.randn_like(x)
        y = torch.empty_like(x)
        M = N * H * W
        st = torch.empty(4, C, device="cuda")
        g5 = torch.empty(5, C, device="cuda")
        w = torch.ones(C, device="cuda")
        b = torch.zeros(C, device="cuda")
        rm, rv = torch.zeros(C, device="cuda"), torch.ones(C, device="cuda")
        ws = torch.zeros(int(lib.detb200_bn_workspace_bytes(C)), dtype=torch.uint8, device="cuda")
        s = _lib.stream_ptr()
        fwd = lambda: lib.detb200_bn_fwd_train(x.data_ptr(), y.data_ptr(), M, C, w.data_ptr(), b.data_ptr(), rm.data_ptr(), rv.data_ptr(), None,  # noqa: E731
                                               st[0].data_ptr(), st[1].data_ptr(), st[2].data_ptr(), st[3].data_ptr(), 0.1, 1e-5, 0, ws.data_ptr(), 1, s)
        bwd = lambda: lib.detb200_bn_bwd(dy.data_ptr(), x.data_ptr(), y.data_ptr(), M, C, w.data_ptr(), st[0].data_ptr(), st[1].data_ptr(),  # noqa: E731
                                         g5[0].data_ptr(), g5[1].data_ptr(), g5[2].data_ptr(), ws.data_ptr(), 1, s)
        bn = torch.nn.BatchNorm2d(C).cuda()
        xr = x.clone().requires_grad_()
        def aten():
            out = bn(xr)
            out.backward(dy)
        nbytes = x.numel() * 2
        t_f, t_b, t_a = timeit(fwd), timeit(bwd), timeit(aten)
        # algorithmic bytes: fwd reads x twice (second pass may hit L2) + writes y; bwd reads x,dy twice + writes dx
        rows.append({"kernel": "bn_fwd_train", "shape": [N, C, H, W], "us": t_f, "algo_GBs": 3 * nbytes / t_f / 1e3, "frac_of_hbm": 3 * nbytes / t_f / 1e3 / hbm, "peak_source": src})
        rows.append({"kernel": "bn_bwd", "shape": [N, C, H, W], "us": t_b, "algo_GBs": 5 * nbytes / t_b / 1e3, "frac_of_hbm": 5 * nbytes / t_b / 1e3 / hbm, "peak_source": src})
        rows.append({"kernel": "aten_bn_fwd+bwd", "shape": [N, C, H, W], "us": t_a, "ours_us": t_f + t_b, "speedup": t_a / (t_f + t_b)})


def bench_norm(rows):
    """Raw kernel launches (no autograd / allocator on the timed path); ATen timed through its
    native functions the same way."""
    from determined_b200.ops import _lib

    lib = _lib.load()
    hbm, _, src = peaks()
    for rows_, cols in [(8192, 768), (8192, 1024), (16384, 2560), (16384, 4096), (4096, 8192)]:
        x = torch.randn(rows_, cols, device="cuda", dtype=torch.bfloat16)
        dy = torch.randn_like(x)
        y, dx = torch.empty_like(x), torch.empty_like(x)
        g = torch.ones(cols, device="cuda", dtype=torch.bfloat16)
        b = torch.zeros(cols, device="cuda", dtype=torch.bfloat16)
        dg, db = torch.empty_like(g), torch.empty_like(g)
        mean = torch.empty(rows_, device="cuda")
        rstd = torch.empty(rows_, device="cuda")
        ws = torch.empty(int(lib.detb200_norm_bwd_workspace_elems(rows_, cols, 0)), device="cuda")
        s = _lib.stream_ptr()
        fwd = lambda: lib.detb200_norm_fwd(x.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(),  # noqa: E731
                                           rows_, cols, 1e-5, 1, 0, s)
        bwd = lambda: lib.detb200_norm_bwd(dy.data_ptr(), x.data_ptr(), g.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dx.data_ptr(),  # noqa: E731
                                           dg.data_ptr(), db.data_ptr(), ws.data_ptr(), rows_, cols, 1, 0, s)
        t_f = timeit(fwd)
        t_b = timeit(bwd)
        _, m_a, r_a = torch.native_layer_norm(x, (cols,), g, b, 1e-5)
        t_rf = timeit(lambda: torch.native_layer_norm(x, (cols,), g, b, 1e-5))
        t_rb = timeit(lambda: torch.ops.aten.native_layer_norm_backward(dy, x, (cols,), m_a, r_a, g, b, [True, True, True]))
        nb = x.numel() * 2
        rows.append({"kernel": "layernorm_fwd", "shape": [rows_, cols], "us": t_f, "algo_GBs": 2 * nb / t_f / 1e3, "frac_of_hbm": 2 * nb / t_f / 1e3 / hbm, "aten_us": t_rf, "peak_source": src})
        rows.append({"kernel": "layernorm_bwd", "shape": [rows_, cols], "us": t_b, "algo_GBs": 3 * nb / t_b / 1e3, "frac_of_hbm": 3 * nb / t_b / 1e3 / hbm, "aten_us": t_rb, "peak_source": src})


def bench_xent(rows):
    from determined_b200.ops import xent

    hbm, _, src = peaks()
    for r, v in [(8192, 50257), (16384, 50304)]:
        logits = torch.randn(r, v, device="cuda", dtype=torch.bfloat16, requires_grad=True)
        tgt = torch.randint(0, v, (r,), device="cuda")
        loss = xent.cross_entropy(logits, tgt)
        t_f = timeit(lambda: xent.cross_entropy(logits, tgt), iters=10)
        t_b = timeit(lambda: loss.backward(retain_graph=True), iters=10)
        lr = torch.nn.functional.cross_entropy(logits.float(), tgt)
        t_rf = timeit(lambda: torch.nn.functional.cross_entropy(logits.float(), tgt), iters=10)
        t_rb = timeit(lambda: lr.backward(retain_graph=True), iters=10)
        nb = logits.numel() * 2
        rows.append({"kernel": "xent_fwd", "shape": [r, v], "us": t_f, "algo_GBs": nb / t_f / 1e3, "frac_of_hbm": nb / t_f / 1e3 / hbm, "aten_us": t_rf, "peak_source": src})
        rows.append({"kernel": "xent_bwd", "shape": [r, v], "us": t_b, "algo_GBs": 2 * nb / t_b / 1e3, "frac_of_hbm": 2 * nb / t_b / 1e3 / hbm, "aten_us": t_rb, "peak_source": src})


def bench_optim(rows):
    from determined_b200.ops import optim
    from determined_b200.parallel.flat import FlatLayout

    hbm, _, src = peaks()
    for n in (11_173_962, 355_000_000):
        p = torch.nn.Parameter(torch.randn(n, device="cuda"))
        opt = torch.optim.AdamW([p], lr=1e-3)
        lay = FlatLayout(opt.param_groups, torch.device("cuda"))
        optim.fuse_optimizer_(opt, lay)
        p.grad.normal_()
        t = timeit(lambda: opt.step(zero_grad=True), iters=10)
        q = torch.nn.Parameter(torch.randn(n, device="cuda"))
        q.grad = torch.randn_like(q)
        ref = torch.optim.AdamW([q], lr=1e-3, fused=True)
        def ref_step():
            ref.step()
            q.grad.zero_()
        t_r = timeit(ref_step, iters=10)
        nb = 32 * n   # read p,g,m,v + write p,m,v,g
        rows.append({"kernel": "adamw_flat(+zero_grad)", "n": n, "us": t, "algo_GBs": nb / t / 1e3, "frac_of_hbm": nb / t / 1e3 / hbm, "torch_fused_us": t_r, "peak_source": src})
        del p, q, opt, ref, lay
        torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", nargs="?", default="all")
    ap.add_argument("--json")
    a = ap.parse_args()
    rows = []
    torch.backends.cudnn.benchmark = True
    for name, fn in (("bn", bench_bn), ("norm", bench_norm), ("xent", bench_xent), ("optim", bench_optim)):
        if a.what in ("all", name):
            fn(rows)
    for r in rows:
        print(json.dumps(r))
    if a.json:
        json.dump(rows, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
