"""Run under torchrun with >= 2 GPUs: checks the in-kernel collectives and the fused engine
against torch.distributed / analytic references.  Prints DIST_CHECK_OK on rank 0 on success."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def check_allreduce(rank, world, dev):
    from determined_b200.ops import comm

    group = comm.PeerGroup.from_torch_distributed(dev)
    n = 4 * world * 250_000 + 4 * world * 3
    buf = comm.SymmetricBuffer(4 * n, dev)
    t = buf.tensor(torch.float32, n)
    peers = group.map(buf)
    stream = torch.cuda.Stream()
    for it in range(3):
        torch.manual_seed(100 * it + rank)
        t.copy_(torch.randn(n, device=dev))
        ref = t.clone()
        dist.all_reduce(ref)
        ref.mul_(0.25 / world)
        group.begin_step()
        torch.cuda.synchronize()
        dist.barrier()
        with torch.cuda.stream(stream):
            # two buckets back to back exercise the epoch logic
            half = (n // 2) // (4 * world) * (4 * world)
            comm.allreduce_fused(group, peers, 0, half, 0.25 / world, found_inf=group.found_inf, stream=stream, timeout_s=20)
            comm.allreduce_fused(group, peers, half, n - half, 0.25 / world, found_inf=group.found_inf, stream=stream, timeout_s=20)
        stream.synchronize()
        torch.testing.assert_close(t, ref, rtol=1e-5, atol=1e-6)
        stats = group.stats_tensor()
        sumsq = float(stats[0, :world].sum())
        torch.testing.assert_close(torch.tensor(sumsq), ref.double().pow(2).sum().float().cpu(), rtol=1e-3, atol=1e-3)
        assert int(group.found_inf.item()) == 0
    # non-finite detection propagates to every rank
    t.zero_()
    if rank == world - 1:
        t[5] = float("inf")
    group.begin_step()
    torch.cuda.synchronize()
    dist.barrier()
    comm.allreduce_fused(group, peers, 0, n, 1.0, found_inf=group.found_inf, timeout_s=20)
    torch.cuda.synchronize()
    assert int(group.found_inf.item()) == 1, "inf not detected"
    # broadcast
    t.fill_(float(rank + 1))
    torch.cuda.synchronize()
    dist.barrier()
    comm.broadcast(group, peers, 0, n, root=0, timeout_s=20)
    torch.cuda.synchronize()
    assert float(t.min()) == float(t.max()) == 1.0
    dist.barrier()
    group.close()


def check_allreduce_nvls(rank, world, dev):
    """VMM symmetric buffers (fd exchange) + the multicast object: in-switch reduction and the bf16 wire format against
    NCCL, over several CTA budgets.  Where the box has no NVLS the same calls run peer-to-peer (and say so)."""
    from determined_b200.ops import comm

    group = comm.PeerGroup.from_torch_distributed(dev)
    assert group.vmm_ok, "VMM file-descriptor handles should be available on the B200 boxes"
    n = 8 * world * 200_000 + 8 * world * 5
    buf = group.alloc(4 * n)
    assert buf.vmm
    t = buf.tensor(torch.float32, n)
    peers = group.map(buf)
    mc = group.multicast(buf)
    if rank == 0:
        print("NVLS multicast mapping:", "yes" if mc else "no (peer-to-peer fallback)", flush=True)
    stage_buf = group.alloc(2 * n)
    stage = stage_buf.tensor(torch.bfloat16, n)
    peer_stage = group.map(stage_buf)
    mc_stage = group.multicast(stage_buf)
    for it, (ctas, use_mc) in enumerate([(0, True), (4, True), (148, True), (8, False), (0, False)]):
        torch.manual_seed(7 * it + rank)
        t.copy_(torch.randn(n, device=dev))
        ref = t.clone()
        dist.all_reduce(ref)
        ref.mul_(0.5 / world)
        group.begin_step()
        group.found_inf.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        half = (n // 2) // (8 * world) * (8 * world)
        m = mc if use_mc else None
        comm.allreduce_fused(group, peers, 0, half, 0.5 / world, found_inf=group.found_inf, timeout_s=20, mc_ptr=m, max_ctas=ctas)
        comm.allreduce_fused(group, peers, half, n - half, 0.5 / world, found_inf=group.found_inf, timeout_s=20, mc_ptr=m, max_ctas=ctas)
        torch.cuda.synchronize()
        torch.testing.assert_close(t, ref, rtol=1e-5, atol=1e-5)
        sumsq = float(group.stats_tensor()[0, :world].sum())
        torch.testing.assert_close(torch.tensor(sumsq), ref.double().pow(2).sum().float().cpu(), rtol=1e-3, atol=1e-3)
        assert int(group.found_inf.item()) == 0
        # every rank holds the same bits
        chk = t[:1024].clone()
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        assert torch.equal(lo, hi), "ranks disagree after the all-reduce"
    # bf16 wire format: narrow -> reduce -> widen, both transports
    for use_mc in (True, False):
        torch.manual_seed(99 + rank)
        t.copy_(torch.randn(n, device=dev))
        ref = t.to(torch.bfloat16).float()
        dist.all_reduce(ref)
        ref.mul_(1.0 / world)
        group.begin_step()
        torch.cuda.synchronize()
        dist.barrier()
        comm.narrow_bf16(t, stage)
        comm.allreduce_bf16(group, peer_stage, 0, n, 1.0 / world, found_inf=group.found_inf, timeout_s=20,
                            mc_ptr=mc_stage if use_mc else None)
        comm.widen_bf16(stage, t)
        torch.cuda.synchronize()
        torch.testing.assert_close(t, ref, rtol=2e-2, atol=2e-2)
        sumsq = float(group.stats_tensor()[0, :world].sum())
        torch.testing.assert_close(torch.tensor(sumsq), ref.double().pow(2).sum().float().cpu(), rtol=2e-2, atol=1e-2)
    # inf on one rank reaches everybody through the switch as well
    t.zero_()
    if rank == world - 1:
        t[13] = float("nan")
    group.begin_step()
    group.found_inf.zero_()
    torch.cuda.synchronize()
    dist.barrier()
    comm.allreduce_fused(group, peers, 0, n, 1.0, found_inf=group.found_inf, timeout_s=20, mc_ptr=mc)
    torch.cuda.synchronize()
    assert int(group.found_inf.item()) == 1, "non-finite value not detected"
    dist.barrier()
    group.close()


def _two_net_setup(dev, seed=0):
    torch.manual_seed(seed)
    gen = torch.nn.Sequential(torch.nn.Linear(16, 64), torch.nn.Tanh(), torch.nn.Linear(64, 32)).to(dev)
    dis = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.Tanh(), torch.nn.Linear(64, 1)).to(dev)
    return gen, dis


def check_gan_two_optimizers(rank, world, dev):
    """The GAN pattern the reference supports through torch DDP: g_loss.backward() flows into the discriminator's
    parameters, then d_loss.backward() adds to them before opt_d steps.  Every backward is reduced; the sums must match
    DDP's.  Also ``backward_passes_per_step=2`` (communicate on the second pass only) on a third optimizer."""
    import copy

    from determined_b200 import core, pytorch

    torch.backends.cuda.matmul.allow_tf32 = False
    for overlap in (True, False):
        _gan_case(rank, world, dev, overlap)
    _bpps_case(rank, world, dev)


def _gan_case(rank, world, dev, overlap):
    import copy

    from determined_b200 import core, pytorch

    gen0, dis0 = _two_net_setup(dev)
    ref_g = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(gen0), device_ids=[dev.index])
    ref_d = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(dis0), device_ids=[dev.index])
    ro_g = torch.optim.SGD(ref_g.parameters(), lr=0.05)
    ro_d = torch.optim.SGD(ref_d.parameters(), lr=0.05)
    dctx = core.DistributedContext.from_torch_distributed()
    with pytorch.init(hparams={"global_batch_size": 8 * world}, distributed=dctx) as ctx:
        ctx.experimental.use_fused_engine(True, bucket_cap_mb=0.004, overlap=overlap)     # several buckets per net
        g = ctx.wrap_model(copy.deepcopy(gen0))
        d = ctx.wrap_model(copy.deepcopy(dis0))
        og = ctx.wrap_optimizer(torch.optim.SGD(g.parameters(), lr=0.05))
        od = ctx.wrap_optimizer(torch.optim.SGD(d.parameters(), lr=0.05))
        assert type(og).__name__ == "FusedSGD" and type(od).__name__ == "FusedSGD"
        rng = torch.Generator().manual_seed(77 + rank)
        for i in range(4):
            ctx._current_batch_idx = i
            z = torch.randn(8, 16, generator=rng).to(dev)
            real = torch.randn(8, 32, generator=rng).to(dev)
            # ours
            g_loss = -d(g(z)).mean()
            ctx.backward(g_loss)                      # also fills the discriminator's gradients
            ctx.step_optimizer(og)
            d_loss = d(real).mean() - d(g(z).detach()).mean()
            ctx.backward(d_loss)                      # second gradient for the same parameters in this step
            ctx.step_optimizer(od)
            # reference
            ro_g.zero_grad()
            rg = -ref_d(ref_g(z)).mean()
            rg.backward()
            ro_g.step()
            rd = ref_d(real).mean() - ref_d(ref_g(z).detach()).mean()
            rd.backward()                             # d grads = g_loss part + d_loss part, both reduced
            ro_d.step()
            ro_d.zero_grad()
        for a, b in zip(list(g.parameters()) + list(d.parameters()), list(ref_g.module.parameters()) + list(ref_d.module.parameters())):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5)



def _bpps_case(rank, world, dev):
    # backward_passes_per_step=2: two backward passes per optimizer step, exchange on the second
    import copy

    from determined_b200 import core, pytorch

    dctx = core.DistributedContext.from_torch_distributed()
    torch.manual_seed(3)
    net0 = torch.nn.Linear(32, 8).to(dev)
    ref = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(net0), device_ids=[dev.index])
    ro = torch.optim.SGD(ref.parameters(), lr=0.1)
    with pytorch.init(hparams={"global_batch_size": 4 * world}, distributed=dctx) as ctx:
        ctx.experimental.use_fused_engine(True, overlap=True)
        net = ctx.wrap_model(copy.deepcopy(net0))
        opt = ctx.wrap_optimizer(torch.optim.SGD(net.parameters(), lr=0.1), backward_passes_per_step=2)
        rng = torch.Generator().manual_seed(5 + rank)
        for i in range(3):
            ctx._current_batch_idx = i
            xa, xb = torch.randn(4, 32, generator=rng).to(dev), torch.randn(4, 32, generator=rng).to(dev)
            ctx.backward(net(xa).pow(2).mean())
            ctx.backward(net(xb).pow(2).mean())
            ctx.step_optimizer(opt)
            ro.zero_grad()
            with ref.no_sync():
                ref(xa).pow(2).mean().backward()
            ref(xb).pow(2).mean().backward()
            ro.step()
        for a, b in zip(net.parameters(), ref.module.parameters()):
            torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)


def check_amp_and_compression(rank, world, dev):
    """K4: with a GradScaler the fused path stays on (unscale + inf check inside the all-reduce, skip flag into the fused
    optimizer, scaler.update() fed by the engine) and never calls scaler.unscale_.  K7: fp16_compression = bf16 wire."""
    import copy

    from determined_b200 import core, pytorch

    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(11)
    net0 = torch.nn.Sequential(torch.nn.Linear(24, 48), torch.nn.ReLU(), torch.nn.Linear(48, 4)).to(dev)
    ref = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(net0), device_ids=[dev.index])
    ro = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9)
    rs = torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=2)
    dctx = core.DistributedContext.from_torch_distributed()
    calls = {"unscale": 0}
    with pytorch.init(hparams={"global_batch_size": 8 * world}, distributed=dctx) as ctx:
        ctx.experimental.use_fused_engine(True, bucket_cap_mb=0.002, overlap=True)
        net = ctx.wrap_model(copy.deepcopy(net0))
        opt = ctx.wrap_optimizer(torch.optim.SGD(net.parameters(), lr=0.05, momentum=0.9))
        scaler = ctx.wrap_scaler(torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=2))
        orig = scaler.unscale_
        def counting(o):  # noqa: E306
            calls["unscale"] += 1
            return orig(o)
        scaler.unscale_ = counting
        assert type(opt).__name__ == "FusedSGD"
        rng = torch.Generator().manual_seed(21 + rank)
        for i in range(6):
            ctx._current_batch_idx = i
            x = torch.randn(8, 24, generator=rng).to(dev)
            if i == 3 and rank == world - 1:
                x[0, 0] = float("inf")           # one rank overflows: every rank must skip this step
            loss = net(x).pow(2).mean()
            ctx.backward(scaler.scale(loss))
            ctx.step_optimizer(opt, clip_grads=pytorch.FusedClipGradNorm(0.5), scaler=scaler)
            scaler.update()
            ro.zero_grad()
            rl = ref(x).pow(2).mean()
            rs.scale(rl).backward()
            rs.unscale_(ro)
            torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.5)
            rs.step(ro)
            rs.update()
            assert float(scaler.get_scale()) == float(rs.get_scale()), (i, scaler.get_scale(), rs.get_scale())
        for a, b in zip(net.parameters(), ref.module.parameters()):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5)
        assert calls["unscale"] == 0, "the fused AMP path must not fall back to scaler.unscale_"
        assert type(opt).__name__ == "FusedSGD"

    # bf16 wire format through the engine
    torch.manual_seed(12)
    net0 = torch.nn.Linear(64, 64).to(dev)
    with pytorch.init(hparams={"global_batch_size": 8 * world}, distributed=dctx) as ctx:
        a = ctx.wrap_model(copy.deepcopy(net0))
        b = ctx.wrap_model(copy.deepcopy(net0))
        oa = ctx.wrap_optimizer(torch.optim.SGD(a.parameters(), lr=0.1))
        ob = ctx.wrap_optimizer(torch.optim.SGD(b.parameters(), lr=0.1), fp16_compression=True)
        assert ctx._get_engine().records[id(ob)].wire_bf16
        rng = torch.Generator().manual_seed(31 + rank)
        for i in range(3):
            ctx._current_batch_idx = i
            x = torch.randn(8, 64, generator=rng).to(dev)
            ctx.backward(a(x).pow(2).mean() + b(x).pow(2).mean())
            ctx.step_optimizer(oa)
            ctx.step_optimizer(ob)
        for pa, pb in zip(a.parameters(), b.parameters()):
            torch.testing.assert_close(pa, pb, rtol=2e-2, atol=2e-3)
            assert not torch.equal(pa, pb) or pa.numel() < 8      # it really went through bf16
        ws = [p.detach().clone() for p in b.parameters()]
        for w in ws:
            lo, hi = w.clone(), w.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            assert torch.equal(lo, hi), "ranks diverged under the bf16 wire format"


def check_ddp_comm_hook(rank, world, dev):
    """A stock torch DDP model (what HuggingFace Trainer builds) with the fused all-reduce registered as its communication
    hook follows the same trajectory as the same model under DDP's own NCCL all-reduce."""
    import copy

    from determined_b200.parallel import ddp as pddp

    torch.backends.cuda.matmul.allow_tf32 = False
    for overlap in (False, True):
        _ddp_hook_case(rank, world, dev, overlap)


def _ddp_hook_case(rank, world, dev, overlap):
    import copy

    from determined_b200.parallel import ddp as pddp

    torch.manual_seed(4)
    net0 = torch.nn.Sequential(torch.nn.Linear(96, 512), torch.nn.GELU(), torch.nn.Linear(512, 512), torch.nn.GELU(),
                               torch.nn.Linear(512, 10)).to(dev)
    a = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(net0), device_ids=[dev.index], bucket_cap_mb=0.5)
    b = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(net0), device_ids=[dev.index], bucket_cap_mb=0.5)
    state = pddp.FusedAllReduceHook(dev, overlap=overlap)
    a.register_comm_hook(state, pddp.fused_allreduce_hook)
    oa = torch.optim.AdamW(a.parameters(), lr=1e-3)
    ob = torch.optim.AdamW(b.parameters(), lr=1e-3)
    rng = torch.Generator().manual_seed(50 + rank)
    for i in range(5):
        x = torch.randn(16, 96, generator=rng).to(dev)
        y = torch.randint(0, 10, (16,), generator=rng).to(dev)
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad()
            torch.nn.functional.cross_entropy(m(x), y).backward()
            o.step()
    # summation order differs between the in-switch reduction and NCCL's ring: last-bit differences after five AdamW steps
    for pa, pb in zip(a.parameters(), b.parameters()):
        torch.testing.assert_close(pa, pb, rtol=1e-3, atol=1e-5)
    assert state.launches >= 5 and len(state.layout) >= 2, (state.launches, len(state.layout))
    dist.barrier()
    state.close()


def check_engine(rank, world, dev):
    """One-variable model, aggregation_frequency=2: weights on every rank must follow the analytic
    trajectory of averaged gradients (reference test_pytorch_trial.py:980-1001,1604-1642)."""
    sys.path.insert(0, os.path.join(ROOT))
    from determined_b200 import core, pytorch
    from tests.fixtures import onevar

    agg = 2
    dctx = core.DistributedContext.from_torch_distributed()
    with pytorch.init(hparams={"learning_rate": 0.1, "batch_size": 4, "dataset_len": 64, "global_batch_size": 4 * world},
                      distributed=dctx, aggregation_frequency=agg) as ctx:
        trial = onevar.OneVarTrial(ctx)
        assert type(trial.opt).__name__ == "FusedSGD"
        os.chdir("/tmp")
        pytorch.Trainer(trial, ctx).fit(max_length=pytorch.Batch(12), checkpoint_policy="none")
        ds = onevar.OnesDataset(64)
        allb = [(ds.x[i:i + 4].numpy(), ds.y[i:i + 4].numpy()) for i in range(0, 64, 4)]
        per_rank = [[allb[r::world][i % len(allb[r::world])] for i in range(12)] for r in range(world)]
        w, acc, exp = 0.0, 0.0, []
        for i in range(12):
            g = np.mean([float(np.mean(2.0 * x * (w * x - y))) for (x, y) in (per_rank[r][i] for r in range(world))])
            acc += g
            if (i + 1) % agg == 0:
                w -= 0.1 * acc / agg
                acc = 0.0
            exp.append(w)
        np.testing.assert_allclose(trial.weights, exp, rtol=1e-4, atol=1e-6)
        ws = ctx.distributed.allgather(trial.weights[-1])
        assert max(ws) - min(ws) < 1e-7, f"ranks diverged: {ws}"


def check_resnet_matches_ddp(rank, world, dev):
    """ResNet-18 for a few steps: fused engine vs torch DDP + SGD on identical data."""
    import copy

    from determined_b200 import core, pytorch
    from determined_b200.models.resnet_cifar import resnet18_cifar

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)
    base = resnet18_cifar().to(dev)
    ddp_model = torch.nn.parallel.DistributedDataParallel(copy.deepcopy(base), device_ids=[dev.index])
    ref_opt = torch.optim.SGD(ddp_model.parameters(), lr=0.05, momentum=0.9)
    dctx = core.DistributedContext.from_torch_distributed()
    with pytorch.init(hparams={"global_batch_size": 8 * world}, distributed=dctx) as ctx:
        ctx.experimental.use_fused_engine(True, bucket_cap_mb=4, overlap=True)
        model = ctx.wrap_model(copy.deepcopy(base))
        opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9))
        g = torch.Generator().manual_seed(1234 + rank)
        for i in range(3):
            ctx._current_batch_idx = i
            x = torch.randn(8, 3, 32, 32, generator=g).to(dev)
            y = torch.randint(0, 10, (8,), generator=g).to(dev)
            loss = torch.nn.functional.cross_entropy(model(x), y)
            ctx.backward(loss)
            ctx.step_optimizer(opt, clip_grads=pytorch.FusedClipGradNorm(1e9))
            rl = torch.nn.functional.cross_entropy(ddp_model(x), y)
            ref_opt.zero_grad()
            rl.backward()
            ref_opt.step()
        # cuDNN's weight-gradient kernels are not deterministic and BatchNorm at 8 samples per rank amplifies the noise:
        # the two trajectories agree to a few 1e-3 after three steps, not better
        for a, b in zip(model.parameters(), ddp_model.module.parameters()):
            torch.testing.assert_close(a, b, rtol=5e-3, atol=3e-3)
        eng = ctx._get_engine()
        assert eng.kernel_launches > 0


def check_gemm_comm(rank, world, dev):
    """all-gather->GEMM and GEMM->reduce-scatter in one kernel each, vs NCCL collectives + fp32 matmul;
    also timed against the unfused (NCCL + our GEMM) sequence."""
    from determined_b200.ops import comm, gemm as G, gemm_comm

    group = comm.PeerGroup.from_torch_distributed(dev)
    tp = gemm_comm.TensorParallelGemm(group)
    rows, K, N = 512, 1024, 1536
    T = rows * world
    for it in range(3):
        torch.manual_seed(1000 + 10 * it + rank)
        x_shard = torch.randn(rows, K, device=dev).to(torch.bfloat16)
        w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
        bias = torch.randn(N, device=dev)
        full, mine, _ = tp.gather_buffer(T, K)
        mine.copy_(x_shard)
        y = tp.all_gather_gemm(T, w, bias=bias, epilogue="gelu", out_dtype=torch.float32)
        xs = [torch.empty_like(x_shard) for _ in range(world)]
        dist.all_gather(xs, x_shard)
        x_full = torch.cat(xs, 0)
        want = torch.nn.functional.gelu(x_full.float() @ w.float().t() + bias)
        torch.cuda.synchronize()
        torch.testing.assert_close(full, x_full)                       # the gather itself is bit exact
        torch.testing.assert_close(y, want, rtol=3e-3, atol=3e-2)
        # reduce-scatter: a [T, K] (this rank's K-slice), b [N, K]
        a = torch.randn(T, K, device=dev).to(torch.bfloat16)
        out = tp.gemm_reduce_scatter(a, w, bias=bias, out_dtype=torch.float32)
        part = (a.float() @ w.float().t()).to(torch.bfloat16).float()     # partials travel as bf16
        dist.all_reduce(part)
        want_rs = part[rank * rows:(rank + 1) * rows] + bias
        torch.cuda.synchronize()
        torch.testing.assert_close(out, want_rs, rtol=1e-2, atol=0.25)
    # autograd: column-parallel -> row-parallel block vs single-device math on gathered tensors
    torch.manual_seed(7)
    H, F4 = 1024, 2048                       # per-rank inner width F4
    w1_all = (torch.randn(world, F4, H, device=dev) * 0.03)
    w2_all = (torch.randn(world, H, F4, device=dev) * 0.03)
    dist.broadcast(w1_all, 0)
    dist.broadcast(w2_all, 0)
    x_all = torch.randn(T, H, device=dev)
    dist.broadcast(x_all, 0)
    w1 = w1_all[rank].clone().requires_grad_(True)
    w2 = w2_all[rank].clone().requires_grad_(True)
    xs = x_all[rank * rows:(rank + 1) * rows].clone().to(torch.bfloat16).requires_grad_(True)
    h = gemm_comm.column_parallel_linear(tp, xs, w1)
    y = gemm_comm.row_parallel_linear(tp, torch.relu(h), w2)
    y.float().sum().backward()
    xr = x_all.to(torch.bfloat16).float().requires_grad_(True)
    w1r = w1_all.to(torch.bfloat16).float().requires_grad_(True)
    w2r = w2_all.to(torch.bfloat16).float().requires_grad_(True)
    yr = sum(torch.relu(xr @ w1r[r].t()) @ w2r[r].t() for r in range(world))
    yr.sum().backward()
    torch.cuda.synchronize()

    def rel(a, b):
        return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-9))

    assert rel(y, yr[rank * rows:(rank + 1) * rows]) < 2e-2, rel(y, yr[rank * rows:(rank + 1) * rows])
    assert rel(w1.grad, w1r.grad[rank]) < 3e-2, rel(w1.grad, w1r.grad[rank])
    assert rel(w2.grad, w2r.grad[rank]) < 3e-2, rel(w2.grad, w2r.grad[rank])
    assert rel(xs.grad, xr.grad[rank * rows:(rank + 1) * rows]) < 3e-2
    # ---- timing: fused kernels vs NCCL collective followed by the same GEMM (device time, max over ranks)
    rows, K, N = 2048, 4096, 4096
    T = rows * world
    full, mine, _ = tp.gather_buffer(T, K)
    w = (torch.randn(N, K, device=dev) * 0.02).to(torch.bfloat16)
    x_shard = torch.randn(rows, K, device=dev).to(torch.bfloat16)
    gathered = torch.empty(T, K, device=dev, dtype=torch.bfloat16)
    a = torch.randn(T, K, device=dev).to(torch.bfloat16)
    part = torch.empty(T, N, device=dev, dtype=torch.bfloat16)
    rs_out = torch.empty(rows, N, device=dev, dtype=torch.bfloat16)

    def fused_ag():
        mine.copy_(x_shard)
        tp.all_gather_gemm(T, w)

    def fused_ag_n(n):
        def f():
            mine.copy_(x_shard)
            tp.all_gather_gemm(T, w, push_ctas=n)
        return f

    def gemm_only():
        G.gemm(gathered, w)

    def nccl_ag():
        dist.all_gather_into_tensor(gathered, x_shard)
        G.gemm(gathered, w)

    def fused_rs():
        tp.gemm_reduce_scatter(a, w)

    def nccl_rs():
        G.gemm(a, w, out=part)
        dist.reduce_scatter_tensor(rs_out, part)

    res = {}
    for name, fn in (("ag_gemm_fused", fused_ag), ("ag_gemm_fused_push2", fused_ag_n(2)), ("ag_gemm_fused_push8", fused_ag_n(8)),
                     ("ag_gemm_fused_push16", fused_ag_n(16)), ("ag_gemm_nccl", nccl_ag), ("gemm_only_full_T", gemm_only),
                     ("gemm_rs_fused", fused_rs), ("gemm_rs_nccl", nccl_rs)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(10):
            fn()
        e.record()
        torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e) / 10], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name] = round(float(t.item()), 4)
    if rank == 0:
        import json
        print("GEMM_COMM_TIMING", json.dumps({"world": world, "rows_per_rank": rows, "K": K, "N": N, "ms": res}), flush=True)


def check_gpt_tp(rank, world, dev):
    """Tensor+sequence-parallel GPT over the fused GEMM+collective kernels == the unsharded native GPT."""
    from determined_b200.models import gpt, gpt_tp
    from determined_b200.ops import comm, gemm_comm

    group = comm.PeerGroup.from_torch_distributed(dev)
    tp = gemm_comm.TensorParallelGemm(group)
    for kw in ({}, {"rotary_pct": 0.25, "parallel_residual": True, "tie_embeddings": False}):
        torch.manual_seed(11)
        cfg = gpt.GPTConfig(n_layer=2, n_head=4 * world // 2 if world > 2 else 4, n_embd=128 * (4 * world // 2 if world > 2 else 4), block_size=256, vocab_size=1024, **kw)
        full = gpt.GPT(cfg).to(dev)
        for p in full.parameters():
            dist.broadcast(p.data, 0)
        model = gpt_tp.TPGPT(cfg, tp).to(dev)
        model.load_from_full(full)
        B, S = 2 * world, 256
        idx = torch.randint(0, 1024, (B, S + 1), device=dev)
        dist.broadcast(idx, 0)
        _, want = full(idx[:, :-1], idx[:, 1:])
        want.backward()
        got = model(idx[:, :-1], idx[:, 1:])
        got.backward()
        model.sync_replicated_grads()
        tot = got.detach().clone()
        dist.all_reduce(tot)
        assert abs(float(tot) - float(want)) < 2e-2, (float(tot), float(want))

        def rel(a, b):
            return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-9))

        E, D, H = cfg.n_embd, cfg.head_dim, cfg.n_head
        hl = H // world
        for mine, theirs in zip(model.blocks, full.blocks):
            g = theirs.qkv.weight.grad.view(3, H, D, E)[:, rank * hl:(rank + 1) * hl].reshape(-1, E)
            assert rel(mine.qkv_w.grad, g) < 5e-2, ("qkv_w", rel(mine.qkv_w.grad, g))
            assert rel(mine.proj_w.grad, theirs.proj.weight.grad[:, rank * hl * D:(rank + 1) * hl * D]) < 5e-2
            F4 = mine.fc1_w.shape[0]
            assert rel(mine.fc1_w.grad, theirs.mlp.fc1.weight.grad[rank * F4:(rank + 1) * F4]) < 5e-2
            assert rel(mine.fc2_w.grad, theirs.mlp.fc2.weight.grad[:, rank * F4:(rank + 1) * F4]) < 5e-2
            assert rel(mine.fc2_b.grad, theirs.mlp.fc2.bias.grad) < 5e-2
            assert rel(mine.ln1.weight.grad, theirs.ln1.weight.grad) < 5e-2
        assert rel(model.wte.weight.grad, full.wte.weight.grad) < 5e-2
        assert rel(model.ln_f.weight.grad, full.ln_f.weight.grad) < 5e-2
    dist.barrier()


def check_zero3_linear(rank, world, dev):
    """ZeRO-3 linear: weight shards gathered inside the fwd and dgrad GEMMs, wgrad reduce-scattered by its epilogue."""
    from determined_b200.ops import comm, gemm_comm

    group = comm.PeerGroup.from_torch_distributed(dev)
    tp = gemm_comm.TensorParallelGemm(group)
    K, N, T = 1024, 512 * world, 768
    for it in range(3):
        torch.manual_seed(50 + it)
        w_full = (torch.randn(N, K, device=dev) * 0.05)
        bias = torch.randn(N, device=dev)
        dist.broadcast(w_full, 0)
        dist.broadcast(bias, 0)
        lin = gemm_comm.Zero3Linear(tp, K, N).to(dev)
        lin.load_full_weight(w_full)
        with torch.no_grad():
            lin.bias.copy_(bias)
        torch.manual_seed(1000 * it + rank)                      # every rank has its own data
        x = torch.randn(T, K, device=dev).to(torch.bfloat16).requires_grad_(True)
        dy = torch.randn(T, N, device=dev).to(torch.bfloat16)
        y = lin(x)
        y.backward(dy)
        wb = w_full.to(torch.bfloat16).float()
        want_y = x.detach().float() @ wb.t() + bias
        torch.testing.assert_close(y.float(), want_y, rtol=2e-2, atol=5e-2)
        want_dx = dy.float() @ wb
        rel = float((x.grad.float() - want_dx).norm() / want_dx.norm())
        assert rel < 2e-2, ("dx", rel)
        dw_full = dy.float().t() @ x.detach().float()
        dist.all_reduce(dw_full)
        dw_full /= world
        rows = N // world
        want_dw = dw_full[rank * rows:(rank + 1) * rows]
        rel = float((lin.weight_shard.grad - want_dw).norm() / want_dw.norm())
        assert rel < 2e-2, ("dw", rel)
        db = dy.float().sum(0)
        dist.all_reduce(db)
        torch.testing.assert_close(lin.bias.grad, db / world, rtol=2e-2, atol=0.5)
    dist.barrier()


def check_zero3_engine(rank, world, dev):
    """GPT under the ZeRO engine: stage 3 (partitioned weights, gather / reduce-scatter inside the GEMMs) follows the same
    loss curve as stage 1 (replicated weights) and consolidates to the same checkpoint keys."""
    import tempfile

    from determined_b200.models import gpt
    from determined_b200.ops import gemm_comm
    from determined_b200.parallel import zero

    def run(stage):
        torch.manual_seed(0)
        E = 512 if world <= 2 else 256 * world             # every projection must split into 256-row shards
        model = gpt.build("tiny", native=True, n_embd=E, n_head=E // 128, n_layer=2, vocab_size=1024, block_size=256, rotary_pct=0.25,
                          parallel_residual=True, tie_embeddings=False)
        cfg = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 2, "gradient_clipping": 1.0, "bf16": {"enabled": True},
               "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "weight_decay": 0.1}},
               "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 1000}}
        eng, _, _, _ = zero.initialize(model=model, config=cfg)
        g = torch.Generator().manual_seed(100 + rank)
        losses = []
        for _ in range(8):
            idx = torch.randint(0, 1024, (4, 256), generator=g).to(dev)
            _, loss = eng(idx, idx)
            eng.backward(loss)
            eng.step()
            losses.append(loss.detach().float())
        t = torch.stack(losses)
        dist.all_reduce(t)
        return eng, (t / world).tolist()

    e1, l1 = run(1)
    e3, l3 = run(3)
    if world <= 2:
        assert e3.partitioned == 9, e3.partitioned           # qkv, proj, fc1, fc2 of both blocks + the LM head
        assert isinstance(e3.module.blocks[0].mlp, gemm_comm.Zero3MLP)
    else:
        assert e3.partitioned >= 5, e3.partitioned
    assert l3[-1] < l3[0], l3
    for a, b in zip(l1, l3):
        assert abs(a - b) < 3e-2 * abs(a), (l1, l3)
    sd1, sd3 = e1.consolidated_state_dict(), e3.consolidated_state_dict()
    if rank == 0:
        assert set(sd1) == set(sd3)
        for k in sd1:
            assert sd1[k].shape == sd3[k].shape, k
        w1, w3 = sd1["blocks.0.mlp.fc1.weight"], sd3["blocks.0.mlp.fc1.weight"]
        assert "blocks.0.mlp.fc1.weight_shard" not in sd3
        assert float((w1 - w3).abs().max()) < 2e-2, float((w1 - w3).abs().max())
    d = [tempfile.mkdtemp() if rank == 0 else None]
    dist.broadcast_object_list(d, 0)
    e3.save_checkpoint(d[0], tag="z3")
    e3b, _ = run(3)
    e3b.load_checkpoint(d[0])
    for (n, a), (_, b) in zip(e3.module.named_parameters(), e3b.module.named_parameters()):
        assert torch.equal(a, b), n
    assert torch.equal(e3.shard_m, e3b.shard_m)
    dist.barrier()


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl")
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "allreduce"):
        check_allreduce(rank, world, dev)
    if which in ("all", "allreduce_nvls"):
        check_allreduce_nvls(rank, world, dev)
    if which in ("all", "gan"):
        check_gan_two_optimizers(rank, world, dev)
    if which in ("all", "amp"):
        check_amp_and_compression(rank, world, dev)
    if which in ("all", "ddp_hook"):
        check_ddp_comm_hook(rank, world, dev)
    if which in ("all", "engine"):
        check_engine(rank, world, dev)
    if which in ("all", "resnet"):
        check_resnet_matches_ddp(rank, world, dev)
    if which in ("all", "gemm_comm"):
        check_gemm_comm(rank, world, dev)
    if which in ("all", "gpt_tp"):
        check_gpt_tp(rank, world, dev)
    if which in ("all", "zero3"):
        check_zero3_linear(rank, world, dev)
    if which in ("all", "zero3_engine"):
        check_zero3_engine(rank, world, dev)
    dist.barrier()
    if rank == 0:
        print("DIST_CHECK_OK", which, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
