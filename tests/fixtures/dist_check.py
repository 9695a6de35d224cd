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
        ctx.experimental.use_fused_engine(True, bucket_cap_mb=4)
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
        for a, b in zip(model.parameters(), ddp_model.module.parameters()):
            torch.testing.assert_close(a, b, rtol=5e-3, atol=1e-3)
        eng = ctx._get_engine()
        assert eng.kernel_launches > 0


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl")
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "allreduce"):
        check_allreduce(rank, world, dev)
    if which in ("all", "engine"):
        check_engine(rank, world, dev)
    if which in ("all", "resnet"):
        check_resnet_matches_ddp(rank, world, dev)
    dist.barrier()
    if rank == 0:
        print("DIST_CHECK_OK", which, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
