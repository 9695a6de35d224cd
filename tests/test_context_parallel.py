"""Ring attention over a sequence sharded across ranks (gloo, 2 and 4 processes) equals attention over the whole sequence:
outputs and the gradients of Q, K and V, causal and not, with grouped-query heads."""
import pathlib
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = str(pathlib.Path(__file__).resolve().parent.parent)
B, H, HKV, S, D = 2, 4, 2, 32, 16


def _full(causal):
    g = torch.Generator().manual_seed(5)
    q = torch.randn(B, H, S, D, generator=g, requires_grad=True)
    k = torch.randn(B, HKV, S, D, generator=g, requires_grad=True)
    v = torch.randn(B, HKV, S, D, generator=g, requires_grad=True)
    w = torch.randn(B, H, S, D, generator=g)
    rep = H // HKV
    o = torch.nn.functional.scaled_dot_product_attention(q, k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1), is_causal=causal)
    (o * w).sum().backward()
    return q, k, v, w, o


def _worker(rank, world, init_file, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from determined_b200.parallel import context_parallel as cp

    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    res = {}
    for causal in (False, True):
        q, k, v, w, o = _full(causal)
        ql, kl, vl = (cp.shard_sequence(t.detach(), rank, world, dim=2).clone().requires_grad_(True) for t in (q, k, v))
        out = cp.ring_attention(ql, kl, vl, causal=causal)
        (out * cp.shard_sequence(w, rank, world, dim=2)).sum().backward()
        res[causal] = {"o": (out - cp.shard_sequence(o.detach(), rank, world, dim=2)).abs().max().item(),
                       "dq": (ql.grad - cp.shard_sequence(q.grad, rank, world, dim=2)).abs().max().item(),
                       "dk": (kl.grad - cp.shard_sequence(k.grad, rank, world, dim=2)).abs().max().item(),
                       "dv": (vl.grad - cp.shard_sequence(v.grad, rank, world, dim=2)).abs().max().item()}
    assert cp.position_ids(S // world, rank, world).tolist() == list(range(rank * S // world, (rank + 1) * S // world))
    torch.save(res, f"{out_dir}/r{rank}.pt")
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_ring_attention_matches_full_attention(tmp_path, world):
    init = tmp_path / "init"
    mp.spawn(_worker, args=(world, str(init), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = torch.load(tmp_path / f"r{r}.pt")
        for causal, errs in res.items():
            for name, e in errs.items():
                assert e < 2e-5, f"rank {r} causal={causal} {name}: {e}"


def test_single_process_is_plain_attention_and_shards_must_divide():
    from determined_b200.parallel import context_parallel as cp

    q, k, v, w, o = _full(True)
    out = cp.ring_attention(q.detach(), k.detach(), v.detach(), causal=True)
    assert (out - o.detach()).abs().max() < 2e-5
    with pytest.raises(ValueError):
        cp.shard_sequence(torch.zeros(1, 10), 0, 3)
    m = cp.ContextParallelAttention(causal=False)
    assert m(q.detach(), k.detach(), v.detach()).shape == (B, H, S, D)


def _gpt_worker(rank, world, init_file, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from determined_b200.models import gpt
    from determined_b200.parallel import context_parallel as cp

    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    out = {}
    for name, over in (("rotary", {"rotary_pct": 0.5, "parallel_residual": True}), ("learned", {})):
        torch.manual_seed(0)
        model = gpt.build("tiny", native=False, **over)
        g = torch.Generator().manual_seed(9)
        idx = torch.randint(0, model.cfg.vocab_size, (2, 32), generator=g)
        tgt = torch.randint(0, model.cfg.vocab_size, (2, 32), generator=g)
        _, full = model(idx, tgt)
        full.backward()
        ref = {n: p.grad.clone() for n, p in model.named_parameters()}
        model.zero_grad()
        model.set_context_parallel(dist.group.WORLD)
        _, loss = model(cp.shard_sequence(idx, rank, world), cp.shard_sequence(tgt, rank, world))
        loss.backward()
        mean = loss.detach().clone()
        dist.all_reduce(mean)
        err = 0.0
        for n, p in model.named_parameters():
            gsum = p.grad.clone()
            dist.all_reduce(gsum)
            err = max(err, (gsum / world - ref[n]).abs().max().item())
        out[name] = (abs(mean.item() / world - full.item()), err)
    torch.save(out, f"{out_dir}/g{rank}.pt")
    dist.destroy_process_group()


def test_gpt_with_a_sharded_sequence_has_the_loss_and_gradients_of_the_whole_sequence(tmp_path):
    mp.spawn(_gpt_worker, args=(2, str(tmp_path / "init"), str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        for name, (dl, dg) in torch.load(tmp_path / f"g{r}.pt").items():
            assert dl < 1e-5 and dg < 1e-5, (r, name, dl, dg)
