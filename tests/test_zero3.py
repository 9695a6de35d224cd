"""ZeRO stage 3 engine: weights partitioned over 2 gloo ranks train like the unpartitioned model, and checkpoints
consolidate back to the full state dict (host twin of the fused kernels; the GPU path is tests/fixtures/dist_check.py)."""
import os
import pathlib
import sys

import torch
import torch.multiprocessing as mp

ROOT = str(pathlib.Path(__file__).resolve().parent.parent)
HID, INNER, STEPS, MB = 64, 512, 3, 8


def _model():
    from determined_b200.ops import gemm as G

    torch.manual_seed(0)

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.inp = torch.nn.Linear(16, HID)                      # too small / indivisible: stays replicated
            self.mlp = G.FusedMLP(HID, INNER, activation="gelu")     # fc1 512x64, fc2: 64 rows do not split -> whole MLP stays
            self.wide = G.FusedLinear(HID, 1024, activation="relu")  # partitioned (512 rows per rank)
            self.mlp2 = G.FusedMLP(1024, 1024, activation="gelu_tanh")   # both partitioned
            self.out = torch.nn.Linear(1024, 4)

        def forward(self, x):
            h = self.inp(x)
            h = h + self.mlp(h)
            h = self.wide(h)
            h = h + self.mlp2(h)
            return self.out(h.float())

    return Net()


def _batches(rank, world):
    g = torch.Generator().manual_seed(7)
    x = torch.randn(STEPS * 2 * world * MB, 16, generator=g)
    y = torch.randn(STEPS * 2 * world * MB, 4, generator=g)
    for k in range(STEPS * 2):
        lo = (k * world + rank) * MB
        yield x[lo:lo + MB], y[lo:lo + MB]


CFG = {"train_micro_batch_size_per_gpu": MB, "gradient_accumulation_steps": 2, "gradient_clipping": 0.05,
       "optimizer": {"type": "AdamW", "params": {"lr": 2e-3, "weight_decay": 0.1}},
       "zero_optimization": {"stage": 3, "partition_parameters": True, "stage3_param_persistence_threshold": 1024}}


def _worker(rank, world, init_file, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from determined_b200.ops import gemm_comm
    from determined_b200.parallel import zero

    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    eng, _, _, _ = zero.initialize(model=_model(), config=CFG)
    assert eng.zero_optimization_stage() == 3 and eng.partitioned == 3, eng.partitioned
    assert isinstance(eng.module.wide, gemm_comm.Zero3Linear) and eng.module.wide.weight_shard.shape == (1024 // world, HID)
    assert isinstance(eng.module.mlp2, gemm_comm.Zero3MLP) and not isinstance(eng.module.mlp, gemm_comm.Zero3MLP)
    losses = []
    for x, y in _batches(rank, world):
        loss = torch.nn.functional.mse_loss(eng(x), y)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss))
    assert eng.global_steps == STEPS
    eng.save_checkpoint(out_dir, tag="t")
    full = eng.consolidated_state_dict()
    if rank == 0:
        torch.save({"losses": losses, "state": full}, os.path.join(out_dir, "rank0.pt"))
    # a fresh engine restores weights (sliced from the consolidated file) and its own optimizer shards
    eng2, _, _, _ = zero.initialize(model=_model(), config=CFG)
    eng2.load_checkpoint(out_dir)
    assert eng2.global_steps == STEPS
    for (n, a), (_, b) in zip(eng.module.named_parameters(), eng2.module.named_parameters()):
        assert torch.equal(a, b), n
    dist.barrier()
    dist.destroy_process_group()


def _reference(world):
    model = _model()
    opt = torch.optim.AdamW(model.parameters(), lr=2e-3, weight_decay=0.1)
    its = [_batches(r, world) for r in range(world)]
    losses = []
    for _ in range(STEPS):
        for _ in range(2):
            for r, it in enumerate(its):
                x, y = next(it)
                loss = torch.nn.functional.mse_loss(model(x), y)
                (loss / (2 * world)).backward()
                if r == 0:
                    losses.append(float(loss))
        torch.nn.utils.clip_grad_norm_(model.parameters(), 0.05)
        opt.step()
        opt.zero_grad()
    return losses, model.state_dict()


def test_zero3_partitioned_training_matches_full_model(tmp_path):
    world = 2
    init = tmp_path / "init"
    mp.spawn(_worker, args=(world, str(init), str(tmp_path)), nprocs=world, join=True)
    got = torch.load(tmp_path / "rank0.pt", weights_only=False)
    ref_losses, ref_state = _reference(world)
    assert set(got["state"]) == set(ref_state)
    # bf16 GEMM operands on both sides; the partitioned path reduces wgrads in a different order
    for a, b in zip(got["losses"], ref_losses):
        assert abs(a - b) < 2e-2 * max(1.0, abs(b)), (got["losses"], ref_losses)
    for k, v in ref_state.items():
        assert got["state"][k].shape == v.shape, k
        assert torch.allclose(got["state"][k].float(), v.float(), atol=2e-2, rtol=0), (k, (got["state"][k].float() - v.float()).abs().max())


def test_partition_pass_single_process():
    """world 1: every eligible layer converts, state dicts round-trip through the full-weight key."""
    sys.path.insert(0, ROOT)
    from determined_b200.ops import gemm_comm

    m = _model()
    ref = {k: v.clone() for k, v in m.state_dict().items()}
    tp = gemm_comm.HostTensorParallel()
    n = gemm_comm.partition_linears_(m, tp, min_params=1024)
    assert n == 3
    assert "wide.weight_shard" in m.state_dict() and "wide.weight" not in m.state_dict()
    full = gemm_comm.consolidated_state_dict(m)
    assert set(full) == set(ref)
    for k, v in ref.items():
        assert torch.equal(full[k], v), k
    with torch.no_grad():
        for p_ in m.parameters():
            p_.zero_()
    m.load_state_dict(ref)             # full ``weight`` keys are sliced on the way in
    for k, v in ref.items():
        assert torch.equal(gemm_comm.consolidated_state_dict(m)[k], v), k
    x = torch.randn(4, 16)
    m(x).sum().backward()
    assert m.wide.weight_shard.grad is not None and m.mlp2.fc1.weight_shard.grad.shape == m.mlp2.fc1.weight_shard.shape


def test_optimizer_shards_are_recut_for_another_world_size():
    """Save at world 8, load at world 4 (and 3 -> 2 with awkward sizes): the full moments rebuilt from the new shards equal the
    full moments the old shards were cut from.  Pure tensor logic of the native path's checkpoint loader, on CPU."""
    from determined_b200.parallel import zero

    def layout(sizes, world):
        align = 4 * world
        slots, off = [], 0
        for n in sizes:
            slots.append((off, n))
            off += (n + align - 1) // align * align
        total = max(off, align)
        # two buckets cut at a parameter boundary, listed last-layers-first like the engine does
        cut = slots[len(slots) // 2][0]
        return slots, total, [(cut, total - cut), (0, cut)]

    def shards_of(full, buckets, world):
        out = []
        for r in range(world):
            out.append(torch.cat([full[off + r * (n // world): off + (r + 1) * (n // world)] for off, n in buckets]))
        return out

    sizes = [5, 1000, 37, 4096, 129, 64]
    for old_w, new_w in ((8, 4), (3, 2), (2, 8)):
        o_slots, o_total, o_buckets = layout(sizes, old_w)
        n_slots, n_total, n_buckets = layout(sizes, new_w)
        per_param = [torch.randn(n) for n in sizes]
        full_old = torch.zeros(o_total)
        for (off, n), t in zip(o_slots, per_param):
            full_old[off:off + n] = t
        old = shards_of(full_old, o_buckets, old_w)
        full_new = torch.zeros(n_total)
        for r in range(new_w):
            mine = zero.reshard_flat_state(old, o_buckets, o_slots, old_w, n_buckets, n_slots, n_total, r, new_w)
            cur = 0
            for off, n in n_buckets:
                per = n // new_w
                full_new[off + r * per: off + (r + 1) * per] = mine[cur:cur + per]
                cur += per
        for (off, n), t in zip(n_slots, per_param):
            assert torch.equal(full_new[off:off + n], t)
    # row-partitioned weights: 2 weights of 16x6 and 8x4 rows x cols, saved by 4 ranks, loaded by 2
    ws = [torch.randn(16, 6), torch.randn(8, 4)]
    def row_shards(world):
        slots, off, per_rank = [], 0, [[] for _ in range(world)]
        for w in ws:
            rows = w.shape[0] // world
            slots.append((off, rows * w.shape[1]))
            off += rows * w.shape[1]
            for r in range(world):
                per_rank[r].append(w[r * rows:(r + 1) * rows].reshape(-1))
        return slots, off, [torch.cat(p) for p in per_rank]
    o_slots, _, old = row_shards(4)
    n_slots, n_total, want = row_shards(2)
    for r in range(2):
        assert torch.equal(zero.reshard_row_shards(old, o_slots, n_slots, n_total, r, 2), want[r])
