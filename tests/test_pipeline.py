"""Pipeline engine: 2 stages x 2 data-parallel replicas over gloo reproduce single-process training."""
import os
import pathlib
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = str(pathlib.Path(__file__).resolve().parent.parent)


def _layers():
    torch.manual_seed(0)
    return [torch.nn.Linear(8, 16), torch.nn.Tanh(), torch.nn.Linear(16, 16), torch.nn.Tanh(), torch.nn.Linear(16, 4)]


def _data(dp_rank, dp, steps, gas, mbs):
    g = torch.Generator().manual_seed(123)
    x = torch.randn(steps * gas * dp * mbs, 8, generator=g)
    y = torch.randn(steps * gas * dp * mbs, 4, generator=g)
    # global micro batch k of replica r = rows [(k*dp + r)*mbs, ...)
    for k in range(steps * gas):
        lo = (k * dp + dp_rank) * mbs
        yield x[lo:lo + mbs], y[lo:lo + mbs]


def _worker(rank, world, init_file, out_dir, stages):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from determined_b200.parallel import pipeline

    dist.init_process_group("gloo", init_method=f"file://{init_file}", rank=rank, world_size=world)
    cfg = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 3, "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "gradient_clipping": 0.5}
    eng = pipeline.PipelineEngine(_layers(), stages, lambda out, y: torch.nn.functional.mse_loss(out, y), cfg)
    it = _data(eng.grid.data_id, eng.grid.dp, steps=4, gas=3, mbs=4) if (eng.is_first_stage() or eng.is_last_stage()) else None
    losses = [float(eng.train_batch(it)) for _ in range(4)]
    eng.save_checkpoint(out_dir, tag="t")
    torch.save({"losses": losses, "range": eng.layer_range, "state": eng.module.state_dict(), "stage": eng.stage_id, "dp": eng.grid.data_id},
               os.path.join(out_dir, f"rank{rank}.pt"))
    # reload into a fresh engine and check counters + weights
    eng2 = pipeline.PipelineEngine(_layers(), stages, lambda out, y: torch.nn.functional.mse_loss(out, y), cfg)
    eng2.load_checkpoint(out_dir)
    assert eng2.global_steps == 4 and eng2.micro_steps == 12
    for a, b in zip(eng.module.parameters(), eng2.module.parameters()):
        assert torch.equal(a, b)
    dist.barrier()
    dist.destroy_process_group()


def _reference(dp):
    model = torch.nn.Sequential(*_layers())
    opt = torch.optim.SGD(model.parameters(), lr=0.1)
    its = [_data(r, dp, 4, 3, 4) for r in range(dp)]
    losses = []
    for _ in range(4):
        tot = 0.0
        for _ in range(3):
            for it in its:
                x, y = next(it)
                loss = torch.nn.functional.mse_loss(model(x), y)
                (loss / (3 * dp)).backward()
                tot += float(loss) / (3 * dp)
        torch.nn.utils.clip_grad_norm_(model.parameters(), 0.5)
        opt.step()
        opt.zero_grad()
        losses.append(tot)
    return model, losses


@pytest.mark.parametrize("world,stages", [(2, 2), (4, 2), (3, 3)])
def test_pipeline_matches_single_process(tmp_path, world, stages):
    init = tmp_path / "init"
    mp.spawn(_worker, args=(world, str(init), str(tmp_path), stages), nprocs=world, join=True)
    model, ref_losses = _reference(world // stages)
    ref_sd = model.state_dict()
    seen = set()
    for r in range(world):
        d = torch.load(tmp_path / f"rank{r}.pt", weights_only=False)
        lo, hi = d["range"]
        seen.update(range(lo, hi))
        for k, v in d["state"].items():
            idx, name = k.split(".", 1)
            torch.testing.assert_close(v, ref_sd[f"{lo + int(idx)}.{name}"], rtol=1e-5, atol=1e-6)
        assert d["losses"] == pytest.approx(ref_losses, rel=1e-5)
    assert seen == set(range(5))
    assert (tmp_path / "latest").read_text() == "t" and (tmp_path / "t" / "layer_00-model_states.pt").exists()


def test_partition_balanced():
    from determined_b200.parallel.pipeline import partition_balanced

    assert partition_balanced([1, 1, 1, 1], 2) == [0, 2, 4]
    assert partition_balanced([10, 1, 1, 1, 1], 2) == [0, 1, 5]
    b = partition_balanced([5, 5, 1, 1, 1, 1, 5, 5], 4)
    assert b[0] == 0 and b[-1] == 8 and all(x < y for x, y in zip(b, b[1:]))
    with pytest.raises(ValueError):
        partition_balanced([1, 1], 3)
