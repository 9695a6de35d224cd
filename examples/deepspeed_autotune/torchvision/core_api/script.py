"""The same model tuned through the Core API: a plain training script that DSAT drives as trials of a custom search.

Every trial gets its candidate engine config in ``hparams["overwrite_deepspeed_args"]``; the script measures a few steps and
reports throughput with ``dsat.report_profile`` inside ``dsat.dsat_reporting_context`` (which turns an out-of-memory error
into an OOM report instead of a failed trial).  Reference counterpart:
examples/deepspeed_autotune/torchvision/core_api/script.py.
"""
import time

import torch
import torchvision

import determined_b200 as det
from determined_b200 import core
from determined_b200.parallel import zero
from determined_b200.pytorch import dsat


def main(core_context: core.Context, hparams: dict) -> None:
    ds_config = dsat.get_ds_config_from_hparams(hparams)
    model = getattr(torchvision.models, hparams.get("model_name", "resnet18"))(num_classes=100)
    engine, _, _, _ = zero.initialize(model=model, config=ds_config)
    mbs, size = engine.train_micro_batch_size_per_gpu(), int(hparams.get("image_size", 64))
    loss_fn = torch.nn.CrossEntropyLoss()
    x = torch.randn(mbs, 3, size, size, device=engine.device)
    y = torch.randint(0, 100, (mbs,), device=engine.device)
    steps = 0
    for op in core_context.searcher.operations():
        with dsat.dsat_reporting_context(core_context, op):
            t0, timed = time.perf_counter(), 0
            while steps < op.length:
                loss = loss_fn(engine(x).float(), y)
                engine.backward(loss)
                engine.step()
                steps += 1
                if steps == 2:                       # skip warm-up in the measurement
                    if engine.device.type == "cuda":
                        torch.cuda.synchronize()
                    t0, timed = time.perf_counter(), 0
                timed += 1
            if engine.device.type == "cuda":
                torch.cuda.synchronize()
            dt = max(time.perf_counter() - t0, 1e-9)
            world = core_context.distributed.size
            dsat.report_profile(core_context, op, steps, throughput=timed * mbs * world / dt, latency_ms=1e3 * dt / max(timed, 1))


if __name__ == "__main__":
    info = det.get_cluster_info()
    hp = info.trial.hparams if info is not None else {"deepspeed_config": "ds_config.json", "model_name": "resnet18"}
    zero.init_distributed()
    distributed = core.DistributedContext.from_torch_distributed() if torch.distributed.is_initialized() else None
    with core.init(distributed=distributed) as ctx:
        main(ctx, hp)
