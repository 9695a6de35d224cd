"""An unmanaged run with several workers: launch it with torchrun yourself; the chief owns the experiment and trial rows,
the others join through the distributed context (reference examples/features/unmanaged/3_torch_distributed.py).

    DET_MASTER=http://127.0.0.1:8080 torchrun --standalone --nproc-per-node 2 3_torch_distributed.py
"""
import torch.distributed as dist

from determined_b200 import core
from determined_b200.experimental import core_v2


def main() -> None:
    dist.init_process_group("gloo")
    distributed = core.DistributedContext.from_torch_distributed()
    core_v2.init(defaults=core_v2.DefaultConfig(name="unmanaged-3-torch-distributed", hparams={"workers": distributed.size}), distributed=distributed)
    for step in range(50):
        # every worker computes its part, the chief reports the mean
        parts = core_v2.distributed.gather(1.0 / (1 + step) + 0.01 * core_v2.distributed.rank)
        if core_v2.distributed.rank == 0 and step % 10 == 9:
            core_v2.train.report_training_metrics(steps_completed=step + 1, metrics={"loss": sum(parts) / len(parts)})
    if core_v2.distributed.rank == 0:
        core_v2.train.report_validation_metrics(steps_completed=50, metrics={"val_loss": sum(parts) / len(parts)})
        print("experiment page:", core_v2.url_reverse_webui_exp_view())
    core_v2.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
