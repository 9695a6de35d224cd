"""Which distributed backend the launcher selected (reference ``harness/determined/_trial_controller.py:10-27``).

The launch layer exports one of three environment switches before the rank processes start
(``launch/_common.py``); trial controllers and user code ask this class instead of reading the environment themselves.
On this build all three lead to the same in-process engines (``torch.distributed`` for bootstrap, the fused kernels for
the gradient exchange): the switches only keep telling *which API family* launched the job."""
import os


class _DistributedBackend:
    HOROVOD = "USE_HOROVOD"
    DEEPSPEED = "USE_DEEPSPEED"
    TORCH = "USE_TORCH_DISTRIBUTED"

    def _on(self, var: str) -> bool:
        return bool(os.environ.get(var))

    def use_horovod(self) -> bool:
        return self._on(self.HOROVOD)

    def use_torch(self) -> bool:
        return self._on(self.TORCH)

    def use_deepspeed(self) -> bool:
        return self._on(self.DEEPSPEED)
