from determined_b200.pytorch.experimental._torch_batch_process import (  # noqa: F401
    TorchBatchProcessor, TorchBatchProcessorContext, get_default_device, torch_batch_process,
)
