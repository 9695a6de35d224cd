from determined_b200.pytorch.experimental._torch_batch_process import (  # noqa: F401
    TorchBatchProcessor, TorchBatchProcessorContext, torch_batch_process,
)
