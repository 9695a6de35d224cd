"""``context.experimental`` (reference ``harness/determined/pytorch/_experimental.py:17-86``)
plus the B200 engine switches."""
import logging
from typing import Any, Optional

import torch

logger = logging.getLogger("determined.pytorch")


class PyTorchExperimentalContext:
    def __init__(self, parent: Any) -> None:
        self._parent = parent
        self._auto_amp = False
        self._amp_dtype: Optional[torch.dtype] = None
        self._data_repro_checks_disabled = False
        self._auto_to_device = True
        # B200 engine options (no counterpart in the reference)
        self._fused_engine: Optional[bool] = None     # None = auto (on when CUDA + native lib)
        self._cuda_graphs = False
        self._bucket_cap_mb = 25.0
        self._grad_overlap: Optional[bool] = None
        self._wire_dtype: Optional[torch.dtype] = None
        self._async_checkpoint = False

    def use_amp(self, dtype: Optional[torch.dtype] = None) -> None:
        """Automatic mixed precision: ``train_batch``/``evaluate_batch`` forward passes run under
        ``torch.autocast``.  With fp16 (the reference's only mode) a ``GradScaler`` is created
        and driven by ``backward``/``step_optimizer``; with ``dtype=torch.bfloat16`` (the B200
        default recommendation) no scaler is needed."""
        self._parent._check_pre_wrap("use_amp()")
        self._auto_amp = True
        self._amp_dtype = dtype or torch.float16
        if self._amp_dtype == torch.float16:
            self._parent.wrap_scaler(torch.amp.GradScaler("cuda", enabled=torch.cuda.is_available()))

    def disable_dataset_reproducibility_checks(self) -> None:
        """Allow ``build_*_data_loader`` to return plain ``torch.utils.data.DataLoader``\\ s; the
        user then owns sharding/skip-on-resume."""
        self._data_repro_checks_disabled = True

    def disable_auto_to_device(self) -> None:
        """Do not move batches to the device automatically; call ``context.to_device`` yourself."""
        self._auto_to_device = False

    # ---- B200 engine -------------------------------------------------------------------------
    def use_fused_engine(self, enabled: bool = True, bucket_cap_mb: float = 25.0,
                         wire_dtype: Optional[torch.dtype] = None, overlap: Optional[bool] = None) -> None:
        """Force the flat-bucket engine on/off.  On: gradients live in one symmetric flat buffer,
        the bucketed all-reduce is our P2P kernel fused with averaging / unscale / inf-check /
        norm partials, and ``torch.optim.{SGD,Adam,AdamW,Adadelta}`` steps run as one fused
        sm_100a kernel per param group."""
        self._parent._check_pre_wrap("use_fused_engine()")
        self._fused_engine = enabled
        self._bucket_cap_mb = bucket_cap_mb
        self._wire_dtype = wire_dtype
        # True: bucketed all-reduces overlapped with backward; False: one all-reduce at step time; None: by gradient volume
        self._grad_overlap = overlap

    def enable_async_checkpoint(self, enabled: bool = True) -> None:
        """Checkpoint fast path: at a checkpoint boundary the chief snapshots models / optimizers / schedulers into pinned host
        memory with non-blocking copies on the training stream and a writer thread serialises, uploads and reports the
        checkpoint while training continues (``pytorch/_async_checkpoint.py``).  The files are identical to the synchronous
        path.  ``on_checkpoint_write_end`` / ``on_checkpoint_end`` then run on the writer thread and
        ``on_checkpoint_upload_end`` is delivered at the next checkpoint or at exit; checkpoints taken while exiting
        (preemption, end of training) are always synchronous."""
        self._async_checkpoint = enabled

    def enable_cuda_graphs(self, enabled: bool = True) -> None:
        """Capture ``train_batch`` (forward + backward + fused all-reduce + optimizer) into CUDA
        graphs (one per aggregation phase) after a few eager warm-up batches.  Requires static
        batch shapes and no host synchronisation inside ``train_batch``."""
        self._cuda_graphs = enabled
