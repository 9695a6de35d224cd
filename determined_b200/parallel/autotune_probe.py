"""What the ZeRO engine does when its config says ``"autotuning": {"enabled": true}``: measure, write a JSON file, exit.

This is the engine half of the autotuner in ``determined_b200.pytorch.dsat`` (the reference delegates it to DeepSpeed's own
autotuning profiler, which its search methods drive through ``overwrite_deepspeed_args``: reference
``harness/determined/pytorch/dsat/defaults.py:18-31`` are the switches, ``_utils.py:293-356`` picks the files up).  Two kinds
of probe, told apart by the config exactly as there:

* ``autotuning.model_info.profile``: one step at micro-batch size 1.  Reports ``num_params`` / ``trainable_num_params``
  (unpartitioned sizes, also under stage 3) and ``activation_mem_per_gpu``: the bytes autograd keeps alive between forward
  and backward, counted with ``saved_tensors_hooks`` over the first forward (parameters excluded, storages counted once), so
  the number is the same on any device and independent of allocator behaviour.  Written to ``autotuning.model_info_path``.
* otherwise: time the steps ``start_profile_step .. end_profile_step`` (device-synchronised at both ends) and report
  ``latency`` (s / step), ``throughput`` (samples / s over all ranks), ``FLOPS_per_gpu`` (6 x trainable parameters x tokens
  this rank processed per step / latency; "tokens" are the elements of the first integer input tensor, else its batch
  dimension) and ``forward`` / ``backward`` (s, means over the window).  Written to ``autotuning_metric.json``.

Either way the process then exits (``SystemExit``), which ``dsat.dsat_reporting_context`` turns into the trial's report.
"""
import json
import logging
import pathlib
import sys
import time
from typing import Any, Dict, Optional, Set

import torch

logger = logging.getLogger("determined.parallel.autotune")

MODEL_INFO_DEFAULT_PATH = "model_info.json"
METRIC_PATH = "autotuning_metric.json"


class EngineProbe:
    def __init__(self, engine: Any, cfg: Dict[str, Any]) -> None:
        self.engine = engine
        self.start = int(cfg.get("start_profile_step", 3))
        self.end = int(cfg.get("end_profile_step", 5))
        if self.end < self.start:
            raise ValueError(f"autotuning: end_profile_step {self.end} < start_profile_step {self.start}")
        mi = cfg.get("model_info") or {}
        self.model_info = bool(mi.get("profile"))
        self.model_info_path = pathlib.Path(cfg.get("model_info_path") or MODEL_INFO_DEFAULT_PATH)
        self.metric_path = pathlib.Path(cfg.get("metric_path") or METRIC_PATH)
        self._t0: Optional[float] = None
        self._activation_bytes: Optional[int] = None
        self._tokens_per_step = 0
        self._tokens_this_step = 0
        self._fwd_s = 0.0
        self._bwd_s = 0.0
        self._fwd_mark: Optional[float] = None
        self._bwd_mark: Optional[float] = None
        self._timed_steps = 0

    # ---- helpers ----------------------------------------------------------------------------------------------------------
    def _sync(self) -> None:
        if self.engine.device.type == "cuda":
            torch.cuda.synchronize(self.engine.device)

    def _count_params(self) -> Dict[str, int]:
        w = self.engine.world
        full = lambda p: p.numel() * (w if getattr(p, "_zero3_shard", False) else 1)     # noqa: E731
        ps = list(self.engine.module.parameters())
        return {"num_params": sum(full(p) for p in ps), "trainable_num_params": sum(full(p) for p in ps if p.requires_grad)}

    @staticmethod
    def _tokens(args: Any, kwargs: Any) -> int:
        for a in list(args) + list(kwargs.values()):
            if isinstance(a, dict):
                a = next((v for v in a.values() if isinstance(v, torch.Tensor)), None)
            if isinstance(a, (list, tuple)):
                a = next((v for v in a if isinstance(v, torch.Tensor)), None)
            if isinstance(a, torch.Tensor) and a.dim() > 0:
                return a.numel() if not a.is_floating_point() else a.shape[0]
        return 1

    # ---- called by the engine -----------------------------------------------------------------------------------------------
    def forward(self, call: Any, args: Any, kwargs: Any) -> Any:
        """Run ``call(*args, **kwargs)`` (the module's forward) under whatever this probe measures."""
        self._tokens_this_step += self._tokens(args, kwargs)
        if self.model_info and self._activation_bytes is None:
            params: Set[int] = {p.untyped_storage().data_ptr() for p in self.engine.module.parameters()}
            seen: Set[int] = set()
            total = [0]

            def pack(t: torch.Tensor) -> torch.Tensor:
                st = t.untyped_storage()
                key = st.data_ptr()
                if key not in params and key not in seen:
                    seen.add(key)
                    total[0] += st.nbytes()
                return t

            with torch.autograd.graph.saved_tensors_hooks(pack, lambda t: t):
                out = call(*args, **kwargs)
            self._activation_bytes = total[0]
            return out
        timing = self._t0 is not None
        if timing:
            self._sync()
            self._fwd_mark = time.perf_counter()
        out = call(*args, **kwargs)
        if timing:
            self._sync()
            self._fwd_s += time.perf_counter() - self._fwd_mark       # type: ignore[operator]
        return out

    def backward_begin(self) -> None:
        if self._t0 is not None:
            self._sync()
            self._bwd_mark = time.perf_counter()

    def backward_end(self) -> None:
        if self._t0 is not None and self._bwd_mark is not None:
            self._sync()
            self._bwd_s += time.perf_counter() - self._bwd_mark
            self._bwd_mark = None

    def step_done(self) -> None:
        """After an optimizer step (``engine.global_steps`` already counts it)."""
        e = self.engine
        tokens, self._tokens_this_step = self._tokens_this_step, 0
        if self.model_info:
            info = dict(self._count_params(), activation_mem_per_gpu=int(self._activation_bytes or 0), rank=e.rank)
            self._finish(self.model_info_path, info)
        g = e.global_steps
        if self._t0 is not None:
            self._timed_steps += 1
            self._tokens_per_step = tokens
        if g == self.start - 1 or (self._t0 is None and g >= self.start):
            self._sync()
            self._t0 = time.perf_counter()          # steps start .. end are the ones after this point
            return
        if g >= self.end and self._t0 is not None and self._timed_steps > 0:
            self._sync()
            n = self._timed_steps
            latency = (time.perf_counter() - self._t0) / n
            trainable = self._count_params()["trainable_num_params"]
            self._finish(self.metric_path, {
                "latency": latency, "throughput": e.train_batch_size() / latency,
                "FLOPS_per_gpu": 6.0 * trainable * self._tokens_per_step / latency,
                "forward": self._fwd_s / n, "backward": self._bwd_s / n})

    def _finish(self, path: pathlib.Path, result: Dict[str, Any]) -> None:
        if self.engine.rank == 0:
            path.write_text(json.dumps(result))
            logger.info("autotuning probe: %s -> %s", result, path)
        if self.engine.world > 1:
            torch.distributed.barrier()          # the file exists before any rank leaves
        sys.exit(0)
