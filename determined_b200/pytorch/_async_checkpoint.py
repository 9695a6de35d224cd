"""Checkpoint fast path: snapshot on the training stream, serialise and upload behind the training loop.

Reference: ``pytorch/_pytorch_trial.py:1287-1360`` (``_save``) runs ``torch.save`` of the live state dicts and the
upload synchronously, so every checkpoint stalls training for the D2H copy, the pickle and the storage write.
Here the state is *snapshotted* into pinned host memory with non-blocking copies queued on the training stream
(they are ordered after the step that produced the weights and before the next one, so the snapshot is exact and
the host does not wait), and one writer thread turns snapshots into checkpoints (``state_dict.pth`` and friends,
byte-for-byte the same layout) while training goes on.  The pinned buffers are reused from checkpoint to
checkpoint.  On CPU the snapshot is a clone.
"""
import logging
import queue
import threading
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch

logger = logging.getLogger("determined.pytorch.checkpoint")


class PinnedPool:
    """Host staging buffers keyed by (shape, dtype), handed out in snapshot order and reused by the next snapshot."""

    def __init__(self) -> None:
        self._free: Dict[Tuple[Any, ...], List[torch.Tensor]] = {}
        self.allocated_bytes = 0

    def take(self, like: torch.Tensor) -> torch.Tensor:
        key = (tuple(like.shape), like.dtype)
        lst = self._free.get(key)
        if lst:
            return lst.pop()
        self.allocated_bytes += like.numel() * like.element_size()
        return torch.empty(like.shape, dtype=like.dtype, device="cpu", pin_memory=torch.cuda.is_available())

    def give(self, bufs: List[torch.Tensor]) -> None:
        for b in bufs:
            self._free.setdefault((tuple(b.shape), b.dtype), []).append(b)


class Snapshot:
    def __init__(self, tree: Any, event: Optional["torch.cuda.Event"], bufs: List[torch.Tensor]) -> None:
        self.tree, self.event, self.bufs = tree, event, bufs

    def wait(self) -> Any:
        """Block (the *calling* thread only) until every device->host copy has landed; returns the host tree."""
        if self.event is not None:
            self.event.synchronize()
        return self.tree


def snapshot(tree: Any, pool: Optional[PinnedPool] = None) -> Snapshot:
    """Deep copy of a nested dict / list / tuple of tensors and plain values.  CUDA tensors go to (pooled) pinned host
    buffers with ``non_blocking`` copies on the current stream; CPU tensors are cloned; storage sharing between tensors of
    the input (tied weights, views of one flat buffer) is not preserved — every tensor is saved dense, like ``torch.save``
    of a state dict of clones."""
    bufs: List[torch.Tensor] = []
    saw_cuda = False

    def walk(x: Any) -> Any:
        nonlocal saw_cuda
        if torch.is_tensor(x):
            if x.is_cuda:
                saw_cuda = True
                src = x.detach()
                dst = pool.take(src) if pool is not None else torch.empty(src.shape, dtype=src.dtype, device="cpu", pin_memory=True)
                dst.copy_(src, non_blocking=True)
                bufs.append(dst)
                return dst
            return x.detach().clone()
        if isinstance(x, dict):
            items = [(k, walk(v)) for k, v in x.items()]
            try:
                return type(x)(items)
            except TypeError:               # defaultdict and friends
                return dict(items)
        if isinstance(x, (list, tuple)):
            vals = [walk(v) for v in x]
            return vals if isinstance(x, list) else (type(x)(*vals) if hasattr(x, "_fields") else tuple(vals))
        return x

    out = walk(tree)
    ev = None
    if saw_cuda:
        ev = torch.cuda.Event()
        ev.record()
    return Snapshot(out, ev, bufs)


class CheckpointWriter:
    """One background thread; ``submit(fn)`` queues a write, ``join()`` waits for everything queued and re-raises the first
    failure.  At most ``max_pending`` snapshots are outstanding (a second checkpoint waits for the first)."""

    def __init__(self, max_pending: int = 1) -> None:
        self._q: "queue.Queue[Optional[Callable[[], None]]]" = queue.Queue()
        self._sem = threading.Semaphore(max_pending)
        self._err: Optional[BaseException] = None
        self._thread = threading.Thread(target=self._run, daemon=True, name="checkpoint-writer")
        self._thread.start()
        self.written = 0

    def _run(self) -> None:
        while True:
            fn = self._q.get()
            if fn is None:
                self._q.task_done()
                return
            try:
                fn()
                self.written += 1
            except BaseException as e:  # noqa: BLE001 - surfaced at join()
                logger.exception("background checkpoint failed")
                if self._err is None:
                    self._err = e
            finally:
                self._sem.release()
                self._q.task_done()

    def submit(self, fn: Callable[[], None]) -> None:
        self._sem.acquire()
        try:
            self._raise()
        except BaseException:
            self._sem.release()
            raise
        self._q.put(fn)

    def _raise(self) -> None:
        if self._err is not None:
            e, self._err = self._err, None
            raise RuntimeError("a background checkpoint failed") from e

    def join(self) -> None:
        self._q.join()
        self._raise()

    def close(self) -> None:
        self._q.join()
        self._q.put(None)
        self._thread.join(timeout=30)
        self._raise()
