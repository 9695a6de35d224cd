"""``core.ProfilerContext`` — system-metrics sampling shipped as ``ReportTrialMetrics`` groups.

Parity: reference ``harness/determined/core/_profiler.py:61-531``: ``on(sampling_interval,
samples_per_report)``/``off()`` ``:61-140``; collectors Network ``:246-275``, Disk ``:278-326``,
Memory ``:329-339``, CPU ``:342-352``, GPU via pynvml ``:355-407``; collector/shipper threads
``:422-531``; metrics keyed by agent id and grouped as ``gpu|cpu|memory|disk|network``.

B200 additions (SURVEY §5.1 "new build"): :class:`StepTimer` — device-side per-step timing
with CUDA events (no host sync inside the step), used by the trial controller and ``bench.py``,
and SM-clock/throttle sampling next to utilisation so a throughput number can be qualified.
"""
import collections
import datetime
import logging
import queue
import threading
import time
from typing import Any, Deque, Dict, List, Optional

logger = logging.getLogger("determined.core")


class _MetricGroupCollector:
    group = ""

    def collect(self) -> Dict[str, Any]:
        raise NotImplementedError


class _Network(_MetricGroupCollector):
    group = "network"

    def __init__(self) -> None:
        import psutil

        self._ps = psutil
        self._last = psutil.net_io_counters()
        self._t = time.time()

    def collect(self) -> Dict[str, Any]:
        now, cur = time.time(), self._ps.net_io_counters()
        dt = max(now - self._t, 1e-6)
        out = {"net_throughput_sent": (cur.bytes_sent - self._last.bytes_sent) / dt,
               "net_throughput_recv": (cur.bytes_recv - self._last.bytes_recv) / dt}
        self._last, self._t = cur, now
        return out


class _Disk(_MetricGroupCollector):
    group = "disk"

    def __init__(self, paths: Optional[List[str]] = None) -> None:
        import psutil

        self._ps = psutil
        self._paths = paths or ["/"]
        self._last = psutil.disk_io_counters()
        self._t = time.time()

    def collect(self) -> Dict[str, Any]:
        now, cur = time.time(), self._ps.disk_io_counters()
        dt = max(now - self._t, 1e-6)
        out: Dict[str, Any] = {}
        if cur is not None and self._last is not None:
            out.update({"disk_iops": ((cur.read_count + cur.write_count) - (self._last.read_count + self._last.write_count)) / dt,
                        "disk_throughput_read": (cur.read_bytes - self._last.read_bytes) / dt,
                        "disk_throughput_write": (cur.write_bytes - self._last.write_bytes) / dt})
        self._last, self._t = cur, now
        for p in self._paths:
            try:
                out.setdefault("disk_util", {})[p] = self._ps.disk_usage(p).percent
            except OSError:
                pass
        return out


class _Memory(_MetricGroupCollector):
    group = "memory"

    def collect(self) -> Dict[str, Any]:
        import psutil

        return {"memory_free": psutil.virtual_memory().available / 1e9}


class _CPU(_MetricGroupCollector):
    group = "cpu"

    def collect(self) -> Dict[str, Any]:
        import psutil

        return {"cpu_util_simple": psutil.cpu_percent()}


class _GPU(_MetricGroupCollector):
    group = "gpu"

    def __init__(self) -> None:
        self._nv = None
        self._handles: Dict[str, Any] = {}
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            for i in range(pynvml.nvmlDeviceGetCount()):
                h = pynvml.nvmlDeviceGetHandleByIndex(i)
                uuid = pynvml.nvmlDeviceGetUUID(h)
                self._handles[uuid.decode() if isinstance(uuid, bytes) else uuid] = h
        except Exception as e:
            logger.debug("pynvml unavailable (%s); GPU metrics disabled", e)

    def collect(self) -> Dict[str, Any]:
        if self._nv is None:
            return {}
        nv = self._nv
        out: Dict[str, Any] = {}
        for uuid, h in self._handles.items():
            try:
                out[uuid] = {
                    "gpu_util": nv.nvmlDeviceGetUtilizationRates(h).gpu,
                    "gpu_free_memory": nv.nvmlDeviceGetMemoryInfo(h).free,
                    "gpu_sm_clock_mhz": nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM),
                    "gpu_power_w": nv.nvmlDeviceGetPowerUsage(h) / 1000.0,
                    "gpu_throttle_reasons": int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                    if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") else 0,
                }
            except Exception:
                continue
        return out


def _mean_nested(samples: List[Dict[str, Any]]) -> Dict[str, Any]:
    out: Dict[str, Any] = {}
    keys = set().union(*[s.keys() for s in samples]) if samples else set()
    for k in keys:
        vals = [s[k] for s in samples if k in s]
        if vals and isinstance(vals[0], dict):
            out[k] = _mean_nested(vals)
        else:
            nums = [v for v in vals if isinstance(v, (int, float))]
            if nums:
                out[k] = sum(nums) / len(nums)
    return out


class _Collector(threading.Thread):
    def __init__(self, collectors: List[_MetricGroupCollector], out: "queue.Queue[Any]",
                 sampling_interval: float, samples_per_report: int) -> None:
        super().__init__(daemon=True, name="profiler-collector")
        self._collectors, self._out = collectors, out
        self._interval, self._n = sampling_interval, max(1, samples_per_report)
        self._quit = threading.Event()

    def run(self) -> None:
        acc: Dict[str, List[Dict[str, Any]]] = collections.defaultdict(list)
        n = 0
        while not self._quit.wait(self._interval):
            for c in self._collectors:
                try:
                    s = c.collect()
                    if s:
                        acc[c.group].append(s)
                except Exception as e:  # never kill training because a sensor failed
                    logger.debug("collector %s failed: %s", c.group, e)
            n += 1
            if n >= self._n:
                ts = datetime.datetime.now(datetime.timezone.utc)
                for g, samples in acc.items():
                    self._out.put((g, _mean_nested(samples), ts))
                acc.clear()
                n = 0
        self._out.put(None)

    def stop(self) -> None:
        self._quit.set()


class _Shipper(threading.Thread):
    def __init__(self, train: Any, agent_id: str, inq: "queue.Queue[Any]") -> None:
        super().__init__(daemon=True, name="profiler-shipper")
        self._train, self._agent, self._q = train, agent_id, inq

    def run(self) -> None:
        while True:
            item = self._q.get()
            if item is None:
                return
            group, metrics, ts = item
            try:
                self._train.report_metrics(group=group, steps_completed=0, metrics={self._agent: metrics}, report_time=ts)
            except Exception as e:
                logger.debug("failed to ship profiler metrics: %s", e)


class ProfilerContext:
    """Collect system metrics while training.  Only the local chief of each node samples."""

    def __init__(self, agent_id: str, distributed: Any, train: Any) -> None:
        self._agent_id = agent_id
        self._distributed = distributed
        self._train = train
        self._collector: Optional[_Collector] = None
        self._shipper: Optional[_Shipper] = None
        self._on = False

    def on(self, sampling_interval: int = 1, samples_per_report: int = 10) -> None:
        if self._on:
            raise RuntimeError("profiler is already on; call .off() first")
        if sampling_interval < 0.1:
            raise ValueError("sampling_interval must be >= 0.1 s")
        if samples_per_report < 1:
            raise ValueError("samples_per_report must be >= 1")
        self._on = True
        if self._distributed.local_rank != 0:
            return
        q: "queue.Queue[Any]" = queue.Queue()
        cols: List[_MetricGroupCollector] = []
        for cls in (_GPU, _CPU, _Memory, _Disk, _Network):
            try:
                cols.append(cls())
            except Exception as e:
                logger.debug("profiler collector %s unavailable: %s", cls.__name__, e)
        self._collector = _Collector(cols, q, sampling_interval, samples_per_report)
        self._shipper = _Shipper(self._train, self._agent_id, q)
        self._collector.start()
        self._shipper.start()

    def off(self) -> None:
        if not self._on:
            return
        self._on = False
        if self._collector is not None:
            self._collector.stop()
            self._collector.join(timeout=10)
            self._collector = None
        if self._shipper is not None:
            self._shipper.join(timeout=10)
            self._shipper = None

    def _close(self) -> None:
        self.off()


class DummyProfilerContext(ProfilerContext):
    def __init__(self) -> None:
        self._on = False

    def on(self, sampling_interval: int = 1, samples_per_report: int = 10) -> None:
        self._on = True

    def off(self) -> None:
        self._on = False

    def _close(self) -> None:
        pass


class StepTimer:
    """Device-side step timing: ``start()``/``stop()`` record CUDA events on the current stream;
    nothing synchronises until :meth:`summary`.  On CPU falls back to ``perf_counter``."""

    def __init__(self, window: int = 1024) -> None:
        self._pairs: Deque[Any] = collections.deque(maxlen=window)
        self._cur: Any = None
        try:
            import torch

            self._torch = torch if torch.cuda.is_available() else None
        except ImportError:
            self._torch = None

    def start(self) -> None:
        if self._torch is not None:
            ev = self._torch.cuda.Event(enable_timing=True)
            ev.record()
            self._cur = ev
        else:
            self._cur = time.perf_counter()

    def stop(self) -> None:
        if self._cur is None:
            return
        if self._torch is not None:
            ev = self._torch.cuda.Event(enable_timing=True)
            ev.record()
            self._pairs.append((self._cur, ev))
        else:
            self._pairs.append((self._cur, time.perf_counter()))
        self._cur = None

    def summary(self) -> Dict[str, float]:
        if not self._pairs:
            return {}
        if self._torch is not None:
            self._torch.cuda.synchronize()
            ms = [a.elapsed_time(b) for a, b in self._pairs]
        else:
            ms = [(b - a) * 1e3 for a, b in self._pairs]
        ms_sorted = sorted(ms)
        return {"step_ms_mean": sum(ms) / len(ms), "step_ms_p50": ms_sorted[len(ms) // 2],
                "step_ms_max": ms_sorted[-1], "steps": float(len(ms))}
