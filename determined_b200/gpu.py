"""GPU discovery for slot detection and the profiler.

Parity: reference ``harness/determined/gpu.py:18-190`` and ``agent/internal/detect/nvidia.go:19-117``
(``nvidia-smi --query-gpu=index,name,uuid``).  B200-first: NVML (pynvml) is the primary source
because it is a library call rather than a subprocess; ``nvidia-smi`` CSV is the fallback; there is
no ROCm branch (single-vendor by design).
"""
import csv
import io
import logging
import subprocess
from typing import List, NamedTuple, Optional, Tuple

logger = logging.getLogger("determined.gpu")


class GPU(NamedTuple):
    id: int
    uuid: str
    name: str = ""
    load: float = 0.0
    memoryUtil: float = 0.0
    memoryTotalMiB: float = 0.0
    sm_clock_mhz: float = 0.0


class GPUProcess(NamedTuple):
    pid: int
    process_name: str
    gpu_uuid: str
    used_memory: str


def _nvml_gpus() -> Optional[List[GPU]]:
    try:
        import pynvml  # provided by nvidia_ml_py

        pynvml.nvmlInit()
    except Exception:
        return None
    out = []
    try:
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            uuid = pynvml.nvmlDeviceGetUUID(h)
            name = pynvml.nvmlDeviceGetName(h)
            if isinstance(uuid, bytes):
                uuid = uuid.decode()
            if isinstance(name, bytes):
                name = name.decode()
            try:
                util = pynvml.nvmlDeviceGetUtilizationRates(h).gpu / 100.0
                mem = pynvml.nvmlDeviceGetMemoryInfo(h)
                clk = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                out.append(GPU(i, uuid, name, util, mem.used / max(mem.total, 1), mem.total / 2**20, clk))
            except Exception:
                out.append(GPU(i, uuid, name))
    except Exception as e:  # pragma: no cover - depends on driver state
        logger.debug("nvml enumeration failed: %s", e)
        return None
    return out


def _smi_gpus() -> List[GPU]:
    try:
        proc = subprocess.run(
            ["nvidia-smi", "--query-gpu=index,uuid,name,utilization.gpu,memory.used,memory.total,clocks.sm",
             "--format=csv,noheader,nounits"],
            stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=20, text=True,
        )
    except (FileNotFoundError, subprocess.TimeoutExpired):
        return []
    if proc.returncode != 0:
        return []
    gpus = []
    for row in csv.reader(io.StringIO(proc.stdout)):
        if len(row) < 3:
            continue
        f = [c.strip() for c in row]

        def num(s: str) -> float:
            try:
                return float(s)
            except ValueError:
                return 0.0

        total = num(f[5]) if len(f) > 5 else 0.0
        gpus.append(GPU(int(f[0]), f[1], f[2], num(f[3]) / 100.0 if len(f) > 3 else 0.0,
                        (num(f[4]) / total) if total else 0.0, total, num(f[6]) if len(f) > 6 else 0.0))
    return gpus


def get_gpus() -> Tuple[List[GPU], str]:
    gpus = _nvml_gpus()
    if gpus is None:
        gpus = _smi_gpus()
    return gpus, ("cuda" if gpus else "cpu")


def get_gpu_uuids() -> List[str]:
    return [g.uuid for g in get_gpus()[0]]


def get_gpu_processes() -> List[GPUProcess]:
    try:
        proc = subprocess.run(
            ["nvidia-smi", "--query-compute-apps=pid,process_name,gpu_uuid,used_memory", "--format=csv,noheader"],
            stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=20, text=True,
        )
    except (FileNotFoundError, subprocess.TimeoutExpired):
        return []
    res = []
    for row in csv.reader(io.StringIO(proc.stdout)):
        if len(row) == 4:
            res.append(GPUProcess(int(row[0]), row[1].strip(), row[2].strip(), row[3].strip()))
    return res
