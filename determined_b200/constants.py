"""Defaults and well-known locations.  Parity: reference ``harness/determined/constants.py:1-60``.

Single-box design: there are no containers, so the "well-known" paths are rooted at
``$DET_RUN_DIR`` (default ``/run/determined``) which the launcher points at a per-allocation
scratch directory.
"""
import os

MAX_SLOTS_PER_AGENT = 16
SLOTS_PER_BOX = 8  # one 8xB200 node

DEFAULT_SEARCHER_CFG = {"name": "single", "max_length": {"batches": 100}}
DEFAULT_RESOURCES_CFG = {"slots_per_trial": 1, "native_parallel": False}
DEFAULT_SCHEDULING_UNIT = 100
DEFAULT_OPTIMIZATIONS = {
    "aggregation_frequency": 1,
    "average_aggregated_gradients": True,
    "average_training_metrics": True,
    "gradient_compression": False,
    "mixed_precision": "O0",
    "tensor_fusion_threshold": 64,
    "tensor_fusion_cycle_time": 1,
    "auto_tune_tensor_fusion": False,
}
DEFAULT_EXP_CFG = {
    "searcher": DEFAULT_SEARCHER_CFG,
    "scheduling_unit": DEFAULT_SCHEDULING_UNIT,
    "resources": DEFAULT_RESOURCES_CFG,
    "optimizations": DEFAULT_OPTIMIZATIONS,
}

DTRAIN_SSH_PORT = int(os.getenv("DTRAIN_SSH_PORT", "12350"))
INTER_TRAIN_PROCESS_COMM_PORT_1 = int(os.getenv("INTER_TRAIN_PROCESS_COMM_PORT_1", "12360"))
INTER_TRAIN_PROCESS_COMM_PORT_2 = int(
    os.getenv("INTER_TRAIN_PROCESS_COMM_PORT_2", str(INTER_TRAIN_PROCESS_COMM_PORT_1 + MAX_SLOTS_PER_AGENT))
)
C10D_PORT = int(os.getenv("C10D_PORT", "29400"))

RUN_DIR = os.getenv("DET_RUN_DIR", "/run/determined")
INFO_DIR = os.path.join(RUN_DIR, "info")
CONTAINER_STDOUT = os.path.join(RUN_DIR, "train/logs/stdout.log")
CONTAINER_STDERR = os.path.join(RUN_DIR, "train/logs/stderr.log")
MANAGED_TRAINING_MODEL_COPY = os.path.join(RUN_DIR, "train/model")

SHARED_FS_CONTAINER_PATH = "/determined_shared_fs"
DEFAULT_CHECKPOINT_PATH = os.path.expanduser(os.getenv("DET_CHECKPOINT_DIR", "~/.local/share/determined_b200/checkpoints"))
DEFAULT_MASTER_PORT = 8080

# kept for configs and scripts that read them; the launch layer here has no Horovod underneath
HOROVOD_STARTUP_TIMEOUT_SECONDS = 1200
HOROVOD_GLOO_TIMEOUT_SECONDS = 240
HOROVOD_AUTOTUNE_LOG_FILEPATH = "/tmp/autotune_log.csv"
