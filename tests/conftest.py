import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "gpu_parallel: needs >= 2 GPUs")
    config.addinivalue_line("markers", "slow: long-running")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
        ngpu = torch.cuda.device_count() if has_gpu else 0
    except Exception:
        has_gpu, ngpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_par = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)
        if "gpu_parallel" in item.keywords and ngpu < 2:
            item.add_marker(skip_par)


@pytest.fixture(autouse=True)
def _isolate_env(tmp_path, monkeypatch):
    # keep tests hermetic: checkpoints and auth caches go to tmp
    monkeypatch.setenv("DET_CONFIG_DIR", str(tmp_path / "cfg"))
    monkeypatch.setenv("DET_INFO_DIR", str(tmp_path / "info"))
    yield
