import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs TWO CUDA devices (NCCL refuses two ranks on one "
                                       "device); run with -m multigpu on a 2-GPU box, not part of -m gpu")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def _gpu_count():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    n = _gpu_count()
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    skip2 = pytest.mark.skip(reason="needs two CUDA devices")
    for it in items:
        if "gpu" in it.keywords and n < 1:
            it.add_marker(skip)
        if "multigpu" in it.keywords and n < 2:
            it.add_marker(skip2)
