import os
import sys

import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 with `pytest -m gpu`)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")
    config.addinivalue_line("markers", "slow: large inputs (GiB scale)")


def pytest_collection_modifyitems(config, items):
    """GPU tests FAIL (not skip) when selected with -m gpu on a box without CUDA, so a silent CPU fallback is
    impossible; without -m gpu they are skipped on CPU-only machines."""
    import torch

    has_cuda = torch.cuda.is_available()
    ngpu = torch.cuda.device_count() if has_cuda else 0
    selected_gpu = "gpu" in (config.getoption("-m") or "")
    for item in items:
        if "gpu" in item.keywords and not has_cuda and not selected_gpu:
            item.add_marker(pytest.mark.skip(reason="no CUDA device"))
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(pytest.mark.skip(reason=f"needs >= 2 GPUs, have {ngpu}"))
