import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "rocm-apex_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # GPU tests are selected with -m gpu; when no GPU is present they are skipped rather than failed
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _same_random_draws_every_run():
    """Every test starts from the same generator state, whatever ran before it: a few tests draw from torch's global
    generators, and a tolerance test whose data depends on the test order is a flake waiting for a different order."""
    import torch
    torch.manual_seed(20240607)      # seeds the CPU and (lazily) every CUDA generator
    yield
