import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    import torch
    has_gpu = torch.cuda.is_available()
    ngpu = torch.cuda.device_count() if has_gpu else 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(skip_multi)


@pytest.fixture(scope="session")
def reference_modules():
    """The reference's own compression.py / utils.py, imported read-only at test time when the
    reference tree is present (it is not on the GPU box)."""
    if not os.path.isdir(REFERENCE):
        pytest.skip("reference tree not available")
    import importlib
    sys.path.insert(0, REFERENCE)
    try:
        saved = {k: sys.modules.pop(k) for k in ("compression", "utils", "settings") if k in sys.modules}
        comp = importlib.import_module("compression")
        utils = importlib.import_module("utils")
        return comp, utils
    finally:
        sys.path.remove(REFERENCE)
