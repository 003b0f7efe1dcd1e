import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "weak-cube-r-cnn_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "ref: needs /root/reference (authoring container only)")


def pytest_collection_modifyitems(config, items):
    from oracle import refshim
    skip_ref = pytest.mark.skip(reason="/root/reference not present (GPU box)")
    for item in items:
        if "ref" in item.keywords and not refshim.available():
            item.add_marker(skip_ref)


@pytest.fixture(scope="session")
def cuda_lib():
    """The product library on a real GPU. Fails loudly if it is not built."""
    import torch
    from wcr import _abi
    assert torch.cuda.is_available(), "gpu test without a GPU"
    lib = _abi.load()
    _abi.check(lib.wcr_device_check(), "wcr_device_check")
    return lib
