"""No GPU needed: the C-ABI library builds, loads and exports every symbol that
include/wcr.h declares; the ctypes mirrors of the structs have the C sizes; calls
fail loudly (no CPU fallback) when there is no B200."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "wcr.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(wcr_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from wcr import _abi
    lib = ctypes.CDLL(_abi.LIB_PATH)
    names = _declared()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/wcr.h but not exported"
    assert set(names) == set(_abi.SIGNATURES), (set(names) ^ set(_abi.SIGNATURES))
    assert lib.wcr_abi_version() == 1


def test_struct_sizes_match_header():
    from wcr import _abi
    from wcr.images import IMAGE_DTYPE
    from wcr.proposals import BOX_DTYPE
    assert IMAGE_DTYPE.itemsize == 48 and BOX_DTYPE.itemsize == 32
    assert ctypes.sizeof(_abi.SampleParams) == 48
    assert ctypes.sizeof(_abi.ScoreParams) == 32
    assert ctypes.sizeof(_abi.LossParams) == 40
    assert ctypes.sizeof(_abi.LossIO) == 8 + 8 * len(_abi._LOSS_IO_PTRS)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from wcr import _abi, synth
    from wcr.images import ImageShard
    from wcr.plane import Plane_torch
    lib = _abi.load()
    assert lib.wcr_device_check() != 0          # no sm_100 device here
    im = synth.make_image(0, 60, 80, n_boxes=1)
    with pytest.raises(_abi.WcrError):
        ImageShard([im["depth"]], None, [im["K"]])
    with pytest.raises(_abi.WcrError):
        Plane_torch().fit_parallel(torch.randn(100, 3))


def test_product_path_never_imports_oracle():
    pkg = os.path.join(ROOT, "weak-cube-r-cnn_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, os.path.join(dp, f)
