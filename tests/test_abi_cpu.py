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
    assert lib.wcr_abi_version() == 2


def test_struct_sizes_match_header():
    from wcr import _abi
    from wcr.images import IMAGE_DTYPE
    from wcr.proposals import BOX_DTYPE
    assert IMAGE_DTYPE.itemsize == 48 and BOX_DTYPE.itemsize == 32
    assert ctypes.sizeof(_abi.SampleParams) == 48
    assert ctypes.sizeof(_abi.ScoreParams) == 32
    lib = _abi.load()
    # sizeof() as compiled into the library vs the ctypes / numpy mirrors
    assert lib.wcr_sizeof(0) == IMAGE_DTYPE.itemsize and lib.wcr_sizeof(1) == BOX_DTYPE.itemsize
    assert lib.wcr_sizeof(2) == ctypes.sizeof(_abi.SampleParams)
    assert lib.wcr_sizeof(3) == ctypes.sizeof(_abi.ScoreParams)
    assert lib.wcr_sizeof(4) == ctypes.sizeof(_abi.LossParams) == 48
    assert lib.wcr_sizeof(5) == ctypes.sizeof(_abi.LossIO)
    assert lib.wcr_sizeof(99) == 0


def test_z_loss_steps_equal_torch_linspace():
    """The table the fused 'z' term uses == values_tensor of z_loss (roi_heads.py:547)."""
    import numpy as np
    import torch
    from wcr import _abi
    out = np.zeros(50, dtype=np.float32)
    _abi.load().wcr_z_loss_steps(out.ctypes.data)
    assert np.array_equal(out, torch.linspace(0.0, (50 - 1) / 10, 50).numpy())


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
