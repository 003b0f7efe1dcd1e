"""TEST INFRASTRUCTURE ONLY (oracle). Builds and loads oracle/csrc/wcr_oracle.c.

The shared object lands in oracle/_build/ (git-ignored, travels to the GPU box).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "csrc", "wcr_oracle.c")
_OUT_DIR = os.path.join(_HERE, "_build")
_OUT = os.path.join(_OUT_DIR, "libwcr_oracle.so")

_lib = None


def build(force: bool = False) -> str:
    os.makedirs(_OUT_DIR, exist_ok=True)
    if (not force and os.path.exists(_OUT)
            and os.path.getmtime(_OUT) >= os.path.getmtime(_SRC)):
        return _OUT
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-std=c11", "-ffp-contract=off",
           "-fno-fast-math", "-fopenmp", "-o", _OUT, _SRC, "-lm"]
    subprocess.check_call(cmd)
    return _OUT


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


def plane_hypotheses(pts: np.ndarray, ids: np.ndarray) -> np.ndarray:
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    ids = np.ascontiguousarray(ids, dtype=np.int64)
    eqs = np.empty((ids.shape[0], 4), dtype=np.float32)
    lib().orc_plane_hypotheses(_p(pts, ctypes.c_float), ctypes.c_int64(pts.shape[0]),
                               _p(ids, ctypes.c_int64), ctypes.c_int64(ids.shape[0]),
                               _p(eqs, ctypes.c_float))
    return eqs


def plane_counts(pts: np.ndarray, eqs: np.ndarray, thresh: float) -> np.ndarray:
    pts = np.ascontiguousarray(pts, dtype=np.float32)
    eqs = np.ascontiguousarray(eqs, dtype=np.float32)
    counts = np.empty(eqs.shape[0], dtype=np.int64)
    lib().orc_plane_counts(_p(pts, ctypes.c_float), ctypes.c_int64(pts.shape[0]),
                           _p(eqs, ctypes.c_float), ctypes.c_int64(eqs.shape[0]),
                           ctypes.c_float(thresh), _p(counts, ctypes.c_int64))
    return counts
