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
           "-fno-fast-math", "-fopenmp", "-mfma", "-mavx2", "-o", _OUT, _SRC, "-lm"]
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


# ---- steps 2+3 --------------------------------------------------------------
c_f, c_i = ctypes.c_float, ctypes.c_int


def box_center_depth(depth: np.ndarray, box: np.ndarray) -> np.float32:
    depth = np.ascontiguousarray(depth, dtype=np.float32)
    box = np.ascontiguousarray(box, dtype=np.float32)
    fn = lib().orc_box_center_depth
    fn.restype = c_f
    return np.float32(fn(_p(depth, c_f), c_i(depth.shape[0]), c_i(depth.shape[1]),
                         c_i(depth.shape[1]), _p(box, c_f)))


def assemble_cubes(K4: np.ndarray, normal: np.ndarray, attrs: np.ndarray) -> np.ndarray:
    """attrs (7, N) = u v z w h l yaw -> cubes (N, 15)."""
    K4 = np.ascontiguousarray(K4, dtype=np.float32)
    normal = np.ascontiguousarray(normal, dtype=np.float32)
    attrs = np.ascontiguousarray(attrs, dtype=np.float32)
    N = attrs.shape[1]
    out = np.empty((N, 15), dtype=np.float32)
    lib().orc_assemble_cubes(_p(K4, c_f), _p(normal, c_f), _p(attrs, c_f), c_i(N), _p(out, c_f))
    return out


def project_cube(cube: np.ndarray, K4: np.ndarray, clamp0: float, clamp1: float):
    cube = np.ascontiguousarray(cube, dtype=np.float32)
    K4 = np.ascontiguousarray(K4, dtype=np.float32)
    box = np.empty(4, dtype=np.float32)
    corners = np.empty((8, 2), dtype=np.float32)
    lib().orc_project_cube(_p(cube, c_f), _p(K4, c_f), c_f(clamp0), c_f(clamp1), _p(box, c_f),
                           _p(corners, c_f))
    return box, corners


def score_box(depth: np.ndarray, K4: np.ndarray, gbox: np.ndarray, clamp0: float, clamp1: float,
              cubes: np.ndarray, stride: int = 1, e_iou: int = 1, e_pts: int = 1, e_seg: int = 0,
              seg_mode: int = 0, seg: np.ndarray | None = None, fast_seg: bool = False):
    """-> dict(score (N,), iou (N,), count (N,) i32, proj (N,4), seg (N,), npts int).
    ``seg``: full-image (H, W) u8 mask of THIS box (only its integer ROI is read).
    ``fast_seg``: evaluate the mask term with orc_seg_term_fast (same result; what the
    timed CPU baseline uses)."""
    depth = np.ascontiguousarray(depth, dtype=np.float32)
    K4 = np.ascontiguousarray(K4, dtype=np.float32)
    gbox = np.ascontiguousarray(gbox, dtype=np.float32)
    cubes = np.ascontiguousarray(cubes, dtype=np.float32)
    N = cubes.shape[0]
    score = np.empty(N, dtype=np.float32)
    iou = np.empty(N, dtype=np.float32)
    count = np.empty(N, dtype=np.int32)
    proj = np.empty((N, 4), dtype=np.float32)
    sseg = np.ones(N, dtype=np.float32)
    if seg is not None:
        seg = np.ascontiguousarray(seg, dtype=np.uint8)
        assert seg.shape == depth.shape
    fn = lib().orc_score_box_seg
    fn.restype = c_i
    npts = fn(_p(depth, c_f), c_i(depth.shape[0]), c_i(depth.shape[1]), c_i(depth.shape[1]),
              _p(K4, c_f), _p(gbox, c_f), c_f(clamp0), c_f(clamp1), _p(cubes, c_f), c_i(N),
              c_i(stride), c_i(e_iou), c_i(e_pts), c_i(e_seg if seg is not None else 0),
              c_i(seg_mode | (256 if fast_seg else 0)), _p(seg, ctypes.c_uint8) if seg is not None else None,
              c_i(depth.shape[1]), _p(score, c_f), _p(iou, c_f), _p(count, ctypes.c_int32),
              _p(proj, c_f), _p(sseg, c_f))
    return {"score": score, "iou": iou, "count": count, "proj": proj, "seg": sseg,
            "npts": int(npts)}


def jarvis_march(points: np.ndarray):
    """utils.py:410-454 -> hull (m,2) f32 in the reference's (flipped) order, or None
    when the march does not close."""
    pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 2)
    hull = np.empty((pts.shape[0], 2), dtype=np.float32)
    fn = lib().orc_jarvis_march
    fn.restype = c_i
    m = fn(_p(pts, c_f), c_i(pts.shape[0]), _p(hull, c_f))
    return None if m < 0 else hull[:m].copy()


def fill_polygon(hull: np.ndarray, H: int, W: int) -> np.ndarray:
    """utils.py:456-488 -> soft mask (H, W) f32."""
    hull = np.ascontiguousarray(hull, dtype=np.float32).reshape(-1, 2)
    out = np.empty((H, W), dtype=np.float32)
    lib().orc_fill_polygon(_p(hull, c_f), c_i(hull.shape[0]), c_i(H), c_i(W), _p(out, c_f))
    return out


def seg_term(corners: np.ndarray, clamp0: float, clamp1: float, seg: np.ndarray, roi, mode: int = 0):
    """-> (s_seg f32, inter, bube) for ONE cube's 8 projected corners (8,2)."""
    corners = np.ascontiguousarray(corners, dtype=np.float32).reshape(16)
    seg = np.ascontiguousarray(seg, dtype=np.uint8)
    roi = np.ascontiguousarray(roi, dtype=np.int32)
    inter, bube = ctypes.c_int(0), ctypes.c_int(0)
    fn = lib().orc_seg_term
    fn.restype = c_f
    v = fn(_p(corners, c_f), c_f(clamp0), c_f(clamp1), _p(seg, ctypes.c_uint8), c_i(seg.shape[0]),
           c_i(seg.shape[1]), c_i(seg.shape[1]), _p(roi, ctypes.c_int32), c_i(-1), c_i(mode),
           ctypes.byref(inter), ctypes.byref(bube))
    return np.float32(v), inter.value, bube.value


def philox4x32_10(ctr, key) -> np.ndarray:
    ctr = np.ascontiguousarray(ctr, dtype=np.uint32)
    key = np.ascontiguousarray(key, dtype=np.uint32)
    out = np.empty(4, dtype=np.uint32)
    lib().orc_philox4x32_10(_p(ctr, ctypes.c_uint32), _p(key, ctypes.c_uint32),
                            _p(out, ctypes.c_uint32))
    return out


def set_num_threads(n: int | None = None) -> int:
    """Use ``n`` OpenMP threads (default: every core this process may run on)."""
    if n is None:
        n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().orc_set_num_threads(c_i(int(n)))
    return num_threads()


def num_threads() -> int:
    fn = lib().orc_num_threads
    fn.restype = c_i
    return int(fn())
