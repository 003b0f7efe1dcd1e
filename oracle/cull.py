"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

numpy float32 restatement of the CULL ARGUMENT of the tile kernel
(weak-cube-r-cnn_b200/csrc/score_tiles.cuh: score_prep_kernel, score_zgrid_kernel). The
reference has no such code (the point-in-cube composite is project spec); what is
pinned here is a claim, checked against the C oracle's counts in tests/test_oracle_cull.py:

    a depth pixel can be inside a cube only if
      (1) its ROI column / row index lies in the cube's cull box (unclamped projected box of
          the 8 corners, grown by one pixel),
      (2) its depth lies in the corner depth range grown by 1 mm per metre,
    and both tests are waived when a corner is closer than 1 cm to the camera plane or a
    projection is not finite. A cube whose cull box and depth range miss every cell of a
    16 x 16 grid of per-cell depth ranges of the ROI holds no point.

The device code uses fmaf where this file uses separate multiplies and adds; the margins (one
pixel, 1 mm/m) are five orders of magnitude above that difference.
"""
from __future__ import annotations

import numpy as np

f32 = np.float32
_SIGNS = [(-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1),
          (-1, -1, 1), (1, -1, 1), (1, 1, 1), (-1, 1, 1)]   # math_util.py:198-217 corner order


def cull_record(cube: np.ndarray, K4: np.ndarray, roi, stride: int = 1):
    """-> (cl, ch, rl, rh, zlo, zhi) in ROI sub-sampled index units (inclusive), or None when the
    proposal is dead (non-positive size, or its box cannot reach the ROI)."""
    u0, v0, u1, v1 = roi
    ncols = max(0, (u1 - u0 + stride - 1) // stride)
    nrows = max(0, (v1 - v0 + stride - 1) // stride)
    cube = np.asarray(cube, dtype=f32)
    fx, fy, cx, cy = (f32(x) for x in K4)
    if nrows == 0 or ncols == 0 or not (cube[3] > 0 and cube[4] > 0 and cube[5] > 0):
        return None
    hx, hy, hz = cube[5] / f32(2), cube[4] / f32(2), cube[3] / f32(2)
    R = cube[6:].reshape(3, 3)
    px, py, zs = [], [], []
    with np.errstate(all="ignore"):
        for sx, sy, sz in _SIGNS:
            loc = np.array([f32(sx) * hx, f32(sy) * hy, f32(sz) * hz], dtype=f32)
            X = f32(R[0] @ loc) + cube[0]
            Y = f32(R[1] @ loc) + cube[1]
            Z = f32(R[2] @ loc) + cube[2]
            px.append(f32(fx * X + cx * Z) / Z)
            py.append(f32(fy * Y + cy * Z) / Z)
            zs.append(Z)
    px, py, zs = np.array(px, f32), np.array(py, f32), np.array(zs, f32)
    cl, ch, rl, rh = 0, ncols - 1, 0, nrows - 1
    zlo, zhi = -np.inf, np.inf
    finite = np.isfinite(px).all() and np.isfinite(py).all() and np.isfinite(zs).all()
    if finite and zs.min() >= f32(0.01):
        big = 1.0e6
        x_lo = int(np.floor(np.clip(px.min(), -big, big))) - 1 - u0
        x_hi = int(np.ceil(np.clip(px.max(), -big, big))) + 1 - u0
        y_lo = int(np.floor(np.clip(py.min(), -big, big))) - 1 - v0
        y_hi = int(np.ceil(np.clip(py.max(), -big, big))) + 1 - v0
        s = stride
        cl = max(cl, 0 if x_lo <= 0 else (x_lo + s - 1) // s)
        ch = min(ch, -1 if x_hi < 0 else x_hi // s)
        rl = max(rl, 0 if y_lo <= 0 else (y_lo + s - 1) // s)
        rh = min(rh, -1 if y_hi < 0 else y_hi // s)
        slack = 1.0e-3 * (abs(float(zs.min())) + abs(float(zs.max()))) + 1.0e-4
        zlo, zhi = float(zs.min()) - slack, float(zs.max()) + slack
    if cl > ch or rl > rh:
        return None
    return cl, ch, rl, rh, zlo, zhi


def zgrid(depth: np.ndarray, roi, stride: int = 1):
    """(16, 16, 2) min / max depth of the valid pixels of every cell of the sub-sampled ROI;
    empty cells hold (+inf, -inf). -> (grid, cell_w, cell_h)."""
    u0, v0, u1, v1 = roi
    sub = np.asarray(depth, dtype=f32)[v0:v1:stride, u0:u1:stride]
    nrows, ncols = sub.shape
    cw, chh = max(1, (ncols + 15) >> 4), max(1, (nrows + 15) >> 4)
    g = np.empty((16, 16, 2), dtype=np.float64)
    g[..., 0], g[..., 1] = np.inf, -np.inf
    for cyi in range(16):
        for cxi in range(16):
            t = sub[cyi * chh:(cyi + 1) * chh, cxi * cw:(cxi + 1) * cw]
            t = t[(t > 0) & np.isfinite(t)]
            if t.size:
                g[cyi, cxi] = (t.min(), t.max())
    return g, cw, chh


def may_hold_a_point(rec, grid, cw: int, chh: int) -> bool:
    if rec is None:
        return False
    cl, ch, rl, rh, zlo, zhi = rec
    sub = grid[rl // chh:min(15, rh // chh) + 1, cl // cw:min(15, ch // cw) + 1]
    return bool((~((zhi < sub[..., 0]) | (zlo > sub[..., 1]))).any())


def masked_depth(depth: np.ndarray, roi, rec, stride: int = 1) -> np.ndarray:
    """The depth map with every ROI pixel the cull record rules out set to 0 (= invalid)."""
    u0, v0, u1, v1 = roi
    out = np.array(depth, dtype=f32, copy=True)
    sub = out[v0:v1:stride, u0:u1:stride]          # view
    if rec is None:
        sub[...] = 0
        return out
    cl, ch, rl, rh, zlo, zhi = rec
    keep = np.zeros(sub.shape, dtype=bool)
    keep[rl:rh + 1, cl:ch + 1] = True
    keep &= (sub >= zlo) & (sub <= zhi)
    sub[~keep] = 0
    return out
