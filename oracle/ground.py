"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

CPU restatement (numpy float32, every rounding explicit) of north_star step 1:
depth -> point cloud unprojection, RANSAC plane fit, up-vector fix-ups.

Follows
  cubercnn/modeling/roi_heads/roi_heads.py:466-528  normal_vector_from_maps
  ProposalNetwork/utils/plane.py:79-134             Plane_torch.fit_parallel
  cubercnn/modeling/roi_heads/roi_heads.py:918-928  K-principal-point lifting (A1.4)

Parity status: PINNED -- tests/test_oracle_ground.py checks this file against
the reference's own functions executed in-process (when /root/reference is
present) and against tests/golden/ground_*.npz generated from them.
"""
from __future__ import annotations

import numpy as np

f32 = np.float32


def unproject_ground_frame(depth: np.ndarray, ground: np.ndarray | None, K: np.ndarray,
                           use_nth: int = 5):
    """roi_heads.py:474-502. Returns (points (M,3) f32 in row-major scan order,
    flat indices (M,) into the strided grid).

    Quirks kept: fx = K[0,0] // use_nth (float floor-div, :479); principal point
    is the centre of the *strided* grid (W'/2, H'/2), not K's (:482);
    x = ((u - cx) * z) / fx evaluated left to right (:484).
    ``ground`` None == the (1,1) dummy: every pixel of the un-padded image (:492-499).
    """
    depth = np.asarray(depth, dtype=f32)
    z = depth[::use_nth, ::use_nth]
    hh, ww = z.shape
    fx = f32(np.floor(f32(K[0, 0]) / f32(use_nth)))
    fy = f32(np.floor(f32(K[1, 1]) / f32(use_nth)))
    u = np.arange(ww, dtype=np.int64)[None, :].repeat(hh, 0)
    v = np.arange(hh, dtype=np.int64)[:, None].repeat(ww, 1)
    cx, cy = ww / 2, hh / 2  # python floats; torch promotes int64 - float -> f32
    x = ((u.astype(f32) - f32(cx)) * z) / fx
    y = ((v.astype(f32) - f32(cy)) * z) / fy
    if ground is not None:
        g = np.asarray(ground)[::use_nth, ::use_nth] > 0
    else:
        g = np.ones_like(z, dtype=bool)
    idx = np.flatnonzero(g.reshape(-1))
    pts = np.stack((x.reshape(-1)[idx], y.reshape(-1)[idx], z.reshape(-1)[idx]), axis=-1)
    return pts.astype(f32), idx


def unproject_K_frame(depth: np.ndarray, K: np.ndarray, stride: int = 1):
    """A1.4: X = z*(u - K02)/K00, Y = z*(v - K12)/K11 (roi_heads.py:927-928
    operation order: (z * (u - cx)) / fx). Invalid depth (<=0 or non-finite)
    -> NaN point. Returns (H', W', 3) f32."""
    depth = np.asarray(depth, dtype=f32)
    z = depth[::stride, ::stride]
    hh, ww = z.shape
    u = (np.arange(ww, dtype=np.int64) * stride).astype(f32)[None, :]
    v = (np.arange(hh, dtype=np.int64) * stride).astype(f32)[:, None]
    with np.errstate(invalid="ignore"):
        x = (z * (u - f32(K[0, 2]))) / f32(K[0, 0])
        y = (z * (v - f32(K[1, 2]))) / f32(K[1, 1])
    out = np.stack((x, y, z), axis=-1).astype(f32)
    bad = ~(np.isfinite(z) & (z > 0))
    out[bad] = np.nan
    return out


def plane_hypotheses(pts: np.ndarray, id_samples: np.ndarray):
    """plane.py:96-113. pts (N,3) f32, id_samples (iters,3) int -> eqs (iters,4) f32.

    Rounding order (established by running the reference on torch-CPU, see
    oracle/csrc/wcr_oracle.c): cross = fma(a1,b2,-(a2*b1)); norm =
    sqrt(fma(z,z,fma(y,y,x*x))); k = -((c0*q0 + c1*q1) + c2*q2) with q = sample
    point **1** (:112)."""
    from . import clib
    return clib.plane_hypotheses(pts, id_samples)


def plane_inlier_counts(pts: np.ndarray, eqs: np.ndarray, thresh: float = 0.05):
    """plane.py:117-126. dist = (((a*x + b*y) + c*z) + d) / sqrt((a^2+b^2)+c^2);
    inlier iff |dist| <= thresh (compared in f32). Returns int64 (iters,)."""
    from . import clib
    return clib.plane_counts(pts, eqs, thresh)


def plane_inlier_mask(pts: np.ndarray, eq: np.ndarray, thresh: float = 0.05):
    a, b, c, d = (f32(v) for v in eq)
    with np.errstate(invalid="ignore", divide="ignore"):
        den = np.sqrt(f32(f32(a * a + b * b) + c * c))
        dist = (((a * pts[:, 0] + b * pts[:, 1]) + c * pts[:, 2]) + d) / den
    return np.abs(dist) <= f32(thresh)


def fit_parallel(pts: np.ndarray, id_samples: np.ndarray, thresh: float = 0.05):
    """plane.py:79-134 with the random.sample draws passed in (replay mode).
    Returns (-eq (4,), inlier indices int64, counts)."""
    eqs = plane_hypotheses(pts, id_samples)
    counts = plane_inlier_counts(pts, eqs, thresh)
    best = int(np.argmax(counts))  # first max, plane.py:128
    eq = eqs[best]
    inl = np.flatnonzero(plane_inlier_mask(pts, eq, thresh))
    return (-eq).astype(f32), inl.astype(np.int64), counts


def fix_up_normal(n: np.ndarray) -> np.ndarray:
    """roi_heads.py:510-525: wall -> floor axis swaps, then force +y."""
    n = np.asarray(n, dtype=f32).copy()
    if abs(n[2]) > abs(n[1]):
        n = np.array([n[0], n[2], -n[1]], dtype=f32)
    if abs(n[0]) > abs(n[1]):
        n = np.array([-n[2], n[0], n[1]], dtype=f32)
    if n[1] < 0:
        n = -n
    return n


def normal_vector_from_maps(grounds, depths, Ks, id_samples_per_image, use_nth: int = 5):
    """roi_heads.py:466-528, replay mode: ``id_samples_per_image[i]`` are the
    (1000,3) triples the reference drew with random.sample for image i.
    ``grounds[i]`` None or shape (1,1) == missing ground map."""
    out = []
    for g, d, K, ids in zip(grounds, depths, Ks, id_samples_per_image):
        if g is not None and tuple(np.shape(g)) == (1, 1):
            g = None
        pts, _ = unproject_ground_frame(d, g, np.asarray(K), use_nth)
        eq, _, _ = fit_parallel(pts, np.asarray(ids), 0.05)
        out.append(fix_up_normal(eq[:3]))
    return np.stack(out).astype(f32)


def draw_id_samples(n_points: int, iters: int, rng):
    """The reference's draw (plane.py:95): ``rng`` is a ``random.Random`` (or the
    ``random`` module) so a seeded stream reproduces the reference exactly."""
    return np.array([rng.sample(range(0, n_points), 3) for _ in range(iters)], dtype=np.int64)
