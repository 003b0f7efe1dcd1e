"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

Float64 numpy restatement of the intersection volume / IoU of two oriented boxes given as
8 corners in pytorch3d's order (pytorch3d/ops/iou_box3d.py: _box_planes), as the reference
uses it through box3d_overlap (ProposalNetwork/utils/utils.py:189-205, roi_heads.py:942).

Third-party dependency: pytorch3d (absent from /root/reference and from this image; the
reference's requirements.txt pins no version). Its published algorithm -- clip the faces of
each box by the planes of the other, sum the resulting polyhedron's volume -- is restated with
quads instead of triangles. Parity status: PARITY UNPINNED against pytorch3d itself (cannot be
run here); anchored instead on closed forms (axis-aligned overlaps, containment, identical and
face-sharing boxes) and on a float64 Monte-Carlo estimate for random oriented pairs
(tests/test_oracle_iou3d.py).
"""
from __future__ import annotations

import numpy as np

FACES = [[0, 1, 2, 3], [3, 2, 6, 7], [0, 1, 5, 4], [0, 3, 7, 4], [1, 2, 6, 5], [4, 5, 6, 7]]


def cube_corners(c, dims, R):
    """math_util.py:142-217 corner order: dims = (w, h, l); x <- l, y <- h, z <- w."""
    w, h, l = dims
    sx = np.array([-1, 1, 1, -1, -1, 1, 1, -1]) * l / 2
    sy = np.array([-1, -1, 1, 1, -1, -1, 1, 1]) * h / 2
    sz = np.array([-1, -1, -1, -1, 1, 1, 1, 1]) * w / 2
    return (np.asarray(R, float) @ np.stack((sx, sy, sz))).T + np.asarray(c, float)


def _planes(c):
    ctr = c.mean(0)
    out = []
    for f in FACES:
        v0, v1, v3 = c[f[0]], c[f[1]], c[f[3]]
        n = np.cross(v1 - v0, v3 - v0)
        ln = np.linalg.norm(n)
        n = n / ln if ln > 0 else n * 0
        if n @ (ctr - v0) > 0:
            n = -n
        out.append((n, n @ v0))
    return out


def _area(p):
    s = np.zeros(3)
    for i in range(1, len(p) - 1):
        s += np.cross(p[i] - p[0], p[i + 1] - p[0])
    return 0.5 * np.linalg.norm(s)


def box_volume(c):
    ctr = c.mean(0)
    return sum(_area([c[k] for k in f]) * (d - n @ ctr) for f, (n, d) in zip(FACES, _planes(c))) / 3


def _clip(poly, n, d):
    out = []
    for k in range(len(poly)):
        s, e = poly[k], poly[(k + 1) % len(poly)]
        ds, de = n @ s - d, n @ e - d
        if ds <= 0:
            out.append(s)
            if de > 0:
                out.append(s + (e - s) * (ds / (ds - de)))
        elif de <= 0:
            out.append(s + (e - s) * (ds / (ds - de)))
    return out


def box3d_overlap_pair(c1, c2):
    c1, c2 = np.asarray(c1, float), np.asarray(c2, float)
    p1, p2 = _planes(c1), _planes(c2)
    ctr1 = c1.mean(0)
    eps = 1e-9 * max(np.linalg.norm(c1[6] - c1[0]), np.linalg.norm(c2[6] - c2[0]))

    def same(f1, g2):      # face f1 of box1 in the plane g2 of box2, same outward normal
        (n1, _), (n2, d2) = p1[f1], p2[g2]
        return n1 @ n2 > 0.999 and max(abs(n2 @ c1[k] - d2) for k in FACES[f1]) <= eps

    vol = 0.0
    for f in range(6):
        poly = [c1[k] for k in FACES[f]]
        for g in range(6):
            if same(f, g) or not poly:
                continue
            poly = _clip(poly, *p2[g])
        if len(poly) >= 3:
            vol += _area(poly) * (p1[f][1] - p1[f][0] @ ctr1) / 3
    for f in range(6):
        if any(same(g, f) for g in range(6)):
            continue
        poly = [c2[k] for k in FACES[f]]
        for g in range(6):
            if not poly:
                break
            poly = _clip(poly, *p1[g])
        if len(poly) >= 3:
            vol += _area(poly) * (p2[f][1] - p2[f][0] @ ctr1) / 3
    v1, v2 = box_volume(c1), box_volume(c2)
    vol = min(max(vol, 0.0), min(v1, v2))
    return vol, vol / (v1 + v2 - vol)


def box3d_overlap(b1, b2):
    vol = np.zeros((len(b1), len(b2)))
    iou = np.zeros_like(vol)
    for i, a in enumerate(b1):
        for j, b in enumerate(b2):
            vol[i, j], iou[i, j] = box3d_overlap_pair(a, b)
    return vol, iou


def monte_carlo_volume(c1, c2, n=400000, seed=0):
    """Independent check: fraction of uniform samples of box1 that fall inside box2."""
    c1, c2 = np.asarray(c1, float), np.asarray(c2, float)
    rng = np.random.default_rng(seed)
    o, ex, ey, ez = c1[0], c1[1] - c1[0], c1[3] - c1[0], c1[4] - c1[0]
    u = rng.random((n, 3))
    pts = o + u[:, :1] * ex + u[:, 1:2] * ey + u[:, 2:] * ez
    inside = np.ones(n, bool)
    for nrm, d in _planes(c2):
        inside &= pts @ nrm <= d
    return inside.mean() * box_volume(c1)
