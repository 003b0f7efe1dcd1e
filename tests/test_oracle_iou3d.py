"""oracle/iou3d.py against closed forms and a Monte-Carlo estimate (pytorch3d itself is not
available: parity unpinned against it, see the oracle's header)."""
import numpy as np
import pytest

from oracle import iou3d as o3


def _rot(rng):
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    return q * np.sign(np.linalg.det(q))


def test_axis_aligned_closed_forms():
    I = np.eye(3)
    a = o3.cube_corners([0, 0, 5], (2, 2, 2), I)
    assert abs(o3.box_volume(a) - 8) < 1e-12
    for shift, want in [((0, 0, 0), 8.0), ((1, 0, 0), 4.0), ((1, 1, 1), 1.0), ((2, 0, 0), 0.0), ((3, 0, 0), 0.0)]:
        b = o3.cube_corners(np.array([0, 0, 5]) + shift, (2, 2, 2), I)
        vol, iou = o3.box3d_overlap_pair(a, b)
        assert abs(vol - want) < 1e-9, (shift, vol)
        assert abs(iou - want / (16 - want)) < 1e-9
    inner = o3.cube_corners([0.2, -0.1, 5.1], (0.5, 0.7, 0.9), _rot(np.random.default_rng(0)))
    vol, iou = o3.box3d_overlap_pair(a, inner)
    assert abs(vol - 0.5 * 0.7 * 0.9) < 1e-9                     # containment
    # shares a face plane (same normal) with the big box, sticks half out of it
    b = o3.cube_corners([0.5, 0, 5], (2, 2, 1), I)      # dims (w,h,l): l = 1 along x, spans x in [0, 1]
    vol, _ = o3.box3d_overlap_pair(a, b)
    assert abs(vol - 4.0) < 1e-9


@pytest.mark.parametrize("seed", range(6))
def test_random_pairs_against_monte_carlo(seed):
    rng = np.random.default_rng(seed)
    a = o3.cube_corners(rng.normal(size=3) * 0.3 + [0, 0, 4], rng.uniform(0.5, 2, 3), _rot(rng))
    b = o3.cube_corners(rng.normal(size=3) * 0.5 + [0, 0, 4], rng.uniform(0.5, 2, 3), _rot(rng))
    vol, iou = o3.box3d_overlap_pair(a, b)
    mc = o3.monte_carlo_volume(a, b, n=300000, seed=seed)
    assert abs(vol - mc) <= 0.02 * max(o3.box_volume(a), 1e-9) + 1e-3
    v2, i2 = o3.box3d_overlap_pair(b, a)                           # symmetry
    assert abs(vol - v2) < 1e-9 and abs(iou - i2) < 1e-9
