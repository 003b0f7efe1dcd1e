"""Pins oracle/hull.py (C restatement of jarvis_march / fill_polygon / mask IoUs,
ProposalNetwork/utils/utils.py:225-245, 373-493) against the REFERENCE: live when
/root/reference is present, and against tests/golden/hull_*.npz (the reference's
outputs on its own five fixture point sets, utils.py:500-544, and on seeded random
clamped corner sets) everywhere."""
import glob
import os

import numpy as np
import pytest

from oracle import hull as oh, refshim

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
CASES = sorted(glob.glob(os.path.join(GOLDEN, "hull_*.npz")))


def test_golden_present():
    assert len(CASES) >= 2


@pytest.mark.parametrize("path", CASES)
def test_hull_matches_reference_golden(path):
    z = np.load(path)
    H, W = int(z["H"]), int(z["W"])
    n = int(z["n_sets"])
    for i in range(n):
        pts = z[f"pts_{i}"]
        hull = oh.jarvis_march(pts)
        assert np.array_equal(hull, z[f"hull_{i}"])                     # bit-exact vertices, same order
        soft = oh.fill_polygon(hull, H, W)
        want = np.unpackbits(z[f"maskbits_{i}"])[:H * W].reshape(H, W).astype(bool)
        assert np.array_equal(soft > 0, want)                           # binarised fill
        # soft values: stored at the pixels where they are fractional
        frac = z[f"frac_idx_{i}"]
        assert np.array_equal(np.flatnonzero((soft > 0) & (soft < 1)), frac)
        assert np.array_equal(soft.reshape(-1)[frac], z[f"frac_val_{i}"])
        seg = z[f"seg_{i}"]
        np.testing.assert_allclose(oh.mask_iou(seg, soft), z[f"miou_{i}"], rtol=1e-6)
        np.testing.assert_allclose(oh.mod_mask_iou(seg, soft), z[f"modiou_{i}"], rtol=1e-5)


@pytest.mark.skipif(not refshim.available(), reason="needs /root/reference")
def test_hull_matches_reference_live():
    import torch
    ns = refshim.load()
    U = ns.utils
    g = np.random.default_rng(7)
    H, W = 96, 128
    for t in range(60):
        pts = g.uniform(-30, 160, size=(8, 2)).astype(np.float32)
        pts[:, 0] = np.clip(pts[:, 0], 0, W - 1)
        pts[:, 1] = np.clip(pts[:, 1], 0, H - 1)
        if t % 3 == 0:
            pts[g.integers(0, 8)] = pts[g.integers(0, 8)]              # forced duplicate
        ref_hull = U.jarvis_march(torch.from_numpy(pts.copy())).numpy()
        hull = oh.jarvis_march(pts)
        assert np.array_equal(hull, ref_hull)
        ref_mask = U.fill_polygon(torch.zeros(H, W), torch.from_numpy(ref_hull)).numpy()
        assert np.array_equal(oh.fill_polygon(hull, H, W), ref_mask)    # soft mask bit-exact
        assert np.array_equal(oh.convex_hull((H, W), pts), U.convex_hull(torch.zeros(H, W), torch.from_numpy(pts.copy())).numpy())
        seg = (g.uniform(size=(H, W)) < 0.3)
        a = U.mask_iou(torch.from_numpy(seg), ref_mask)
        np.testing.assert_allclose(oh.mask_iou(seg, ref_mask), float(a), rtol=1e-6)
        b = U.mod_mask_iou(torch.from_numpy(seg).float(), ref_mask)
        np.testing.assert_allclose(oh.mod_mask_iou(seg, ref_mask), float(b), rtol=1e-5)


def test_seg_term_equals_full_image_fill():
    """The composite's row-restricted 'all edges strictly left' rule == binarised
    reference fill over the whole image, on non-degenerate hulls."""
    g = np.random.default_rng(3)
    H, W = 96, 128
    for t in range(40):
        corners = g.uniform(-20, 150, size=(8, 2)).astype(np.float32)
        seg = np.zeros((H, W), dtype=np.uint8)
        roi = (20, 10, 100, 80)
        seg[roi[1]:roi[3], roi[0]:roi[2]] = g.uniform(size=(roi[3] - roi[1], roi[2] - roi[0])) < 0.5
        s, inter, bube = oh.seg_term(corners, (W, H), seg, roi, 0)   # clamp tuple: x by clamp0
        c2 = corners.copy()
        c2[:, 0] = np.clip(c2[:, 0], 0, W - 1)
        c2[:, 1] = np.clip(c2[:, 1], 0, H - 1)
        full = oh.convex_hull((H, W), c2) > 0
        assert bube == int(full.sum())
        assert inter == int((full & (seg > 0)).sum())
        want = oh.mask_iou(seg, full)
        assert s == want


def test_fast_seg_term_equals_definition():
    """orc_seg_term_fast (row intervals + prefix sums; the timed CPU baseline) ==
    orc_seg_term (per-pixel definition) through score_box, bit for bit."""
    import torch
    from oracle import cubes as oc
    g = torch.Generator().manual_seed(9)
    H, W = 96, 128
    depth = (1.0 + 3.0 * torch.rand(H, W, generator=g)).numpy()
    K = np.array([[110.0, 0, 64], [0, 110, 48], [0, 0, 1]], np.float32)
    N = 300
    attrs = np.zeros((7, N), np.float32)
    attrs[0] = 64 + 40 * torch.randn(N, generator=g).numpy()
    attrs[1] = 48 + 30 * torch.randn(N, generator=g).numpy()
    attrs[2] = 1.0 + 3 * torch.rand(N, generator=g).numpy()
    attrs[3:6] = 0.2 + 1.5 * torch.rand(3, N, generator=g).numpy()
    attrs[6] = np.pi * torch.rand(N, generator=g).numpy()
    cubes = oc.assemble_cubes(K, np.array([0.0, 1.0, 0.0], np.float32), attrs)
    gbox = np.array([20.3, 10.1, 100.9, 80.2], np.float32)
    seg = np.zeros((H, W), np.uint8)
    seg[15:70, 30:90] = (torch.rand(55, 60, generator=g) < 0.6).numpy()
    for mode in (0, 1):
        for clamp in ((H, W), (W, H)):
            a = oc.score_box(depth, K, gbox, clamp, cubes, 1, 1, 1, 1, mode, seg)
            b = oc.score_box(depth, K, gbox, clamp, cubes, 1, 1, 1, 1, mode, seg, fast_seg=True)
            assert np.array_equal(a["seg"], b["seg"]) and np.array_equal(a["score"], b["score"])
            assert (a["seg"] > 0).any()
