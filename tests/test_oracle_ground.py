"""Pin oracle/ground.py (step 1 restatement) to the reference: committed goldens
(always) and the reference's own code run live (authoring container, `ref`)."""
import glob
import os
import random

import numpy as np
import pytest
import torch

from oracle import ground as og, refshim

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "ground_*.npz")))


@pytest.mark.parametrize("path", GOLDEN)
def test_ground_oracle_matches_reference_golden(path):
    z = np.load(path)
    normals = []
    for i in range(int(z["n"])):
        depth, ground, K, ids = z[f"depth{i}"], z[f"ground{i}"], z[f"K{i}"], z[f"ids{i}"]
        g = None if ground.shape == (1, 1) else ground
        pts, _ = og.unproject_ground_frame(depth, g, K, 5)
        assert np.array_equal(pts, z[f"pts{i}"])
        eq, inl, _ = og.fit_parallel(pts, ids.astype(np.int64), 0.05)
        assert np.array_equal(eq, z[f"eq{i}"])          # bit-exact plane (negated, plane.py:134)
        assert np.array_equal(inl, z[f"inl{i}"])        # same inlier set
        normals.append(og.fix_up_normal(eq[:3]))
    assert np.array_equal(np.stack(normals), z["normals"])


@pytest.mark.ref
def test_ground_oracle_matches_reference_live():
    ns = refshim.load()
    g = torch.Generator().manual_seed(3)
    pts = torch.randn(2500, 3, generator=g) * torch.tensor([2.0, 0.04, 3.0]) + torch.tensor([0.0, 1.3, 4.0])
    random.seed(9)
    eq_ref, inl_ref = ns.plane.Plane_torch().fit_parallel(pts, thresh=0.05, maxIteration=1000)
    random.seed(9)
    ids = og.draw_id_samples(2500, 1000, random)
    eq, inl, _ = og.fit_parallel(pts.numpy(), ids)
    assert np.array_equal(eq, eq_ref.numpy()) and np.array_equal(inl, inl_ref.numpy())


def test_fix_up_normal_cases():
    assert np.array_equal(og.fix_up_normal(np.array([0.1, 0.2, 0.9], np.float32)),
                          np.array([0.1, 0.9, -0.2], np.float32))          # back wall
    assert np.array_equal(og.fix_up_normal(np.array([0.9, 0.2, 0.1], np.float32)),
                          np.array([-0.1, 0.9, 0.2], np.float32))          # side wall
    assert np.array_equal(og.fix_up_normal(np.array([0.0, -1.0, 0.0], np.float32)),
                          np.array([0.0, 1.0, 0.0], np.float32))           # flipped
