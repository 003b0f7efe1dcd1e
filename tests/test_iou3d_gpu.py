"""3D IoU on the GPU through the drop-in names (wcr.utils.iou_3d /
box3d_overlap) against oracle/iou3d.py (float64). Tolerance: 2e-4 of the smaller box's volume
(fp32 clipping of O(1)-sized boxes), IoU 2e-4 absolute."""
import numpy as np
import pytest
import torch

from oracle import iou3d as o3

pytestmark = pytest.mark.gpu


def _rot(rng, det=1.0):
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    return q * np.sign(np.linalg.det(q)) * det


def _cubes(rng, n, spread=0.6, det=1.0):
    return [(rng.normal(size=3) * spread + [0, 0, 4], rng.uniform(0.4, 2.0, 3), _rot(rng, det)) for _ in range(n)]


def test_box3d_overlap_matches_oracle(cuda_lib):
    from wcr.utils import box3d_overlap
    rng = np.random.default_rng(0)
    A = np.stack([o3.cube_corners(*c) for c in _cubes(rng, 6)])
    B = np.stack([o3.cube_corners(*c) for c in _cubes(rng, 40)])
    vol, iou = box3d_overlap(torch.tensor(A, dtype=torch.float32).cuda(), torch.tensor(B, dtype=torch.float32).cuda())
    wv, wi = o3.box3d_overlap(A, B)
    vmin = np.minimum(np.array([o3.box_volume(a) for a in A])[:, None], np.array([o3.box_volume(b) for b in B])[None])
    assert np.all(np.abs(vol.cpu().numpy() - wv) <= 2e-4 * vmin + 1e-6)
    assert np.all(np.abs(iou.cpu().numpy() - wi) <= 2e-4)
    assert (wi > 0.05).sum() > 20 and (wi == 0).sum() > 5        # both regimes are exercised


def test_closed_forms_and_special_positions(cuda_lib):
    from wcr.utils import box3d_overlap
    I = np.eye(3)
    a = o3.cube_corners([0, 0, 5], (2, 2, 2), I)
    others = [o3.cube_corners(np.array([0, 0, 5.0]) + s, d, I) for s, d in
              [((0, 0, 0), (2, 2, 2)), ((1, 0, 0), (2, 2, 2)), ((1, 1, 1), (2, 2, 2)), ((2, 0, 0), (2, 2, 2)),
               ((5, 0, 0), (2, 2, 2)), ((0.5, 0, 0), (2, 2, 1)), ((0.1, 0.2, -0.1), (0.5, 0.5, 0.5))]]
    want = np.array([8.0, 4.0, 1.0, 0.0, 0.0, 4.0, 0.125])
    vol, iou = box3d_overlap(torch.tensor(a[None], dtype=torch.float32).cuda(),
                             torch.tensor(np.stack(others), dtype=torch.float32).cuda())
    assert np.allclose(vol.cpu().numpy()[0], want, atol=2e-5)
    assert abs(float(iou[0, 0]) - 1.0) < 1e-5
    with pytest.raises(ValueError):                                # pytorch3d rejects flat boxes
        flat = o3.cube_corners([0, 0, 5], (0.0, 1, 1), I)
        box3d_overlap(torch.tensor(a[None], dtype=torch.float32).cuda(), torch.tensor(flat[None], dtype=torch.float32).cuda())


def test_iou_3d_drop_in_on_cubes(cuda_lib):
    """utils.py:189-205 signature: one GT cube vs P proposal cubes (the reference's det -1 bases too)."""
    from wcr.spaces import Cubes
    from wcr.utils import iou_3d
    rng = np.random.default_rng(3)
    gt = _cubes(rng, 1, det=-1.0)[0]
    props = _cubes(rng, 64, spread=0.4, det=-1.0)
    def row(c):
        return np.concatenate((c[0], c[1], np.asarray(c[2]).reshape(-1)))
    G = Cubes(torch.tensor(row(gt)[None, None], dtype=torch.float32).cuda())
    P = Cubes(torch.tensor(np.stack([row(c) for c in props])[None], dtype=torch.float32).cuda())
    got = iou_3d(G, P).cpu().numpy()
    want = np.array([o3.box3d_overlap_pair(o3.cube_corners(*gt), o3.cube_corners(*c))[1] for c in props])
    assert got.shape == (64,) and np.all(np.abs(got - want) <= 2e-4)
