"""The cull argument of the tile kernel, checked on the CPU against the C oracle's counts
(oracle/cull.py restates it): removing every pixel the cull record rules out must not change a
single point-in-cube count, and a proposal the depth grid declares empty must count 0. Cubes:
the sampler's kind plus adversarial ones (arbitrary rotations, near the camera plane, huge, tiny)."""
import numpy as np
import pytest
import torch

from oracle import cubes as oc, cull


def _cubes(rng, im, box, n):
    H, W = im["depth"].shape
    f = float(im["K"][0, 0])
    c = np.zeros((n, 15), dtype=np.float32)
    u = rng.uniform(box[0] - 40, box[2] + 40, n)
    v = rng.uniform(box[1] - 40, box[3] + 40, n)
    z = np.exp(rng.uniform(-2.5, 2.7, n))                     # 0.08 .. 15 m
    c[:, 0], c[:, 1], c[:, 2] = z * (u - W / 2) / f, z * (v - H / 2) / f, z
    c[:, 3:6] = np.exp(rng.uniform(-4.0, 2.0, (n, 3)))        # 0.02 .. 7.4 m
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    w_, x_, y_, z_ = q.T
    c[:, 6:] = np.stack([1 - 2 * (y_ * y_ + z_ * z_), 2 * (x_ * y_ - z_ * w_), 2 * (x_ * z_ + y_ * w_),
                         2 * (x_ * y_ + z_ * w_), 1 - 2 * (x_ * x_ + z_ * z_), 2 * (y_ * z_ - x_ * w_),
                         2 * (x_ * z_ - y_ * w_), 2 * (y_ * z_ + x_ * w_), 1 - 2 * (x_ * x_ + y_ * y_)], 1)
    half = n // 2                                             # the other half: yaw-only, sampler-like
    yaw = rng.uniform(0, np.pi, half)
    c[:half, 6:] = np.stack([np.cos(yaw), 0 * yaw, np.sin(yaw), 0 * yaw, 1 + 0 * yaw, 0 * yaw,
                             -np.sin(yaw), 0 * yaw, np.cos(yaw)], 1)
    c[0, 0:3], c[0, 3:6] = (0.0, 0.0, 0.5), 3.0               # swallows the camera
    c[1, 2] = -1.0                                            # behind the camera
    c[2, 2], c[2, 3:6] = 0.0105, 0.001                        # a hair in front of the 1 cm guard
    return c


@pytest.mark.parametrize("stride", [1, 2])
def test_cull_never_removes_an_inside_point(stride):
    from wcr import synth
    rng = np.random.default_rng(3)
    H, W = 200, 280
    checked = culled_some = proved_empty = inside_total = 0
    for i in (21, 22):
        im = synth.make_image(i, H, W, n_boxes=3, dummy_ground=False)
        depth = im["depth"].numpy().copy()
        depth[::9, ::7] *= 0.15                               # a lattice of near points
        K = im["K"].numpy()
        K4 = oc.K4_of(K)
        for b in range(3):
            box = im["boxes"][b].numpy()
            roi = oc.box_roi(box, H, W)
            grid, cw, chh = cull.zgrid(depth, roi, stride)
            cubes = _cubes(rng, {"depth": depth, "K": K}, box, 60)
            full = oc.score_box(depth, K, box, (H, W), cubes, stride)["count"]
            for j in range(len(cubes)):
                rec = cull.cull_record(cubes[j], K4, roi, stride)
                masked = cull.masked_depth(depth, roi, rec, stride)
                got = oc.score_box(masked, K, box, (H, W), cubes[j:j + 1], stride)["count"][0]
                assert got == full[j], (i, b, j, rec)
                if not cull.may_hold_a_point(rec, grid, cw, chh):
                    assert full[j] == 0, (i, b, j, rec)
                    proved_empty += 1
                culled_some += int((masked != depth).any())
                inside_total += int(full[j])
                checked += 1
    assert checked == 360 and culled_some > 200 and proved_empty > 20 and inside_total > 2000
