"""Pin oracle/cubes.py (steps 2+3 blocks) to the reference's own outputs
(tests/golden/cubes_0.npz from orthobasis_from_normal_t, Cubes.get_all_corners /
get_bube_corners, cubes_to_box, iou_2d) and check the composite's invariants."""
import os

import numpy as np
import pytest
import torch

from oracle import clib, cubes as oc, refshim

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "cubes_0.npz"))
RTOL = 1e-5


def test_orthobasis_matches_reference():
    K = np.eye(3, dtype=np.float32)
    N = len(G["yaw"])
    attrs = np.zeros((7, N), np.float32)
    attrs[2] = 1.0; attrs[3:6] = 1.0; attrs[6] = G["yaw"]
    for normal, want in ((G["normal"], G["R"]), (np.array([0, 1, 0], np.float32), G["R0"])):
        cubes = oc.assemble_cubes(K, normal, attrs)
        np.testing.assert_allclose(cubes[:, 6:].reshape(N, 3, 3), want, rtol=0, atol=3e-7)


def test_projection_box_iou_match_reference():
    cubes, K = G["cubes"], G["K"]
    I, N = cubes.shape[:2]
    for i in range(I):
        for j in range(N):
            box, corners = oc.project_cube(cubes[i, j], K, (480, 640))
            np.testing.assert_allclose(corners, G["bube"][i, j], rtol=RTOL, atol=1e-3)
            np.testing.assert_allclose(box, G["boxes"][i, j], rtol=RTOL, atol=1e-3)
        depth = np.ones((480, 640), np.float32)
        r = oc.score_box(depth, K, G["gt"][i], (480, 640), cubes[i], e_pts=0)
        np.testing.assert_allclose(r["iou"], G["ious"][i], rtol=RTOL, atol=1e-7)
        np.testing.assert_allclose(r["score"], G["ious"][i], rtol=RTOL, atol=1e-7)   # e_pts=0: score == IoU


def test_point_in_cube_vs_plain_composition():
    """The kernel's factored unit-cube test vs the textbook q = R^T (p - c), |q| <= half:
    counts may differ only on points within a few ulp of a face (north_star's allowance)."""
    g = torch.Generator().manual_seed(5)
    H, W = 60, 80
    depth = (1.0 + 3.0 * torch.rand(H, W, generator=g)).numpy()
    depth[3, 4] = 0.0
    K = np.array([[70.0, 0, 40], [0, 70, 30], [0, 0, 1]], np.float32)
    N = 64
    attrs = np.zeros((7, N), np.float32)
    attrs[0] = 40 + 10 * torch.randn(N, generator=g).numpy(); attrs[1] = 30 + 8 * torch.randn(N, generator=g).numpy()
    attrs[2] = 2.5 + torch.rand(N, generator=g).numpy()
    attrs[3:6] = 0.3 + 1.5 * torch.rand(3, N, generator=g).numpy()
    attrs[6] = np.pi * torch.rand(N, generator=g).numpy()
    normal = np.array([0.02, 0.999, 0.03], np.float32); normal /= np.linalg.norm(normal)
    cubes = oc.assemble_cubes(K, normal, attrs)
    gbox = np.array([5.3, 4.1, 70.9, 55.2], np.float32)
    r = oc.score_box(depth, K, gbox, (H, W), cubes)
    v, u = np.meshgrid(np.arange(4, 55), np.arange(5, 70), indexing="ij")
    z = depth[v, u].astype(np.float64)
    ok = z > 0
    P = np.stack((z * (u - 40.0) / 70.0, z * (v - 30.0) / 70.0, z), -1)[ok]
    assert r["npts"] == ok.sum()
    for j in range(N):
        c = cubes[j].astype(np.float64)
        q = (P - c[:3]) @ c[6:].reshape(3, 3)
        half = np.array([c[5], c[4], c[3]]) / 2
        m = np.abs(q) / half
        inside = (m <= 1).all(1).sum()
        near = ((np.abs(m - 1) < 1e-5).any(1)).sum()
        assert abs(int(r["count"][j]) - int(inside)) <= near
    assert r["count"].max() > 50


def test_topk_order_and_padding():
    s = np.array([0.5, 0.9, 0.9, 0.1, 0.0, 0.9], np.float32)
    assert list(oc.topk(s, 4)) == [1, 2, 5, 0]
    assert list(oc.topk(s[:2], 4)) == [1, 0, -1, -1]


def test_philox_known_answer_vectors():
    """Random123 kat_vectors, philox4x32-10."""
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        assert tuple(int(x) for x in clib.philox4x32_10(ctr, key)) == want


@pytest.mark.ref
def test_blocks_match_reference_live():
    """Same checks against the reference imported live (other seed than the golden)."""
    ns = refshim.load()
    g = torch.Generator().manual_seed(77)
    N = 25
    normal = torch.nn.functional.normalize(torch.tensor([-0.1, 0.97, 0.08]), dim=0)
    yaw = torch.rand(N, generator=g) * 3.14159
    R = ns.utils.orthobasis_from_normal_t(normal, yaw)
    attrs = np.zeros((7, N), np.float32)
    attrs[0] = 300 + 50 * torch.randn(N, generator=g).numpy(); attrs[1] = 250 + 40 * torch.randn(N, generator=g).numpy()
    attrs[2] = 2 + 3 * torch.rand(N, generator=g).numpy()
    attrs[3:6] = 0.3 + torch.rand(3, N, generator=g).numpy(); attrs[6] = yaw.numpy()
    K = torch.tensor([[500.0, 0, 310], [0, 505, 245], [0, 0, 1]])
    cubes = oc.assemble_cubes(K.numpy(), normal.numpy(), attrs)
    np.testing.assert_allclose(cubes[:, 6:].reshape(N, 3, 3), R.numpy(), rtol=0, atol=3e-7)
    C = ns.spaces.Cubes(torch.from_numpy(cubes).unsqueeze(0))
    want = ns.conversions.cubes_to_box(C, K, [480, 640])[0].tensor.numpy()
    for j in range(N):
        box, _ = oc.project_cube(cubes[j], K.numpy(), (480, 640))
        np.testing.assert_allclose(box, want[j], rtol=RTOL, atol=1e-3)
