"""Round-2 block-level drop-ins on the GPU (through the C ABI) against tests/golden/blocks2_0.npz
-- outputs and AUTOGRAD gradients of the reference's own functions -- and against oracle/blocks2.py:
custom_mapping, randn_orthobasis_torch, gt_in_norm_range, is_gt_included, differentiable
Cubes.get_all_corners / get_bube_corners / cubes_to_box, stand-alone z_loss / depth_range_loss /
normal_to_rotation, cube_to_box, get_cuboids_verts_faces."""
import os

import numpy as np
import pytest
import torch

from oracle import blocks2 as ob, loss_inputs as li

pytestmark = pytest.mark.gpu
Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "blocks2_0.npz"))


def _grad_close(a, b, rtol=1e-4, tiny=1e-3):
    a, b = a.reshape(-1).double(), b.reshape(-1).double()
    scale = b.abs().max().item()
    assert (a - b).abs().max().item() <= rtol * scale
    big = b.abs() >= tiny * scale
    assert ((a - b).abs()[big] / b.abs()[big]).max().item() <= rtol


def test_custom_mapping(cuda_lib):
    from wcr.utils import custom_mapping
    for bi, beta in enumerate(Z["cm_betas"]):
        got = custom_mapping([float(v) for v in Z["cm_x"]], float(beta))
        assert isinstance(got, list) and len(got) == len(Z["cm_x"])
        np.testing.assert_allclose(np.asarray(got), Z["cm_out"][bi], rtol=1e-6, atol=1e-12)
        t = custom_mapping(torch.from_numpy(Z["cm_x"]).cuda(), float(beta))
        assert t.is_cuda and np.allclose(t.cpu().numpy(), Z["cm_out"][bi], rtol=1e-6, atol=1e-12)
    assert custom_mapping([0.0, -3.0], 1.7) == [0.0, 0.0]


def test_randn_orthobasis_replay_and_native(cuda_lib):
    from wcr.utils import randn_orthobasis_torch
    z = torch.from_numpy(Z["ro_z"]).cuda()
    got = randn_orthobasis_torch(7, 3, z=z)
    assert tuple(got.shape) == (3, 7, 3, 3)
    np.testing.assert_allclose(got.cpu().numpy(), Z["ro_out"], rtol=0, atol=3e-7)
    a = randn_orthobasis_torch(1000, 2, seed=4)
    b = randn_orthobasis_torch(1000, 2, seed=4)
    c = randn_orthobasis_torch(1000, 2, seed=5)
    assert torch.equal(a, b) and not torch.equal(a, c)
    R = a.reshape(-1, 3, 3)
    assert torch.allclose(R @ R.transpose(1, 2), torch.eye(3, device="cuda").expand_as(R), atol=1e-5)
    assert abs(float(R[:, 0, 0].mean())) < 0.05 and 0.28 < float((R[:, 0, 0] ** 2).mean()) < 0.39   # isotropic


def test_gt_in_norm_range(cuda_lib):
    from wcr.utils import gt_in_norm_range
    lo, hi, gt = (torch.from_numpy(Z[k]).cuda() for k in ("gn_lo", "gn_hi", "gn_gt"))
    got = gt_in_norm_range((lo, hi), gt)
    assert np.array_equal(got.cpu().numpy(), Z["gn_out"])
    one = gt_in_norm_range((float(lo[3]), float(hi[3])), gt[3:4])
    assert float(one) == float(Z["gn_out"][3])


def test_is_gt_included(cuda_lib, capsys):
    from wcr.utils import is_gt_included, is_gt_included_batch
    n = len(Z["ig_included"])
    cubes = torch.zeros(n, 15)
    cubes[:, 0:3] = torch.from_numpy(Z["ig_center"]); cubes[:, 3:6] = torch.from_numpy(Z["ig_dims"])
    r, p = torch.from_numpy(Z["ig_ranges"]), torch.from_numpy(Z["ig_priors"])
    bits, by = is_gt_included_batch(cubes.cuda(), r[:, 0].cuda(), r[:, 1].cuda(), r[:, 2].cuda(),
                                    p[:, 0].cuda(), p[:, 1].cuda(), p[:, 2].cuda())
    assert np.array_equal(bits.cpu().numpy() == 0, Z["ig_included"])
    for i in range(n):
        wb, wby = ob.is_gt_included_bits(Z["ig_center"][i], Z["ig_dims"][i], *Z["ig_ranges"][i], *Z["ig_priors"][i])
        assert int(bits[i]) == wb and np.allclose(by[i].cpu().numpy(), wby, rtol=0, atol=0)
    import types
    i = int(np.flatnonzero(~Z["ig_included"])[0])
    cube = types.SimpleNamespace(center=torch.from_numpy(Z["ig_center"][i]), dimensions=torch.from_numpy(Z["ig_dims"][i]))
    assert is_gt_included(cube, r[i, 0], r[i, 1], r[i, 2], p[i, 0], p[i, 1], p[i, 2]) is False
    assert "GT cannot be found due to" in capsys.readouterr().out
    j = int(np.flatnonzero(Z["ig_included"])[0])
    cube = types.SimpleNamespace(center=torch.from_numpy(Z["ig_center"][j]), dimensions=torch.from_numpy(Z["ig_dims"][j]))
    assert is_gt_included(cube, r[j, 0], r[j, 1], r[j, 2], p[j, 0], p[j, 1], p[j, 2]) is True


def test_cubes_geometry_is_differentiable_like_the_reference(cuda_lib):
    """Gradients the reference's autograd produced through get_all_corners / get_bube_corners /
    cubes_to_box (spaces.py:192-245, conversions.py:25-48) vs the backward kernels."""
    from wcr.spaces import Cubes
    from wcr.conversions import cubes_to_box, cube_to_box
    cubes = torch.from_numpy(Z["ag_cubes"]).cuda()
    K, clamp = torch.from_numpy(Z["ag_K"]), [int(v) for v in Z["ag_clamp"]]
    ct = cubes.clone().requires_grad_(True)
    (Cubes(ct).get_all_corners() * torch.from_numpy(Z["ag_up3"]).cuda()).sum().backward()
    _grad_close(ct.grad.cpu(), torch.from_numpy(Z["ag_g3"]))
    ct = cubes.clone().requires_grad_(True)
    (Cubes(ct).get_bube_corners(K, clamp) * torch.from_numpy(Z["ag_up2"]).cuda()).sum().backward()
    _grad_close(ct.grad.cpu(), torch.from_numpy(Z["ag_g2"]))
    assert float(ct.grad[0, 0].abs().sum()) >= 0.0
    ct = cubes.clone().requires_grad_(True)
    boxes = torch.stack([b.tensor for b in cubes_to_box(Cubes(ct), K, clamp)])
    (boxes * torch.from_numpy(Z["ag_up4"]).cuda()).sum().backward()
    _grad_close(ct.grad.cpu(), torch.from_numpy(Z["ag_g4"]))
    # a leaf BEHIND further torch ops keeps its graph (the call site builds cubes with torch.cat)
    base = cubes.clone().requires_grad_(True)
    (Cubes(base * 1.0).get_all_corners()).sum().backward()
    assert base.grad is not None and float(base.grad.abs().sum()) > 0
    with pytest.raises(NameError):                 # the reference's cube_to_box calls get_bube_corners(K) without clamp
        cube_to_box(Cubes(cubes[0, 0]), K)
    verts, faces = Cubes(cubes).get_cuboids_verts_faces()
    assert tuple(verts.shape) == (2, 9, 8, 3) and tuple(faces.shape) == (2, 9, 12, 3)
    assert faces[1, 3].tolist()[:2] == [[0.0, 1.0, 2.0], [2.0, 3.0, 0.0]]


def _sa_inputs():
    seed, nb, H, W = int(Z["sa_seed"]), [int(v) for v in Z["sa_nb"]], int(Z["sa_H"]), int(Z["sa_W"])
    return li.add_masks(li.make_loss_inputs(seed, nb, H=H, W=W), seed, H=H, W=W), nb


def test_standalone_z_loss_and_normal_to_rotation(cuda_lib):
    from wcr.roi_heads import z_loss, normal_to_rotation
    from wcr.spaces import Cubes
    from wcr.conversions import Boxes
    inp, nb = _sa_inputs()
    cubes = torch.from_numpy(Z["sa_cubes"]).cuda()
    n = cubes.shape[0]
    K4 = inp["K4"]
    Ks = torch.zeros(n, 3, 3)
    Ks[:, 0, 0], Ks[:, 1, 1], Ks[:, 0, 2], Ks[:, 1, 2], Ks[:, 2, 2] = K4[:, 0], K4[:, 1], K4[:, 2], K4[:, 3], 1
    im_sizes = [[int(c[0]), int(c[1])] for c in inp["clamp"]]
    got = z_loss(Boxes(inp["gt_boxes"].cuda()), Cubes(cubes), [Ks[i] for i in range(n)], im_sizes,
                 Boxes(torch.from_numpy(Z["sa_proj"]).cuda()))
    assert np.array_equal(got.cpu().numpy(), Z["sa_z_loss"])
    rot = normal_to_rotation(torch.from_numpy(Z["sa_normals"]).cuda())
    np.testing.assert_allclose(rot.cpu().numpy(), Z["sa_rot"], rtol=0, atol=2e-7)


@pytest.mark.parametrize("tag", ["", "_pad"])
def test_standalone_depth_range_loss(cuda_lib, tag):
    """Values 1e-5, gradients 1e-4 vs the reference's own depth_range_loss; `_pad`: masks at a
    padded size, so every depth map goes through the align_corners bilinear resize first."""
    from wcr.roi_heads import depth_range_loss
    from wcr.spaces import Cubes
    from wcr.conversions import Boxes
    inp, nb = _sa_inputs()
    pad = [int(v) for v in Z["sa_pad"]] if tag else [0, 0]
    masks = [torch.nn.functional.pad(m, (0, pad[1], 0, pad[0])).cuda() for m in inp["gt_masks"]]
    ct = torch.from_numpy(Z["sa_cubes"]).cuda().requires_grad_(True)
    d = depth_range_loss(masks, inp["at_which_mask_idx"], [x.cuda() for x in inp["depth_maps"]], Cubes(ct),
                         Boxes(inp["gt_boxes"].cuda()), nb)
    d.sum().backward()
    np.testing.assert_allclose(d.detach().cpu().numpy(), Z["sa_depth" + tag], rtol=1e-5, atol=2e-6)
    _grad_close(ct.grad.cpu(), torch.from_numpy(Z["sa_depth_grad" + tag]))
