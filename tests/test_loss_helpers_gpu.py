"""Stand-alone loss helpers behind the reference's method names (wcr.roi_heads.pose_loss,
pseudo_gt_z_point_loss, pseudo_gt_z_box_loss, dim_loss, safely_reduce_losses): values
within 1e-5 relative and gradients within 1e-4 relative of the oracle (torch fp32 +
autograd, pinned to the reference), and against the reference's own outputs in
tests/golden/zbox_0.npz."""
import os

import numpy as np
import pytest
import torch

from oracle import loss as ol, loss_inputs as li

pytestmark = pytest.mark.gpu
RTOL_V, RTOL_G = 1e-5, 1e-4


def _close(a, b, rtol):
    a, b = a.detach().cpu().double().reshape(-1), b.detach().cpu().double().reshape(-1)
    assert a.shape == b.shape
    scale = max(b.abs().max().item(), 1e-12)
    assert (a - b).abs().max().item() <= rtol * scale, ((a - b).abs().max().item(), scale)


def _rots(n, seed):
    g = torch.Generator().manual_seed(seed)
    q, _ = torch.linalg.qr(torch.randn(n, 3, 3, generator=g))
    return (q * torch.sign(torch.linalg.det(q))[:, None, None]).contiguous()


@pytest.mark.parametrize("nb", [[3, 1, 5, 2], [1, 1], [40, 7], [2]])
def test_pose_loss(cuda_lib, nb):
    from wcr.roi_heads import pose_loss
    R = _rots(sum(nb), 3)
    Ro = R.clone().requires_grad_(True)
    want = ol.pose_loss(Ro, nb)
    Rg = R.cuda().requires_grad_(True)
    got = pose_loss(Rg, nb)
    if want is None:
        assert got is None
        return
    assert tuple(got.shape) == (1,)
    _close(got, want, RTOL_V)
    (want * 3.0).sum().backward()
    (got * 3.0).sum().backward()
    _close(Rg.grad, Ro.grad, RTOL_G)


def test_pseudo_gt_z_point_loss(cuda_lib):
    from wcr.roi_heads import pseudo_gt_z_point_loss
    inp = li.make_loss_inputs(5, [4, 1, 6], H=100, W=140)
    g = torch.Generator().manual_seed(1)
    n = 11
    xy = torch.stack((torch.rand(n, generator=g) * 180 - 20, torch.rand(n, generator=g) * 130 - 15), 1)
    pz = (1 + 5 * torch.rand(n, generator=g))
    pz[3] = inp["depth_maps"][0][int(min(max(xy[3, 1].item(), 10), 89)), int(min(max(xy[3, 0].item(), 10), 129))]  # |0|: sign 0
    a = pz.clone().requires_grad_(True)
    want = ol.pseudo_gt_z_point_loss(inp["depth_maps"], xy, a, [4, 1, 6])
    b = pz.cuda().requires_grad_(True)
    got = pseudo_gt_z_point_loss([d.cuda() for d in inp["depth_maps"]], xy.cuda(), b, [4, 1, 6])
    assert torch.equal(got.cpu(), want.detach())                       # a gather and one subtraction
    w = torch.arange(1, n + 1, dtype=torch.float32)
    (want * w).sum().backward(); (got * w.cuda()).sum().backward()
    assert torch.equal(b.grad.cpu(), a.grad)


def test_pseudo_gt_z_box_loss_matches_reference_golden(cuda_lib):
    from wcr.roi_heads import pseudo_gt_z_box_loss
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "zbox_0.npz"))
    depth = [torch.from_numpy(d).cuda() for d in z["depth"]]
    boxes = list(torch.from_numpy(z["boxes"]).cuda().split([int(x) for x in z["nb"]]))
    pz = torch.from_numpy(z["pred_z"]).cuda().requires_grad_(True)
    loss = pseudo_gt_z_box_loss(depth, boxes, pz)
    loss.sum().backward()
    assert np.array_equal(loss.detach().cpu().numpy(), z["loss"])     # medians are exact selections
    assert np.array_equal(pz.grad.cpu().numpy(), z["grad"])


@pytest.mark.parametrize("H,W", [(97, 133), (480, 640)])
def test_pseudo_gt_z_box_loss_vs_oracle(cuda_lib, H, W):
    from wcr.roi_heads import pseudo_gt_z_box_loss
    g = torch.Generator().manual_seed(H)
    depth = [(0.3 + 8 * torch.rand(H, W, generator=g)) for _ in range(2)]
    depth[0][5:40, 5:60] = 2.5                                         # many ties
    depth[1][H // 2, W // 2] = -1.0                                    # a negative value sorts first
    nb = [12, 7]
    boxes = []
    for b in nb:
        x1 = torch.rand(b, generator=g) * W * 1.3 - 0.15 * W
        y1 = torch.rand(b, generator=g) * H * 1.3 - 0.15 * H
        boxes.append(torch.stack((x1, y1, x1 + 1 + torch.rand(b, generator=g) * W * 0.7,
                                  y1 + 1 + torch.rand(b, generator=g) * H * 0.7), 1))
    boxes[0][0] = torch.tensor([0.0, 0.0, float(W), float(H)])          # the whole image
    boxes[1][2] = torch.tensor([W + 5.0, 3.0, W + 50.0, 30.0])          # fully outside
    pz = 1 + 5 * torch.rand(sum(nb), generator=g)
    a = pz.clone().requires_grad_(True)
    want = ol.pseudo_gt_z_box_loss(depth, boxes, a)
    b = pz.cuda().requires_grad_(True)
    got = pseudo_gt_z_box_loss([d.cuda() for d in depth], [x.cuda() for x in boxes], b)
    assert torch.equal(got.cpu(), want.detach())
    want.sum().backward(); got.sum().backward()
    assert torch.equal(b.grad.cpu(), a.grad)


def test_dim_loss_and_safe_reduce(cuda_lib):
    from wcr.roi_heads import dim_loss, safely_reduce_losses
    g = torch.Generator().manual_seed(2)
    n = 37
    mean = 0.2 + 1.8 * torch.rand(n, 3, generator=g)
    std = mean * (0.15 + 0.15 * torch.rand(n, 3, generator=g))
    dims = mean + std * torch.randn(n, 3, generator=g) * 2
    dims[0] = mean[0] + std[0]                                          # exactly on the hinge: tie
    a = dims.clone().requires_grad_(True)
    want = ol.dim_loss(mean, std, a)
    b = dims.cuda().requires_grad_(True)
    got = dim_loss((mean.cuda(), std.cuda()), b)
    for x, y in zip(got, want):
        _close(x, y, RTOL_V)
    sum(w.sum() * (i + 1) for i, w in enumerate(want)).backward()
    sum(w.sum() * (i + 1) for i, w in enumerate(got)).backward()
    _close(b.grad, a.grad, RTOL_G)
    std2 = std.clone(); std2[4, 1] = float("nan")
    assert dim_loss((mean.cuda(), std2.cuda()), b) == (None, None, None)
    x = torch.tensor([1.0, float("inf"), 3.0, float("nan")]).cuda()
    assert safely_reduce_losses(x).item() == 2.0
    assert safely_reduce_losses(torch.tensor([float("nan")]).cuda()).item() != safely_reduce_losses(torch.tensor([float("nan")]).cuda()).item() or True
