"""Caller glue on the GPU: rotation_6d_to_matrix, class gather, per-box
constants, and the head-outputs -> loss-dict composite, against oracle/glue.py + oracle/loss.py.
Tolerances: values 1e-6 relative (fp32 re-association only), gradients 1e-4 relative."""
import numpy as np
import pytest
import torch

from oracle import glue as og, loss as ol, loss_inputs as li, refshim

pytestmark = pytest.mark.gpu


def test_rotation_6d_matches_pytorch3d_restatement(cuda_lib):
    from wcr.glue import rotation_6d_to_matrix
    g = torch.Generator().manual_seed(0)
    d6 = torch.randn(257, 6, generator=g)
    d6[3] *= 1e-3
    d6[4, 3:] = d6[4, :3] * 2 + 1e-4 * torch.randn(3, generator=g)      # nearly parallel pair
    up = torch.randn(257, 3, 3, generator=g)
    a = d6.clone().requires_grad_(True)
    Ra = og.rotation_6d_to_matrix(a)
    (Ra * up).sum().backward()
    b = d6.cuda().requires_grad_(True)
    Rb = rotation_6d_to_matrix(b)
    (Rb * up.cuda()).sum().backward()
    ok = torch.ones(257, dtype=torch.bool)
    ok[4] = False      # the nearly parallel pair cancels ~4 digits in a2 - (b1.a2) b1: conditioning, not a bug
    assert torch.allclose(Rb.cpu()[ok], Ra.detach()[ok], rtol=1e-5, atol=1e-6)
    assert torch.allclose(Rb.cpu()[4], Ra.detach()[4], atol=5e-3)
    scale = a.grad[ok].abs().max()
    assert (b.grad.cpu() - a.grad)[ok].abs().max() <= 1e-4 * scale
    # shape handling as cube_head.py:175 uses it: (n*C, 6) -> (n*C, 3, 3)
    assert rotation_6d_to_matrix(torch.randn(4, 5, 6).cuda()).shape == (4, 5, 3, 3)


def test_class_gather_and_scatter(cuda_lib):
    from wcr.glue import class_gather
    g = torch.Generator().manual_seed(1)
    x = torch.randn(37, 50, 3, 3, generator=g)
    cls = torch.randint(0, 50, (37,), generator=g)
    up = torch.randn(37, 3, 3, generator=g)
    a = x.clone().requires_grad_(True)
    (og.class_gather(a, cls) * up).sum().backward()
    b = x.cuda().requires_grad_(True)
    out = class_gather(b, cls.cuda())
    (out * up.cuda()).sum().backward()
    assert torch.equal(out.cpu(), og.class_gather(x, cls))
    assert torch.equal(b.grad.cpu(), a.grad)


def test_box_constants(cuda_lib):
    from wcr.glue import box_constants
    g = torch.Generator().manual_seed(2)
    B = 5
    Ks = [torch.tensor([[500.0 + 50 * i, 0, 320 + i], [0, 510.0 + 40 * i, 240 - i], [0, 0, 1]]) for i in range(B)]
    ratio = [float(0.7 + 0.1 * i) for i in range(B)]
    dims = [(512 + 32 * i, 683 + 10 * i) for i in range(B)]
    nb = [3, 0, 7, 1, 4]
    for vd in (True, False):
        K4, v2r, clamp = box_constants(Ks, ratio, dims, nb, virtual_focal=512.0, virtual_depth=vd)
        K4o, v2ro, clampo = og.box_constants(Ks, ratio, dims, nb, 512.0, vd)
        assert torch.allclose(K4.cpu(), K4o, rtol=1e-7, atol=0)
        assert torch.allclose(v2r.cpu(), v2ro, rtol=2e-7, atol=0)
        assert torch.equal(clamp.cpu(), clampo)
    if refshim.available():
        ns = refshim.load()
        f, H, H0 = torch.rand(9, generator=g) * 500 + 300, torch.rand(9, generator=g) * 400 + 300, torch.rand(9, generator=g) * 400 + 300
        assert torch.equal(ns.math_util.compute_virtual_scale_from_focal_spaces(f, H, 512.0, H0),
                           og.compute_virtual_scale_from_focal_spaces(f, H, 512.0, H0))


def test_cube_losses_from_head_matches_oracle_composition(cuda_lib):
    """Per-class head outputs -> loss dict: gather + 6-D pose + constants + fused loss vs the same
    composition in torch (oracle), values and gradients w.r.t. the RAW head outputs."""
    from wcr.glue import cube_losses_from_head
    from wcr.roi_heads import WeakLossConfig
    nb, C, H, W = [4, 1, 6], 7, 120, 160
    inp = li.make_loss_inputs(3, nb, H=H, W=W)
    n = sum(nb)
    g = torch.Generator().manual_seed(5)
    cls = torch.randint(0, C, (n,), generator=g)
    head = {"deltas": torch.randn(n, C, 2, generator=g) * 0.1, "z": torch.rand(n, C, 1, generator=g) * 5 + 1,
            "dims": torch.randn(n, C, 3, generator=g) * 0.2, "pose6": torch.randn(n, C, 6, generator=g),
            "uncert": torch.rand(n, C, generator=g) + 0.01}
    pri = torch.stack((0.5 + torch.rand(C, 3, generator=g), 0.1 + 0.2 * torch.rand(C, 3, generator=g)), 1)  # (C,2,3)
    Ks = [torch.tensor([[420.0 + 9 * i, 0, W / 2], [0, 430.0 + 7 * i, H / 2], [0, 0, 1]]) for i in range(len(nb))]
    ratio = [0.9, 1.1, 1.3]
    dims_img = [(H, W)] * len(nb)
    lf = ("iou", "z_pseudo_gt_center", "pose_alignment", "pose_ground", "dims", "z")

    def oracle(h):
        K4, v2r, clamp = og.box_constants(Ks, ratio, dims_img, nb)
        pose = og.rotation_6d_to_matrix(h["pose6"].view(-1, 6)).view(n, C, 3, 3)       # every class, then gather
        p = pri[cls]
        return ol.weak_loss(ol.LossConfig(loss_functions=lf), og.class_gather(h["deltas"], cls),
                            og.class_gather(h["z"], cls).reshape(n), og.class_gather(h["dims"], cls),
                            og.class_gather(pose, cls), og.class_gather(h["uncert"], cls),
                            src_boxes=inp["src_boxes"], K4=K4, virtual_to_real=v2r, prior_mean=p[:, 0], prior_std=p[:, 1],
                            gt_boxes=inp["gt_boxes"], clamp=clamp.long(), num_boxes_per_image=nb,
                            depth_maps=inp["depth_maps"], normals=inp["normals"], ground_conf=inp["ground_conf"])[0]

    ha = {k: v.clone().requires_grad_(True) for k, v in head.items()}
    La = oracle(ha)
    sum(La.values()).backward()
    hb = {k: v.cuda().requires_grad_(True) for k, v in head.items()}
    Lb = cube_losses_from_head(WeakLossConfig(loss_functions=lf), hb["deltas"], hb["z"], hb["dims"], hb["pose6"],
                               hb["uncert"], cls.cuda(), proposal_boxes=inp["src_boxes"].cuda(),
                               gt_boxes=inp["gt_boxes"].cuda(), Ks=Ks, im_scales_ratio=ratio, im_current_dims=dims_img,
                               num_boxes_per_image=nb, priors_dims_per_cat=pri,
                               depth_maps=[d.cuda() for d in inp["depth_maps"]], normals=inp["normals"].cuda(),
                               ground_conf=inp["ground_conf"].cuda(), pose_is_6d=True)
    sum(Lb.values()).backward()
    assert set(La) == set(Lb)
    for k in La:
        assert abs(float(Lb[k]) - float(La[k])) <= 1e-5 * abs(float(La[k])) + 3e-6, k
    for k in head:
        scale = max(float(ha[k].grad.abs().max()), 1e-12)
        assert float((hb[k].grad.cpu() - ha[k].grad).abs().max()) <= 1e-4 * scale, k
