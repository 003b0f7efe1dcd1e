"""Pin oracle/loss.py (step 4 restatement) to the reference: committed goldens
(values + autograd gradients produced by the reference's own functions) and, in
the authoring container, the reference run live on further seeds."""
import glob
import os

import numpy as np
import pytest

from oracle import loss as ol, loss_inputs as li, refshim
from tests.golden.golden_io import load_loss_case

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "loss_*.npz")))


def _close(La, ga, Lb, gb, lr=1e-6, gr=1e-5):
    assert set(La) == set(Lb)
    for k in Lb:
        assert abs(float(La[k]) - float(Lb[k])) <= lr * max(abs(float(Lb[k])), 1e-6), k
    for k in gb:
        d = (ga[k].reshape(-1) - gb[k].reshape(-1)).abs().max().item()
        assert d <= gr * max(gb[k].abs().max().item(), 1e-12), (k, d)


@pytest.mark.parametrize("path", GOLDEN)
def test_loss_oracle_matches_reference_golden(path):
    cfg_kw, inp, L_ref, g_ref = load_loss_case(path)
    L, _, g = li.run_with_grads(ol.weak_loss, ol.LossConfig(**cfg_kw), inp)
    _close(L, g, L_ref, g_ref)


@pytest.mark.ref
@pytest.mark.parametrize("seed,nb", [(21, [4, 1, 6]), (22, [2, 2, 2, 2])])
def test_loss_oracle_matches_reference_live(seed, nb):
    from oracle import ref_loss
    inp = li.make_loss_inputs(seed, nb, H=100, W=140)
    for cfg in (ol.LossConfig(), ol.LossConfig(z_type="sigmoid", dims_priors_func="sigmoid")):
        La, _, ga = li.run_with_grads(ol.weak_loss, cfg, inp)
        Lb, _, gb = li.run_with_grads(ref_loss.reference_weak_loss, cfg, inp)
        _close(La, ga, Lb, gb)


def test_giou_matches_torchvision():
    import torch
    from torchvision.ops import generalized_box_iou_loss
    g = torch.Generator().manual_seed(0)
    a = torch.rand(50, 4, generator=g) * 100; a[:, 2:] += a[:, :2] + 1
    b = torch.rand(50, 4, generator=g) * 100; b[:, 2:] += b[:, :2] + 1
    assert torch.equal(ol.giou_loss(a, b), generalized_box_iou_loss(a, b, reduction="none"))


def test_zbox_oracle_matches_reference_golden():
    """oracle.pseudo_gt_z_box_loss vs ROIHeads3DScore.pseudo_gt_z_box_loss run as is
    (tests/golden/zbox_0.npz): values and the gradient w.r.t. pred_z."""
    import torch
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "zbox_0.npz"))
    depth = [torch.from_numpy(d) for d in z["depth"]]
    boxes = list(torch.from_numpy(z["boxes"]).split([int(x) for x in z["nb"]]))
    pz = torch.from_numpy(z["pred_z"]).requires_grad_(True)
    loss = ol.pseudo_gt_z_box_loss(depth, boxes, pz)
    loss.sum().backward()
    assert np.array_equal(loss.detach().numpy(), z["loss"])
    assert np.array_equal(pz.grad.numpy(), z["grad"])
