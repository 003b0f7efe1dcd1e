"""oracle/loss_tier2.py (segment_loss, z_loss, depth_range_loss, pose_ground2) pinned to the
reference's own methods: live when /root/reference is present, else tests/golden/tier2_*.npz
(values and autograd gradients, tests/golden/make_golden.py)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import loss as ol, loss_inputs as li, loss_tier2 as t2, refshim

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
T2 = ("iou", "z", "depth", "pose_ground2", "dims")


def _close(La, ga, Lb, gb, lr=1e-6, gr=1e-5):
    assert set(La) == set(Lb), (sorted(La), sorted(Lb))
    for k in Lb:
        a, b = float(La[k]), float(Lb[k])
        assert (np.isnan(a) and np.isnan(b)) or abs(a - b) <= lr * max(abs(b), 1e-3), (k, a, b)
    for k in gb:
        scale = max(float(gb[k].abs().max()), 1e-12)
        assert float((ga[k] - gb[k]).abs().max()) <= gr * scale, k


def _inputs(seed, nb, H, W):
    return li.add_masks(li.make_loss_inputs(seed, nb, H=H, W=W), seed, H=H, W=W)


@pytest.mark.skipif(not refshim.available(), reason="needs /root/reference")
@pytest.mark.parametrize("seed,nb", [(0, [3, 1, 5, 2]), (3, [6, 9])])
@pytest.mark.parametrize("conf", [1.0, 0.0])
def test_tier2_oracle_matches_reference_live(seed, nb, conf):
    from oracle import ref_loss
    H, W = 96, 128
    inp = _inputs(seed, nb, H, W)
    cfg = ol.LossConfig(loss_functions=T2, use_confidence=conf)
    La, pa, ga = li.run_with_grads(ol.weak_loss, cfg, inp)
    Lb, pb, gb = li.run_with_grads(ref_loss.reference_weak_loss, cfg, inp)
    for k in ("raw_z", "raw_depth", "raw_normal_vec"):
        assert torch.allclose(pa[k], pb[k], rtol=1e-6, atol=1e-7), k
    _close(La, ga, Lb, gb)


@pytest.mark.skipif(not refshim.available(), reason="needs /root/reference")
@pytest.mark.parametrize("seed,nb", [(0, [3, 1, 5, 2]), (6, [4, 7])])
def test_segment_loss_oracle_matches_reference_live(seed, nb):
    """The soft raster is differentiable in the reference (utils.py:479-483): values AND gradients."""
    from oracle import ref_loss
    H, W = 96, 128
    inp = _inputs(seed, nb, H, W)
    cfg = ol.LossConfig(loss_functions=("segmentation",), use_confidence=1.0)
    La, pa, ga = li.run_with_grads(ol.weak_loss, cfg, inp)
    Lb, pb, gb = li.run_with_grads(ref_loss.reference_weak_loss, cfg, inp)
    assert torch.allclose(pa["raw_seg"], pb["raw_seg"], rtol=1e-6, atol=1e-8)
    assert any(float(g.abs().max()) > 0 for g in gb.values())
    _close(La, ga, Lb, gb)


@pytest.mark.skipif(not refshim.available(), reason="needs /root/reference")
def test_normal_to_rotation_and_linspace_quirks():
    head = refshim.roi_head_instance()
    g = torch.Generator().manual_seed(1)
    nv = torch.nn.functional.normalize(torch.randn(7, 3, generator=g), dim=1)
    assert torch.equal(t2.normal_to_rotation(nv), head.normal_to_rotation(nv))
    ex = torch.tensor([[1.0, 0, 0]]).repeat(3, 1)       # every t0 is exactly zero -> y-axis branch
    assert torch.equal(t2.normal_to_rotation(ex), head.normal_to_rotation(ex))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "tier2_*.npz"))))
def test_tier2_oracle_matches_reference_golden(path):
    z = np.load(path, allow_pickle=False)
    nb = [int(v) for v in z["nb"]]
    inp = _inputs(int(z["seed"]), nb, int(z["H"]), int(z["W"]))
    cfg = ol.LossConfig(loss_functions=T2, use_confidence=float(z["use_confidence"]))
    La, pa, ga = li.run_with_grads(ol.weak_loss, cfg, inp)
    Lb = {k[5:]: torch.tensor(z[k]) for k in z.files if k.startswith("loss:")}
    gb = {k[5:]: torch.tensor(z[k]) for k in z.files if k.startswith("grad:")}
    for k in ("raw_z", "raw_depth", "raw_normal_vec"):
        assert np.allclose(pa[k].numpy(), z["box:" + k], rtol=1e-6, atol=1e-7), k
    _close(La, ga, Lb, gb)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "tier2seg_*.npz"))))
def test_segment_loss_oracle_matches_reference_golden(path):
    z = np.load(path, allow_pickle=False)
    nb = [int(v) for v in z["nb"]]
    inp = _inputs(int(z["seed"]), nb, int(z["H"]), int(z["W"]))
    cfg = ol.LossConfig(loss_functions=("segmentation", "iou"), use_confidence=float(z["use_confidence"]))
    La, pa, ga = li.run_with_grads(ol.weak_loss, cfg, inp)
    Lb = {k[5:]: torch.tensor(z[k]) for k in z.files if k.startswith("loss:")}
    gb = {k[5:]: torch.tensor(z[k]) for k in z.files if k.startswith("grad:")}
    assert np.allclose(pa["raw_seg"].numpy(), z["box:raw_seg"], rtol=1e-6, atol=1e-8)
    _close(La, ga, Lb, gb)
