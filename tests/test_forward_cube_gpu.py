"""``ROIHeads3DScoreB200._forward_cube`` (the ROI-head drop-in whose training
loss block is one call into the fused kernels) against tests/golden/forward_cube_*.npz -- the
Cube/loss_* dict and the autograd gradients w.r.t. the five per-class cube-head outputs that
the REFERENCE's own ``ROIHeads3DScore._forward_cube`` (roi_heads.py:698-1144) produced for the
same seeded step (oracle/ref_forward_cube.py through the import shim, make_golden.py).

The class is built on a minimal stand-in base here (the reference and detectron2 do not exist
on the GPU box); with both importable, ``wcr.roi_heads.register_roi_heads_b200()`` builds it on
the reference's class and registers it."""
import ast
import glob
import os
import random

import numpy as np
import pytest
import torch

from oracle import ref_forward_cube as rf, standins

pytestmark = pytest.mark.gpu
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "forward_cube_*.npz")))


# ---- what the subclass takes from the reference's module / base class -----------------------
def select_foreground_proposals(instances, num_classes):
    return instances, None


Instances = rf._Inst


class FakeROIHeads3DScore:
    scale_roi_boxes = 0

    def scale_proposals(self, proposal_boxes):
        return proposal_boxes

    def _forward_cube(self, *a, **k):
        raise AssertionError("the reference's torch path must not run in training")


def _head(case, leaves):
    from wcr.roi_heads import make_roi_heads_b200
    n, C = sum(case["nb"]), case["C"]
    h = make_roi_heads_b200(FakeROIHeads3DScore)()
    h.training, h.in_features, h.num_classes = True, ["p2"], C
    h.cube_pooler = lambda feats, boxes: rf._Pooled(n)
    h.cube_head = lambda feats: (leaves["deltas"], leaves["z"], leaves["dims"], leaves["pose"], leaves["uncert"])
    h.cluster_bins, h.use_confidence = 1, case["use_confidence"]
    h.dims_priors_enabled, h.dims_priors_func, h.priors_dims_per_cat = True, "exp", case["priors"]
    h.allocentric_pose, h.z_type, h.virtual_depth, h.virtual_focal = True, "direct", True, 512.0
    h.loss_functions = list(case["loss_functions"])
    h.loss_w_3d, h.loss_w_iou, h.loss_w_seg, h.loss_w_pose = 1.0, 1.0, 2.5, 7.0
    h.loss_w_normal_vec, h.loss_w_z, h.loss_w_dims, h.loss_w_depth = 20.0, 1.0, 20.0, 1.0
    h.wcr_replay_rng = random.Random(case["rng_seed"])      # the stream plane.py:95 consumed in the golden run
    return h


def _instances(case):
    insts, first_occ, k = [], {}, 0
    for b in case["nb"]:
        sl = slice(k, k + b)
        gt3 = torch.zeros(b, 6)
        gt3[:, 0] = torch.arange(k, k + b, dtype=torch.float32)
        for j in range(b):
            first_occ[float(k + j)] = int(case["at_which_mask_idx"][k + j])
        insts.append(rf._Inst(gt_classes=case["cls"][sl], proposal_boxes=standins.Boxes(case["src_boxes"][sl]),
                              pred_boxes=standins.Boxes(case["src_boxes"][sl]), gt_boxes3D=gt3,
                              gt_boxes=standins.Boxes(case["gt_boxes"][sl])))
        k += b
    return insts, first_occ


@pytest.mark.parametrize("path", GOLDEN)
def test_forward_cube_matches_the_reference(cuda_lib, path):
    z = np.load(path)
    kw = ast.literal_eval(bytes(z["kw"]).decode())
    case = rf.make_case(**kw)
    leaves = {k: v.cuda().requires_grad_(True) for k, v in case["head"].items()}
    h = _head(case, leaves)
    insts, first_occ = _instances(case)
    depth = standins.ImageList.from_tensors(case["depth_maps"])
    gts = case["ground_maps"]
    ground = standins.ImageList([None] * len(gts), [tuple(g.shape) for g in gts])
    ground.__class__ = type("IL", (standins.ImageList,), {"__getitem__": lambda self, i: gts[i],
                                                          "__iter__": lambda self: iter(gts)})
    pred, losses = h._forward_cube({"p2": None}, insts, case["Ks"], case["dims_img"], case["ratio"],
                                   case["gt_masks"], first_occ, ground, depth)
    sum(losses.values()).backward()
    want = {k[5:]: float(z[k]) for k in z.files if k.startswith("loss:")}
    assert set(losses) == set(want), (sorted(losses), sorted(want))
    for k, v in want.items():
        assert abs(float(losses[k].detach()) - v) <= 1e-5 * abs(v) + 3e-6, (k, float(losses[k].detach()), v)
    seg = "segmentation" in case["loss_functions"]
    for k, v in leaves.items():          # a head output the configuration does not use (uncert without USE_CONFIDENCE)
        if v.grad is None:
            v.grad = torch.zeros_like(v)
    ga = torch.cat([leaves[k].grad.reshape(-1).double().cpu() for k in sorted(leaves)])
    gb = torch.cat([torch.from_numpy(z["grad:" + k]).reshape(-1).double() for k in sorted(leaves)])
    if seg:   # the soft raster's gradient is discontinuous in the corners (tests/test_loss_gpu.py::_check_seg)
        cos = float((ga * gb).sum() / (ga.norm() * gb.norm()))
        assert cos >= 0.98, cos
    else:
        for k in leaves:
            a, b = leaves[k].grad.reshape(-1).double().cpu(), torch.from_numpy(z["grad:" + k]).reshape(-1).double()
            scale = b.abs().max().item()
            assert (a - b).abs().max().item() <= 1e-4 * max(scale, 1e-12), k
            big = b.abs() >= 1e-3 * scale
            if scale > 0 and bool(big.any()):
                assert ((a - b).abs()[big] / b.abs()[big]).max().item() <= 1e-4, k
    # the per-image prediction instances of the training branch (roi_heads.py:1153-1195)
    assert len(pred) == len(case["nb"])
    for i, b in enumerate(case["nb"]):
        assert tuple(pred[i].pred_bbox3D.shape) == (b, 8, 3) and tuple(pred[i].pred_pose.shape) == (b, 3, 3)
        assert tuple(pred[i].pred_center_cam.shape) == (b, 3) and len(pred[i].pred_classes) == b


def test_forward_cube_rejects_configs_off_the_fused_path(cuda_lib):
    from wcr import _abi
    case = rf.make_case(1, [2, 2])
    leaves = {k: v.cuda() for k, v in case["head"].items()}
    h = _head(case, leaves)
    h.loss_functions = ["iou", "chamfer"]
    insts, first_occ = _instances(case)
    with pytest.raises(_abi.WcrError, match="not on the fused path"):
        h._forward_cube({"p2": None}, insts, case["Ks"], case["dims_img"], case["ratio"], None, first_occ, None, None)
