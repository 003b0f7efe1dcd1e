"""TEST INFRASTRUCTURE ONLY (oracle; authoring container). Drives the REFERENCE's own
``ROIHeads3DScore._forward_cube`` training branch (cubercnn/modeling/roi_heads/roi_heads.py:
698-1144: class gather :823-840, decode :786-905, the Python loops :947-960, every loss term,
uncertainty weighting and safe reductions :1075-1144) through oracle/refshim.py, with the pooler
and the cube head replaced by stubs that return given tensors. tests/golden/make_golden.py
records its loss dict and the autograd gradients w.r.t. the five head outputs
(forward_cube_*.npz); tests/test_forward_cube_gpu.py compares ``wcr.roi_heads.
ROIHeads3DScoreB200`` / ``wcr.glue.cube_losses_from_head`` with them on the GPU.
"""
from __future__ import annotations

import random
import types

import torch

from . import refshim, standins


class _Inst:
    """Just enough of detectron2.structures.Instances for _forward_cube."""

    def __init__(self, image_size=None, **fields):
        self.image_size = image_size
        for k, v in fields.items():
            setattr(self, k, v)

    def __len__(self):
        for k in ("gt_classes", "pred_classes", "proposal_boxes"):
            if k in self.__dict__:
                return len(self.__dict__[k])
        return 0


class _Pooled:
    def __init__(self, n):
        self.n = n

    def flatten(self, _):
        return torch.zeros(self.n, 4)


def make_case(seed: int, nb, H=96, W=128, C=5, loss_functions=None, use_confidence=1.0, pad=(0, 0)):
    """Seeded inputs of one training step: head outputs per class, proposals, GT, maps, masks."""
    from . import loss_inputs as li
    g = torch.Generator().manual_seed(1000 + seed)
    n = sum(nb)
    inp = li.add_masks(li.make_loss_inputs(seed, nb, H=H, W=W), seed, H=H, W=W)
    cls = torch.randint(0, C, (n,), generator=g)
    head = {"deltas": torch.randn(n, C, 2, generator=g) * 0.1, "z": torch.rand(n, C, 1, generator=g) * 4 + 1.5,
            "dims": torch.randn(n, C, 3, generator=g) * 0.2, "pose": torch.randn(n, C, 3, 3, generator=g) * 0.3,
            "uncert": torch.rand(n, C, generator=g) + 0.01}
    # allocentric pose per class: a perturbed rotation (the reference feeds rotation matrices here)
    q, _ = torch.linalg.qr(torch.eye(3).expand(n, C, 3, 3) + head["pose"])
    head["pose"] = q.contiguous()
    pri = torch.stack((0.5 + torch.rand(C, 3, generator=g), 0.1 + 0.2 * torch.rand(C, 3, generator=g)), 1)
    Ks = [torch.tensor([[105.0 + 3 * i, 0, W / 2], [0, 104.0 + 2 * i, H / 2], [0, 0, 1]]) for i in range(len(nb))]
    ratio = [0.9, 1.1, 1.3, 1.0, 0.8][:len(nb)]
    ground = []
    for i in range(len(nb)):
        gm = torch.zeros(H, W, dtype=torch.uint8)
        gm[H // 2 + 4:, :] = 1
        ground.append(gm if i != 1 else torch.ones(1, 1, dtype=torch.uint8))
    masks = [torch.nn.functional.pad(m, (0, pad[1], 0, pad[0])) for m in inp["gt_masks"]]
    lf = tuple(loss_functions or ("dims", "pose_alignment", "pose_ground", "iou", "z", "z_pseudo_gt_patch", "depth"))
    return {"nb": list(nb), "H": H, "W": W, "C": C, "cls": cls, "head": head, "priors": pri, "Ks": Ks,
            "ratio": ratio, "dims_img": [(H, W)] * len(nb), "src_boxes": inp["src_boxes"],
            "gt_boxes": inp["gt_boxes"], "depth_maps": inp["depth_maps"], "ground_maps": ground,
            "gt_masks": masks, "at_which_mask_idx": inp["at_which_mask_idx"], "loss_functions": lf,
            "use_confidence": use_confidence, "rng_seed": 4000 + seed}


def run_reference(case, head_leaves):
    """-> loss dict of the reference's _forward_cube for ``case`` with the cube head returning
    ``head_leaves`` (dict of tensors that may require grad)."""
    ns = refshim.load()
    rh = ns.roi_heads
    rh.select_foreground_proposals = lambda instances, num_classes: (instances, None)
    rh.Instances = _Inst
    h = object.__new__(rh.ROIHeads3DScore)
    n, C = sum(case["nb"]), case["C"]
    h.training = True
    h.in_features = ["p2"]
    h.num_classes = C
    h.scale_roi_boxes = 0
    h.cube_pooler = lambda feats, boxes: _Pooled(n)
    h.cube_head = lambda feats: (head_leaves["deltas"], head_leaves["z"], head_leaves["dims"],
                                 head_leaves["pose"], head_leaves["uncert"])
    h.cluster_bins = 1
    h.use_confidence = case["use_confidence"]
    h.dims_priors_enabled, h.dims_priors_func = True, "exp"
    h.priors_dims_per_cat = case["priors"]
    h.allocentric_pose = True
    h.z_type = "direct"
    h.virtual_depth, h.virtual_focal = True, 512.0
    h.loss_functions = list(case["loss_functions"])
    h.loss_w_3d, h.loss_w_iou, h.loss_w_seg, h.loss_w_pose = 1.0, 1.0, 2.5, 7.0
    h.loss_w_normal_vec, h.loss_w_z, h.loss_w_dims, h.loss_w_depth = 20.0, 1.0, 20.0, 1.0
    # per-image instances; gt_boxes3D[:, 0] doubles as the key into first_occurrence_indices
    insts, first_occ, k = [], {}, 0
    for i, b in enumerate(case["nb"]):
        sl = slice(k, k + b)
        gt3 = torch.zeros(b, 6)
        gt3[:, 0] = torch.arange(k, k + b, dtype=torch.float32)       # unique object ids
        gt3[:, 2] = 3.0
        gt3[:, 3:6] = 1.0
        for j in range(b):
            first_occ[float(k + j)] = int(case["at_which_mask_idx"][k + j])
        insts.append(_Inst(gt_classes=case["cls"][sl], proposal_boxes=standins.Boxes(case["src_boxes"][sl]),
                           pred_boxes=standins.Boxes(case["src_boxes"][sl]), gt_boxes3D=gt3,
                           gt_poses=torch.eye(3).expand(b, 3, 3).contiguous(),
                           gt_boxes=standins.Boxes(case["gt_boxes"][sl])))
        k += b
    depth = standins.ImageList.from_tensors(case["depth_maps"])
    gts = case["ground_maps"]
    ground = standins.ImageList([None] * len(gts), [tuple(g.shape) for g in gts])
    ground.__class__ = type("IL", (standins.ImageList,), {
        "__getitem__": lambda self, i: gts[i], "device": property(lambda self: torch.device("cpu"))})
    random.seed(case["rng_seed"])                   # plane.py:95 draws from Python's global generator
    _, losses = h._forward_cube({"p2": None}, insts, case["Ks"], case["dims_img"], case["ratio"],
                                case["gt_masks"], first_occ, ground, depth)
    return losses


def run_reference_with_grads(case):
    leaves = {k: v.clone().requires_grad_(True) for k, v in case["head"].items()}
    losses = run_reference(case, leaves)
    total = sum(losses.values())
    total.backward()
    grads = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in leaves.items()}
    return {k: v.detach() for k, v in losses.items()}, grads
