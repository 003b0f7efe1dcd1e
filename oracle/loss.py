"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

Torch-fp32 restatement (CPU, differentiable through torch autograd) of north_star
step 4: the weak-loss block of ``ROIHeads3DScore._forward_cube`` -- decode of the
cube head outputs, projected-cube-to-2D-box GIoU loss, pose-alignment, ground-normal
(pseudo orientation), pseudo-depth and dimension-prior terms, uncertainty scaling
and the safe reductions. This is a floating-point path, so the oracle is a torch
fp32 reference; gradients come from autograd over these very ops, which keeps the
reference's sub-gradient conventions (torch.min/max over a dim -> first
arg-extremum; clamp passes gradient on the closed interval; binary max/min split
ties; abs has gradient 0 at 0).

Follows (paths relative to /root/reference):
  cubercnn/modeling/roi_heads/roi_heads.py:786-905   decode
  cubercnn/modeling/roi_heads/roi_heads.py:927-960   cubes, projected boxes
  cubercnn/modeling/roi_heads/roi_heads.py:973-1031  loss terms
  cubercnn/modeling/roi_heads/roi_heads.py:1075-1144 uncertainty + reduction
  cubercnn/modeling/roi_heads/roi_heads.py:445-464   pose_loss
  cubercnn/modeling/roi_heads/roi_heads.py:612-656   dim_loss, pseudo_gt_z_point_loss
  cubercnn/modeling/roi_heads/roi_heads.py:1307-1315 safely_reduce_losses
  cubercnn/util/math_util.py:142-217, 732-743, 802-830, 1132-1237
  ProposalNetwork/utils/spaces.py:224-245, conversions.py:25-48
  torchvision.ops.generalized_box_iou_loss (third party, 0.26.0 here)
  pytorch3d.transforms.axis_angle_to_matrix (third party; quaternion route)

Parity status: PINNED -- tests/test_oracle_loss.py runs the reference's own methods
(ROIHeads3DScore.pose_loss / pseudo_gt_z_point_loss / dim_loss /
safely_reduce_losses, Cubes, cubes_to_box, R_from_allocentric, torchvision GIoU)
on the same seeded inputs when /root/reference is present, and
tests/golden/loss_*.npz holds their outputs (values and autograd gradients) for the
GPU box.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch
import torch.nn.functional as F

SQRT_2_CONSTANT = 1.41421356  # roi_heads.py:43


@dataclass
class LossConfig:
    """cfg.MODEL.ROI_CUBE_HEAD.* / cfg.loss_functions (config.py:52-127, 186)."""
    loss_functions: tuple = ("iou", "z_pseudo_gt_center", "pose_alignment", "pose_ground", "dims")
    z_type: str = "direct"              # 'direct' | 'sigmoid' | 'log'
    dims_priors_func: str = "exp"       # 'exp' | 'sigmoid'
    use_confidence: float = 1.0
    loss_w_3d: float = 1.0
    loss_w_iou: float = 1.0
    loss_w_pose: float = 7.0
    loss_w_normal_vec: float = 20.0
    loss_w_z: float = 1.0
    loss_w_dims: float = 20.0
    loss_w_seg: float = 2.5             # config.py:124
    loss_w_depth: float = 1.0           # config.py:127


def axis_angle_to_matrix(aa: torch.Tensor) -> torch.Tensor:
    """pytorch3d: axis-angle -> quaternion -> matrix."""
    angles = torch.norm(aa, p=2, dim=-1, keepdim=True)
    half = angles * 0.5
    small = angles.abs() < 1e-6
    s = torch.where(small, 0.5 - (angles * angles) / 48, torch.sin(half) / torch.where(small, torch.ones_like(angles), angles))
    q = torch.cat([torch.cos(half), aa * s], dim=-1)
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def R_from_allocentric(K4, R_view, u, v):
    """math_util.py:802-830 (tensor branch). K4 (n,4) = fx fy cx cy."""
    fx, fy, sx, sy = K4.unbind(-1)
    oray = torch.stack(((u - sx) / fx, (v - sy) / fy, torch.ones_like(u))).T
    oray = oray / torch.linalg.norm(oray, dim=1).unsqueeze(1)
    angle = torch.acos(oray[:, -1])
    axis = torch.zeros_like(oray)
    axis[:, 0] = axis[:, 0] - oray[:, 1]
    axis[:, 1] = axis[:, 1] + oray[:, 0]
    norms = torch.linalg.norm(axis, dim=1)
    valid = angle > 0
    M = axis_angle_to_matrix(angle.unsqueeze(1) * axis / norms.unsqueeze(1))
    R = R_view.clone()
    R[valid] = torch.bmm(M[valid], R_view[valid])
    return R


def cuboid_corners(box6, R):
    """math_util.py:142-217: (n,6)=[x y z w h l], (n,3,3) -> (n,8,3)."""
    n = box6.shape[0]
    w3, h3, l3 = box6[:, 3:4], box6[:, 4:5], box6[:, 5:6]
    verts = torch.zeros((n, 3, 8), dtype=box6.dtype)
    verts[:, 0, [0, 3, 4, 7]] = -l3 / 2
    verts[:, 0, [1, 2, 5, 6]] = l3 / 2
    verts[:, 1, [0, 1, 4, 5]] = -h3 / 2
    verts[:, 1, [2, 3, 6, 7]] = h3 / 2
    verts[:, 2, [0, 1, 2, 3]] = -w3 / 2
    verts[:, 2, [4, 5, 6, 7]] = w3 / 2
    verts = R @ verts
    verts[:, 0, :] += box6[:, 0:1]
    verts[:, 1, :] += box6[:, 1:2]
    verts[:, 2, :] += box6[:, 2:3]
    return verts.transpose(1, 2)


def bube_corners(box6, R, K4, clamp):
    """spaces.py:224-245 with one K and one clamp per box: (n,8,2) clamped projected corners."""
    corners = cuboid_corners(box6, R)                         # (n,8,3)
    fx, fy, cx, cy = (K4[:, i:i + 1] for i in range(4))
    X, Y, Z = corners[..., 0], corners[..., 1], corners[..., 2]
    px = (fx * X + cx * Z) / Z                                  # K @ corner, rows 0 and 1 over row 2
    py = (fy * Y + cy * Z) / Z
    c0, c1 = clamp[:, 0:1].to(px.dtype), clamp[:, 1:2].to(px.dtype)
    lo0, hi0 = torch.trunc(-c0 / 2 + 1), torch.trunc(c0 - 1 + c0)
    lo1, hi1 = torch.trunc(-c1 / 2 + 1), torch.trunc(c1 - 1 + c1)
    x = _clamp_closed(px, lo0, hi0)      # torch.clamp(x, int, int) forward and backward rule
    y = _clamp_closed(py, lo1, hi1)
    return torch.stack((x, y), -1)


def projected_boxes(box6, R, K4, clamp):
    """conversions.py:25-48 over bube_corners (roi_heads.py:957-960 loops box by box).
    clamp (n,2) integers."""
    b = bube_corners(box6, R, K4, clamp)
    x, y = b[..., 0], b[..., 1]
    return torch.stack((x.min(1).values, y.min(1).values, x.max(1).values, y.max(1).values), 1)


class _ClampClosed(torch.autograd.Function):
    """torch.clamp with tensor bounds and clamp's own backward (gradient passes
    where lo <= x <= hi, closed)."""

    @staticmethod
    def forward(ctx, x, lo, hi):
        ctx.save_for_backward((x >= lo) & (x <= hi))
        return torch.where(x != x, x, torch.max(torch.min(x, hi), lo))

    @staticmethod
    def backward(ctx, g):
        (m,) = ctx.saved_tensors
        return g * m.to(g.dtype), None, None


def _clamp_closed(x, lo, hi):
    return _ClampClosed.apply(x, lo.expand_as(x), hi.expand_as(x))


def giou_loss(b1, b2, eps=1e-7):
    """torchvision.ops.generalized_box_iou_loss(boxes1=b1, boxes2=b2, 'none')."""
    x1, y1, x2, y2 = b1.unbind(-1)
    x1g, y1g, x2g, y2g = b2.unbind(-1)
    xk1, yk1 = torch.max(x1, x1g), torch.max(y1, y1g)
    xk2, yk2 = torch.min(x2, x2g), torch.min(y2, y2g)
    inter = torch.zeros_like(x1)
    mask = (yk2 > yk1) & (xk2 > xk1)
    inter[mask] = (xk2[mask] - xk1[mask]) * (yk2[mask] - yk1[mask])
    union = (x2 - x1) * (y2 - y1) + (x2g - x1g) * (y2g - y1g) - inter
    iou = inter / (union + eps)
    xc1, yc1 = torch.min(x1, x1g), torch.min(y1, y1g)
    xc2, yc2 = torch.max(x2, x2g), torch.max(y2, y2g)
    area_c = (xc2 - xc1) * (yc2 - yc1)
    return 1 - (iou - ((area_c - union) / (area_c + eps)))


def pose_loss(cube_pose, num_boxes_per_image):
    """roi_heads.py:445-464 + math_util.py:1132-1181, 1222-1231 (cos_angle=True).
    Pairs (i, j<i) in the reference's order; returns None when every image has a
    single box; divides by (#single-box images + 1) as the reference does."""
    loss = torch.zeros(1)
    fail = 0
    for R in cube_pose.split(list(num_boxes_per_image)):
        if len(R) == 1:
            fail += 1
            continue
        if len(R) == 0:   # torch.mean of an empty tensor is NaN in the reference too
            loss = loss + torch.mean(torch.zeros(0))
            continue
        i, j = torch.tril_indices(len(R), len(R), -1)
        R12 = torch.matmul(R[i], R[j].permute(0, 2, 1))
        tr = R12[:, 0, 0] + R12[:, 1, 1] + R12[:, 2, 2]
        cos = (tr - 1.0) * 0.5
        loss = loss + torch.mean(1 - cos.abs())
    if fail == len(num_boxes_per_image):
        return None
    return loss * 1 / (fail + 1)


def pseudo_gt_z_point_loss(depth_maps, pred_xy, pred_z, num_boxes_per_image):
    """roi_heads.py:635-656. depth_maps: list of un-padded (h,w) tensors."""
    gt = []
    for d, xy in zip(depth_maps, pred_xy.split(list(num_boxes_per_image))):
        h, w = d.shape
        x = torch.clamp(xy[:, 0], 10, w - 11)
        y = torch.clamp(xy[:, 1], 10, h - 11)
        gt.append(d[y.long(), x.long()])
    return F.smooth_l1_loss(pred_z, torch.cat(gt), reduction="none", beta=0.0)


def pseudo_gt_z_box_loss(depth_maps, proposal_boxes, pred_z):
    """roi_heads.py:575-610 ('z_pseudo_gt_patch'). depth_maps: list of un-padded (h,w)
    tensors; proposal_boxes: list of (b_i,4) tensors. Per image: clip to the image
    (Boxes.clip), boxes with positive clipped area -> torch.median (LOWER median) of the
    patch sliced with .long() bounds, appended first; the others -> centre lookup with
    the 10-px clamp, appended after them. The list is paired with pred_z by position."""
    gt = []
    for d, b in zip(depth_maps, proposal_boxes):
        h, w = d.shape
        b = torch.stack((b[:, 0].clamp(0, w), b[:, 1].clamp(0, h), b[:, 2].clamp(0, w), b[:, 3].clamp(0, h)), 1)
        area = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
        m = area > 0
        for box in b[m]:
            gt.append(torch.median(d[box[1].long():box[3].long(), box[0].long():box[2].long()]).unsqueeze(0))
        out = b[~m]
        if len(out) == 0:
            continue
        c = (out[:, :2] + out[:, 2:]) / 2
        x = torch.clamp(c[:, 0], 10, w - 11)
        y = torch.clamp(c[:, 1], 10, h - 11)
        gt.append(d[y.long(), x.long()])
    return F.smooth_l1_loss(pred_z, torch.cat(gt), reduction="none", beta=0.0)


def dim_loss(prior_mean, prior_std, dims):
    """roi_heads.py:612-633."""
    if torch.isnan(prior_std).any():
        return None, None, None
    s = (dims - prior_mean).abs() / prior_std
    s = torch.max(s - 1.0, torch.zeros_like(s))
    return s[:, 0], s[:, 1], s[:, 2]


def safely_reduce_losses(loss):
    """roi_heads.py:1307-1315."""
    valid = (~loss.isinf()) & (~loss.isnan())
    if valid.any():
        return loss[valid].mean()
    return loss.mean() * 0.0


def decode(cfg: LossConfig, deltas, z_raw, dims_raw, pose, src_boxes, K4, virtual_to_real,
           prior_mean, prior_std):
    """roi_heads.py:786-905, 927-931 -> cube_x, cube_y, cube_z, dims, R_ego, box6."""
    sw = src_boxes[:, 2] - src_boxes[:, 0]
    sh = src_boxes[:, 3] - src_boxes[:, 1]
    scx = src_boxes[:, 0] + 0.5 * sw
    scy = src_boxes[:, 1] + 0.5 * sh
    cube_x = scx + sw * deltas[:, 0]
    cube_y = scy + sh * deltas[:, 1]
    if cfg.dims_priors_func == "sigmoid":
        dmin = (prior_mean - 3 * prior_std).clip(0.0)
        dmax = prior_mean + 3 * prior_std
        dims = dmin + (dmax - dmin) * torch.sigmoid(dims_raw)
    else:
        dims = torch.exp(dims_raw.clip(max=5)) * prior_mean
    R = R_from_allocentric(K4, pose, cube_x.detach(), cube_y.detach())
    if cfg.z_type == "sigmoid":
        z = torch.sigmoid(z_raw) * 100
    elif cfg.z_type == "log":
        z = torch.exp(z_raw)
    else:
        z = z_raw
    z = z * virtual_to_real
    x3d = z * (cube_x - K4[:, 2]) / K4[:, 0]
    y3d = z * (cube_y - K4[:, 3]) / K4[:, 1]
    box6 = torch.cat((x3d[:, None], y3d[:, None], z[:, None], dims), 1)
    return cube_x, cube_y, z, dims, R, box6


def weak_loss(cfg: LossConfig, deltas, z_raw, dims_raw, pose, uncert, *, src_boxes, K4,
              virtual_to_real, prior_mean, prior_std, gt_boxes, clamp, num_boxes_per_image,
              depth_maps, normals, ground_conf, gt_masks=None, at_which_mask_idx=None):
    """The loss block. Per-box tensors are (n, ...); normals (B,3) and ground_conf
    (B,) are per image (0.1 where the ground map is the (1,1) dummy, roi_heads.py:988).
    -> (dict of scalar losses keyed like the reference, dict of per-box raw terms)."""
    n = deltas.shape[0]
    lf = cfg.loss_functions
    cube_x, cube_y, cube_z, dims, R, box6 = decode(cfg, deltas, z_raw, dims_raw, pose, src_boxes, K4,
                                                   virtual_to_real, prior_mean, prior_std)
    proj = projected_boxes(box6, R, K4, clamp)
    per_box = {"proj": proj, "cube_z": cube_z, "dims": dims, "R": R}
    nb = torch.tensor(list(num_boxes_per_image))
    terms = {}
    if "iou" in lf:
        terms["iou"] = giou_loss(gt_boxes, proj).view(n, -1).mean(dim=1)
    if "pose_alignment" in lf:
        lp = pose_loss(R, num_boxes_per_image)
        if lp is not None:
            terms["pose"] = lp.repeat(n)
    if "pose_ground" in lf:
        nv = normals.repeat_interleave(nb, 0)
        conf = ground_conf.repeat_interleave(nb, 0)
        pred_normal = R[:, 1, :]
        terms["normal_vec"] = (1 - F.cosine_similarity(nv, pred_normal, dim=1).abs()) * conf
    if "pose_ground2" in lf:         # roi_heads.py:997-1007, overrides 'pose_ground'
        from . import loss_tier2 as t2
        terms["normal_vec"] = t2.pose_ground2(R, normals, ground_conf, num_boxes_per_image)
    if "z_pseudo_gt_patch" in lf:    # if / elif of roi_heads.py:1010-1013
        terms["pseudo_gt_z"] = pseudo_gt_z_box_loss(depth_maps, proj.split(list(num_boxes_per_image)), cube_z)
    elif "z_pseudo_gt_center" in lf:
        cube_xy = torch.cat((cube_x.unsqueeze(1), cube_y.unsqueeze(1)), dim=1)
        terms["pseudo_gt_z"] = pseudo_gt_z_point_loss(depth_maps, cube_xy, cube_z, num_boxes_per_image)
    if "segmentation" in lf:         # roi_heads.py:948-954, 1016-1017
        from . import loss_tier2 as t2
        bube = torch.stack([t2.clamp_bube_corners(bube_corners(box6[i:i + 1], R[i:i + 1], K4[i:i + 1],
                                                               clamp[i:i + 1])[0], clamp[i]) for i in range(n)])
        terms["seg"] = t2.segment_loss(gt_masks, bube, at_which_mask_idx)
    if "z" in lf:                    # :1019-1021
        from . import loss_tier2 as t2
        terms["z"] = t2.z_loss(gt_boxes, box6, R, K4, clamp.float(), proj)
    if "dims" in lf:
        lw, lh, ll = dim_loss(prior_mean, prior_std, dims)
        if lw is not None:
            terms["dims_w"], terms["dims_h"], terms["dims_l"] = lw, lh, ll
    if "depth" in lf:                # :1027-1029
        from . import loss_tier2 as t2
        terms["depth"], gts = t2.depth_range_loss(gt_masks, at_which_mask_idx, depth_maps, box6, R, gt_boxes,
                                                  num_boxes_per_image)
        per_box["depth_gt"] = gts
    per_box.update({"raw_" + k: v.detach().clone() for k, v in terms.items()})
    losses = {}
    if cfg.use_confidence > 0:
        sf = SQRT_2_CONSTANT * torch.exp(-uncert)
        terms = {k: v * sf for k, v in terms.items()}
        losses["Cube/uncert"] = cfg.use_confidence * safely_reduce_losses(uncert.clone())
    w = {"iou": cfg.loss_w_iou, "pose": cfg.loss_w_pose, "normal_vec": cfg.loss_w_normal_vec,
         "pseudo_gt_z": cfg.loss_w_z, "dims_w": cfg.loss_w_dims, "dims_h": cfg.loss_w_dims,
         "dims_l": cfg.loss_w_dims, "seg": cfg.loss_w_seg, "z": cfg.loss_w_z, "depth": cfg.loss_w_depth}
    for k, v in terms.items():
        losses["Cube/loss_" + k] = safely_reduce_losses(v) * w[k] * cfg.loss_w_3d
    return losses, per_box
