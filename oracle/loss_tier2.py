"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

Torch-fp32 restatement of the tier-2 weak-loss terms of ``ROIHeads3DScore``
(SURVEY 8(f) rank 1), each following the reference's method line by line including
its quirks:

  segment_loss        roi_heads.py:420-443  (convex_hull raster vs GT mask; the MASK is
                                             passed as logits and the RASTER as target of
                                             torchvision.ops.sigmoid_focal_loss)
  z_loss              roi_heads.py:530-573  (50 depth offsets of 0.1 m, argmin of the
                                             projected-area mismatch; zero gradient)
  depth_range_loss    roi_heads.py:658-684  (q90 - q10 of the masked depth vs the depth
                                             extent of the cube corners)
  normal_to_rotation  roi_heads.py:686-696  (Frobenius norm over the WHOLE batch, and
                                             ``.any() < 0.001`` evaluated on a bool)
  'pose_ground2'      roi_heads.py:997-1007 (1 - cos of the relative angle, pytorch3d
                                             so3_relative_angle(cos_angle=True))
  bube-corner clamp   roi_heads.py:948-954

Third party restated: torchvision.ops.sigmoid_focal_loss 0.26.0 (fvcore focal loss,
alpha 0.25, gamma 2), pytorch3d.transforms.so3.so3_relative_angle (trace form; the
reference's own copy is math_util.py:1184-1237).

Parity status: PINNED -- tests/test_oracle_loss_tier2.py calls the reference's own
methods on the same seeded inputs when /root/reference is present and
tests/golden/tier2_*.npz holds their outputs (and autograd gradients) otherwise.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from . import hull as ohull
from .loss import cuboid_corners


def l1_loss(vals, target):
    """roi_heads.py:352-353: smooth_l1 with beta=0 == |vals - target|."""
    return F.smooth_l1_loss(vals, target, reduction="none", beta=0.0)


def sigmoid_focal_loss(inputs, targets, alpha=0.25, gamma=2.0):
    p = torch.sigmoid(inputs)
    ce = F.binary_cross_entropy_with_logits(inputs, targets, reduction="none")
    p_t = p * targets + (1 - p) * (1 - targets)
    loss = ce * ((1 - p_t) ** gamma)
    if alpha >= 0:
        loss = (alpha * targets + (1 - alpha) * (1 - targets)) * loss
    return loss


def clamp_bube_corners(bube, clamp):
    """roi_heads.py:950-953: x into [0, int(clamp[0]-1)], y into [0, int(clamp[1]-1)]."""
    x = torch.clamp(bube[..., 0], 0, int(clamp[0] - 1))
    y = torch.clamp(bube[..., 1], 0, int(clamp[1] - 1))
    return torch.stack((x, y), dim=-1)


def hull_indices(points):
    """jarvis_march (utils.py:410-454) on float32 values, returning WHICH points form the hull
    (in the reference's flipped order) and the duplicate bump added to them, so that the hull can
    be rebuilt differentiably as (points + bump)[idx]. None when the march does not close within
    len(points) steps (the reference would loop forever)."""
    f = np.float32
    pts = np.asarray(points, dtype=f).copy()
    n = len(pts)
    dup = sorted({i for i in range(n - 1) for j in range(i + 1, n) if np.all(pts[i] == pts[j])})
    bump = np.zeros_like(pts)
    for k, d in enumerate(dup):
        bump[d] += f(k + 1)
    pts = (pts + bump).astype(f)
    min_x = pts[:, 0].min()
    cand = np.nonzero(pts[:, 0] == min_x)[0]
    index = int(cand[np.argmax(pts[cand][:, 1])]) if len(cand) > 1 else int(cand[0])

    def direction(p1, p2, p3):
        return f(f(f(p2[0] - p1[0]) * f(p3[1] - p1[1])) - f(f(p2[1] - p1[1]) * f(p3[0] - p1[0])))

    def dist(p1, p2):
        return f(f(f(p2[0] - p1[0]) ** 2) + f(f(p2[1] - p1[1]) ** 2))

    l, result = index, [index]
    for _ in range(n + 1):
        q = (l + 1) % n
        for i in range(n):
            if i == l:
                continue
            d = direction(pts[l], pts[i], pts[q])
            if d > 0 or (d == 0 and dist(pts[i], pts[l]) > dist(pts[q], pts[l])):
                q = i
        l = q
        if l == index:
            return result[::-1], bump
        result.append(q)
    return None, bump


def fill_polygon_torch(H, W, hull):
    """utils.py:456-488, differentiable in the hull vertices exactly like the reference."""
    Y, X = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    gx, gy = X.flatten().float(), Y.flatten().float()
    new_mask = torch.ones(H, W)
    zeros, ones = torch.zeros(H, W), torch.ones(H, W)
    for i in range(len(hull)):
        v1, v2 = hull[i], hull[(i + 1) % len(hull)]
        e = v2 - v1
        points = (gx - v1[0]) * e[1] - (gy - v1[1]) * e[0]
        new_mask = new_mask * torch.min(torch.max(points.view(H, W), zeros), ones)
    return new_mask


def convex_hull_torch(H, W, coords):
    idx, bump = hull_indices(coords.detach().numpy())
    if idx is None:
        return torch.zeros(H, W)
    return fill_polygon_torch(H, W, (coords + torch.from_numpy(bump))[idx])


def segment_loss(gt_mask, bube_corners, at_which_mask_idx, loss="focal"):
    """gt_mask: list of (1, H, W) bool; bube_corners (n, 8, 2) already clamped (roi_heads.py
    :948-954); differentiable w.r.t. the corners through the soft raster, as in the reference."""
    n = len(bube_corners)
    y_hat, y = [], []
    H, W = gt_mask[0].squeeze().shape
    for i in range(n):
        gt_i = gt_mask[at_which_mask_idx[i]][0]
        y.append((gt_i * 1.0).float())
        y_hat.append(convex_hull_torch(H, W, bube_corners[i]))
    y, y_hat = torch.stack(y), torch.stack(y_hat)
    if loss == "bce":
        return F.binary_cross_entropy_with_logits(y, y_hat, reduction="none").mean((1, 2))
    if loss == "dice":
        inter = (y * y_hat).sum()
        return 1 - (2.0 * inter + 1) / (y.sum() + y_hat.sum() + 1)
    return sigmoid_focal_loss(y, y_hat).mean((1, 2))


def _proj_area(box6, R, K4, clamp):
    """Area of cubes_to_box for a batch sharing one K/clamp (conversions.py:25-48)."""
    from .loss import projected_boxes
    m = box6.shape[0]
    pb = projected_boxes(box6, R.expand(m, 3, 3), K4.expand(m, 4), clamp.expand(m, 2))
    return (pb[:, 2] - pb[:, 0]) * (pb[:, 3] - pb[:, 1])


def z_loss(gt_boxes, box6, R, K4, clamp, proj):
    """gt_boxes (n,4), box6 (n,6) = xyz + dims, R (n,3,3), K4 (n,4), clamp (n,2),
    proj (n,4) projected boxes -> (n,) scores (already halved)."""
    max_count = 50
    n = box6.shape[0]
    scores = torch.zeros(n)
    gt_area = (gt_boxes[:, 2] - gt_boxes[:, 0]) * (gt_boxes[:, 3] - gt_boxes[:, 1])
    pc = (proj[:, :2] + proj[:, 2:]) / 2
    pred_area = (proj[:, 2] - proj[:, 0]) * (proj[:, 3] - proj[:, 1])
    # the chained comparison of :545-546 compares a BOOL with the upper bound
    within = ((gt_boxes[:, 0] - max_count <= pc[:, 0]) <= gt_boxes[:, 2] + max_count) & \
             ((gt_boxes[:, 1] - max_count <= pc[:, 1]) <= gt_boxes[:, 3] + max_count)
    values = torch.linspace(0.0, (max_count - 1) / 10, max_count)
    smaller = gt_area < pred_area
    for i in range(n):
        if within[i]:
            mod = box6[i].clone().unsqueeze(0).repeat((max_count, 1))
            if smaller[i]:
                mod[:, 2] += values
            else:
                mod[:, 2] -= values
            areas = _proj_area(mod, R[i:i + 1], K4[i:i + 1], clamp[i:i + 1])
            areas = areas + (areas == 0) * 10000000
            idx = torch.argmin(l1_loss(gt_area[i].repeat(max_count), areas))
            scores[i] = l1_loss(box6[i, 2], mod[idx, 2])
        else:
            scores[i] = torch.tensor(0.1 * max_count)
    return scores / 2


def depth_range_loss(gt_mask, at_which_mask_idx, depth_maps, box6, R, gt_boxes, num_instances):
    """depth_maps: list of (H, W) f32; -> (n,) |gt_range - pred_range|."""
    corners_z = cuboid_corners(box6, R)[:, :, 2]
    pred = torch.max(corners_z, dim=1)[0] - torch.min(corners_z, dim=1)[0]
    gts, counter = [], 0
    for depth_map, ni in zip(depth_maps, num_instances):
        for _ in range(ni):
            seg = gt_mask[at_which_mask_idx[counter]][0]
            depth_map = F.interpolate(depth_map[None, None], size=seg.shape, mode="bilinear",
                                      align_corners=True).squeeze()
            rng = depth_map[seg]
            if rng.numel() == 0:
                b = gt_boxes[counter]
                rng = depth_map[b[1].long():b[3].long(), b[0].long():b[2].long()]
            gts.append(torch.quantile(rng, 0.9) - torch.quantile(rng, 0.1))
            counter += 1
    return l1_loss(torch.stack(gts), pred), torch.stack(gts)


def normal_to_rotation(normal):
    x1 = torch.tensor([1.0, 0, 0]).repeat(normal.shape[0], 1)
    t0 = torch.cross(normal, x1, dim=1)
    if torch.bmm(t0.view(-1, 1, 3), t0.view(-1, 3, 1)).flatten().any() < 0.001:
        y1 = torch.tensor([0, 1.0, 0]).repeat(normal.shape[0], 1)
        t0 = torch.cross(normal, y1, dim=1)
    t0 = t0 / torch.norm(t0)
    t1t = torch.cross(normal, t0, dim=1)
    t1 = t1t / torch.norm(t1t)
    return torch.cat([t0, t1, normal], dim=1).reshape((normal.shape[0], 3, 3))


def so3_relative_angle_cos(R1, R2, eps=1e-4):
    R12 = torch.bmm(R1, R2.permute(0, 2, 1))
    tr = R12[:, 0, 0] + R12[:, 1, 1] + R12[:, 2, 2]
    if ((tr < -1.0 - eps) + (tr > 3.0 + eps)).any():
        raise ValueError("A matrix has trace outside valid range [-1-eps,3+eps].")
    return (tr - 1.0) * 0.5


def pose_ground2(R, normals, ground_conf, num_boxes_per_image):
    nb = torch.tensor(list(num_boxes_per_image))
    nv = normals.repeat_interleave(nb, 0)
    conf = ground_conf.repeat_interleave(nb, 0)
    gt = normal_to_rotation(nv)
    return (1 - so3_relative_angle_cos(R, gt)) * conf
