"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

The proposal composites (SURVEY A2.7 `propose`, A3.12 `score` + top-k) written the
plain way, in torch, from the reference's own building blocks in the reference's own
style -- an independent cross-check of oracle/csrc/wcr_oracle.c (whose rounding order
is what the CUDA kernels are bit-compared with).

Blocks and where they come from (reference paths):
  sample_normal_in_range          ProposalNetwork/utils/utils.py:41-59
  orthobasis_from_normal_t        ProposalNetwork/utils/utils.py:119-131 (+ :111-117, :133-145)
  lifting through K               cubercnn/modeling/roi_heads/roi_heads.py:927-928
  pseudo depth at the box centre  cubercnn/modeling/roi_heads/roi_heads.py:645-652
  Cubes / get_bube_corners        ProposalNetwork/utils/spaces.py:95-245
  cubes_to_box                    ProposalNetwork/utils/conversions.py:25-48
  iou_2d                          ProposalNetwork/utils/utils.py:181-187
  convex_hull / mask_iou          ProposalNetwork/utils/utils.py:225-245, 410-493

When /root/reference is importable (authoring container) the reference's functions are
called directly (`blocks = refshim.load()`); otherwise the oracle's pinned restatements
are used. Parity status: composite level UNPINNED -- the reference snapshot ships no
driver that chains these blocks; this file and the C restatement
are two statements of this project's spec and are tested against each other
(tests/test_oracle_compose.py).
"""
from __future__ import annotations

import math

import numpy as np
import torch

from . import cubes as oc, hull as oh, refshim, standins


def _ref():
    return refshim.load() if refshim.available() else None


def draw_attributes(box, z0, prior, N, *, sigma_xy=0.25, sigma_z=0.15, z_lo=0.5, z_hi=1.5, dim_lo=0.05,
                    dim_hi_sigmas=3.0, yaw_lo=0.0, yaw_hi=math.pi, generator_seed=0):
    """(7, N) = (u v z w h l yaw) for ONE box with the reference's sampler
    (sample_normal_in_range) under ``torch.manual_seed``; yaw ~ U[yaw_lo, yaw_hi).
    ``prior`` (2,3) = [[w h l mean], [w h l std]]. Defaults == wcr.proposals.ProposalConfig."""
    ns = _ref()
    assert ns is not None, "drawing with the reference's sampler needs /root/reference"
    torch.manual_seed(generator_seed)
    box = torch.as_tensor(box, dtype=torch.float32)
    prior = torch.as_tensor(prior, dtype=torch.float32)
    bw, bh = box[2] - box[0], box[3] - box[1]
    snr = ns.utils.sample_normal_in_range
    u = snr(((box[0] + box[2]) / 2).view(1), (sigma_xy * bw).view(1), N, float(box[0]), box[2].view(1))[0]
    v = snr(((box[1] + box[3]) / 2).view(1), (sigma_xy * bh).view(1), N, float(box[1]), box[3].view(1))[0]
    z0 = torch.as_tensor(z0, dtype=torch.float32)
    z = snr(z0.view(1), (sigma_z * z0).view(1), N, float(z0 * z_lo), (z0 * z_hi).view(1))[0]
    dims = snr(prior[0], prior[1], N, dim_lo, prior[0] + dim_hi_sigmas * prior[1])
    yaw = yaw_lo + (yaw_hi - yaw_lo) * torch.rand(N)
    return torch.stack((u, v, z, dims[0], dims[1], dims[2], yaw)).numpy().astype(np.float32)


def assemble(K, normal, attrs):
    """(N, 15) cubes from (7, N) attributes: lift the centre through K, yaw -> R by
    orthobasis_from_normal_t (columns [x, n, y])."""
    K = torch.as_tensor(K, dtype=torch.float32)
    a = torch.as_tensor(attrs, dtype=torch.float32)
    ns = _ref()
    if ns is not None:
        R = ns.utils.orthobasis_from_normal_t(torch.as_tensor(normal, dtype=torch.float32), a[6])
    else:
        cub = oc.assemble_cubes(K.numpy(), np.asarray(normal, np.float32), a.numpy())
        R = torch.from_numpy(cub[:, 6:]).view(-1, 3, 3)
    x = a[2] * (a[0] - K[0, 2]) / K[0, 0]
    y = a[2] * (a[1] - K[1, 2]) / K[1, 1]
    return torch.cat((torch.stack((x, y, a[2], a[3], a[4], a[5]), 1), R.reshape(-1, 9)), 1)


def score(depth, K, gbox, clamp, cubes, seg=None, e_iou=1, e_pts=1, e_seg=0, seg_mode=0, stride=1):
    """dict(score, iou, count, npts, seg) for ONE box, the textbook way:
    q = R^T (p - c), inside iff |q| <= (l/2, h/2, w/2); IoU of the projected box;
    mask term = mask_iou(seg in the ROI, convex_hull(second-clamped corners) > 0)."""
    depth = torch.as_tensor(depth, dtype=torch.float32)
    K = torch.as_tensor(K, dtype=torch.float32)
    cubes = torch.as_tensor(cubes, dtype=torch.float32)
    H, W = depth.shape
    u0, v0, u1, v1 = oc.box_roi(gbox, H, W)
    vv, uu = torch.meshgrid(torch.arange(v0, v1, stride), torch.arange(u0, u1, stride), indexing="ij")
    z = depth[vv, uu]
    ok = (z > 0) & torch.isfinite(z)
    P = torch.stack((z * (uu - K[0, 2]) / K[0, 0], z * (vv - K[1, 2]) / K[1, 1], z), -1)[ok].double()
    npts = int(ok.sum())
    ns = _ref()
    N = cubes.shape[0]
    if ns is not None:
        C = ns.spaces.Cubes(cubes.unsqueeze(0))
        pbox = ns.conversions.cubes_to_box(C, K, list(clamp))[0].tensor
        corners = C.get_bube_corners(K, list(clamp))[0]
        iou = ns.utils.iou_2d(standins.Boxes(torch.as_tensor(gbox, dtype=torch.float32).view(1, 4)), standins.Boxes(pbox))
    else:
        pb = [oc.project_cube(c.numpy(), K.numpy(), clamp) for c in cubes]
        pbox = torch.from_numpy(np.stack([p[0] for p in pb]))
        corners = torch.from_numpy(np.stack([p[1] for p in pb]))
        iou = standins.pairwise_iou(standins.Boxes(torch.as_tensor(gbox, dtype=torch.float32).view(1, 4)),
                                    standins.Boxes(pbox)).flatten()
    count = torch.zeros(N, dtype=torch.int64)
    near = torch.zeros(N, dtype=torch.int64)      # points within 1e-5 (relative) of a face
    for j in range(N):
        c = cubes[j].double()
        if not (c[3] > 0 and c[4] > 0 and c[5] > 0) or npts == 0:
            continue
        q = (P - c[:3]) @ c[6:].view(3, 3)
        m = q.abs() / (torch.stack((c[5], c[4], c[3])) / 2)
        count[j] = int((m <= 1).all(1).sum())
        near[j] = int(((m - 1).abs() < 1e-5).any(1).sum())
    s_pts = count.float() / max(npts, 1) if npts else torch.zeros(N)
    s_seg = torch.ones(N)
    if e_seg > 0 and seg is not None:
        seg_roi = np.zeros((H, W), dtype=bool)
        seg_roi[v0:v1, u0:u1] = np.asarray(seg)[v0:v1, u0:u1] != 0
        hi0, hi1 = int(clamp[0] - 1), int(clamp[1] - 1)
        for j in range(N):
            cj = corners[j]
            if not torch.isfinite(cj).all():
                s_seg[j] = 0.0
                continue
            c2 = torch.stack((cj[:, 0].clamp(0, hi0), cj[:, 1].clamp(0, hi1)), 1)   # roi_heads.py:950-952
            if ns is not None:
                bube = ns.utils.convex_hull(torch.zeros(H, W), c2.clone()).numpy() > 0
            else:
                bube = oh.convex_hull((H, W), c2.numpy()) > 0
            s_seg[j] = float(oh.mod_mask_iou(seg_roi, bube) if seg_mode == 1 else oh.mask_iou(seg_roi, bube))
    sc = iou ** e_iou * s_pts ** e_pts * s_seg ** e_seg
    return {"score": sc.numpy(), "iou": iou.numpy(), "count": count.numpy(), "near": near.numpy(),
            "npts": npts, "seg": s_seg.numpy(), "proj": pbox.numpy()}
