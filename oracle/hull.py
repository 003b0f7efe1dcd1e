"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

CPU restatement of the reference's convex-hull rasteriser and mask IoUs over
oracle/csrc/wcr_oracle.c.

Follows
  ProposalNetwork/utils/utils.py:373-408  direction / distance_sq / findDuplicates
  ProposalNetwork/utils/utils.py:410-454  jarvis_march (duplicate bump, start rule, flip)
  ProposalNetwork/utils/utils.py:456-488  fill_polygon (product of clamped half-plane values)
  ProposalNetwork/utils/utils.py:490-493  convex_hull
  ProposalNetwork/utils/utils.py:225-245  mask_iou / mod_mask_iou
  cubercnn/modeling/roi_heads/roi_heads.py:428, 950-952  call site: second clamp of the corners

Parity status: block level PINNED -- tests/test_oracle_hull.py runs the reference's
own convex_hull / mask_iou on its own five fixture point sets (utils.py:500-544) and
on seeded random corner sets when /root/reference is present;
tests/golden/hull_*.npz holds the reference's outputs otherwise. The composite mask
term (seg_term: which pixels, which ROI, binarisation) is THIS PROJECT'S SPEC
(SURVEY A3.12): parity unpinned at composite level.
"""
from __future__ import annotations

import numpy as np

from . import clib

f32 = np.float32


def jarvis_march(points):
    return clib.jarvis_march(points)


def fill_polygon(hull, H, W):
    return clib.fill_polygon(hull, H, W)


def convex_hull(mask_shape, coords):
    """utils.py:490-493: soft mask (H, W) f32 of the hull of ``coords``; an unclosed
    march (the reference would not terminate) gives zeros."""
    H, W = mask_shape
    hull = jarvis_march(coords)
    if hull is None:
        return np.zeros((H, W), dtype=f32)
    return fill_polygon(hull, H, W)


def mask_iou(seg, bube):
    """utils.py:225-234 with fp32 accumulation of exact small integers/fractions:
    intersection = sum(seg * bube); 0 -> 0.0; union = count(seg | bube)."""
    seg = np.asarray(seg)
    bube = np.asarray(bube)
    inter = (seg.astype(f32) * bube.astype(f32)).sum(dtype=np.float64)
    if inter == 0:
        return f32(0.0)
    union = np.logical_or(seg != 0, bube != 0).sum()
    return f32(f32(inter) / f32(union))


def mod_mask_iou(seg, bube):
    """utils.py:236-245: intersection**5 / union."""
    seg = np.asarray(seg)
    bube = np.asarray(bube)
    inter = (seg.astype(f32) * bube.astype(f32)).sum(dtype=np.float64)
    if inter == 0:
        return f32(0.0)
    union = np.logical_or(seg != 0, bube != 0).sum()
    return f32(f32(inter) ** 5 / f32(union))


def seg_term(corners, clamp, seg, roi, mode=0):
    """Composite mask term of one cube (see orc_seg_term): -> (s_seg, inter, bube)."""
    return clib.seg_term(corners, float(clamp[0]), float(clamp[1]), seg, roi, mode)
