"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

CPU restatement of north_star steps 2+3 (cube assembly, projection, 2D IoU,
point-in-oriented-cube count, combined score, top-k) over oracle/csrc/wcr_oracle.c.

Follows
  ProposalNetwork/utils/utils.py:111-145   vec_perp_t / rotate_vector_t / orthobasis_from_normal_t
  cubercnn/util/math_util.py:198-217       get_cuboid_verts_faces (corner order, axis mapping)
  ProposalNetwork/utils/spaces.py:224-245  get_bube_corners (projection + clamp)
  ProposalNetwork/utils/conversions.py:25-48 cubes_to_box
  ProposalNetwork/utils/utils.py:181-187   iou_2d (detectron2 pairwise_iou formula)
  cubercnn/modeling/roi_heads/roi_heads.py:590-597, 645-652, 927-928  ROI slicing, depth lookup, lifting

Parity status: block level PINNED (tests/test_oracle_cubes.py runs the reference's
own functions when /root/reference is present; tests/golden/cubes_*.npz otherwise).
Composite level (which block feeds which, point-in-cube, top-k) is THIS PROJECT'S
SPEC -- the reference snapshot has no such function: "parity
unpinned at composite level", cross-checked against oracle/compose.py.
"""
from __future__ import annotations

import numpy as np

from . import clib

f32 = np.float32


def K4_of(K) -> np.ndarray:
    K = np.asarray(K, dtype=f32)
    return np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], dtype=f32)


def box_center_depth(depth, box):
    return clib.box_center_depth(depth, box)


def assemble_cubes(K, normal, attrs):
    """attrs (7,N) = (u v z w h l yaw) -> (N,15) cubes [c, w h l, R row-major]."""
    return clib.assemble_cubes(K4_of(K), normal, attrs)


def project_cube(cube, K, clamp):
    return clib.project_cube(cube, K4_of(K), float(clamp[0]), float(clamp[1]))


def score_box(depth, K, gbox, clamp, cubes, stride=1, e_iou=1, e_pts=1, e_seg=0, seg_mode=0,
              seg=None, fast_seg=False):
    return clib.score_box(depth, K4_of(K), gbox, float(clamp[0]), float(clamp[1]), cubes, stride,
                          e_iou, e_pts, e_seg, seg_mode, seg, fast_seg)


def box_roi(gbox, H, W):
    """Integer ROI of a GT box: Boxes.clip((H, W)) then .long() (roi_heads.py:590-597)."""
    g = np.asarray(gbox, dtype=f32)
    u0, u1 = int(min(max(g[0], 0), W)), int(min(max(g[2], 0), W))
    v0, v1 = int(min(max(g[1], 0), H)), int(min(max(g[3], 0), H))
    return u0, v0, u1, v1


def topk(scores: np.ndarray, k: int):
    """Descending, ties -> lower proposal index (stable sort), -1 padded."""
    order = np.argsort(-scores.astype(np.float64), kind="stable")[:k]
    idx = np.full(k, -1, dtype=np.int32)
    idx[:len(order)] = order
    return idx


def score_images(images, N_cubes_per_box, cfg=None, stride=1, e_iou=1, e_pts=1, k=16):
    """images: list of dict(depth, K, boxes (b,4), clamp (c0,c1)); N_cubes_per_box:
    list (per image) of (b, N, 15). -> list of per-box result dicts (+ 'topk')."""
    out = []
    for im, cubes in zip(images, N_cubes_per_box):
        depth = np.asarray(im["depth"], dtype=f32)
        for b in range(len(im["boxes"])):
            r = score_box(depth, im["K"], np.asarray(im["boxes"][b], dtype=f32), im["clamp"],
                          np.asarray(cubes[b], dtype=f32), stride, e_iou, e_pts)
            r["topk"] = topk(r["score"], k)
            out.append(r)
    return out
