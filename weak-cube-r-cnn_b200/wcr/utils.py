"""Drop-ins for the helpers of ProposalNetwork/utils/utils.py that sit on the hot
path, each one kernel launch in libwcr_b200 (csrc/blocks.cu):

  sample_normal_in_range (:41-59)   orthobasis_from_normal_t (:119-131)
  vectorized_linspace   (:169-176)  iou_2d (:181-187)
  mask_iou (:225-234)  mod_mask_iou (:236-245)  convex_hull (:490-493)

Same positional signatures and return shapes. CUDA tensors only (no CPU fallback).
"""
from __future__ import annotations

import torch

from . import _abi


def _f32c(t, device=None):
    t = torch.as_tensor(t)
    return t.to(device=device or t.device, dtype=torch.float32).contiguous()


def sample_normal_in_range(means, stds, count, threshold_low=None, threshold_high=None, *,
                           seed: int | None = None, max_tries: int = 10000):
    """(M,) means, (M,) stds -> (M, count) normal samples; when both thresholds are
    given (``threshold_low`` scalar, ``threshold_high`` (M,) tensor) every sample is
    redrawn until it lies in range (utils.py:47-57; there: whole-tensor redraws, up
    to 10 000 of them -- here: per-element redraws from the in-kernel Philox
    generator, at most ``max_tries``, then clamped).

    ``seed`` None => one 63-bit seed is taken from torch's global CPU generator, so
    ``torch.manual_seed`` controls the stream as it does in the reference."""
    _abi.require_cuda(means, stds)
    lib = _abi.load()
    dev = means.device
    means, stds = _f32c(means), _f32c(stds, dev)
    M = means.shape[0]
    has = threshold_low is not None and threshold_high is not None
    hi = _f32c(threshold_high, dev).expand(M).contiguous() if has else None
    lo = float(threshold_low) if has else 0.0
    if seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    out = torch.empty((M, int(count)), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_sample_normal_in_range(means.data_ptr(), stds.data_ptr(), M, int(count),
                                              int(has), lo, _abi.ptr(hi), int(seed), int(max_tries),
                                              out.data_ptr(), _abi.stream_ptr(dev)),
               "wcr_sample_normal_in_range")
    return out


def orthobasis_from_normal_t(normal: torch.Tensor, yaw_angles=0):
    """(3,) normal, (T,) yaw -> (T, 3, 3) with columns ``[x, normal, y]`` (det -1),
    utils.py:119-131."""
    _abi.require_cuda(normal)
    lib = _abi.load()
    dev = normal.device
    normal = _f32c(normal)
    yaw = _f32c(yaw_angles, dev).reshape(-1)
    out = torch.empty((yaw.shape[0], 3, 3), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_orthobasis_from_normal(normal.data_ptr(), yaw.data_ptr(), yaw.shape[0],
                                              out.data_ptr(), _abi.stream_ptr(dev)),
               "wcr_orthobasis_from_normal")
    return out


def vectorized_linspace(start_tensor, end_tensor, number_of_steps):
    """(M,), (M,) -> (M, steps), utils.py:169-176."""
    _abi.require_cuda(start_tensor, end_tensor)
    lib = _abi.load()
    dev = start_tensor.device
    a, b = _f32c(start_tensor), _f32c(end_tensor, dev)
    out = torch.empty((a.shape[0], int(number_of_steps)), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_vectorized_linspace(a.data_ptr(), b.data_ptr(), a.shape[0],
                                           int(number_of_steps), out.data_ptr(),
                                           _abi.stream_ptr(dev)), "wcr_vectorized_linspace")
    return out


def iou_2d(gt_box, proposal_boxes):
    """``gt_box``: Boxes holding ONE box; ``proposal_boxes``: Boxes (P) -> (P,) IoUs,
    utils.py:181-187 (detectron2 pairwise_iou, flattened)."""
    g = gt_box.tensor if hasattr(gt_box, "tensor") else gt_box
    p = proposal_boxes.tensor if hasattr(proposal_boxes, "tensor") else proposal_boxes
    _abi.require_cuda(g, p)
    lib = _abi.load()
    dev = p.device
    g, p = _f32c(g, dev).reshape(-1, 4), _f32c(p).reshape(-1, 4)
    outs = []
    for i in range(g.shape[0]):      # pairwise_iou is (n_gt, P); the reference passes n_gt = 1
        out = torch.empty((p.shape[0],), dtype=torch.float32, device=dev)
        _abi.check(lib.wcr_iou_2d(g[i].data_ptr(), p.data_ptr(), p.shape[0], out.data_ptr(),
                                  _abi.stream_ptr(dev)), "wcr_iou_2d")
        outs.append(out)
    return outs[0] if len(outs) == 1 else torch.cat(outs)


def convex_hull(mask: torch.Tensor, coords: torch.Tensor):
    """Soft mask (H, W) of the convex hull of ``coords`` (n, 2), n <= 32: Jarvis march
    (utils.py:410-454) then the product of clamped half-plane values (:456-488).
    ``mask`` only supplies the shape and device, as in the reference."""
    _abi.require_cuda(mask)
    lib = _abi.load()
    dev = mask.device
    H, W = mask.shape[-2], mask.shape[-1]
    coords = _f32c(coords, dev).reshape(-1, 2)
    out = torch.empty((H, W), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_convex_hull(coords.data_ptr(), coords.shape[0], H, W, out.data_ptr(), W,
                                   _abi.stream_ptr(dev)), "wcr_convex_hull")
    return out


def _mask_iou(segmentation_mask, bube_mask, mode):
    _abi.require_cuda(segmentation_mask)
    lib = _abi.load()
    dev = segmentation_mask.device
    seg = segmentation_mask
    if seg.dtype in (torch.bool, torch.uint8):
        seg, is_float = seg.to(torch.uint8).contiguous(), 0
    else:
        seg, is_float = seg.to(torch.float32).contiguous(), 1
    bube = _f32c(torch.as_tensor(bube_mask), dev)
    assert bube.numel() == seg.numel()
    out = torch.empty((1,), dtype=torch.float32, device=dev)
    ws = torch.empty((4,), dtype=torch.float64, device=dev)
    _abi.check(lib.wcr_mask_iou(seg.data_ptr(), is_float, bube.data_ptr(), seg.numel(), mode,
                                out.data_ptr(), ws.data_ptr(), _abi.stream_ptr(dev)), "wcr_mask_iou")
    return out[0]


def mask_iou(segmentation_mask, bube_mask):
    """sum(seg * bube) / count(seg | bube), 0 when the intersection is empty (utils.py:225-234)."""
    return _mask_iou(segmentation_mask, bube_mask, 0)


def mod_mask_iou(segmentation_mask, bube_mask):
    """intersection**5 / union (utils.py:236-245)."""
    return _mask_iou(segmentation_mask, bube_mask, 1)


def box3d_overlap(boxes1: torch.Tensor, boxes2: torch.Tensor, eps: float = 1e-4):
    """pytorch3d.ops.box3d_overlap: (N,8,3), (M,8,3) corner lists in pytorch3d's corner order
    -> (vol (N,M), iou (N,M)). Exact clipping on the device (csrc/iou3d.cu). Like pytorch3d it
    raises ``ValueError`` for a box of ~zero volume (one small host read)."""
    _abi.require_cuda(boxes1, boxes2)
    lib = _abi.load()
    dev = boxes1.device
    b1, b2 = _f32c(boxes1, dev).reshape(-1, 8, 3), _f32c(boxes2, dev).reshape(-1, 8, 3)
    N, M = b1.shape[0], b2.shape[0]
    both = torch.cat((b1, b2))
    vols = torch.empty((N + M,), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_box3d_volume(both.data_ptr(), N + M, vols.data_ptr(), _abi.stream_ptr(dev)),
               "wcr_box3d_volume")
    if N + M and bool((vols < eps).any()):
        raise ValueError("Planes have zero areas")          # pytorch3d _check_nonzero's message
    vol = torch.empty((N, M), dtype=torch.float32, device=dev)
    iou = torch.empty((N, M), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_box3d_overlap(b1.data_ptr(), N, b2.data_ptr(), M, vol.data_ptr(), iou.data_ptr(),
                                     _abi.stream_ptr(dev)), "wcr_box3d_overlap")
    return vol, iou


def iou_3d(gt_cube, proposal_cubes):
    """utils.py:189-205: IoU of ONE GT cube with every proposal cube -> (P,)."""
    gt_corners = gt_cube.get_all_corners()[0]
    proposal_corners = proposal_cubes.get_all_corners()[0]
    vol, iou = box3d_overlap(gt_corners, proposal_corners)
    return iou[0]
