"""Drop-ins for the helpers of ProposalNetwork/utils/utils.py that sit on the hot
path, each one kernel launch in libwcr_b200 (csrc/blocks.cu):

  sample_normal_in_range (:41-59)   orthobasis_from_normal_t (:119-131)
  vectorized_linspace   (:169-176)  iou_2d (:181-187)
  mask_iou (:225-234)  mod_mask_iou (:236-245)  convex_hull (:490-493)
  randn_orthobasis_torch (:61-68)   gt_in_norm_range (:148-152)
  custom_mapping (:207-223)         is_gt_included (:257-298)

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


def randn_orthobasis_torch(num_samples=1, num_instances=1, *, z: torch.Tensor | None = None,
                           seed: int | None = None, device="cuda"):
    """utils.py:61-68 -> (num_instances, num_samples, 3, 3) random orthonormal bases. ``z``: the
    (num_instances, num_samples, 3, 3) ``torch.randn`` draws of the reference (replay mode, exact
    parity); None => in-kernel Philox (``seed`` None: taken from torch's global generator)."""
    lib = _abi.load()
    n = int(num_instances) * int(num_samples)
    if z is not None:
        _abi.require_cuda(z)
        device = z.device
        z = _f32c(z).reshape(n, 9)
    device = torch.device(device)
    if device.type != "cuda":
        raise _abi.WcrError("wcr: CUDA device required (there is no CPU fallback on this path)")
    if seed is None:
        seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    out = torch.empty((n, 9), dtype=torch.float32, device=device)
    _abi.check(lib.wcr_randn_orthobasis(_abi.ptr(z), n, int(seed), out.data_ptr(), _abi.stream_ptr(device)),
               "wcr_randn_orthobasis")
    return out.view(int(num_instances), int(num_samples), 3, 3)


def gt_in_norm_range(range, gt):  # noqa: A002 - the reference's parameter name
    """utils.py:148-152: ``(gt - range[0]) / abs(range[1] - range[0])``. ``range``: a pair of
    scalars or tensors broadcastable against ``gt``; one launch for all elements."""
    gt_t = torch.as_tensor(gt)
    _abi.require_cuda(gt_t)
    lib = _abi.load()
    dev = gt_t.device
    g = _f32c(gt_t)
    lo = _f32c(torch.as_tensor(range[0]), dev).expand_as(g).contiguous()
    hi = _f32c(torch.as_tensor(range[1]), dev).expand_as(g).contiguous()
    out = torch.empty_like(g)
    _abi.check(lib.wcr_gt_in_norm_range(lo.data_ptr(), hi.data_ptr(), g.data_ptr(), g.numel(),
                                        out.data_ptr(), _abi.stream_ptr(dev)), "wcr_gt_in_norm_range")
    return out


def custom_mapping(x, beta=1.7):
    """utils.py:207-223: S-curve ``1 / (1 + (x / (1 - x)) ** -beta)`` (0 for x <= 0). A list in, a
    list of Python floats out, like the reference; a CUDA tensor in, a tensor out (no sync)."""
    lib = _abi.load()
    as_list = not isinstance(x, torch.Tensor)
    t = torch.as_tensor(x, dtype=torch.float32)
    if as_list:
        t = t.cuda()
    _abi.require_cuda(t)
    t = t.contiguous()
    out = torch.empty_like(t)
    _abi.check(lib.wcr_custom_mapping(t.data_ptr(), t.numel(), float(beta), out.data_ptr(),
                                      _abi.stream_ptr(t.device)), "wcr_custom_mapping")
    return [float(v) for v in out.cpu().tolist()] if as_list else out


_REASONS = ("x", "y", "z", "w-", "w+", "h-", "h+", "l-", "l+")


def is_gt_included_batch(gt_cubes, x_range, y_range, z_range, w_prior, h_prior, l_prior):
    """``is_gt_included`` for n GT cubes in one launch. ``gt_cubes`` (n, 15) or a ``Cubes``;
    ranges (n, 2) each; priors (n, 2) = (mean, std) each. -> (bits (n,) int32, by (n, 3)):
    bits == 0 <=> included (bit order: x y z w- w+ h- h+ l- l+)."""
    lib = _abi.load()
    c = gt_cubes.tensor if hasattr(gt_cubes, "tensor") else gt_cubes
    _abi.require_cuda(c)
    dev = c.device
    c = _f32c(c).reshape(-1, 15)
    n = c.shape[0]
    ranges = torch.cat([_f32c(torch.as_tensor(r), dev).reshape(-1, 2).expand(n, 2)
                        for r in (x_range, y_range, z_range)], dim=1).contiguous()
    priors = torch.cat([_f32c(torch.as_tensor(p), dev).reshape(-1, 2).expand(n, 2)
                        for p in (w_prior, h_prior, l_prior)], dim=1).contiguous()
    bits = torch.empty((n,), dtype=torch.int32, device=dev)
    by = torch.empty((n, 3), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_is_gt_included(c.data_ptr(), ranges.data_ptr(), priors.data_ptr(), n, bits.data_ptr(),
                                      by.data_ptr(), _abi.stream_ptr(dev)), "wcr_is_gt_included")
    return bits, by


def is_gt_included(gt_cube, x_range, y_range, z_range, w_prior, h_prior, l_prior):
    """utils.py:257-298 for ONE GT cube (an object with ``center`` / ``dimensions`` like the
    reference's ``Cube``, a ``Cubes`` of one cube, or a (15,) tensor): True when the GT lies inside
    the search ranges and within 1.5 std of the dimension priors, else prints the reference's
    message and returns False. One small host read (the reference is host code throughout)."""
    if hasattr(gt_cube, "center") and hasattr(gt_cube, "dimensions"):
        c = torch.zeros(15)
        c[0:3] = torch.as_tensor(gt_cube.center, dtype=torch.float32).reshape(-1)[:3]
        c[3:6] = torch.as_tensor(gt_cube.dimensions, dtype=torch.float32).reshape(-1)[:3]
        c = c.cuda()
    else:
        c = (gt_cube.tensor if hasattr(gt_cube, "tensor") else torch.as_tensor(gt_cube)).reshape(-1)[:15]
        c = c if c.is_cuda else c.cuda()
    dev = c.device
    mk = lambda v: torch.as_tensor([float(v[0]), float(v[1])], dtype=torch.float32, device=dev)  # noqa: E731
    bits, by = is_gt_included_batch(c.view(1, 15), mk(x_range), mk(y_range), mk(z_range), mk(w_prior),
                                    mk(h_prior), mk(l_prior))
    b, dist = int(bits.item()), by[0].cpu().tolist()
    if b == 0:
        return True
    because_of = [f"{_REASONS[i]} by {dist[i]:.1f}" for i in range(3) if b >> i & 1]
    because_of += [_REASONS[i] for i in range(3, 9) if b >> i & 1]
    print("GT cannot be found due to", because_of)
    return False
