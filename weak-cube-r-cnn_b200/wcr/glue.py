"""Caller glue around the fused weak-loss block.

The reference prepares the loss inputs with a page of small torch ops and Python loops
(roi_heads.py:748-846) and converts the pose head's 6-D output for EVERY class before
throwing all but one class away (cube_head.py:174-176, roi_heads.py:836). Here:

* ``rotation_6d_to_matrix``  pytorch3d's name and semantics, one fused fwd and one bwd kernel
* ``class_gather``           ``x[fg_inds, box_classes, ...]`` with a scatter backward
* ``box_constants``          Ks_scaled_per_box / virtual_to_real / im_sizes per box in one launch
* ``cube_losses_from_head``  NEW: head outputs (per class) -> the reference's loss dict; the
                             6-D pose is converted for the gathered class only.
No CPU fallback.
"""
from __future__ import annotations

import torch

from . import _abi
from .roi_heads import WeakLossConfig, weak_loss


class _Rot6dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, d6):
        lib = _abi.load()
        x = d6.detach().to(torch.float32).contiguous().view(-1, 6)
        R = torch.empty((x.shape[0], 9), dtype=torch.float32, device=x.device)
        _abi.check(lib.wcr_rotation_6d_fwd(x.data_ptr(), x.shape[0], R.data_ptr(), _abi.stream_ptr(x.device)),
                   "wcr_rotation_6d_fwd")
        ctx.save_for_backward(x)
        ctx.shape = d6.shape
        return R.view(*d6.shape[:-1], 3, 3)

    @staticmethod
    def backward(ctx, g):
        lib = _abi.load()
        (x,) = ctx.saved_tensors
        g = g.to(torch.float32).contiguous().view(-1, 9)
        d = torch.empty_like(x)
        _abi.check(lib.wcr_rotation_6d_bwd(x.data_ptr(), g.data_ptr(), x.shape[0], d.data_ptr(),
                                           _abi.stream_ptr(x.device)), "wcr_rotation_6d_bwd")
        return d.view(ctx.shape)


def rotation_6d_to_matrix(d6: torch.Tensor) -> torch.Tensor:
    """pytorch3d.transforms.rotation_6d_to_matrix (cube_head.py:175): (..., 6) -> (..., 3, 3)."""
    _abi.require_cuda(d6)
    return _Rot6dFn.apply(d6)


class _ClassGatherFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, classes):
        lib = _abi.load()
        n, C = src.shape[0], src.shape[1]
        k = int(src[0, 0].numel()) if n else int(torch.tensor(src.shape[2:]).prod()) if src.dim() > 2 else 1
        x = src.detach().to(torch.float32).contiguous()
        cls = classes.to(device=src.device, dtype=torch.int64).contiguous()
        out = torch.empty((n,) + tuple(src.shape[2:]), dtype=torch.float32, device=src.device)
        _abi.check(lib.wcr_class_gather_fwd(x.data_ptr(), n, C, k, cls.data_ptr(), out.data_ptr(),
                                            _abi.stream_ptr(src.device)), "wcr_class_gather_fwd")
        ctx.save_for_backward(cls)
        ctx.meta = (tuple(src.shape), k)
        return out

    @staticmethod
    def backward(ctx, g):
        lib = _abi.load()
        (cls,) = ctx.saved_tensors
        shape, k = ctx.meta
        g = g.to(torch.float32).contiguous()
        d = torch.empty(shape, dtype=torch.float32, device=g.device)
        _abi.check(lib.wcr_class_gather_bwd(g.data_ptr(), shape[0], shape[1], k, cls.data_ptr(), d.data_ptr(),
                                            _abi.stream_ptr(g.device)), "wcr_class_gather_bwd")
        return d, None


def class_gather(x: torch.Tensor, box_classes: torch.Tensor) -> torch.Tensor:
    """``x[torch.arange(n), box_classes]`` (roi_heads.py:823-840) for x of shape (n, C, ...)."""
    _abi.require_cuda(x)
    return _ClassGatherFn.apply(x, box_classes)


def box_constants(Ks, im_scales_ratio, im_current_dims, num_boxes_per_image, *, virtual_focal=512.0,
                  virtual_depth=True, device=None):
    """roi_heads.py:748-781, 810-814 in one launch.

    Ks: list/tensor of (3,3) per image; im_scales_ratio: list of floats; im_current_dims: list of
    (h, w). -> K4 (n,4) = Ks_scaled_per_box's fx fy cx cy, virtual_to_real (n,), clamp (n,2) =
    im_sizes."""
    lib = _abi.load()
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    B = len(num_boxes_per_image)
    K = torch.stack([torch.as_tensor(k, dtype=torch.float32).reshape(3, 3) for k in Ks]).reshape(B, 9).to(dev)
    ratio = torch.as_tensor([float(r) for r in im_scales_ratio], dtype=torch.float32).to(dev)
    dims = torch.as_tensor([[float(d[0]), float(d[1])] for d in im_current_dims], dtype=torch.float32).to(dev)
    off = torch.zeros(B + 1, dtype=torch.int32)
    off[1:] = torch.cumsum(torch.as_tensor(list(num_boxes_per_image), dtype=torch.int32), 0)
    n = int(off[-1])
    off = off.to(dev)
    K4 = torch.empty((n, 4), device=dev)
    v2r = torch.empty((n,), device=dev)
    clamp = torch.empty((n, 2), device=dev)
    _abi.check(lib.wcr_box_constants(K.data_ptr(), ratio.data_ptr(), dims.data_ptr(), off.data_ptr(), B, n,
                                     float(virtual_focal), int(bool(virtual_depth)), K4.data_ptr(),
                                     v2r.data_ptr(), clamp.data_ptr(), _abi.stream_ptr(dev)), "wcr_box_constants")
    return K4, v2r, clamp


def cube_losses_from_head(cfg: WeakLossConfig, cube_2d_deltas, cube_z, cube_dims, cube_pose, cube_uncert,
                          box_classes, *, proposal_boxes, gt_boxes, Ks, im_scales_ratio, im_current_dims,
                          num_boxes_per_image, priors_dims_per_cat, depth_maps, normals, ground_conf,
                          virtual_focal=512.0, virtual_depth=True, pose_is_6d=False, gt_masks=None,
                          at_which_mask_idx=None, return_per_box=False, host_sync=True):
    """NEW (no reference counterpart as one call): roi_heads.py:748-1144 from the cube head's
    per-class outputs to the loss dict.

    cube_2d_deltas (n,C,2), cube_z (n,C) or (n,C,1), cube_dims (n,C,3), cube_pose (n,C,3,3) or, with
    ``pose_is_6d``, the raw (n,C,6) head output (converted for the box's own class only),
    cube_uncert (n,C) or None, box_classes (n,), priors_dims_per_cat (C,2,3) = (mean, std)."""
    n = cube_2d_deltas.shape[0]
    dev = cube_2d_deltas.device
    deltas = class_gather(cube_2d_deltas, box_classes)
    z = class_gather(cube_z.reshape(n, cube_z.shape[1], -1), box_classes).reshape(n)
    dims = class_gather(cube_dims, box_classes)
    if pose_is_6d:
        pose = rotation_6d_to_matrix(class_gather(cube_pose.reshape(n, -1, 6), box_classes))
    else:
        pose = class_gather(cube_pose.reshape(n, -1, 9), box_classes).view(n, 3, 3)
    if cube_uncert is not None:
        uncert = class_gather(cube_uncert.reshape(n, -1, 1), box_classes).reshape(n)
    else:
        uncert = torch.zeros(n, device=dev)
    pri = priors_dims_per_cat.detach().to(dev)[box_classes.to(dev)]          # (n, 2, 3): plain index, no grad
    K4, v2r, clamp = box_constants(Ks, im_scales_ratio, im_current_dims, num_boxes_per_image,
                                   virtual_focal=virtual_focal, virtual_depth=virtual_depth, device=dev)
    return weak_loss(cfg, deltas, z, dims, pose.reshape(n, 9), uncert, src_boxes=proposal_boxes, K4=K4,
                     virtual_to_real=v2r, prior_mean=pri[:, 0, :].contiguous(), prior_std=pri[:, 1, :].contiguous(),
                     gt_boxes=gt_boxes, clamp=clamp, num_boxes_per_image=num_boxes_per_image,
                     depth_maps=depth_maps, normals=normals, ground_conf=ground_conf, gt_masks=gt_masks,
                     at_which_mask_idx=at_which_mask_idx, return_per_box=return_per_box, host_sync=host_sync)
