"""Weak-loss block of ``ROIHeads3DScore._forward_cube`` on B200 (north_star step 4).

``weak_loss`` replaces roi_heads.py:843-1144 (decode of the gathered head outputs
through the reduced loss dict) with one fused forward launch chain and one fused
backward launch chain behind a ``torch.autograd.Function``; the reference runs O(n)
Python iterations with ~15 torch kernels each plus an O(n_i^2) CPU pair loop.

The loss helpers keep the reference's method names and signatures
(``pose_loss``, ``pseudo_gt_z_point_loss``, ``dim_loss``, ``safely_reduce_losses``)
so ``roi_heads.py`` can import them instead of defining them. No CPU fallback.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import torch

from . import _abi
from .images import ImageShard

TERM_KEYS = ("Cube/uncert", "Cube/loss_iou", "Cube/loss_pose", "Cube/loss_normal_vec",
             "Cube/loss_pseudo_gt_z", "Cube/loss_dims_w", "Cube/loss_dims_h", "Cube/loss_dims_l",
             "Cube/loss_z", "Cube/loss_depth", "Cube/loss_seg")
N_TERMS, N_STATS = len(TERM_KEYS), 32          # WCR_LOSS_TERMS, WCR_LOSS_STATS
STAT_TRACE_ERR, STAT_SHAPE_ERR = 18, 19
_FLAG = {"iou": 1, "pose_alignment": 2, "pose_ground": 4, "z_pseudo_gt_center": 8, "dims": 16,
         "z_pseudo_gt_patch": 32, "z": 64, "depth": 128, "pose_ground2": 256, "segmentation": 512}
_ZTYPE = {"direct": 0, "sigmoid": 1, "log": 2}
_DIMS = {"exp": 0, "sigmoid": 1}


@dataclass
class WeakLossConfig:
    """cfg.MODEL.ROI_CUBE_HEAD.* and cfg.loss_functions (config.py:52-127, 186)."""
    loss_functions: tuple = ("iou", "z_pseudo_gt_center", "pose_alignment", "pose_ground", "dims")
    z_type: str = "direct"
    dims_priors_func: str = "exp"
    use_confidence: float = 1.0
    loss_w_3d: float = 1.0
    loss_w_iou: float = 1.0
    loss_w_pose: float = 7.0
    loss_w_normal_vec: float = 20.0
    loss_w_z: float = 1.0
    loss_w_dims: float = 20.0
    loss_w_depth: float = 1.0
    loss_w_seg: float = 2.5

    def c_struct(self) -> _abi.LossParams:
        unknown = [f for f in self.loss_functions if f not in _FLAG]
        if unknown:
            raise _abi.WcrError(f"wcr.weak_loss: loss functions {unknown} are not on the fused path "
                                "(SURVEY.md §8f)")
        flags = 0
        for f in self.loss_functions:
            flags |= _FLAG[f]
        return _abi.LossParams(flags, _ZTYPE[self.z_type], _DIMS[self.dims_priors_func],
                               self.use_confidence, self.loss_w_3d, self.loss_w_iou, self.loss_w_pose,
                               self.loss_w_normal_vec, self.loss_w_z, self.loss_w_dims, self.loss_w_depth,
                               self.loss_w_seg)


def _c(t, dev, dtype=torch.float32):
    return t.detach().to(device=dev, dtype=dtype).contiguous()


def resized_depth_shard(shard: ImageShard, mask_hw):
    """depth_range_loss resizes every depth map to the mask shape with
    ``F.interpolate(mode='bilinear', align_corners=True)`` (roi_heads.py:672) -- the masks have the
    padded batch-tensor size, the depth maps are iterated un-padded. -> an ``ImageShard`` of the
    resized maps (one ``wcr_resize_bilinear_ac`` launch per image whose shape differs), or None when
    every depth map already has the mask shape (the resize is the identity then)."""
    mh, mw = mask_hw
    sizes = shard.sizes()
    if all((h, w) == (mh, mw) for h, w in sizes):
        return None
    lib = _abi.load()
    dev = shard.device
    planes = torch.empty((len(sizes), mh, mw), dtype=torch.float32, device=dev)
    rec = shard.host_table
    for i, (h, w) in enumerate(sizes):
        _abi.check(lib.wcr_resize_bilinear_ac(int(rec["depth"][i]), h, w, int(rec["depth_pitch"][i]),
                                              planes[i].data_ptr(), mh, mw, mw, _abi.stream_ptr(dev)),
                   "wcr_resize_bilinear_ac")
    out = ImageShard([planes[i] for i in range(len(sizes))], None, [torch.eye(3)] * len(sizes), device=dev)
    out._planes = planes
    return out


class _Ctx:
    """Everything one fwd/bwd pair shares: contiguous device inputs + saved buffers."""

    def __init__(self, cfg, deltas, z_raw, dims_raw, pose, uncert, src_boxes, K4, virtual_to_real,
                 prior_mean, prior_std, gt_boxes, clamp, num_boxes_per_image, shard, normals,
                 ground_conf, gt_masks=None, at_which_mask_idx=None):
        _abi.require_cuda(deltas, z_raw, dims_raw, pose, uncert)
        dev = deltas.device
        n = deltas.shape[0]
        self.n, self.dev, self.cfg = n, dev, cfg
        self.params = cfg.c_struct()
        self.shard = shard
        B = len(num_boxes_per_image)
        assert len(shard) == B, "one depth map per image"
        off = torch.zeros(B + 1, dtype=torch.int32)
        off[1:] = torch.cumsum(torch.as_tensor(list(num_boxes_per_image), dtype=torch.int32), 0)
        assert int(off[-1]) == n
        t = {
            "deltas": _c(deltas, dev).view(n, 2), "z_raw": _c(z_raw, dev).view(n),
            "dims_raw": _c(dims_raw, dev).view(n, 3), "pose": _c(pose, dev).view(n, 9),
            "uncert": _c(uncert, dev).view(n), "src_boxes": _c(src_boxes, dev), "K4": _c(K4, dev),
            "v2r": _c(virtual_to_real, dev).view(n), "prior_mean": _c(prior_mean, dev),
            "prior_std": _c(prior_std, dev), "gt_boxes": _c(gt_boxes, dev), "clamp": _c(clamp, dev),
            "img_off": off.to(dev), "images": shard.table, "normals": _c(normals, dev),
            "ground_conf": _c(ground_conf, dev),
            "losses": torch.zeros(N_TERMS, device=dev),
            "present": torch.zeros(N_TERMS + 2, dtype=torch.int32, device=dev),
            "terms": torch.empty((N_TERMS - 1, max(n, 1)), device=dev),
            "R_ego": torch.empty((max(n, 1), 9), device=dev),
            "proj": torch.empty((max(n, 1), 4), device=dev), "stats": torch.zeros(N_STATS, device=dev),
            "img_pose": torch.zeros(B, device=dev),
        }
        if "z_pseudo_gt_patch" in cfg.loss_functions:
            t.update({"z_dec": torch.empty(max(n, 1), device=dev), "z_gt": torch.empty(max(n, 1), device=dev),
                      "z_inside": torch.empty(max(n, 1), dtype=torch.int32, device=dev),
                      "z_sign": torch.zeros(max(n, 1), device=dev)})
        self.mask_hw = (0, 0)
        if "depth" in cfg.loss_functions or "segmentation" in cfg.loss_functions:
            if gt_masks is None or at_which_mask_idx is None:
                raise _abi.WcrError("wcr.weak_loss: 'depth' / 'segmentation' need gt_masks and at_which_mask_idx")
            # masks_all_images (roi_heads.py:273): list of (1, H, W) bool, or one (M, H, W) tensor
            m = gt_masks if isinstance(gt_masks, torch.Tensor) else torch.stack(
                [torch.as_tensor(x).reshape(x.shape[-2], x.shape[-1]) for x in gt_masks])
            m = m.reshape(-1, m.shape[-2], m.shape[-1]).to(device=dev)
            m = (m != 0).to(torch.uint8).contiguous() if m.dtype not in (torch.uint8, torch.bool) else m.contiguous()
            t.update({"masks": m.view(torch.uint8) if m.dtype == torch.bool else m,
                      "mask_idx": torch.as_tensor(list(at_which_mask_idx), dtype=torch.int32).to(dev),
                      "depth_gt": torch.empty(max(n, 1), device=dev)})
            assert t["mask_idx"].numel() == n
            self.mask_hw = (int(m.shape[-2]), int(m.shape[-1]))
        if "depth" in cfg.loss_functions:
            t["depth_images"] = None
            self.depth_shard = resized_depth_shard(shard, self.mask_hw)
            if self.depth_shard is not None:
                t["depth_images"] = self.depth_shard.table
        if "pose_ground2" in cfg.loss_functions:
            t["gt_rot"] = torch.empty((B, 9), device=dev)
        if "depth" in cfg.loss_functions or "z_pseudo_gt_patch" in cfg.loss_functions:
            t["select_ws"] = torch.empty(int(_abi.load().wcr_select_ws_bytes(max(n, 1))), dtype=torch.uint8,
                                         device=dev)
        self.t = t
        self.n_images = B

    def io(self, extra=None) -> _abi.LossIO:
        io = _abi.LossIO()
        io.n, io.n_images = self.n, self.n_images
        tt = dict(self.t)
        if extra:
            tt.update(extra)
        for name in _abi._LOSS_IO_PTRS + _abi._LOSS_IO_PTRS2 + ["masks"]:
            v = tt.get(name)
            setattr(io, name, None if v is None else v.data_ptr())
        io.mask_h, io.mask_w = self.mask_hw
        self._keep = tt
        return io


class _WeakLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, c: _Ctx, deltas, z_raw, dims_raw, pose, uncert):
        lib = _abi.load()
        io = c.io()
        _abi.check(lib.wcr_weak_loss_fwd(ctypes.byref(io), ctypes.byref(c.params),
                                         _abi.stream_ptr(c.dev)), "wcr_weak_loss_fwd")
        ctx.c = c
        ctx.shapes = (deltas.shape, z_raw.shape, dims_raw.shape, pose.shape, uncert.shape)
        return c.t["losses"].clone()

    @staticmethod
    def backward(ctx, grad_losses):
        lib = _abi.load()
        c = ctx.c
        n, dev = c.n, c.dev
        g = {"grad_losses": grad_losses.to(torch.float32).contiguous(),
             "d_deltas": torch.zeros((n, 2), device=dev), "d_z_raw": torch.zeros((n,), device=dev),
             "d_dims_raw": torch.zeros((n, 3), device=dev), "d_pose": torch.zeros((n, 9), device=dev),
             "d_uncert": torch.zeros((n,), device=dev), "d_R_pose": torch.zeros((max(n, 1), 9), device=dev)}
        if "segmentation" in c.cfg.loss_functions:
            g["seg_dcorner"] = torch.zeros((max(n, 1), 16), device=dev)
        io = c.io(g)
        _abi.check(lib.wcr_weak_loss_bwd(ctypes.byref(io), ctypes.byref(c.params),
                                         _abi.stream_ptr(dev)), "wcr_weak_loss_bwd")
        s = ctx.shapes
        return (None, g["d_deltas"].view(s[0]), g["d_z_raw"].view(s[1]), g["d_dims_raw"].view(s[2]),
                g["d_pose"].view(s[3]), g["d_uncert"].view(s[4]))


def weak_loss(cfg: WeakLossConfig, deltas, z_raw, dims_raw, pose, uncert, *, src_boxes, K4,
              virtual_to_real, prior_mean, prior_std, gt_boxes, clamp, num_boxes_per_image,
              depth_maps, normals, ground_conf, gt_masks=None, at_which_mask_idx=None,
              return_per_box: bool = False, host_sync: bool = True):
    """Fused weak-loss block. Per-box tensors follow the reference's torch.cat order
    (boxes of one image contiguous). ``depth_maps``: an ``ImageShard`` or anything
    ``ImageShard`` accepts (ImageList-like / list of (h,w) CUDA tensors); ``normals``
    (B,3) from ``normal_vector_from_maps``; ``ground_conf`` (B,) = 0.1 where the
    ground map is the (1,1) dummy else 1.0 (roi_heads.py:988).

    -> dict keyed like the reference ('Cube/loss_iou', ...); keys the reference would
    not emit (pose term when every image has one box, dims when a prior std is NaN,
    'Cube/uncert' when USE_CONFIDENCE == 0) are absent.

    ``host_sync=True`` (default): ONE 52-byte device -> host read decides the keys exactly as
    the reference does (it syncs many times here: .item(), :1061-1073).
    ``host_sync=False``: NO host synchronisation. The keys are decided on the host from what the
    host knows (loss list, USE_CONFIDENCE, boxes per image); the two data-dependent cases -- a NaN
    prior std (the reference drops the dims keys) and the error flags -- are copied to pinned
    memory asynchronously and checked ONE CALL LATE (``poll_deferred_flags()``, also run at the
    start of the next call): a violation raises then, naming the step it belongs to."""
    shard = depth_maps if isinstance(depth_maps, ImageShard) else ImageShard(
        depth_maps, None, [torch.eye(3)] * len(num_boxes_per_image), device=deltas.device)
    c = _Ctx(cfg, deltas, z_raw, dims_raw, pose, uncert, src_boxes, K4, virtual_to_real, prior_mean,
             prior_std, gt_boxes, clamp, num_boxes_per_image, shard, normals, ground_conf, gt_masks,
             at_which_mask_idx)
    if not host_sync:
        poll_deferred_flags()
    out = _WeakLossFn.apply(c, deltas, z_raw, dims_raw, pose, uncert)
    if host_sync:
        present = c.t["present"].cpu().tolist()     # the one host sync: term flags + two error flags
        _raise_flags(present)
    else:
        present = _host_side_present(cfg, num_boxes_per_image, c.n)
        _defer_flags(c.t["present"], present)
    losses = {k: out[i] for i, k in enumerate(TERM_KEYS) if present[i]}
    if return_per_box:
        t = c.t
        per_box = {"proj": t["proj"][:c.n], "R": t["R_ego"][:c.n].view(-1, 3, 3), "terms": t["terms"][:, :c.n],
                   "stats": t["stats"]}
        return losses, per_box
    return losses


def _raise_flags(present):
    if present[N_TERMS]:
        raise ValueError("A matrix has trace outside valid range [-1-eps,3+eps].")   # pytorch3d so3_rotation_angle
    if present[N_TERMS + 1]:
        raise _abi.WcrError("wcr.weak_loss: 'depth' needs masks of the same shape as their depth maps")


_TERM_OF = {"iou": (1,), "pose_alignment": (2,), "pose_ground": (3,), "pose_ground2": (3,),
            "z_pseudo_gt_center": (4,), "z_pseudo_gt_patch": (4,), "dims": (5, 6, 7), "z": (8,),
            "depth": (9,), "segmentation": (10,)}
_deferred = []      # (step, pinned flags, event, keys assumed present)
_step_counter = [0]


def _host_side_present(cfg, num_boxes_per_image, n):
    """Which keys the reference emits, from host-side knowledge only (data-dependent NaNs aside)."""
    p = [0] * (N_TERMS + 2)
    if n == 0:
        return p
    for f in cfg.loss_functions:
        for t in _TERM_OF[f]:
            p[t] = 1
    if all(int(b) == 1 for b in num_boxes_per_image):
        p[2] = 0                                    # pose_loss returns None (roi_heads.py:462-463)
    p[0] = 1 if cfg.use_confidence > 0 else 0
    return p


def _defer_flags(present_dev, assumed):
    h = torch.empty(present_dev.shape, dtype=present_dev.dtype).pin_memory()
    h.copy_(present_dev, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(present_dev.device))
    _step_counter[0] += 1
    _deferred.append((_step_counter[0], h, ev, list(assumed)))


def poll_deferred_flags(block: bool = False):
    """Check the flags of earlier ``weak_loss(host_sync=False)`` calls whose copy has completed
    (all of them with ``block=True``). Raises what the synchronous path would have raised, one
    call late."""
    while _deferred:
        step, h, ev, assumed = _deferred[0]
        if not block and not ev.query():
            return
        ev.synchronize()
        _deferred.pop(0)
        got = h.tolist()
        _raise_flags(got)
        wrong = [TERM_KEYS[i] for i in range(N_TERMS) if bool(got[i]) != bool(assumed[i])]
        if wrong:
            raise _abi.WcrError(f"wcr.weak_loss(host_sync=False), call #{step}: the keys {wrong} were decided on the "
                                "host but the data says otherwise (a NaN prior std drops the dims terms, "
                                "roi_heads.py:620-623); rerun that step with host_sync=True")


# ---- drop-in helpers with the reference's names (roi_heads.py:445, 575, 612, 635, 1307) ----
def safely_reduce_losses(loss: torch.Tensor) -> torch.Tensor:
    valid = (~(loss.isinf())) & (~(loss.isnan()))
    if valid.any():
        return loss[valid].mean()
    return loss.mean() * 0.0


def _img_off(num_boxes_per_image, dev):
    off = torch.zeros(len(num_boxes_per_image) + 1, dtype=torch.int32)
    off[1:] = torch.cumsum(torch.as_tensor(list(num_boxes_per_image), dtype=torch.int32), 0)
    return off.to(dev), int(off[-1])


def _as_shard(depth_maps, dev) -> ImageShard:
    if isinstance(depth_maps, ImageShard):
        return depth_maps
    n = len(depth_maps.image_sizes) if hasattr(depth_maps, "image_sizes") else len(depth_maps)
    return ImageShard(depth_maps, None, [torch.eye(3)] * n, device=dev)


class _PoseLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cube_pose, off, n_images):
        lib = _abi.load()
        dev = cube_pose.device
        R = cube_pose.detach().to(torch.float32).contiguous().view(-1, 9)
        img_pose = torch.empty((n_images,), dtype=torch.float32, device=dev)
        out3 = torch.empty((3,), dtype=torch.float32, device=dev)
        _abi.check(lib.wcr_pose_loss_fwd(R.data_ptr(), off.data_ptr(), n_images, img_pose.data_ptr(),
                                         out3.data_ptr(), _abi.stream_ptr(dev)), "wcr_pose_loss_fwd")
        ctx.save_for_backward(R, off, out3)
        ctx.n_images, ctx.shape = n_images, cube_pose.shape
        return out3[:1].clone(), out3

    @staticmethod
    def backward(ctx, grad, _grad_aux):
        lib = _abi.load()
        R, off, out3 = ctx.saved_tensors
        d_R = torch.zeros_like(R)
        g = grad.to(torch.float32).contiguous()
        _abi.check(lib.wcr_pose_loss_bwd(R.data_ptr(), off.data_ptr(), ctx.n_images, out3.data_ptr(),
                                         g.data_ptr(), d_R.data_ptr(), _abi.stream_ptr(R.device)),
                   "wcr_pose_loss_bwd")
        return d_R.view(ctx.shape), None, None


def pose_loss(cube_pose: torch.Tensor, num_boxes_per_image):
    """roi_heads.py:445-464. (n,3,3) egocentric poses -> (1,) tensor, or ``None`` when
    every image holds exactly one box. Differentiable (fused backward kernel). The
    ``None`` decision needs no device sync: it depends on ``num_boxes_per_image`` only."""
    _abi.require_cuda(cube_pose)
    nb = list(num_boxes_per_image)
    if all(b == 1 for b in nb):
        return None
    off, n = _img_off(nb, cube_pose.device)
    assert n == cube_pose.shape[0]
    loss, _ = _PoseLossFn.apply(cube_pose, off, len(nb))
    return loss


class _L1GatherFn(torch.autograd.Function):
    """|pred_z - pseudo depth| with the pseudo depth looked up by a kernel."""

    @staticmethod
    def forward(ctx, pred_z, launch):
        dev = pred_z.device
        pz = pred_z.detach().to(torch.float32).contiguous().view(-1)
        loss = torch.empty_like(pz)
        sign = torch.empty_like(pz)
        launch(pz, loss, sign)
        ctx.save_for_backward(sign)
        ctx.shape = pred_z.shape
        return loss.view(pred_z.shape)

    @staticmethod
    def backward(ctx, grad):
        lib = _abi.load()
        (sign,) = ctx.saved_tensors
        g = grad.to(torch.float32).contiguous().view(-1)
        d = torch.empty_like(sign)
        _abi.check(lib.wcr_l1_loss_bwd(sign.data_ptr(), g.data_ptr(), sign.numel(), d.data_ptr(),
                                       _abi.stream_ptr(sign.device)), "wcr_l1_loss_bwd")
        return d.view(ctx.shape), None


def pseudo_gt_z_point_loss(depth_maps, pred_xy, pred_z, num_boxes_per_image):
    """roi_heads.py:635-656: L1 between ``pred_z`` and the depth at the predicted centre
    (clamped 10 px into the image, truncated). Gradient flows to ``pred_z`` only."""
    _abi.require_cuda(pred_xy, pred_z)
    lib = _abi.load()
    dev = pred_z.device
    shard = _as_shard(depth_maps, dev)
    off, n = _img_off(num_boxes_per_image, dev)
    xy = pred_xy.detach().to(torch.float32).contiguous().view(-1, 2)
    assert n == xy.shape[0]

    def launch(pz, loss, sign):
        _abi.check(lib.wcr_pseudo_z_point_loss(shard.table.data_ptr(), off.data_ptr(), len(shard), n,
                                               xy.data_ptr(), pz.data_ptr(), loss.data_ptr(),
                                               sign.data_ptr(), _abi.stream_ptr(dev)),
                   "wcr_pseudo_z_point_loss")
    return _L1GatherFn.apply(pred_z, launch)


def pseudo_gt_z_box_loss(depth_maps, proposal_boxes, pred_z):
    """roi_heads.py:575-610 ('z_pseudo_gt_patch'): L1 between ``pred_z`` and the lower
    median of the depth patch under each predicted 2D box. ``proposal_boxes``: one (b_i,4)
    tensor per image. Reference quirk kept: within an image, boxes with zero area after
    clipping are moved behind the others before pairing with ``pred_z`` by position."""
    _abi.require_cuda(pred_z)
    lib = _abi.load()
    dev = pred_z.device
    shard = _as_shard(depth_maps, dev)
    boxes = [torch.as_tensor(b.tensor if hasattr(b, "tensor") else b) for b in proposal_boxes]
    off, n = _img_off([b.shape[0] for b in boxes], dev)
    bx = torch.cat([b.detach().to(device=dev, dtype=torch.float32).view(-1, 4) for b in boxes]).contiguous()
    gt = torch.empty((max(n, 1),), dtype=torch.float32, device=dev)
    inside = torch.empty((max(n, 1),), dtype=torch.int32, device=dev)

    sel_ws = torch.empty(int(lib.wcr_select_ws_bytes(n)), dtype=torch.uint8, device=dev)

    def launch(pz, loss, sign):
        assert pz.numel() == n
        _abi.check(lib.wcr_pseudo_z_box_loss(shard.table.data_ptr(), off.data_ptr(), len(shard), n,
                                             bx.data_ptr(), pz.data_ptr(), loss.data_ptr(),
                                             sign.data_ptr(), gt.data_ptr(), inside.data_ptr(),
                                             sel_ws.data_ptr(), _abi.stream_ptr(dev)), "wcr_pseudo_z_box_loss")
    return _L1GatherFn.apply(pred_z, launch)


class _DimLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dims, mean, std):
        lib = _abi.load()
        dev = dims.device
        d = dims.detach().to(torch.float32).contiguous().view(-1, 3)
        n = d.shape[0]
        out = torch.empty((3, max(n, 1)), dtype=torch.float32, device=dev)
        flag = torch.zeros((1,), dtype=torch.int32, device=dev)
        _abi.check(lib.wcr_dim_loss_fwd(mean.data_ptr(), std.data_ptr(), d.data_ptr(), n, out.data_ptr(),
                                        flag.data_ptr(), _abi.stream_ptr(dev)), "wcr_dim_loss_fwd")
        ctx.save_for_backward(d, mean, std)
        ctx.shape = dims.shape
        return out[:, :n], flag

    @staticmethod
    def backward(ctx, grad, _gflag):
        lib = _abi.load()
        d, mean, std = ctx.saved_tensors
        n = d.shape[0]
        g = grad.to(torch.float32).contiguous()
        dd = torch.zeros_like(d)
        _abi.check(lib.wcr_dim_loss_bwd(mean.data_ptr(), std.data_ptr(), d.data_ptr(), g.data_ptr(), n,
                                        dd.data_ptr(), _abi.stream_ptr(d.device)), "wcr_dim_loss_bwd")
        return dd.view(ctx.shape), None, None


def dim_loss(priors, dimensions, *, host_sync: bool = True):
    """roi_heads.py:612-633: ``priors = (prior_mean, prior_std)`` (n,3) each ->
    three (n,) hinge losses ``max(|d - mean|/std - 1, 0)``, or ``(None, None, None)``
    when a prior std is NaN (one 4-byte host read, the reference's ``mask.all()``).
    ``host_sync=False`` skips that read: the priors are taken to be finite (category statistics
    are; a NaN std then shows up as a NaN loss instead of a dropped key)."""
    prior_mean, prior_std = priors
    _abi.require_cuda(dimensions, prior_mean, prior_std)
    dev = dimensions.device
    mean, std = _c(prior_mean, dev).view(-1, 3), _c(prior_std, dev).view(-1, 3)
    out, flag = _DimLossFn.apply(dimensions, mean, std)
    if host_sync and int(flag.item()) != 0:
        return None, None, None
    return out[0], out[1], out[2]


def _stack_masks(gt_mask, dev):
    m = gt_mask if isinstance(gt_mask, torch.Tensor) else torch.stack(
        [torch.as_tensor(x).reshape(x.shape[-2], x.shape[-1]) for x in gt_mask])
    m = m.reshape(-1, m.shape[-2], m.shape[-1]).to(device=dev)
    if m.dtype == torch.bool:
        return m.contiguous().view(torch.uint8)
    return (m != 0).to(torch.uint8).contiguous() if m.dtype != torch.uint8 else m.contiguous()


class _SegmentLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, corners, masks, idx):
        lib = _abi.load()
        c = corners.detach().to(torch.float32).contiguous().view(-1, 16)
        n = c.shape[0]
        out = torch.empty((n,), dtype=torch.float32, device=c.device)
        _abi.check(lib.wcr_segment_loss_fwd(masks.data_ptr(), masks.shape[1], masks.shape[2], idx.data_ptr(),
                                            c.data_ptr(), n, out.data_ptr(), _abi.stream_ptr(c.device)),
                   "wcr_segment_loss_fwd")
        ctx.save_for_backward(c, masks, idx)
        ctx.shape = corners.shape
        return out

    @staticmethod
    def backward(ctx, g):
        lib = _abi.load()
        c, masks, idx = ctx.saved_tensors
        n = c.shape[0]
        d = torch.empty_like(c)
        _abi.check(lib.wcr_segment_loss_bwd(masks.data_ptr(), masks.shape[1], masks.shape[2], idx.data_ptr(),
                                            c.data_ptr(), n, d.data_ptr(), _abi.stream_ptr(c.device)),
                   "wcr_segment_loss_bwd")
        return (d * g.to(torch.float32).view(n, 1)).view(ctx.shape), None, None


def segment_loss(gt_mask, bube_corners, at_which_mask_idx, loss: str = "focal"):
    """roi_heads.py:420-443 (``loss='focal'``, the only mode the reference calls, :1017):
    ``gt_mask`` list of (1,H,W) masks (or one (M,H,W) tensor), ``bube_corners`` (n,8,2) clamped
    projected corners, ``at_which_mask_idx`` (n,) -> (n,) losses, differentiable in the corners."""
    if loss != "focal":
        raise _abi.WcrError("wcr.segment_loss: only loss='focal' is on the device path")
    _abi.require_cuda(bube_corners)
    dev = bube_corners.device
    masks = _stack_masks(gt_mask, dev)
    idx = torch.as_tensor(list(at_which_mask_idx) if not isinstance(at_which_mask_idx, torch.Tensor)
                          else at_which_mask_idx, dtype=torch.int32).to(dev)
    return _SegmentLossFn.apply(bube_corners, masks, idx)


# ---- stand-alone tier-2 helpers with the reference's method names (roi_heads.py:530, 658, 686) ----
def _boxes_tensor(b, dev):
    t = b.tensor if hasattr(b, "tensor") else torch.as_tensor(b)
    return t.detach().to(device=dev, dtype=torch.float32).reshape(-1, 4).contiguous()


def z_loss(gt_boxes, cubes, Ks, im_sizes, proj_boxes):
    """roi_heads.py:530-573: per predicted cube, 50 offsets of 0.1 m along z, arg-min of the
    mismatch between the GT box area and the projected area, score = |z - z_best| / 2 (2.5 when
    the projected centre is outside the GT box grown by 50 px -- with the reference's chained
    comparison kept). ``cubes``: a ``Cubes`` of shape (n, 1, 15); ``Ks``: n (3,3) intrinsics;
    ``im_sizes``: n pairs handed to ``cubes_to_box``. One launch instead of n x 50 ``cubes_to_box``
    calls; value only, like the reference's graph (the selected offset is a constant)."""
    lib = _abi.load()
    ct = cubes.tensor if hasattr(cubes, "tensor") else cubes
    _abi.require_cuda(ct)
    dev = ct.device
    flat = ct.detach().to(torch.float32).reshape(-1, 15).contiguous()
    n = flat.shape[0]
    Kt = torch.stack([torch.as_tensor(k, dtype=torch.float32).reshape(3, 3) for k in Ks]) \
        if not isinstance(Ks, torch.Tensor) else Ks.to(torch.float32).reshape(-1, 3, 3)
    K4 = torch.stack((Kt[:, 0, 0], Kt[:, 1, 1], Kt[:, 0, 2], Kt[:, 1, 2]), dim=1).to(dev).contiguous()
    clamp = torch.as_tensor([[float(s[0]), float(s[1])] for s in im_sizes], dtype=torch.float32).to(dev) \
        if not isinstance(im_sizes, torch.Tensor) else im_sizes.to(device=dev, dtype=torch.float32).reshape(-1, 2)
    gt, pj = _boxes_tensor(gt_boxes, dev), _boxes_tensor(proj_boxes, dev)
    assert K4.shape[0] == n and clamp.shape[0] == n and gt.shape[0] == n and pj.shape[0] == n
    out = torch.empty((n,), dtype=torch.float32, device=dev)
    _abi.check(lib.wcr_z_loss(gt.data_ptr(), flat.data_ptr(), K4.data_ptr(), clamp.contiguous().data_ptr(),
                              pj.data_ptr(), n, out.data_ptr(), _abi.stream_ptr(dev)), "wcr_z_loss")
    return out


def masked_depth_range(gt_mask, at_which_mask_idx, depth_maps, gt_boxes, num_instances):
    """The GT side of ``depth_range_loss``: (n,) q90 - q10 of the depth under each object's mask
    (GT box patch where the mask is empty), depth maps resized to the mask shape first where the
    shapes differ (roi_heads.py:668-676)."""
    lib = _abi.load()
    gt_src = gt_boxes.tensor if hasattr(gt_boxes, "tensor") else torch.as_tensor(gt_boxes)
    _abi.require_cuda(gt_src)
    dev = gt_src.device
    gt = _boxes_tensor(gt_src, dev)
    masks = _stack_masks(gt_mask, dev)
    shard = _as_shard(depth_maps, dev)
    rs = resized_depth_shard(shard, (int(masks.shape[1]), int(masks.shape[2])))
    use = rs if rs is not None else shard
    off, n = _img_off(num_instances, dev)
    assert n == gt.shape[0]
    idx = torch.as_tensor(list(at_which_mask_idx) if not isinstance(at_which_mask_idx, torch.Tensor)
                          else at_which_mask_idx, dtype=torch.int32).to(dev)
    out = torch.empty((max(n, 1),), dtype=torch.float32, device=dev)
    stats = torch.zeros((N_STATS,), dtype=torch.float32, device=dev)
    ws = torch.empty(int(lib.wcr_select_ws_bytes(max(n, 1))), dtype=torch.uint8, device=dev)
    _abi.check(lib.wcr_masked_depth_range(use.table.data_ptr(), off.data_ptr(), len(use), n, masks.data_ptr(),
                                          int(masks.shape[1]), int(masks.shape[2]), idx.data_ptr(),
                                          gt.data_ptr(), out.data_ptr(), stats.data_ptr(), ws.data_ptr(),
                                          _abi.stream_ptr(dev)), "wcr_masked_depth_range")
    return out[:n]


def depth_range_loss(gt_mask, at_which_mask_idx, depth_maps, cubes, gt_boxes, num_instances):
    """roi_heads.py:658-684: |(q90 - q10 of the depth under the object's mask) - (depth extent of
    the cube's 8 corners)| per box. Differentiable in ``cubes.tensor`` through
    ``Cubes.get_all_corners`` (first arg-max / arg-min corner, like torch.max / torch.min)."""
    corner_depths = cubes.get_all_corners()[:, 0, :, 2]
    pred_depth = torch.max(corner_depths, dim=1)[0] - torch.min(corner_depths, dim=1)[0]
    gt_depths = masked_depth_range(gt_mask, at_which_mask_idx, depth_maps, gt_boxes, num_instances)
    return torch.nn.functional.smooth_l1_loss(gt_depths, pred_depth, reduction="none", beta=0.0)


def normal_to_rotation(normal):
    """roi_heads.py:686-696: (n, 3) normals -> (n, 3, 3) with rows [t0, t1, normal] (quirks of the
    reference kept: whole-matrix Frobenius norms, bool-compared fallback test; see wcr.h)."""
    _abi.require_cuda(normal)
    lib = _abi.load()
    nrm = normal.detach().to(torch.float32).reshape(-1, 3).contiguous()
    out = torch.empty((nrm.shape[0], 3, 3), dtype=torch.float32, device=nrm.device)
    _abi.check(lib.wcr_normal_to_rotation(nrm.data_ptr(), nrm.shape[0], out.data_ptr(),
                                          _abi.stream_ptr(nrm.device)), "wcr_normal_to_rotation")
    return out


# ---- the ROI head drop-in -------------------------------------------------
_FUSED_LOSSES = frozenset(_FLAG)


def make_roi_heads_b200(base):
    """Subclass ``base`` -- the reference's ``ROIHeads3DScore`` (roi_heads.py:52) -- so that the
    training loss block of ``_forward_cube`` (roi_heads.py:786-1144: class gather, decode, the
    per-box Python loops, every weak-loss term, uncertainty weighting, safe reductions) is ONE
    call into the fused kernels (``wcr.glue.cube_losses_from_head``). Everything else -- the box
    branch, proposal sampling, the pooler, the cube head, inference -- stays the reference's code.

    Configurations outside the fused path (z clusters, no dimension priors, egocentric head,
    a loss name the kernels do not know) raise ``WcrError`` in training: there is no silent
    fallback to the torch path. Attributes a caller may set on the instance:
      ``wcr_replay_rng``  a ``random.Random`` / the ``random`` module: the RANSAC triples are drawn
                          from it exactly as plane.py:95 does (parity mode); None: in-kernel Philox
      ``wcr_seed``        seed of the in-kernel generator (default 0)
      ``wcr_host_sync``   False: no host synchronisation in the loss block (see ``weak_loss``)"""
    import sys

    class ROIHeads3DScoreB200(base):
        wcr_replay_rng = None
        wcr_seed = 0
        wcr_host_sync = True

        def _wcr_check_config(self):
            bad = []
            if getattr(self, "cluster_bins", 1) > 1 or self.z_type == "clusters":
                bad.append("z clusters")
            if not self.dims_priors_enabled:
                bad.append("DIMS_PRIORS_ENABLED = False")
            if not self.allocentric_pose:
                bad.append("ALLOCENTRIC_POSE = False")
            unknown = [f for f in self.loss_functions if f not in _FUSED_LOSSES]
            if unknown:
                bad.append(f"loss functions {unknown}")
            if bad:
                raise _abi.WcrError("ROIHeads3DScoreB200: not on the fused path: " + ", ".join(bad))

        def _forward_cube(self, features, instances, Ks, im_current_dims, im_scales_ratio,
                          masks_all_images, first_occurrence_indices, ground_maps, depth_maps):
            if not self.training:
                return super()._forward_cube(features, instances, Ks, im_current_dims, im_scales_ratio,
                                             masks_all_images, first_occurrence_indices, ground_maps, depth_maps)
            self._wcr_check_config()
            from . import glue, ground as wground
            from .spaces import Cubes as WCubes
            ref_mod = sys.modules[base.__module__]
            feats = [features[f] for f in self.in_features]
            self.normalize_factor = max(sum(i.gt_classes.numel() for i in instances), 1.0)
            proposals, _ = ref_mod.select_foreground_proposals(instances, self.num_classes)
            proposal_boxes = [x.proposal_boxes for x in proposals]
            pred_boxes = [x.pred_boxes for x in proposals]
            cube_features = self.cube_pooler(feats, self.scale_proposals(proposal_boxes)).flatten(1)
            n = cube_features.shape[0]
            if n == 0:
                return instances, {}
            nb = [len(p) for p in proposals]
            heads = self.cube_head(cube_features)
            dev = heads[0].device
            if dev.type != "cuda":
                raise _abi.WcrError("ROIHeads3DScoreB200: the cube head must run on a CUDA device")
            box_classes = torch.cat([p.gt_classes for p in proposals]).to(dev)
            at_which = None
            if masks_all_images is not None:
                ids = torch.cat([p.gt_boxes3D[:, 0] for p in proposals]).tolist()
                at_which = [first_occurrence_indices[e] for e in ids]
            src = torch.cat([b.tensor for b in proposal_boxes]).to(dev)
            gt = torch.cat([p.gt_boxes.tensor for p in proposals]).to(dev)
            # normal_vector_from_maps zips Ks_scaled_per_box (one K per BOX) against IMAGES
            # (roi_heads.py:472, :990): image i is unprojected with the K of box i. Kept.
            B = len(nb)
            im_of_box = [i for i, c in enumerate(nb) for _ in range(c)]
            Kq = []
            for i in range(B):
                j = im_of_box[i] if i < n else i
                Kb = torch.as_tensor(Ks[j], dtype=torch.float32).detach().cpu().clone() / float(im_scales_ratio[j])
                Kb[-1, -1] = 1
                Kq.append(Kb)
            need_normals = any(f in self.loss_functions for f in ("pose_ground", "pose_ground2"))
            if need_normals:
                normals = wground.normal_vector_from_maps(
                    [g.to(dev) for g in ground_maps], [d.to(dev) for d in depth_maps], Kq,
                    replay_rng=self.wcr_replay_rng, seed=self.wcr_seed)
            else:
                normals = torch.tensor([[0.0, 1.0, 0.0]], device=dev).repeat(B, 1)
            conf = torch.tensor([0.1 if tuple(s) == (1, 1) else 1.0 for s in ground_maps.image_sizes], device=dev)
            cfg = WeakLossConfig(
                loss_functions=tuple(self.loss_functions), z_type=self.z_type, dims_priors_func=self.dims_priors_func,
                use_confidence=float(self.use_confidence), loss_w_3d=self.loss_w_3d, loss_w_iou=self.loss_w_iou,
                loss_w_pose=self.loss_w_pose, loss_w_normal_vec=self.loss_w_normal_vec, loss_w_z=self.loss_w_z,
                loss_w_dims=self.loss_w_dims, loss_w_depth=self.loss_w_depth, loss_w_seg=self.loss_w_seg)
            deltas_c, z_c, dims_c, pose_c, uncert_c = heads
            losses, per_box = glue.cube_losses_from_head(
                cfg, deltas_c, z_c, dims_c, pose_c, uncert_c if self.use_confidence else None, box_classes,
                proposal_boxes=src, gt_boxes=gt, Ks=Ks, im_scales_ratio=im_scales_ratio,
                im_current_dims=im_current_dims, num_boxes_per_image=nb,
                priors_dims_per_cat=self.priors_dims_per_cat, depth_maps=[d.to(dev) for d in depth_maps],
                normals=normals, ground_conf=conf, virtual_focal=self.virtual_focal,
                virtual_depth=self.virtual_depth, gt_masks=masks_all_images, at_which_mask_idx=at_which,
                return_per_box=True, host_sync=self.wcr_host_sync)
            pred = self._wcr_pred_instances(ref_mod, heads, box_classes, src, per_box, Ks, im_current_dims,
                                            im_scales_ratio, nb, pred_boxes, WCubes)
            return pred, losses

        @torch.no_grad()
        def _wcr_pred_instances(self, ref_mod, heads, box_classes, src, per_box, Ks, im_dims, ratios, nb,
                                pred_boxes, WCubes):
            """The per-image ``Instances`` the reference returns from the training branch
            (roi_heads.py:1153-1195): a handful of vectorised tensor ops, no gradient."""
            dev = src.device
            n = src.shape[0]
            ar = torch.arange(n, device=dev)
            deltas = heads[0][ar, box_classes]
            z = heads[1][ar, box_classes].reshape(n)
            dims_n = heads[2][ar, box_classes]
            from . import glue
            K4, v2r, _ = glue.box_constants(Ks, ratios, im_dims, nb, virtual_focal=self.virtual_focal,
                                            virtual_depth=self.virtual_depth, device=dev)
            w, h = src[:, 2] - src[:, 0], src[:, 3] - src[:, 1]
            cx, cy = src[:, 0] + 0.5 * w + w * deltas[:, 0], src[:, 1] + 0.5 * h + h * deltas[:, 1]
            pri = self.priors_dims_per_cat.detach().to(dev)[box_classes]
            if self.dims_priors_func == "sigmoid":
                lo, hi = (pri[:, 0] - 3 * pri[:, 1]).clip(0.0), pri[:, 0] + 3 * pri[:, 1]
                dims = lo + (hi - lo) * torch.sigmoid(dims_n)
            else:
                dims = torch.exp(dims_n.clip(max=5)) * pri[:, 0]
            if self.z_type == "sigmoid":
                z = torch.sigmoid(z) * 100
            elif self.z_type == "log":
                z = torch.exp(z)
            z = z * v2r
            x3d, y3d = z * (cx - K4[:, 2]) / K4[:, 0], z * (cy - K4[:, 3]) / K4[:, 1]
            ratio_pb = torch.cat([torch.full((c,), float(ratios[i])) for i, c in enumerate(nb)]).to(dev)
            cube_3d = torch.cat((torch.stack((x3d, y3d, z)).T, dims, torch.stack((cx, cy), 1) * ratio_pb[:, None]), 1)
            if self.use_confidence:
                cube_3d = torch.cat((cube_3d, torch.exp(-heads[4][ar, box_classes])[:, None]), 1)
            R = per_box["R"]
            verts = WCubes(torch.cat((cube_3d[:, :6], R.reshape(n, 9)), 1).unsqueeze(1)).get_all_corners()[:, 0]
            out = []
            for i, (c3, Ri, vi, cls_i) in enumerate(zip(cube_3d.split(nb), R.split(nb), verts.split(nb),
                                                        box_classes.split(nb))):
                inst = ref_mod.Instances(im_dims[i])
                inst.scores = c3[:, -1]
                inst.pred_classes = cls_i
                inst.pred_boxes = pred_boxes[i]
                inst.pred_bbox3D = vi
                inst.pred_center_cam, inst.pred_center_2D, inst.pred_dimensions = c3[:, :3], c3[:, 6:8], c3[:, 3:6]
                inst.pred_pose = Ri
                out.append(inst)
            return out

    ROIHeads3DScoreB200.__qualname__ = "ROIHeads3DScoreB200"
    return ROIHeads3DScoreB200


def register_roi_heads_b200():
    """With detectron2 and the reference importable: build ``ROIHeads3DScoreB200`` on the
    reference's ``ROIHeads3DScore`` and register it in ``ROI_HEADS_REGISTRY`` (select it with
    ``MODEL.ROI_HEADS.NAME: ROIHeads3DScoreB200``, configs/Omni_combined.yaml:25-36). Returns the
    class; raises ImportError where detectron2 is absent (then ``make_roi_heads_b200(base)``
    builds the class on whatever ``ROIHeads3DScore`` the caller imported)."""
    from detectron2.modeling.roi_heads import ROI_HEADS_REGISTRY          # noqa: F401 (ImportError without it)
    from cubercnn.modeling.roi_heads.roi_heads import ROIHeads3DScore
    cls = make_roi_heads_b200(ROIHeads3DScore)
    try:
        ROI_HEADS_REGISTRY.register(cls)
    except AssertionError:
        pass                                                               # already registered
    return cls
