"""ctypes binding of include/wcr.h (libwcr_b200.so). Plumbing only.

The product path has NO CPU fallback: if the shared library is missing or the
device is not sm_100, calls raise. torch is used for device memory and streams;
every pointer handed to C is ``tensor.data_ptr()``.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libwcr_b200.so")

c_void_p, c_int, c_float, c_i64, c_u64 = (ctypes.c_void_p, ctypes.c_int, ctypes.c_float,
                                          ctypes.c_int64, ctypes.c_uint64)
c_u32 = ctypes.c_uint32

# name -> (restype, argtypes); must list every symbol include/wcr.h declares.
SIGNATURES = {
    "wcr_abi_version": (c_int, []),
    "wcr_last_error": (ctypes.c_char_p, []),
    "wcr_device_check": (c_int, []),
    "wcr_launch_count": (c_i64, []),
    "wcr_reset_launch_count": (None, []),
    "wcr_unproject_ground": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_unproject_cloud": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_i64, c_void_p]),
    "wcr_ransac_ground": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_float, c_u64,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p]),
    "wcr_ransac_ground_static": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_float,
                                         c_u64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_void_p]),
    "wcr_probe_ffma": (c_i64, [c_void_p, c_i64, c_int, c_void_p]),
    "wcr_philox_block": (c_int, [c_u64, c_u32, c_u32, c_u32, c_u32, c_void_p, c_void_p]),
    "wcr_box_center_depth": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_sample_cubes": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_int, c_void_p, c_u64, c_void_p, c_void_p]),
    "wcr_score_workspace_bytes": (c_i64, [c_int, c_int, c_int]),
    "wcr_score_bounds_offsets": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "wcr_score_topk": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_u64, c_void_p, c_void_p, c_void_p, c_int,
                               c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                               c_void_p, c_void_p, c_void_p]),
    "wcr_score_set_reserved_ctas": (c_int, [c_int]),
    "wcr_score_topk_after": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_u64, c_void_p, c_void_p, c_void_p, c_int,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p]),
    "wcr_pack_box_masks": (c_int, [c_void_p, c_i64, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                                   c_void_p, c_void_p]),
    "wcr_h2d_box_mask_rois": (c_int, [c_void_p, c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "wcr_h2d_box_mask_rows": (c_int, [c_void_p, c_void_p, c_i64, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_cube_corners": (c_int, [c_void_p, c_i64, c_void_p, c_void_p]),
    "wcr_bube_corners": (c_int, [c_void_p, c_i64, c_float, c_float, c_float, c_float, c_float,
                                 c_float, c_void_p, c_void_p, c_void_p]),
    "wcr_iou_2d": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "wcr_convex_hull": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_i64, c_void_p]),
    "wcr_mask_iou": (c_int, [c_void_p, c_int, c_void_p, c_i64, c_int, c_void_p, c_void_p, c_void_p]),
    "wcr_sample_normal_in_range": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_float,
                                           c_void_p, c_u64, c_int, c_void_p, c_void_p]),
    "wcr_orthobasis_from_normal": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_vectorized_linspace": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "wcr_weak_loss_fwd": (c_int, [c_void_p, c_void_p, c_void_p]),
    "wcr_weak_loss_bwd": (c_int, [c_void_p, c_void_p, c_void_p]),
    "wcr_pose_loss_fwd": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "wcr_pose_loss_bwd": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "wcr_pseudo_z_point_loss": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p]),
    "wcr_pseudo_z_box_loss": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "wcr_select_ws_bytes": (c_i64, [c_int]),
    "wcr_l1_loss_bwd": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_dim_loss_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "wcr_dim_loss_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_z_loss_steps": (None, [c_void_p]),
    "wcr_segment_loss_fwd": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_segment_loss_bwd": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_box3d_overlap": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "wcr_box3d_volume": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_rotation_6d_fwd": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_rotation_6d_bwd": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_class_gather_fwd": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "wcr_class_gather_bwd": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "wcr_box_constants": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_int,
                                  c_void_p, c_void_p, c_void_p, c_void_p]),
    "wcr_pil_nearest_table": (None, [c_int, c_int, c_void_p]),
    "wcr_resize_nearest": (c_int, [c_void_p, c_int, c_int, c_int, c_i64, c_void_p, c_void_p, c_void_p,
                                   c_int, c_int, c_i64, c_void_p]),
    "wcr_sizeof": (c_i64, [c_int]),
    "wcr_resize_bilinear_ac": (c_int, [c_void_p, c_int, c_int, c_i64, c_void_p, c_int, c_int, c_i64, c_void_p]),
    "wcr_cube_corners_bwd": (c_int, [c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "wcr_bube_corners_bwd": (c_int, [c_void_p, c_i64, c_float, c_float, c_float, c_float, c_float, c_float,
                                     c_void_p, c_void_p, c_void_p, c_void_p]),
    "wcr_custom_mapping": (c_int, [c_void_p, c_i64, ctypes.c_double, c_void_p, c_void_p]),
    "wcr_randn_orthobasis": (c_int, [c_void_p, c_i64, c_u64, c_void_p, c_void_p]),
    "wcr_gt_in_norm_range": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p]),
    "wcr_is_gt_included": (c_int, [c_void_p, c_void_p, c_void_p, c_i64, c_void_p, c_void_p, c_void_p]),
    "wcr_z_loss": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_normal_to_rotation": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "wcr_masked_depth_range": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p,
                                       c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
}
_lib = None


class WcrError(RuntimeError):
    pass


def load():
    """Load libwcr_b200.so; raise loudly if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise WcrError(
            f"{LIB_PATH} is missing: build it with `python weak-cube-r-cnn_b200/build.py` "
            "(there is no CPU or eager fallback for this path)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().wcr_last_error()
        raise WcrError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t) -> int | None:
    if t is None:
        return None
    return t.data_ptr()


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise WcrError("wcr: CUDA tensors required (there is no CPU fallback on this path)")


def launch_count() -> int:
    return int(load().wcr_launch_count())


def reset_launch_count():
    load().wcr_reset_launch_count()


class SampleParams(ctypes.Structure):
    """include/wcr.h WcrSampleParams (host struct, passed by pointer)."""
    _fields_ = [("sigma_xy", c_float), ("sigma_z", c_float), ("z_lo", c_float), ("z_hi", c_float),
                ("dim_lo", c_float), ("dim_hi_sigmas", c_float), ("yaw_lo", c_float),
                ("yaw_hi", c_float), ("max_tries", ctypes.c_int32), ("reserved0", ctypes.c_int32),
                ("reserved1", ctypes.c_int32), ("reserved2", ctypes.c_int32)]


class ScoreParams(ctypes.Structure):
    """include/wcr.h WcrScoreParams."""
    _fields_ = [("e_iou", ctypes.c_int32), ("e_pts", ctypes.c_int32), ("e_seg", ctypes.c_int32),
                ("point_stride", ctypes.c_int32), ("seg_mode", ctypes.c_int32),
                ("map_beta", c_float), ("seg_eval", ctypes.c_int32),
                ("pts_eval", ctypes.c_int32)]


class LossParams(ctypes.Structure):
    """include/wcr.h WcrLossParams."""
    _fields_ = [("loss_flags", ctypes.c_int32), ("z_type", ctypes.c_int32),
                ("dims_func", ctypes.c_int32), ("use_confidence", c_float), ("w_3d", c_float),
                ("w_iou", c_float), ("w_pose", c_float), ("w_normal", c_float), ("w_z", c_float),
                ("w_dims", c_float), ("w_depth", c_float), ("w_seg", c_float)]


_LOSS_IO_PTRS = ["deltas", "z_raw", "dims_raw", "pose", "uncert", "src_boxes", "K4", "v2r",
                 "prior_mean", "prior_std", "gt_boxes", "clamp", "img_off", "images", "normals",
                 "ground_conf", "losses", "present", "terms", "R_ego", "proj", "stats", "img_pose",
                 "grad_losses", "d_deltas", "d_z_raw", "d_dims_raw", "d_pose", "d_uncert", "d_R_pose",
                 "z_dec", "z_gt", "z_inside", "z_sign"]


_LOSS_IO_PTRS2 = ["mask_idx", "depth_gt", "gt_rot", "select_ws", "seg_dcorner",   # after the mask shape (tier-2 terms)
                  "depth_images"]


class LossIO(ctypes.Structure):
    """include/wcr.h WcrLossIO: two sizes, device pointers, then the tier-2 block."""
    _fields_ = [("n", ctypes.c_int32), ("n_images", ctypes.c_int32)] + \
               [(name, c_void_p) for name in _LOSS_IO_PTRS] + \
               [("masks", c_void_p), ("mask_h", ctypes.c_int32), ("mask_w", ctypes.c_int32)] + \
               [(name, c_void_p) for name in _LOSS_IO_PTRS2]
