"""Cube proposal generation and fused scoring (north_star steps 2 and 3).

NEW entry points: the reference snapshot ships only the building blocks of its
proposal pipeline; ``propose`` / ``score_topk`` are this project's
composition of them (A2.7 / A3.12) -- ``oracle/compose.py`` is the same composition
written with the reference's own functions.

Host code here is plumbing: it builds the device box table and hands pointers to
libwcr_b200.so (include/wcr.h). No CPU fallback.
"""
from __future__ import annotations

import ctypes
import math
from dataclasses import dataclass

import numpy as np
import torch

from . import _abi
from .images import ImageShard

BOX_DTYPE = np.dtype([("x1", np.float32), ("y1", np.float32), ("x2", np.float32), ("y2", np.float32),
                      ("image", np.int32), ("cls", np.int32),
                      ("clamp0", np.float32), ("clamp1", np.float32)])
assert BOX_DTYPE.itemsize == 32

CUBE_FLOATS, RECORD_FLOATS, MAX_TOPK = 15, 20, 32


@dataclass
class ProposalConfig:
    """Free parameters of the composite (defaults are this project's spec)."""
    sigma_xy: float = 0.25       # std of the 2D centre / box width (height)
    sigma_z: float = 0.15        # std of z / pseudo depth z0
    z_lo: float = 0.5            # z in [z0*z_lo, z0*z_hi]
    z_hi: float = 1.5
    dim_lo: float = 0.05         # dims in [dim_lo, mean + dim_hi_sigmas*std]
    dim_hi_sigmas: float = 3.0
    yaw_lo: float = 0.0          # yaw ~ U[0, pi)
    yaw_hi: float = math.pi
    max_tries: int = 32

    def c_struct(self) -> _abi.SampleParams:
        return _abi.SampleParams(self.sigma_xy, self.sigma_z, self.z_lo, self.z_hi, self.dim_lo,
                                 self.dim_hi_sigmas, self.yaw_lo, self.yaw_hi, self.max_tries, 0, 0, 0)


@dataclass
class ScoreConfig:
    e_iou: int = 1
    e_pts: int = 1
    e_seg: int = 0
    point_stride: int = 1
    seg_mode: int = 0
    k: int = 16
    map_beta: float = 0.0        # > 0: custom_mapping(x, beta) (utils.py:207-223) shapes every term; 0 = off
    seg_eval: int = 0            # mask term: 0 = library decides, 1 = every hull, 2 = on demand (same results)
    pts_eval: int = 0            # point term: 0 = library decides, 1 = every proposal, 2 = on demand (same results)

    def c_struct(self) -> _abi.ScoreParams:
        return _abi.ScoreParams(self.e_iou, self.e_pts, self.e_seg, self.point_stride,
                                self.seg_mode, float(self.map_beta), int(self.seg_eval), int(self.pts_eval))


class BoxTable:
    """Device table of ``WcrBox`` records for the GT boxes of a shard.

    ``boxes_per_image``: list of (b_i, 4) tensors; ``classes_per_image``: list of
    (b_i,) integer tensors. ``clamp`` is the tuple handed to
    ``Cubes.get_bube_corners`` per image: default ``[H, W]`` as the ROI head does
    (roi_heads.py:801-806, 950 -- the reference's own call-site convention)."""

    def __init__(self, shard: ImageShard, boxes_per_image, classes_per_image, clamp=None):
        recs = []
        sizes = shard.sizes()
        for i, (bx, cl) in enumerate(zip(boxes_per_image, classes_per_image)):
            bx = torch.as_tensor(bx, dtype=torch.float32).cpu().numpy().reshape(-1, 4)
            cl = torch.as_tensor(cl).cpu().numpy().reshape(-1)
            c0, c1 = clamp[i] if clamp is not None else sizes[i]
            r = np.zeros(len(bx), dtype=BOX_DTYPE)
            r["x1"], r["y1"], r["x2"], r["y2"] = bx[:, 0], bx[:, 1], bx[:, 2], bx[:, 3]
            r["image"], r["cls"] = i, cl
            r["clamp0"], r["clamp1"] = c0, c1
            recs.append(r)
        self.host = np.concatenate(recs) if recs else np.zeros(0, dtype=BOX_DTYPE)
        self.n_boxes = len(self.host)
        self.device = shard.device
        self.table = torch.from_numpy(self.host.view(np.uint8).copy()).to(shard.device)
        self.num_boxes_per_image = [len(r) for r in recs]

    @classmethod
    def from_host(cls, host: np.ndarray, device, table: torch.Tensor | None = None,
                  num_boxes_per_image=None):
        """Table from an already built ``BOX_DTYPE`` array. ``table``: an existing device byte
        buffer (>= 32 * len(host)) the caller uploads into (pipeline slot); None => uploaded here."""
        self = cls.__new__(cls)
        self.host = host
        self.n_boxes = len(host)
        self.device = torch.device(device)
        self.table = table if table is not None else \
            torch.from_numpy(host.view(np.uint8).copy()).to(self.device, non_blocking=True)
        self.num_boxes_per_image = num_boxes_per_image
        return self

    def __len__(self):
        return self.n_boxes


def box_center_depth(shard: ImageShard, boxes: BoxTable) -> torch.Tensor:
    """Pseudo depth z0 per box (roi_heads.py:645-652 on the GT box centre)."""
    lib = _abi.load()
    z0 = torch.empty((len(boxes),), dtype=torch.float32, device=shard.device)
    _abi.check(lib.wcr_box_center_depth(shard.table.data_ptr(), boxes.table.data_ptr(), len(boxes),
                                        z0.data_ptr(), _abi.stream_ptr(shard.device)),
               "wcr_box_center_depth")
    return z0


def _f32(t, device):
    return t.to(device=device, dtype=torch.float32).contiguous()


def propose(shard: ImageShard, boxes: BoxTable, priors: torch.Tensor, normals: torch.Tensor,
            N: int = 1000, cfg: ProposalConfig | None = None, *, z0=None, replay=None,
            seed: int = 0) -> torch.Tensor:
    """N cubes per box -> (n_boxes, N, 15) f32, the ``Cubes.tensor`` layout
    (spaces.py:130-150). ``replay`` (n_boxes, 7, N): the (u v z w h l yaw) samples
    drawn by the reference's ``sample_normal_in_range``; None => in-kernel Philox."""
    lib = _abi.load()
    cfg = cfg or ProposalConfig()
    dev = shard.device
    priors, normals = _f32(priors, dev), _f32(normals, dev)
    _abi.require_cuda(priors, normals)
    if z0 is None:
        z0 = box_center_depth(shard, boxes)
    if replay is not None:
        replay = _f32(replay, dev)
        assert tuple(replay.shape) == (len(boxes), 7, N)
    out = torch.empty((len(boxes), N, CUBE_FLOATS), dtype=torch.float32, device=dev)
    sp = cfg.c_struct()
    _abi.check(lib.wcr_sample_cubes(shard.table.data_ptr(), boxes.table.data_ptr(), len(boxes),
                                    priors.data_ptr(), normals.data_ptr(), z0.data_ptr(),
                                    ctypes.byref(sp), N, _abi.ptr(replay), int(seed),
                                    out.data_ptr(), _abi.stream_ptr(dev)),
               "wcr_sample_cubes")
    return out


def box_rois(boxes: "BoxTable", shard: ImageShard) -> np.ndarray:
    """(n_boxes, 4) int32 [u0, v0, u1, v1): Boxes.clip((H, W)) then .long(), the way the
    reference slices depth patches (roi_heads.py:590-597). Same arithmetic as
    csrc/score.cu box_roi()."""
    h = boxes.host
    sizes = np.asarray(shard.sizes(), dtype=np.float32).reshape(-1, 2)
    Hs, Ws = sizes[h["image"], 0], sizes[h["image"], 1]
    u0 = np.minimum(np.maximum(h["x1"], 0), Ws).astype(np.int32)
    u1 = np.minimum(np.maximum(h["x2"], 0), Ws).astype(np.int32)
    v0 = np.minimum(np.maximum(h["y1"], 0), Hs).astype(np.int32)
    v1 = np.minimum(np.maximum(h["y2"], 0), Hs).astype(np.int32)
    return np.stack([u0, v0, u1, v1], axis=1)


def pack_box_masks(rois: np.ndarray, masks_per_box) -> list:
    """Bit-pack one (H, W) mask per box over the box's integer ROI (rows padded to 32
    bits, bit c & 31 of word c >> 5). -> list of uint32 arrays (nrows * wpr words each).
    Host-side set-up (numpy packbits)."""
    out = []
    for (u0, v0, u1, v1), m in zip(rois, masks_per_box):
        m = torch.as_tensor(m).cpu().numpy()
        if u1 <= u0 or v1 <= v0:
            out.append(np.zeros(0, dtype=np.uint32))
            continue
        sub = m[v0:v1, u0:u1] != 0
        nrows, ncols = sub.shape
        wpr = (ncols + 31) // 32
        padded = np.zeros((nrows, wpr * 32), dtype=np.uint8)
        padded[:, :ncols] = sub
        out.append(np.packbits(padded, axis=1, bitorder="little").view("<u4").reshape(-1).copy())
    return out


class SegMasks:
    """Bit-packed per-box segmentation masks (include/wcr.h ``seg_bits``/``seg_off``).

    ``masks_per_box``: one (H, W) mask per box (bool / u8 / float, non-zero == object,
    host or device); only the box's integer ROI is kept -- the mask term of the
    composite is defined on the ROI. ``packed``: the same already
    packed by ``pack_box_masks`` (lets a caller stream images without holding every
    dense mask). Packing is host-side set-up, done once per shard."""

    def __init__(self, shard: ImageShard, boxes: "BoxTable", masks_per_box=None, *, packed=None):
        rois = box_rois(boxes, shard)
        words = packed if packed is not None else pack_box_masks(rois, masks_per_box)
        assert len(words) == len(rois)
        off = [0]
        for (u0, v0, u1, v1), w in zip(rois, words):
            expect = max(v1 - v0, 0) * ((max(u1 - u0, 0) + 31) // 32)
            assert len(w) == expect, "packed mask does not match the box ROI"
            off.append(off[-1] + len(w))
        allw = np.concatenate(words) if len(words) and off[-1] else np.zeros(1, dtype=np.uint32)
        self.host_bits, self.host_off = allw.astype(np.uint32), np.asarray(off, dtype=np.int64)
        self.bits = torch.from_numpy(self.host_bits.view(np.int32).copy()).to(shard.device)
        self.off = torch.from_numpy(self.host_off.copy()).to(shard.device)
        self.n_boxes = len(rois)


def _segmasks_from_dense(shard: ImageShard, boxes: "BoxTable", masks: torch.Tensor, mask_idx=None):
    """SegMasks packed ON THE DEVICE from dense masks (``wcr_pack_box_masks``): ``masks``
    (n_masks, Hm, Wm) u8/bool CUDA tensor, the reference's ``masks_all_images``
    (roi_heads.py:273); box b uses mask ``mask_idx[b]`` (default b). Only the word offsets are
    computed on the host (from the box ROIs). ``repack(masks)`` refreshes the bits in place."""
    self = SegMasks.__new__(SegMasks)
    rois = box_rois(boxes, shard)
    off = np.zeros(len(rois) + 1, dtype=np.int64)
    if len(rois):
        nrows = np.maximum(rois[:, 3] - rois[:, 1], 0)
        ncols = np.maximum(rois[:, 2] - rois[:, 0], 0)
        off[1:] = np.cumsum(np.where(ncols > 0, nrows, 0).astype(np.int64) * ((ncols + 31) // 32))
    self.host_off, self.host_bits = off, None
    self.roi_bytes = int((np.where(ncols > 0, nrows, 0).astype(np.int64) * ncols).sum()) if len(rois) else 0
    self.bits = torch.zeros((max(int(off[-1]), 1),), dtype=torch.int32, device=shard.device)
    self.off = torch.from_numpy(off.copy()).to(shard.device)
    self.n_boxes = len(rois)
    self._shard, self._boxes = shard, boxes
    self._mask_idx = None if mask_idx is None else \
        torch.as_tensor(mask_idx, dtype=torch.int32).to(shard.device).contiguous()
    self.repack(masks)
    return self


def _segmasks_repack(self, masks: torch.Tensor):
    """``masks``: a CUDA tensor, or a PINNED host tensor -- pinned memory is mapped into the
    device's address space (unified virtual addressing), so the pack kernel then reads just the
    ROI bytes it needs straight from host memory over PCIe and the dense maps are never copied."""
    if not masks.is_cuda and not masks.is_pinned():
        raise _abi.WcrError("wcr: CUDA or pinned host tensor required (there is no CPU fallback on this path)")
    if masks.dtype == torch.bool:
        masks = masks.view(torch.uint8)
    assert masks.dtype == torch.uint8 and masks.dim() == 3 and masks.stride(2) == 1
    lib = _abi.load()
    _abi.check(lib.wcr_pack_box_masks(masks.data_ptr(), masks.stride(0), masks.stride(1),
                  _abi.ptr(self._mask_idx), self._shard.table.data_ptr(),
                  self._boxes.table.data_ptr(), self.n_boxes, self.off.data_ptr(),
                  self.bits.data_ptr(), _abi.stream_ptr(self.bits.device)),
               "wcr_pack_box_masks")


def _segmasks_from_arrays(host_bits: np.ndarray, host_off: np.ndarray, device, bits=None, off=None):
    """SegMasks over prebuilt arrays (wcr.job); ``bits`` / ``off``: existing device buffers the
    caller uploads into, None => uploaded here."""
    self = SegMasks.__new__(SegMasks)
    self.host_bits, self.host_off = host_bits, host_off
    self.bits = bits if bits is not None else \
        torch.from_numpy(host_bits.view(np.int32).copy()).to(device, non_blocking=True)
    self.off = off if off is not None else torch.from_numpy(host_off.copy()).to(device, non_blocking=True)
    self.n_boxes = len(host_off) - 1
    return self


SegMasks.from_arrays = staticmethod(_segmasks_from_arrays)
SegMasks.from_dense = staticmethod(_segmasks_from_dense)
SegMasks.repack = _segmasks_repack


class ScoreWorkspace:
    """Caller-owned buffers of ``score_topk`` so a steady-state loop allocates
    nothing (and so the top-k records can be written straight into a registered
    NCCL send buffer, see wcr.dist)."""

    def __init__(self, n_boxes: int, N: int, k: int, device, records: torch.Tensor | None = None,
                 ws: torch.Tensor | None = None):
        """``ws``: share the byte buffer of another workspace (>= the size this one needs);
        launches that share it must be ordered on one stream."""
        lib = _abi.load()
        nbytes = int(lib.wcr_score_workspace_bytes(n_boxes, N, k))
        if ws is not None:
            assert ws.numel() >= nbytes, "shared score workspace too small"
            self.ws = ws
        else:
            self.ws = torch.empty((max(nbytes, 256),), dtype=torch.uint8, device=device)  # 256-B aligned by the allocator
        self.records = records if records is not None else torch.empty(
            (n_boxes, k, RECORD_FLOATS), dtype=torch.float32, device=device)
        self.n_boxes, self.N, self.k = n_boxes, N, k

    def executed_tests(self) -> int:
        """Point tests the last ``score_topk`` on this workspace actually issued (after the
        projected-box cull; padding lanes included). The library keeps the counter at byte 8
        of the workspace (include/wcr.h). Synchronises."""
        return int(self.ws[8:16].view(torch.int64).item())

    def hulls_filled(self) -> int:
        """Hulls the last ``score_topk`` filled for the mask term (all of them with
        ``want_all=True``; only those whose bound could still reach the top-k otherwise).
        Synchronises."""
        return int(self.ws[16:24].view(torch.int64).item())

    def bounds(self):
        """-> (ubcnt (n_boxes, N) int32, segb (n_boxes, N, 2) int32, segwin (n_boxes, N) int32,
        mint (n_boxes, 33, 33) int32): views of the per-proposal bounds the last on-demand
        ``score_topk`` on this workspace pruned with (upper bound of the inside-count; max
        |mask & hull| and min |hull| pixels; the scan window's cell range in the 32 x 32 mask grid
        and that grid's integral tables). Test hook."""
        lib = _abi.load()
        o = [ctypes.c_int64() for _ in range(4)]
        _abi.check(lib.wcr_score_bounds_offsets(self.n_boxes, self.N, self.k, *[ctypes.byref(x) for x in o]),
                   "wcr_score_bounds_offsets")
        n = self.n_boxes * self.N
        ub = self.ws[o[0].value:o[0].value + 4 * n].view(torch.int32).view(self.n_boxes, self.N)
        sb = self.ws[o[1].value:o[1].value + 8 * n].view(torch.int32).view(self.n_boxes, self.N, 2)
        sw = self.ws[o[2].value:o[2].value + 4 * n].view(torch.int32).view(self.n_boxes, self.N)
        mi = self.ws[o[3].value:o[3].value + 4 * 1089 * self.n_boxes].view(torch.int32).view(self.n_boxes, 33, 33)
        return ub, sb, sw, mi

    def proposals_pruned(self) -> int:
        """Proposals the last ``score_topk`` on this workspace never tested against the depth
        points (point term on demand: their score bound was below the k-th best score already
        known). 0 when the point term was evaluated for every proposal. Synchronises."""
        return int(self.ws[40:48].view(torch.int64).item())


def score_topk(shard: ImageShard, boxes: BoxTable, cubes: torch.Tensor | None, N: int | None = None,
               cfg: ScoreConfig | None = None, *, priors=None, normals=None, z0=None,
               sample_cfg: ProposalConfig | None = None, seed: int = 0, seg=None,
               workspace: ScoreWorkspace | None = None, want_all: bool = False, seg_ready=None):
    """Fused scoring + per-box top-k. ``cubes`` (n_boxes, N, 15) or None to sample
    in-kernel (then priors/normals are required; same generator as ``propose``).
    ``seg_ready``: a recorded ``torch.cuda.Event`` the scoring kernel waits for before it reads
    ``seg`` (``wcr_score_topk_after``): the masks may still be on their way on another stream.

    -> dict(records (n_boxes,k,20) f32 [idx bits, score, cube15, s_iou, s_pts, s_seg],
            idx (n_boxes,k) i32 view, score (n_boxes,k);
            with want_all: scores, iou, seg (n_boxes,N) f32, count (n_boxes,N) i32,
            proj (n_boxes,N,4), npts (n_boxes,) i32)."""
    lib = _abi.load()
    cfg = cfg or ScoreConfig()
    dev = shard.device
    nb = len(boxes)
    if cubes is not None:
        _abi.require_cuda(cubes)
        cubes = _f32(cubes, dev)
        assert cubes.dim() == 3 and cubes.shape[0] == nb and cubes.shape[2] == CUBE_FLOATS
        N = cubes.shape[1]
    else:
        assert N is not None and priors is not None and normals is not None
        priors, normals = _f32(priors, dev), _f32(normals, dev)
        if z0 is None:
            z0 = box_center_depth(shard, boxes)
    sp = (sample_cfg or ProposalConfig()).c_struct()
    sc = cfg.c_struct()
    if workspace is None:
        workspace = ScoreWorkspace(nb, N, cfg.k, dev)
    assert workspace.n_boxes >= nb and workspace.N == N and workspace.k == cfg.k
    rec = workspace.records
    outs = {}
    outs["workspace"] = workspace
    if want_all:
        outs["scores"] = torch.empty((nb, N), dtype=torch.float32, device=dev)
        outs["iou"] = torch.empty((nb, N), dtype=torch.float32, device=dev)
        outs["seg"] = torch.empty((nb, N), dtype=torch.float32, device=dev)
        outs["count"] = torch.empty((nb, N), dtype=torch.int32, device=dev)
        outs["proj"] = torch.empty((nb, N, 4), dtype=torch.float32, device=dev)
        outs["npts"] = torch.empty((nb,), dtype=torch.int32, device=dev)
    seg_bits = seg_off = None
    if seg is not None:
        seg_bits, seg_off = seg.bits, seg.off
        assert seg.n_boxes == nb
    elif cfg.e_seg > 0:
        raise _abi.WcrError("score_topk: e_seg > 0 needs seg=SegMasks(...)")
    _abi.check(lib.wcr_score_topk_after(
        shard.table.data_ptr(), boxes.table.data_ptr(), nb, _abi.ptr(cubes), N,
        _abi.ptr(priors), _abi.ptr(normals), _abi.ptr(z0), ctypes.byref(sp), int(seed),
        _abi.ptr(seg_bits), _abi.ptr(seg_off), ctypes.byref(sc), cfg.k, rec.data_ptr(),
        _abi.ptr(outs.get("scores")), _abi.ptr(outs.get("iou")), _abi.ptr(outs.get("seg")),
        _abi.ptr(outs.get("count")),
        _abi.ptr(outs.get("proj")), _abi.ptr(outs.get("npts")), workspace.ws.data_ptr(),
        None if seg_ready is None else ctypes.c_void_p(int(seg_ready.cuda_event)),
        _abi.stream_ptr(dev)), "wcr_score_topk")
    rec = rec[:nb]
    outs["records"] = rec
    outs["idx"] = rec[:, :, 0].contiguous().view(torch.int32)
    outs["score"] = rec[:, :, 1]
    return outs
