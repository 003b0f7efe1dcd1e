"""On-disk formats of the path's inputs.

  depth   np.savez_compressed(depth = float32 (H, W))   generate_depth_maps.py:85-86
  ground  np.savez_compressed(mask = bool (H, W))       generate_ground_segmentations.py:200
  no_ground_idx.csv: one ``img_id`` column               generate_ground_segmentations.py:205-206,
                                                         filter_ground.py:21-26
The mapper resizes both maps to the augmented image size with PIL's NEAREST filter and does NOT
flip them (dataset_mapper.py:126-145). Here the zip/deflate decode stays on the host (zlib), the
maps go to the device at their stored size and the resize is a device gather through PIL's own
index tables, bit-identical to ``Image.resize(size, Image.NEAREST)``. No CPU fallback for the
device part.
"""
from __future__ import annotations

import csv
import os

import numpy as np
import torch

from . import _abi


def pil_nearest_table(in_size: int, out_size: int) -> np.ndarray:
    """Source index of every output pixel along one axis, as PIL computes it (host)."""
    tab = np.empty(out_size, dtype=np.int32)
    _abi.load().wcr_pil_nearest_table(int(in_size), int(out_size), tab.ctypes.data)
    return tab


_TABLES: dict = {}


def _device_table(n_in: int, n_out: int, dev) -> torch.Tensor:
    key = (n_in, n_out, str(dev))
    if key not in _TABLES:            # a data loader sees a handful of (stored, augmented) size pairs
        _TABLES[key] = torch.from_numpy(pil_nearest_table(n_in, n_out)).to(dev)
    return _TABLES[key]


def resize_nearest(t: torch.Tensor, size) -> torch.Tensor:
    """``np.array(Image.fromarray(a).resize((w, h), Image.NEAREST))`` for a (H, W) float32 / bool /
    uint8 CUDA tensor; ``size`` = (h, w) of the augmented image (``image.shape[:2]``)."""
    _abi.require_cuda(t)
    lib = _abi.load()
    assert t.dim() == 2
    oh, ow = int(size[0]), int(size[1])
    src = t.contiguous()
    if src.dtype == torch.bool:
        view = src.view(torch.uint8)
    elif src.dtype in (torch.uint8, torch.float32):
        view = src
    else:
        raise _abi.WcrError(f"resize_nearest: unsupported dtype {src.dtype}")
    ih, iw = src.shape
    if (ih, iw) == (oh, ow):
        return src.clone()          # PIL copies when the size is unchanged
    dev = src.device
    xt, yt = _device_table(iw, ow, dev), _device_table(ih, oh, dev)
    out = torch.empty((oh, ow), dtype=view.dtype, device=dev)
    _abi.check(lib.wcr_resize_nearest(view.data_ptr(), view.element_size(), ih, iw, iw, xt.data_ptr(),
                                      yt.data_ptr(), out.data_ptr(), oh, ow, ow, _abi.stream_ptr(dev)),
               "wcr_resize_nearest")
    return out.view(torch.bool) if src.dtype == torch.bool else out


def read_no_ground_idx(path: str) -> set:
    """``datasets/no_ground_idx.csv`` -> set of image ids without a usable ground map."""
    if not os.path.exists(path):
        return set()
    with open(path, newline="") as f:
        rows = list(csv.DictReader(f))
    return {int(r["img_id"]) for r in rows if r.get("img_id") not in (None, "")}


def load_depth_npz(path: str) -> torch.Tensor:
    """Host decode of one depth file -> pinned (H, W) float32 tensor."""
    a = np.load(path)["depth"]
    t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
    return t.pin_memory() if torch.cuda.is_available() else t


def load_ground_npz(path: str) -> torch.Tensor:
    a = np.load(path)["mask"]
    t = torch.from_numpy(np.ascontiguousarray(a).astype(np.bool_))
    return t.pin_memory() if torch.cuda.is_available() else t


def load_maps(img_ids, image_sizes, *, depth_dir="datasets/depth_maps", ground_dir="datasets/ground_maps",
              no_ground=None, device="cuda"):
    """The mapper's depth / ground loading for a batch (dataset_mapper.py:124-147):
    -> (depth_maps, ground_maps) lists of CUDA tensors at ``image_sizes[i]`` = (h, w); an image
    listed in ``no_ground`` (or without a ground file) gets the reference's ``(1, 1)`` dummy
    (rcnn3d.py builds a zeros((1,1)) map for None)."""
    dev = torch.device(device)
    no_ground = no_ground or set()
    depths, grounds = [], []
    for img_id, hw in zip(img_ids, image_sizes):
        d = load_depth_npz(os.path.join(depth_dir, f"{img_id}.npz")).to(dev, non_blocking=True)
        depths.append(resize_nearest(d, hw))
        gp = os.path.join(ground_dir, f"{img_id}.npz")
        if int(img_id) in no_ground or not os.path.exists(gp):
            grounds.append(None)
        else:
            g = load_ground_npz(gp).to(dev, non_blocking=True)
            grounds.append(resize_nearest(g, hw))
    return depths, grounds
