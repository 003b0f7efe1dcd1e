"""Image shard descriptors: the device-side table of ``WcrImage`` records
(include/wcr.h) that every kernel of the path indexes by image.

The reference iterates detectron2 ``ImageList``s image by image
(roi_heads.py:472, :585, :645); here the whole shard is described once and each
kernel covers all images of the shard in one launch.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _abi

IMAGE_DTYPE = np.dtype([
    ("depth", np.uint64), ("ground", np.uint64),
    ("H", np.int32), ("W", np.int32),
    ("depth_pitch", np.int32), ("ground_pitch", np.int32),
    ("fx", np.float32), ("fy", np.float32), ("cx", np.float32), ("cy", np.float32),
])
assert IMAGE_DTYPE.itemsize == 48


def _iter_images(maps):
    """Yield un-padded 2-D views from an ImageList-like, a list of tensors or a
    (B,H,W) tensor."""
    if maps is None:
        return
    if hasattr(maps, "image_sizes") and hasattr(maps, "tensor"):
        for i, (h, w) in enumerate(maps.image_sizes):
            t = maps.tensor[i]
            while t.dim() > 2:
                t = t[0]
            yield t[:h, :w]
        return
    for t in maps:
        if t is None:          # no ground map for this image (no_ground_idx.csv): same as the (1,1) dummy
            yield None
            continue
        while t.dim() > 2:
            t = t[0]
        yield t


class ImageShard:
    """Depth maps, ground masks and intrinsics of the images one rank owns."""

    def __init__(self, depth_maps, ground_maps, Ks, device=None):
        depths = list(_iter_images(depth_maps))
        if device is None:
            device = depths[0].device
        device = torch.device(device)
        if device.type != "cuda":
            raise _abi.WcrError("wcr: CUDA tensors required (there is no CPU fallback on this path)")
        grounds = list(_iter_images(ground_maps)) if ground_maps is not None else [None] * len(depths)
        assert len(grounds) == len(depths), "one ground map (or (1,1) dummy) per depth map"
        self.device = device
        self.depths, self.grounds = [], []
        rec = np.zeros(len(depths), dtype=IMAGE_DTYPE)
        if len(depths) and all(isinstance(Ks[i], torch.Tensor) and Ks[i].is_cuda for i in range(len(depths))):
            # the reference hands Ks_scaled_per_box on the device: one transfer, not one sync per image
            Ks = torch.stack([Ks[i].detach().reshape(3, 3) for i in range(len(depths))]).cpu().numpy()
        for i, (d, g) in enumerate(zip(depths, grounds)):
            d = d.to(device=device, dtype=torch.float32)
            if d.stride(-1) != 1:
                d = d.contiguous()
            if g is not None and tuple(g.shape) == (1, 1):
                g = None  # the "(1,1) dummy" of rcnn3d.py:375-378 / roi_heads.py:473
            if g is not None:
                g = g.to(device)
                if g.dtype != torch.uint8:
                    g = (g > 0).to(torch.uint8)
                if g.stride(-1) != 1:
                    g = g.contiguous()
                if g.shape[0] < d.shape[0] or g.shape[1] < d.shape[1]:
                    raise _abi.WcrError(
                        f"wcr: ground map {tuple(g.shape)} of image {i} is smaller than its depth map "
                        f"{tuple(d.shape)} (only the (1,1) dummy may be)")
            K = Ks[i]
            K = K.detach().cpu().numpy() if isinstance(K, torch.Tensor) else np.asarray(K)
            rec[i] = (d.data_ptr(), 0 if g is None else g.data_ptr(), d.shape[0], d.shape[1],
                      d.stride(0), 0 if g is None else g.stride(0),
                      K[0, 0], K[1, 1], K[0, 2], K[1, 2])
            self.depths.append(d)
            self.grounds.append(g)
        self.host_table = rec
        self.table = torch.from_numpy(rec.view(np.uint8).copy()).to(device, non_blocking=True)
        self.n_images = len(depths)
        self.max_H = int(rec["H"].max()) if len(depths) else 0
        self.max_W = int(rec["W"].max()) if len(depths) else 0
        self.has_ground = [g is not None for g in self.grounds]

    @classmethod
    def from_host_table(cls, rec: np.ndarray, device, table: torch.Tensor | None = None):
        """Shard over planes the CALLER keeps alive: ``rec`` is an ``IMAGE_DTYPE`` array whose
        pointer fields already hold device addresses (wcr.job builds these for views into one
        big resident plane). ``table``: an existing device tensor the records will be copied
        into by the caller (a pipeline slot); None => uploaded here."""
        self = cls.__new__(cls)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _abi.WcrError("wcr: CUDA device required (there is no CPU fallback on this path)")
        self.depths, self.grounds = [], []
        self.host_table = rec
        self.table = table if table is not None else \
            torch.from_numpy(rec.view(np.uint8).copy()).to(self.device, non_blocking=True)
        self.n_images = len(rec)
        self.max_H = int(rec["H"].max()) if len(rec) else 0
        self.max_W = int(rec["W"].max()) if len(rec) else 0
        self.has_ground = [int(g) != 0 for g in rec["ground"]]
        return self

    def __len__(self):
        return self.n_images

    def sizes(self):
        return [(int(r["H"]), int(r["W"])) for r in self.host_table]
