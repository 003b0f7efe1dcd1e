"""Step 1 host side: ground-frame unprojection + RANSAC plane + up vector.

Drop-in for ``ROIHeads3DScore.normal_vector_from_maps`` (reference
cubercnn/modeling/roi_heads/roi_heads.py:466-528) -- one launch chain for the
whole shard instead of a Python loop of ~25 torch kernels per image, no host
synchronisation (the reference syncs on boolean-mask indexing, :489-491).
"""
from __future__ import annotations

import random as _random

import torch

from . import _abi
from .images import ImageShard


def unproject_ground(shard: ImageShard, use_nth: int = 5):
    """-> (pts (B, cap, 4) f32, count (B,) i32). Row-major scan order of the
    strided grid, exactly the order of roi_heads.py:489-502."""
    lib = _abi.load()
    B = len(shard)
    cap = max(1, -(-shard.max_H // use_nth) * -(-shard.max_W // use_nth))
    pts = torch.empty((B, cap, 4), dtype=torch.float32, device=shard.device)
    count = torch.empty((B,), dtype=torch.int32, device=shard.device)
    _abi.check(lib.wcr_unproject_ground(shard.table.data_ptr(), B, use_nth, pts.data_ptr(), cap,
                                        count.data_ptr(), _abi.stream_ptr(shard.device)),
               "wcr_unproject_ground")
    return pts, count


def unproject_cloud(shard: ImageShard, stride: int = 1):
    """Dense K-frame cloud: (B, Hs_max*Ws_max, 4) f32; image i is
    packed with its own row pitch Ws_i = ceil(W_i/stride). NaN == invalid depth."""
    lib = _abi.load()
    B = len(shard)
    Hs, Ws = -(-shard.max_H // stride), -(-shard.max_W // stride)
    out = torch.empty((B, Hs * Ws, 4), dtype=torch.float32, device=shard.device)
    _abi.check(lib.wcr_unproject_cloud(shard.table.data_ptr(), B, stride, shard.max_H, shard.max_W,
                                       out.data_ptr(), Hs * Ws, _abi.stream_ptr(shard.device)),
               "wcr_unproject_cloud")
    return out


def ransac_ground(pts, count, sample_idx=None, iters: int = 1000, thresh: float = 0.05,
                  seed: int = 0, want_mask: bool = False):
    """All images in one call. ``sample_idx`` (B, iters, 3) int32 replays the
    reference's ``random.sample`` draws (plane.py:95); None => in-kernel Philox.
    -> dict(eq (B,4) negated, normal (B,3), best (B,), counts (B,iters), mask)."""
    lib = _abi.load()
    _abi.require_cuda(pts, count, sample_idx)
    B, cap = pts.shape[0], pts.shape[1]
    dev = pts.device
    eq = torch.empty((B, 4), dtype=torch.float32, device=dev)
    normal = torch.empty((B, 3), dtype=torch.float32, device=dev)
    best = torch.empty((B,), dtype=torch.int32, device=dev)
    counts = torch.empty((B, iters), dtype=torch.int32, device=dev)
    ws = torch.empty((B, iters, 4), dtype=torch.float32, device=dev)
    mask = torch.zeros((B, cap), dtype=torch.uint8, device=dev) if want_mask else None
    if sample_idx is not None:
        sample_idx = sample_idx.to(device=dev, dtype=torch.int32).contiguous()
        assert tuple(sample_idx.shape) == (B, iters, 3)
    _abi.check(lib.wcr_ransac_ground(pts.data_ptr(), count.data_ptr(), cap, B,
                                     _abi.ptr(sample_idx), iters, float(thresh), int(seed),
                                     eq.data_ptr(), normal.data_ptr(), best.data_ptr(),
                                     counts.data_ptr(), _abi.ptr(mask), ws.data_ptr(),
                                     _abi.stream_ptr(dev)),
               "wcr_ransac_ground")
    return {"eq": eq, "normal": normal, "best": best, "counts": counts, "mask": mask, "eqs": ws}


def draw_reference_samples(counts_host, iters: int = 1000, rng=_random):
    """Replay helper: the exact draw of plane.py:95 for each image, in image
    order, from Python's ``random`` stream (host side; needs the point counts,
    hence one D2H sync -- used for parity runs, not for throughput runs)."""
    out = torch.empty((len(counts_host), iters, 3), dtype=torch.int32)
    for i, n in enumerate(counts_host):
        if int(n) < 3:  # the reference's random.sample raises here; the kernel yields NaN
            out[i] = 0
            continue
        out[i] = torch.tensor([rng.sample(range(0, int(n)), 3) for _ in range(iters)],
                              dtype=torch.int32)
    return out


def normal_vector_from_maps(ground_maps, depth_maps, Ks, use_nth: int = 5, *, replay_rng=None,
                            seed: int = 0, shard: ImageShard | None = None):
    """Same positional signature and return as the reference method
    (roi_heads.py:466): (B,3) up vectors.

    Call-site quirk kept: ``Ks`` is zipped against *images*, so
    image i uses ``Ks[i]`` even when the caller passes one K per box.
    ``replay_rng``: a ``random.Random``/the ``random`` module => consume the same
    Mersenne-Twister stream as plane.py:95 (bit-exact parity mode, one host sync);
    None => Philox in-kernel (throughput mode, no sync)."""
    if shard is None:
        shard = ImageShard(depth_maps, ground_maps, Ks)
    pts, count = unproject_ground(shard, use_nth)
    sample_idx = None
    if replay_rng is not None:
        sample_idx = draw_reference_samples(count.cpu().tolist(), 1000, replay_rng).to(shard.device)
    res = ransac_ground(pts, count, sample_idx, iters=1000, thresh=0.05, seed=seed)
    return res["normal"]
