"""Drop-in for ProposalNetwork/utils/plane.py: ``Plane_torch().fit_parallel``.

Same signature and return as the reference (plane.py:79): ``(-equation (4,),
inlier indices int64)``. The (maxIteration x N) distance matrix is never
materialised: wcr_ransac_ground keeps one hypothesis per thread and streams the
points through shared memory.
"""
from __future__ import annotations

import random

import torch

from . import _abi
from .ground import ransac_ground


class Plane_torch:
    def __init__(self):
        self.inliers = []
        self.equation = []

    def fit_parallel(self, pts: torch.Tensor, thresh=0.05, minPoints=100, maxIteration=1000,
                     *, id_samples=None, seed=None):
        """``pts`` (N,3) CUDA tensor. By default the 3-point samples come from
        Python's ``random.sample`` exactly as in the reference (plane.py:95), so a
        seeded ``random`` gives the reference's stream. ``id_samples`` overrides
        them; ``seed`` (int) switches to the in-kernel Philox generator."""
        _abi.require_cuda(pts)
        n = pts.shape[0]
        dev = pts.device
        p4 = torch.zeros((1, max(n, 1), 4), dtype=torch.float32, device=dev)
        p4[0, :n, :3] = pts.to(torch.float32)
        count = torch.tensor([n], dtype=torch.int32, device=dev)
        if seed is None and id_samples is None:
            id_samples = torch.tensor(
                [random.sample(range(0, n), 3) for _ in range(maxIteration)], dtype=torch.int32)
        if id_samples is not None:
            id_samples = torch.as_tensor(id_samples).to(device=dev, dtype=torch.int32).reshape(
                1, maxIteration, 3)
        res = ransac_ground(p4, count, id_samples, iters=maxIteration, thresh=thresh,
                            seed=0 if seed is None else seed, want_mask=True)
        self.equation = -res["eq"][0]  # the reference stores the un-negated plane
        self.inliers = torch.where(res["mask"][0, :n] > 0)[0]
        return res["eq"][0], self.inliers
