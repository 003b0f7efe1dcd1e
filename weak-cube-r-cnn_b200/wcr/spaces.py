"""Drop-in for ProposalNetwork/utils/spaces.py: the ``Cubes`` container.

Same constructor, properties and methods as the reference class (spaces.py:95-328);
the geometry (``get_all_corners``, ``get_bube_corners``) runs in libwcr_b200
(wcr_cube_corners / wcr_bube_corners) instead of a Python loop over instances of
torch ops. Container logic (views, indexing, cat, split) is plain tensor plumbing.

Differences: tensors must live on a CUDA device for the geometry methods (no CPU
fallback). The geometry methods are DIFFERENTIABLE in ``Cubes.tensor`` like the reference's
torch ops (it calls them inside autograd: spaces.py:224-245, roi_heads.py:947-960): each is a
``torch.autograd.Function`` over a forward and a backward kernel (wcr_cube_corners[_bwd],
wcr_bube_corners[_bwd]); the fused ``wcr.roi_heads.weak_loss`` remains the fast path.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _abi


def _K4(K):
    K = K.detach().cpu() if isinstance(K, torch.Tensor) else torch.as_tensor(np.asarray(K))
    return float(K[0, 0]), float(K[1, 1]), float(K[0, 2]), float(K[1, 2])


# faces of get_cuboid_verts_faces (math_util.py:220-238): vertex indices per triangle
_FACES = [[0, 1, 2], [2, 3, 0], [1, 5, 6], [6, 2, 1], [4, 0, 3], [3, 7, 4],
          [5, 4, 7], [7, 6, 5], [4, 5, 1], [1, 0, 4], [3, 2, 6], [6, 7, 3]]


class _CornersFn(torch.autograd.Function):
    """Cubes.get_all_corners: (M, 15) -> (M, 8, 3)."""

    @staticmethod
    def forward(ctx, flat):
        lib = _abi.load()
        f = flat.detach().contiguous()
        out = torch.empty((f.shape[0], 8, 3), dtype=torch.float32, device=f.device)
        _abi.check(lib.wcr_cube_corners(f.data_ptr(), f.shape[0], out.data_ptr(),
                                        _abi.stream_ptr(f.device)), "wcr_cube_corners")
        ctx.save_for_backward(f)
        return out

    @staticmethod
    def backward(ctx, g):
        lib = _abi.load()
        (f,) = ctx.saved_tensors
        g = g.to(torch.float32).contiguous()
        d = torch.empty_like(f)
        _abi.check(lib.wcr_cube_corners_bwd(f.data_ptr(), g.data_ptr(), f.shape[0], d.data_ptr(),
                                            _abi.stream_ptr(f.device)), "wcr_cube_corners_bwd")
        return d


class _BubeFn(torch.autograd.Function):
    """get_bube_corners (which == 0: (M, 8, 2)) or the cubes_to_box min / max (which == 1: (M, 4))."""

    @staticmethod
    def forward(ctx, flat, K4, clamp, which):
        lib = _abi.load()
        f = flat.detach().contiguous()
        M = f.shape[0]
        out = torch.empty((M, 8, 2) if which == 0 else (M, 4), dtype=torch.float32, device=f.device)
        _abi.check(lib.wcr_bube_corners(f.data_ptr(), M, *K4, float(clamp[0]), float(clamp[1]),
                                        out.data_ptr() if which == 0 else None,
                                        out.data_ptr() if which == 1 else None,
                                        _abi.stream_ptr(f.device)), "wcr_bube_corners")
        ctx.save_for_backward(f)
        ctx.K4, ctx.clamp, ctx.which = K4, (float(clamp[0]), float(clamp[1])), which
        return out

    @staticmethod
    def backward(ctx, g):
        lib = _abi.load()
        (f,) = ctx.saved_tensors
        g = g.to(torch.float32).contiguous()
        d = torch.empty_like(f)
        _abi.check(lib.wcr_bube_corners_bwd(f.data_ptr(), f.shape[0], *ctx.K4, *ctx.clamp,
                                            g.data_ptr() if ctx.which == 0 else None,
                                            g.data_ptr() if ctx.which == 1 else None,
                                            d.data_ptr(), _abi.stream_ptr(f.device)), "wcr_bube_corners_bwd")
        return d, None, None, None


class Cubes:
    """3D boxes ``[cx, cy, cz, w, h, l, R (9, row-major)]``, tensor shape
    (num_instances, num_proposals, 15) -- spaces.py:95-150."""

    def __init__(self, tensor, scores=None, labels=None) -> None:
        if scores is not None:
            assert scores.ndim == 2, \
                f"scores.shape must be (n_instances, n_proposals), but was {scores.shape}"
        self.scores, self.labels = scores, labels
        if isinstance(tensor, torch.Tensor):
            tensor = tensor.to(torch.float32)
        else:
            tensor = torch.as_tensor(np.asarray(tensor), dtype=torch.float32, device="cpu")
        if tensor.numel() == 0:
            tensor = tensor.reshape((-1, 15))
        while tensor.dim() < 3:          # (15,) -> (1,15) -> (1,1,15); (P,15) -> (1,P,15)
            tensor = tensor.unsqueeze(0)
        self.tensor = tensor

    # ---- views (spaces.py:152-180) -------------------------------------------
    @property
    def centers(self):
        return self.tensor[:, :, :3]

    @property
    def dimensions(self):
        return self.tensor[:, :, 3:6]

    @property
    def rotations(self):
        s = self.tensor.shape
        return self.tensor[:, :, 6:].reshape(s[0], s[1], 3, 3)

    @property
    def device(self):
        return self.tensor.device

    @property
    def num_instances(self):
        return self.tensor.shape[0]

    @property
    def shape(self):
        return self.tensor.shape

    def clone(self) -> "Cubes":
        return Cubes(self.tensor.clone())

    # ---- geometry: one kernel launch each -------------------------------------
    def _flat(self):
        _abi.require_cuda(self.tensor)
        return self.tensor.contiguous().view(-1, 15)      # keeps the autograd graph of self.tensor

    def get_all_corners(self) -> torch.Tensor:
        """(I, P, 8, 3) camera-space vertices, ``get_cuboid_verts_faces`` order
        (math_util.py:198-217) -- spaces.py:192-204. Differentiable."""
        out = _CornersFn.apply(self._flat())
        return out.view(self.tensor.shape[0], self.tensor.shape[1], 8, 3)

    def get_cuboids_verts_faces(self):
        """spaces.py:206-222: ``(verts (I, P, 8, 3), faces (I, P, 12, 3))``; the faces are the
        constant triangle table of math_util.py:220-238 (as floats, like the reference)."""
        verts = self.get_all_corners()
        faces = torch.tensor(_FACES, dtype=torch.float32, device=verts.device)
        return verts, faces.expand(verts.shape[0], verts.shape[1], 12, 3).contiguous()

    def get_bube_corners(self, K, clamp: tuple = None) -> torch.Tensor:
        """(I, P, 8, 2) projected corners clamped to
        ``[int(-clamp/2+1), int(clamp-1+clamp)]`` -- spaces.py:224-245. Like the
        reference, ``clamp=None`` is an error (there: NameError at :243). Differentiable
        (clamp passes the gradient on the closed interval, like torch.clamp)."""
        if clamp is None:
            raise NameError("get_bube_corners: clamp is required (the reference leaves x, y "
                            "undefined without it, spaces.py:240-243)")
        out = _BubeFn.apply(self._flat(), _K4(K), clamp, 0)
        return out.view(self.tensor.shape[0], self.tensor.shape[1], 8, 2)

    def projected_boxes(self, K, clamp) -> torch.Tensor:
        """(I, P, 4) min/max of the clamped projected corners: the tensor behind
        ``cubes_to_box`` (conversions.py:36-42). Differentiable (first arg-min / arg-max)."""
        out = _BubeFn.apply(self._flat(), _K4(K), clamp, 1)
        return out.view(self.tensor.shape[0], self.tensor.shape[1], 4)

    def get_volumes(self):
        return self.dimensions.prod(-1)

    # ---- container protocol (spaces.py:247-327) --------------------------------
    def __len__(self) -> int:
        return self.tensor.shape[0]

    def __repr__(self) -> str:
        return f"Cubes({self.tensor})"

    def to(self, device):
        sc = self.scores.to(device=device) if isinstance(self.scores, torch.Tensor) else self.scores
        lb = self.labels.to(device=device) if isinstance(self.labels, torch.Tensor) else self.labels
        self.scores, self.labels = sc, lb
        return Cubes(self.tensor.to(device=device), sc, lb)

    def __getitem__(self, item) -> "Cubes":
        if isinstance(item, int):
            return Cubes(self.tensor[item].view(1, self.tensor.shape[1], -1))
        if isinstance(item, tuple):
            return Cubes(self.tensor[item[0], item[1]].view(1, 1, -1))
        sub = self.tensor[item]
        assert sub.dim() == 2, f"Indexing on Cubes with {item} failed to return a matrix!"
        return Cubes(sub)

    @classmethod
    def cat(cls, cubes_list) -> "Cubes":
        assert isinstance(cubes_list, (list, tuple))
        if len(cubes_list) == 0:
            return cls(torch.empty(0))
        assert all(isinstance(c, Cubes) for c in cubes_list)
        return cls(torch.cat([c.tensor for c in cubes_list], dim=0))

    def __iter__(self):
        yield from self.tensor

    def split(self, split_size: int, dim=1):
        return tuple(Cubes(x) for x in self.tensor.split(split_size, dim=dim))

    def reshape(self, *args) -> "Cubes":
        return Cubes(self.tensor.reshape(*args), self.scores, self.labels)
