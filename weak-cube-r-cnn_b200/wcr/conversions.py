"""Drop-in for ProposalNetwork/utils/conversions.py (``cubes_to_box``) plus the
minimal ``Boxes`` it returns when detectron2 is not importable.

``Boxes`` restates the small part of detectron2.structures.Boxes the hot path
touches (tensor, area, clip, get_centers, len, indexing, iteration); when
detectron2 is installed its own class is used so that callers get the type they
expect.
"""
from __future__ import annotations

import torch

try:  # pragma: no cover - detectron2 is absent in this environment
    from detectron2.structures import Boxes  # type: ignore
except Exception:  # noqa: BLE001
    class Boxes:
        def __init__(self, tensor: torch.Tensor):
            if not isinstance(tensor, torch.Tensor):
                tensor = torch.as_tensor(tensor, dtype=torch.float32)
            tensor = tensor.to(torch.float32)
            if tensor.numel() == 0:
                tensor = tensor.reshape((-1, 4))
            assert tensor.dim() == 2 and tensor.size(-1) == 4, tensor.size()
            self.tensor = tensor

        def area(self):
            b = self.tensor
            return (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])

        def clip(self, box_size):
            h, w = box_size
            self.tensor = torch.stack((self.tensor[:, 0].clamp(0, w), self.tensor[:, 1].clamp(0, h),
                                       self.tensor[:, 2].clamp(0, w), self.tensor[:, 3].clamp(0, h)), dim=-1)

        def get_centers(self):
            return (self.tensor[:, :2] + self.tensor[:, 2:]) / 2

        def to(self, device):
            return Boxes(self.tensor.to(device=device))

        @property
        def device(self):
            return self.tensor.device

        def __getitem__(self, item):
            if isinstance(item, int):
                return Boxes(self.tensor[item].view(1, -1))
            return Boxes(self.tensor[item])

        def __len__(self):
            return self.tensor.shape[0]

        def __iter__(self):
            yield from self.tensor

        def __repr__(self):
            return f"Boxes({self.tensor})"


def cube_to_box(cube, K):
    """conversions.py:5-23: ONE cube -> a ``Boxes`` of one box. The reference calls
    ``cube.get_bube_corners(K)`` without a clamp there, which raises (spaces.py:240-243 leaves x, y
    undefined); the same ``NameError`` surfaces here, so a caller ported from the reference behaves
    identically. (With a clamp the box is ``cubes_to_box(cube, K, clamp)[0]``.)"""
    bube_corners = cube.get_bube_corners(K)        # raises NameError, like the reference
    min_x, max_x = bube_corners[..., 0].min(), bube_corners[..., 0].max()
    min_y, max_y = bube_corners[..., 1].min(), bube_corners[..., 1].max()
    return Boxes(torch.stack((min_x, min_y, max_x, max_y)).view(1, 4))


def cubes_to_box(cubes, K, im_shape):
    """conversions.py:25-48: per instance, the (P, 4) boxes ``(min_x, min_y, max_x,
    max_y)`` over the 8 clamped projected corners. One kernel launch
    (wcr_bube_corners) for all instances."""
    boxes = cubes.projected_boxes(K, im_shape)
    return [Boxes(boxes[i]) for i in range(cubes.num_instances)]
