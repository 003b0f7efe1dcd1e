"""TEST INFRASTRUCTURE ONLY (oracle). Never imported by the product path.

Small, real stand-ins for the third-party arithmetic the reference calls but
does not vendor:

* detectron2 (requirements.txt:9, unpinned git HEAD): ``Boxes``,
  ``pairwise_iou`` -- call sites ProposalNetwork/utils/utils.py:186,
  ProposalNetwork/utils/conversions.py:44-48, roi_heads.py:575-610.
* pytorch3d (requirements.txt:8, ``@stable``): ``axis_angle_to_matrix`` used by
  cubercnn/util/math_util.py:827, ``rotation_6d_to_matrix`` (cube_head.py:176).
* an ``ImageList`` look-alike: iteration yields the un-padded per-image view and
  ``.image_sizes`` lists (h, w) -- behaviour relied on at roi_heads.py:472,585,645,988.

These restate the libraries' *published* formulas; the reference pins none of
them with tests, so parity at these boundaries is
"pinned by this project", and each is cross-checked in tests/ against an
independent formula.
"""
from __future__ import annotations

import torch


class Boxes:
    """Axis-aligned boxes (x1, y1, x2, y2); detectron2.structures.Boxes subset."""

    def __init__(self, tensor):
        if not isinstance(tensor, torch.Tensor):
            tensor = torch.as_tensor(tensor, dtype=torch.float32)
        else:
            tensor = tensor.to(torch.float32)
        if tensor.numel() == 0:
            tensor = tensor.reshape((-1, 4)).to(dtype=torch.float32)
        assert tensor.dim() == 2 and tensor.size(-1) == 4, tensor.size()
        self.tensor = tensor

    def area(self):
        b = self.tensor
        return (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])

    def clip(self, box_size):
        h, w = box_size
        x1 = self.tensor[:, 0].clamp(min=0, max=w)
        y1 = self.tensor[:, 1].clamp(min=0, max=h)
        x2 = self.tensor[:, 2].clamp(min=0, max=w)
        y2 = self.tensor[:, 3].clamp(min=0, max=h)
        self.tensor = torch.stack((x1, y1, x2, y2), dim=-1)

    def get_centers(self):
        return (self.tensor[:, :2] + self.tensor[:, 2:]) / 2

    def __getitem__(self, item):
        if isinstance(item, int):
            return Boxes(self.tensor[item].view(1, -1))
        b = self.tensor[item]
        assert b.dim() == 2
        return Boxes(b)

    def __len__(self):
        return self.tensor.shape[0]

    def __iter__(self):
        yield from self.tensor

    def to(self, device):
        return Boxes(self.tensor.to(device=device))

    @property
    def device(self):
        return self.tensor.device


def pairwise_intersection(boxes1: Boxes, boxes2: Boxes):
    b1, b2 = boxes1.tensor, boxes2.tensor
    wh = torch.min(b1[:, None, 2:], b2[:, 2:]) - torch.max(b1[:, None, :2], b2[:, :2])
    wh.clamp_(min=0)
    return wh.prod(dim=2)


def pairwise_iou(boxes1: Boxes, boxes2: Boxes):
    """inter / (a1 + a2 - inter) where inter > 0 else 0 (detectron2 formula)."""
    area1 = boxes1.area()
    area2 = boxes2.area()
    inter = pairwise_intersection(boxes1, boxes2)
    iou = torch.where(
        inter > 0,
        inter / (area1[:, None] + area2 - inter),
        torch.zeros(1, dtype=inter.dtype, device=inter.device),
    )
    return iou


class ImageList:
    """Padded batch + true sizes; iterating yields un-padded views."""

    def __init__(self, tensor, image_sizes):
        self.tensor = tensor
        self.image_sizes = [tuple(s) for s in image_sizes]

    @classmethod
    def from_tensors(cls, tensors):
        sizes = [tuple(t.shape[-2:]) for t in tensors]
        hm = max(s[0] for s in sizes)
        wm = max(s[1] for s in sizes)
        out = tensors[0].new_zeros((len(tensors), hm, wm))
        for i, t in enumerate(tensors):
            out[i, : t.shape[-2], : t.shape[-1]] = t
        return cls(out, sizes)

    def __len__(self):
        return len(self.image_sizes)

    def __getitem__(self, idx):
        h, w = self.image_sizes[idx]
        return self.tensor[idx, ..., :h, :w]

    def __iter__(self):
        for i in range(len(self)):
            yield self[i]

    @property
    def device(self):
        return self.tensor.device


def axis_angle_to_matrix(axis_angle: torch.Tensor) -> torch.Tensor:
    """Rodrigues via quaternion, as pytorch3d.transforms.axis_angle_to_matrix
    (axis_angle -> quaternion -> matrix)."""
    angles = torch.norm(axis_angle, p=2, dim=-1, keepdim=True)
    half = angles * 0.5
    eps = 1e-6
    small = angles.abs() < eps
    s = torch.empty_like(angles)
    s[~small] = torch.sin(half[~small]) / angles[~small]
    s[small] = 0.5 - (angles[small] * angles[small]) / 48
    q = torch.cat([torch.cos(half), axis_angle * s], dim=-1)
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack(
        (
            1 - two_s * (j * j + k * k),
            two_s * (i * j - k * r),
            two_s * (i * k + j * r),
            two_s * (i * j + k * r),
            1 - two_s * (i * i + k * k),
            two_s * (j * k - i * r),
            two_s * (i * k - j * r),
            two_s * (j * k + i * r),
            1 - two_s * (i * i + j * j),
        ),
        -1,
    )
    return o.reshape(q.shape[:-1] + (3, 3))


def rotation_6d_to_matrix(d6: torch.Tensor) -> torch.Tensor:
    """Gram-Schmidt of two 3-vectors (Zhou et al.), rows = b1, b2, b3."""
    import torch.nn.functional as F

    a1, a2 = d6[..., :3], d6[..., 3:]
    b1 = F.normalize(a1, dim=-1)
    b2 = a2 - (b1 * a2).sum(-1, keepdim=True) * b1
    b2 = F.normalize(b2, dim=-1)
    b3 = torch.cross(b1, b2, dim=-1)
    return torch.stack((b1, b2, b3), dim=-2)


def so3_relative_angle(R1, R2, cos_angle: bool = False, cos_bound: float = 1e-4, eps: float = 1e-4):
    """pytorch3d.transforms.so3.so3_relative_angle, cos_angle=True branch (the only one the
    hot path uses, roi_heads.py:1006): cos = (tr(R1 R2^T) - 1) / 2 with the trace-range check."""
    R12 = torch.bmm(R1, R2.permute(0, 2, 1))
    tr = R12[:, 0, 0] + R12[:, 1, 1] + R12[:, 2, 2]
    if ((tr < -1.0 - eps) + (tr > 3.0 + eps)).any():
        raise ValueError("A matrix has trace outside valid range [-1-eps,3+eps].")
    phi_cos = (tr - 1.0) * 0.5
    if cos_angle:
        return phi_cos
    raise NotImplementedError("stand-in covers cos_angle=True only")


class _EventSink:
    iter = 1

    def put_scalar(self, *a, **k):
        pass


def get_event_storage():
    return _EventSink()
