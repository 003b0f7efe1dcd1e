"""Synthetic SUN RGB-D-shaped inputs. Seeded, host-side, torch-CPU.

There is no dataset in this environment; bench.py and the tests use these.
Image i uses ``torch.Generator().manual_seed(1234 + i)``.
"""
from __future__ import annotations

import math

import torch


def make_image(i: int, H: int = 480, W: int = 640, n_boxes: int | None = None,
               dummy_ground: bool | None = None, n_classes: int = 50):
    """One synthetic indoor frame.

    depth  (H,W) f32 metres: floor plane below the horizon at camera height
           h~U[1.0,1.6], back wall at z~U[3,8], 3-8 box-shaped bumps, clamp
           [0.3,20] (Depth-Anything indoor max_depth=20), noise N(1,0.01).
    K      fx=fy=525*W/640, cx=W/2, cy=H/2.
    ground (H,W) u8 floor mask, or the (1,1) dummy for ~10% of images.
    boxes  (b,4) f32 x1,y1,x2,y2 fully inside the image, w~U[40,0.6W], h~U[40,0.6H].
    classes (b,) int64 in [0,50).  seg (b,H,W) is produced lazily by seg_masks().
    """
    g = torch.Generator().manual_seed(1234 + i)
    f = 525.0 * W / 640.0
    cam_h = 1.0 + 0.6 * torch.rand(1, generator=g).item()
    wall = 3.0 + 5.0 * torch.rand(1, generator=g).item()
    v = torch.arange(H, dtype=torch.float32)[:, None].expand(H, W)
    u = torch.arange(W, dtype=torch.float32)[None, :].expand(H, W)
    floor_z = cam_h * f / (v - H / 2).clamp(min=1e-3)
    below = v > H / 2 + 2
    depth = torch.where(below, torch.minimum(floor_z, torch.full_like(v, wall)),
                        torch.full_like(v, wall))
    is_floor = below & (floor_z < wall)
    n_bumps = int(torch.randint(3, 9, (1,), generator=g).item())
    for _ in range(n_bumps):
        bw = int(torch.randint(30, max(31, W // 4), (1,), generator=g).item())
        bh = int(torch.randint(30, max(31, H // 3), (1,), generator=g).item())
        bw, bh = min(bw, W - 1), min(bh, H - H // 4 - 1)
        x0 = int(torch.randint(0, W - bw, (1,), generator=g).item())
        y0 = int(torch.randint(H // 4, H - bh, (1,), generator=g).item())
        zb = depth[min(y0 + bh, H - 1), x0 + bw // 2].item()  # stands on what is below it
        zb = max(0.5, zb - 0.2 * torch.rand(1, generator=g).item())
        patch = depth[y0:y0 + bh, x0:x0 + bw]
        closer = patch > zb
        patch[closer] = zb
        is_floor[y0:y0 + bh, x0:x0 + bw] &= ~closer
    depth = depth * (1.0 + 0.01 * torch.randn(H, W, generator=g))
    depth = depth.clamp(0.3, 20.0).contiguous()
    K = torch.tensor([[f, 0.0, W / 2.0], [0.0, f, H / 2.0], [0.0, 0.0, 1.0]], dtype=torch.float32)
    if dummy_ground is None:
        dummy_ground = torch.rand(1, generator=g).item() < 0.1
    ground = torch.ones((1, 1), dtype=torch.uint8) if dummy_ground else is_floor.to(torch.uint8)
    if n_boxes is None:
        n_boxes = int(torch.poisson(torch.tensor([20.0]), generator=g).clamp(1, 64).item())
    bw = 40 + (0.6 * W - 40) * torch.rand(n_boxes, generator=g)
    bh = 40 + (0.6 * H - 40) * torch.rand(n_boxes, generator=g)
    x1 = (W - bw) * torch.rand(n_boxes, generator=g)
    y1 = (H - bh) * torch.rand(n_boxes, generator=g)
    boxes = torch.stack((x1, y1, x1 + bw, y1 + bh), dim=1).float()
    classes = torch.randint(0, n_classes, (n_boxes,), generator=g)
    return {"depth": depth, "K": K, "ground": ground, "boxes": boxes, "classes": classes,
            "H": H, "W": W}


def make_priors(n_classes: int = 50, seed: int = 99):
    """(C,2,3): [[w,h,l] mean ~U[0.2,2.0] m, std = mean*U[0.15,0.3]]
    (stand-in for priors['priors_dims_per_cat'], math_util.py:419)."""
    g = torch.Generator().manual_seed(seed)
    mean = 0.2 + 1.8 * torch.rand(n_classes, 3, generator=g)
    std = mean * (0.15 + 0.15 * torch.rand(n_classes, 3, generator=g))
    return torch.stack((mean, std), dim=1).float()


def seg_masks(boxes: torch.Tensor, H: int, W: int):
    """(b,H,W) u8: ellipse inscribed in each GT box (stand-in for SAM masks)."""
    v = torch.arange(H, dtype=torch.float32)[None, :, None]
    u = torch.arange(W, dtype=torch.float32)[None, None, :]
    cx = ((boxes[:, 0] + boxes[:, 2]) * 0.5)[:, None, None]
    cy = ((boxes[:, 1] + boxes[:, 3]) * 0.5)[:, None, None]
    rx = ((boxes[:, 2] - boxes[:, 0]) * 0.5)[:, None, None]
    ry = ((boxes[:, 3] - boxes[:, 1]) * 0.5)[:, None, None]
    return ((((u - cx) / rx) ** 2 + ((v - cy) / ry) ** 2) <= 1.0).to(torch.uint8)


def make_batch(n_images: int, H: int = 480, W: int = 640, n_boxes: int | None = None,
               first: int = 0):
    return [make_image(first + i, H, W, n_boxes) for i in range(n_images)]
