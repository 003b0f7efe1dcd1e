"""TEST INFRASTRUCTURE ONLY (oracle). Seeded synthetic inputs of the weak-loss
block: head outputs at init-like scale, proposal boxes
jittered around GT boxes, per-image depth maps and up vectors."""
from __future__ import annotations

import torch


def make_loss_inputs(seed: int, boxes_per_image, H=480, W=640, n_classes=50, big=False):
    g = torch.Generator().manual_seed(seed)
    B = len(boxes_per_image)
    n = sum(boxes_per_image)
    r = lambda *s: torch.rand(*s, generator=g)
    rn = lambda *s: torch.randn(*s, generator=g)
    bw = 40 + (0.5 * W - 40) * r(n)
    bh = 40 + (0.5 * H - 40) * r(n)
    x1 = (W - bw) * r(n)
    y1 = (H - bh) * r(n)
    gt = torch.stack((x1, y1, x1 + bw, y1 + bh), 1)
    src = gt + rn(n, 4) * 4.0
    f = 525.0 * W / 640.0
    ratio = 0.8 + 0.4 * r(B)                       # im_scales_ratio per image
    per_img = lambda t: torch.cat([t[i:i + 1].expand(b, *t.shape[1:]) for i, b in enumerate(boxes_per_image)])
    K4 = per_img(torch.stack((f / ratio, f / ratio, (W / 2) / ratio, (H / 2) / ratio), 1)).contiguous()
    v2r = per_img((0.8 + 0.4 * r(B))[:, None])[:, 0].contiguous()
    mean = 0.2 + 1.8 * r(n_classes, 3)
    std = mean * (0.15 + 0.15 * r(n_classes, 3))
    cls = torch.randint(0, n_classes, (n,), generator=g)
    scale = 1.0 if big else 0.05
    deltas = rn(n, 2) * scale * 2
    z_raw = 1.0 + r(n) * 5.0
    dims_raw = rn(n, 3) * (1.0 if big else 0.2)
    if big:
        dims_raw[::7, 0] = 5.5                       # exercises clip(max=5)
    q, _ = torch.linalg.qr(rn(n, 3, 3))
    pose = q * torch.sign(torch.linalg.det(q))[:, None, None]
    uncert = 0.01 + r(n) * (5.0 if not big else 1.0)
    depth_maps = [(0.5 + 6 * r(H, W)).contiguous() for _ in range(B)]
    normals = torch.nn.functional.normalize(torch.tensor([[0.0, 1.0, 0.0]]) + 0.1 * rn(B, 3), dim=1)
    conf = torch.where(r(B) < 0.3, torch.tensor(0.1), torch.tensor(1.0))
    return dict(deltas=deltas, z_raw=z_raw, dims_raw=dims_raw, pose=pose.contiguous(), uncert=uncert,
                src_boxes=src, K4=K4, virtual_to_real=v2r, prior_mean=mean[cls].contiguous(),
                prior_std=std[cls].contiguous(), gt_boxes=gt,
                clamp=torch.tensor([[H, W]]).repeat(n, 1), num_boxes_per_image=list(boxes_per_image),
                depth_maps=depth_maps, normals=normals, ground_conf=conf)


GRAD_KEYS = ("deltas", "z_raw", "dims_raw", "pose", "uncert")


def run_with_grads(fn, cfg, inp):
    """-> (losses dict of floats, per_box dict, grads dict) with d(sum of losses)."""
    leaves = {k: inp[k].clone().requires_grad_(True) for k in GRAD_KEYS}
    rest = {k: v for k, v in inp.items() if k not in GRAD_KEYS}
    losses, per_box = fn(cfg, **leaves, **rest)
    total = sum(losses.values())
    total.backward()
    return ({k: v.detach() for k, v in losses.items()}, per_box,
            {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in leaves.items()})


def add_masks(inp: dict, seed: int, H=480, W=640, share_every=4, empty_every=7) -> dict:
    """Per-object GT masks for the tier-2 terms ('segmentation', 'depth'): one filled ellipse
    inside each GT box as a (1, H, W) bool tensor; every ``share_every``-th box re-uses
    the previous box's mask (several proposals matched to one object,
    roi_heads.py:262-270) and every ``empty_every``-th mask is empty (segmentation failed:
    depth_range_loss falls back to the GT box, roi_heads.py:674-675)."""
    g = torch.Generator().manual_seed(seed + 991)
    gt = inp["gt_boxes"]
    n = gt.shape[0]
    ys = torch.arange(H, dtype=torch.float32)[:, None]
    xs = torch.arange(W, dtype=torch.float32)[None, :]
    masks, idx = [], []
    for i in range(n):
        if i % share_every == share_every - 1 and masks:
            idx.append(len(masks) - 1)
            continue
        x1, y1, x2, y2 = gt[i].tolist()
        cx, cy = 0.5 * (x1 + x2), 0.5 * (y1 + y2)
        rx = 0.5 * (x2 - x1) * (0.5 + 0.5 * torch.rand((), generator=g).item())
        ry = 0.5 * (y2 - y1) * (0.5 + 0.5 * torch.rand((), generator=g).item())
        m = ((xs - cx) / rx) ** 2 + ((ys - cy) / ry) ** 2 <= 1.0
        if len(masks) % empty_every == empty_every - 1:
            m = torch.zeros_like(m)
        masks.append(m[None])
        idx.append(len(masks) - 1)
    out = dict(inp)
    out["gt_masks"] = masks
    out["at_which_mask_idx"] = idx
    return out
