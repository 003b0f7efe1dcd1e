"""Diagnostic (GPU box): how much of the point-test work at C2 belongs to proposals that an upper
bound of their score rules out of the top-k before a single point is tested?

Per box (a sample of them), in torch: the proposals (wcr.proposals.propose, the generator the
scoring kernel uses), their conservative cull box + depth range (the arithmetic of
score_prep_kernel), and two upper bounds of the point count: `cell` = valid points of the 16 x 16
ROI cells the cull box overlaps whose depth range meets the cube's; `rect` = valid points inside the
cull box with a depth in the cube's range. A proposal survives when bound-score >= T, T the k-th
best true score of a phase A = the PA proposals with the largest bound. Prints the share of
proposals and of point-test work (cull-box area) that remains, phase A included."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import numpy as np
import torch
import bench
from wcr import synth
from wcr.pipeline import ShardRunner
from wcr.proposals import ScoreConfig, score_topk, box_center_depth, propose, box_rois

n_img = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 64
every = 8                                       # study every 8th box
ims = bench.make_images(n_img, 0, True, None, dense=False)
priors = synth.make_priors()
K, N = 16, 1000
r = ShardRunner(ims, N, "cuda", priors=priors, score_cfg=ScoreConfig(k=K, e_seg=1), seed=1234)
r.upload(); r.step(); torch.cuda.synchronize()
z0 = box_center_depth(r.shard, r.boxes)
outs = {}
for use_seg in (True, False):
    outs[use_seg] = score_topk(r.shard, r.boxes, None, N, ScoreConfig(k=K, e_seg=1 if use_seg else 0),
                               priors=r.priors, normals=r.normals, z0=z0, seed=1234,
                               seg=r.seg if use_seg else None, want_all=True)
cubes = propose(r.shard, r.boxes, r.priors, r.normals, N, z0=z0, seed=1234)
rois = box_rois(r.boxes, r.shard)
h = r.boxes.host
tab = r.shard.host_table
SGN = torch.tensor([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1],
                    [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=torch.float32, device="cuda")
acc = {}
def add(key, num, den):
    a = acc.setdefault(key, [0.0, 0.0]); a[0] += float(num); a[1] += float(den)

for b in range(0, len(h), every):
    img = int(h["image"][b]); u0, v0, u1, v1 = [int(x) for x in rois[b]]
    if u1 <= u0 or v1 <= v0:
        continue
    fx, fy, cx, cy = [float(tab[k][img]) for k in ("fx", "fy", "cx", "cy")]
    D = r.shard.depths[img][v0:v1, u0:u1]
    valid = (D > 0) & torch.isfinite(D)
    nrows, ncols = D.shape
    cw, chh = max(1, (ncols + 15) >> 4), max(1, (nrows + 15) >> 4)
    rr = torch.arange(nrows, device="cuda")[:, None].expand(nrows, ncols)
    cc = torch.arange(ncols, device="cuda")[None, :].expand(nrows, ncols)
    cell = (rr // chh) * 16 + (cc // cw)
    ccnt = torch.zeros(256, device="cuda").index_add_(0, cell[valid], torch.ones(int(valid.sum()), device="cuda"))
    zmin = torch.full((256,), float("inf"), device="cuda").scatter_reduce_(0, cell[valid], D[valid], "amin")
    zmax = torch.full((256,), float("-inf"), device="cuda").scatter_reduce_(0, cell[valid], D[valid], "amax")
    c = cubes[b]
    half = torch.stack([c[:, 5], c[:, 4], c[:, 3]], 1) / 2
    R = c[:, 6:15].reshape(N, 3, 3)
    loc = SGN[None] * half[:, None]                                  # (N, 8, 3)
    P = torch.einsum("nij,nkj->nki", R, loc) + c[:, None, 0:3]       # (N, 8, 3)
    Z = P[..., 2]
    px = (cx * Z + fx * P[..., 0]) / Z; py = (cy * Z + fy * P[..., 1]) / Z
    live = (c[:, 3] > 0) & (c[:, 4] > 0) & (c[:, 5] > 0) & (Z.min(1).values >= 0.01) & \
           torch.isfinite(px).all(1) & torch.isfinite(py).all(1)
    x_lo = torch.floor(px.min(1).values.clamp(-1e6, 1e6)) - 1 - u0
    x_hi = torch.ceil(px.max(1).values.clamp(-1e6, 1e6)) + 1 - u0
    y_lo = torch.floor(py.min(1).values.clamp(-1e6, 1e6)) - 1 - v0
    y_hi = torch.ceil(py.max(1).values.clamp(-1e6, 1e6)) + 1 - v0
    cl = x_lo.clamp(min=0); ch = x_hi.clamp(max=ncols - 1); rl = y_lo.clamp(min=0); rh = y_hi.clamp(max=nrows - 1)
    slack = 1e-3 * (Z.min(1).values.abs() + Z.max(1).values.abs()) + 1e-4
    zlo, zhi = Z.min(1).values - slack, Z.max(1).values + slack
    # unlive proposals are never culled by the kernel; keep them out of the study (bound = all points)
    reach = (cl <= ch) & (rl <= rh)
    gx = torch.arange(16, device="cuda")
    ox = (gx[None] >= (cl // cw)[:, None]) & (gx[None] <= (ch // cw).clamp(max=15)[:, None])      # (N,16)
    oy = (gx[None] >= (rl // chh)[:, None]) & (gx[None] <= (rh // chh).clamp(max=15)[:, None])
    oz = (zmax[None] >= zlo[:, None]) & (zmin[None] <= zhi[:, None])                                  # (N,256)
    over = (oy[:, :, None] & ox[:, None, :]).reshape(N, 256) & oz
    ub_cell = (over.float() * ccnt[None]).sum(1)
    npts = float(valid.sum())
    ub_cell = torch.where(live, torch.where(reach, ub_cell, torch.zeros_like(ub_cell)), torch.full_like(ub_cell, npts))
    # rect bound on a pixel subsample (every 2nd row / column) scaled back: study only
    Ds, vs = D[::2, ::2], valid[::2, ::2]
    rs = torch.arange(0, nrows, 2, device="cuda"); cs = torch.arange(0, ncols, 2, device="cuda")
    inr = (rs[None] >= rl[:, None]) & (rs[None] <= rh[:, None])       # (N, R)
    inc = (cs[None] >= cl[:, None]) & (cs[None] <= ch[:, None])       # (N, C)
    ub_rect = torch.zeros(N, device="cuda")
    for s in range(0, N, 100):
        e = s + 100
        inz = vs[None] & (Ds[None] >= zlo[s:e, None, None]) & (Ds[None] <= zhi[s:e, None, None])
        ub_rect[s:e] = (inz & inr[s:e, :, None] & inc[s:e, None, :]).flatten(1).sum(1).float() * 4.0
    ub_rect = torch.where(live, torch.where(reach, ub_rect * 1.05 + 4, torch.zeros_like(ub_rect)), torch.full_like(ub_rect, npts))
    work = torch.where(live & reach, (ch - cl + 1) * (rh - rl + 1), torch.zeros_like(cl)).clamp(min=0)
    work = torch.where(live, work, torch.full_like(work, float(nrows * ncols)))
    for use_seg in (True, False):
        o = outs[use_seg]
        sc = torch.nan_to_num(o["scores"][b], nan=0.0)
        iou = torch.nan_to_num(o["iou"][b], nan=1.0)
        cnt = o["count"][b].float()
        tag = "seg" if use_seg else "noseg"
        add(f"{tag}/viol_cell", (cnt > ub_cell + 0.5).sum(), N)
        for name, ub in (("cell", ub_cell), ("rect", ub_rect), ("oracle1.0", cnt)):
            ubs = iou * (ub / max(npts, 1.0)).clamp(max=1)
            for PA in (64, 128, 256):
                order = ubs.argsort(descending=True)
                inA = torch.zeros(N, dtype=torch.bool, device="cuda"); inA[order[:PA]] = True
                TA = sc[inA].topk(K).values[-1]
                keepB = (~inA) & (ubs >= TA)
                add(f"{tag}/{name}/PA{PA}/count", inA.sum() + keepB.sum(), N)
                add(f"{tag}/{name}/PA{PA}/work", work[inA].sum() + work[keepB].sum(), work.sum())
                add(f"{tag}/{name}/PA{PA}/workA", work[inA].sum(), work.sum())
print(json.dumps({k: round(v[0] / max(v[1], 1e-9), 4) for k, v in acc.items()}, indent=0))
