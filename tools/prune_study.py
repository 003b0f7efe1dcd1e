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
every = 16                                      # study every 16th box
ims = bench.make_images(n_img, 0, True, None, dense="both")
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
all_masks = [m for im in ims for m in im["seg"]]
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
    # slab bound: a cell (pixel rectangle x depth range) against the cube's own three slab pairs,
    # interval arithmetic on q_k = z * (A_k a + B_k b + C_k) - O_k, inside iff |q_k| <= 1
    cyi, cxi = torch.arange(256, device="cuda") // 16, torch.arange(256, device="cuda") % 16
    a_lo = ((u0 + cxi * cw).float() - cx) / fx; a_hi = ((u0 + torch.minimum((cxi + 1) * cw, torch.tensor(ncols, device="cuda")) - 1).float() - cx) / fx
    b_lo = ((v0 + cyi * chh).float() - cy) / fy; b_hi = ((v0 + torch.minimum((cyi + 1) * chh, torch.tensor(nrows, device="cuda")) - 1).float() - cy) / fy
    Rm = c[:, 6:15].reshape(N, 3, 3)
    Ak, Bk, Ck = Rm[:, 0, :] / half, Rm[:, 1, :] / half, Rm[:, 2, :] / half          # (N,3)
    Ok = (Ak * c[:, None, 0] + Bk * c[:, None, 1] + Ck * c[:, None, 2])                 # (N,3)
    def iv(coef, lo, hi):      # (N,3) x (256,) -> (N,256,3) interval of coef * x
        p1, p2 = coef[:, None, :] * lo[None, :, None], coef[:, None, :] * hi[None, :, None]
        return torch.minimum(p1, p2), torch.maximum(p1, p2)
    al, ah = iv(Ak, a_lo, a_hi); bl, bh = iv(Bk, b_lo, b_hi)
    g_lo = al + bl + Ck[:, None, :]; g_hi = ah + bh + Ck[:, None, :]
    zl, zh = zmin[None, :, None], zmax[None, :, None]
    q_lo = torch.minimum(zl * g_lo, zh * g_lo) - Ok[:, None, :]
    q_hi = torch.maximum(zl * g_hi, zh * g_hi) - Ok[:, None, :]
    slab_ok = ((q_lo <= 1.001) & (q_hi >= -1.001)).all(2)                                 # (N,256)
    ub_slab = ((over & slab_ok).float() * ccnt[None]).sum(1)
    # variants: two axes only (the cube's x and z axes), and 2 x 2 blocks of cells
    slab2_ok = ((q_lo <= 1.001) & (q_hi >= -1.001))[:, :, [0, 2]].all(2)
    ub_slab2 = ((over & slab2_ok).float() * ccnt[None]).sum(1)
    def blk(x, red):   # (256,) -> (64,) over 2 x 2 blocks
        x4 = x.reshape(8, 2, 8, 2)
        return red(red(x4, 3), 1).reshape(64)
    mn = lambda t, d: t.min(d).values
    mx = lambda t, d: t.max(d).values
    zmin8, zmax8 = blk(zmin, mn), blk(zmax, mx)
    a_lo8, a_hi8, b_lo8, b_hi8 = blk(a_lo, mn), blk(a_hi, mx), blk(b_lo, mn), blk(b_hi, mx)
    cnt8 = blk(ccnt, lambda t, d: t.sum(d))
    over8 = over.reshape(N, 8, 2, 8, 2).any(4).any(2).reshape(N, 64)
    al8, ah8 = iv(Ak, a_lo8, a_hi8); bl8, bh8 = iv(Bk, b_lo8, b_hi8)
    g_lo8 = al8 + bl8 + Ck[:, None, :]; g_hi8 = ah8 + bh8 + Ck[:, None, :]
    zl8 = torch.maximum(zmin8[None, :, None], zlo[:, None, None]); zh8 = torch.minimum(zmax8[None, :, None], zhi[:, None, None])
    q_lo8 = torch.minimum(zl8 * g_lo8, zh8 * g_lo8) - Ok[:, None, :]
    q_hi8 = torch.maximum(zl8 * g_hi8, zh8 * g_hi8) - Ok[:, None, :]
    slab8_ok = ((q_lo8 <= 1.001) & (q_hi8 >= -1.001)).all(2)
    # block passes -> its z-overlapping cells count
    pass8 = (over8 & slab8_ok).reshape(N, 8, 1, 8, 1).expand(N, 8, 2, 8, 2).reshape(N, 256)
    ub_slab8 = ((over & pass8).float() * ccnt[None]).sum(1)
    add("cost/blocks8_tested_per_proposal", over8.float().sum(), N)
    for nm_, u_ in (("ub_slab2", ub_slab2), ("ub_slab8", ub_slab8)):
        u_[:] = torch.where(live, torch.where(reach, u_, torch.zeros_like(u_)), torch.full_like(u_, float(valid.sum())))
    ub_slab = torch.where(live, torch.where(reach, ub_slab, torch.zeros_like(ub_slab)), torch.full_like(ub_slab, float(valid.sum())))
    # executed-work proxy under tile culling by cull box + depth range vs by the slab mask
    lr = live & reach
    add("cull/slab_over_bbox_cells", ((over & slab_ok).float().sum(1) * lr).sum(), (over.float().sum(1) * lr).sum())
    oxy = (oy[:, :, None] & ox[:, None, :]).reshape(N, 256)
    add("cull/slab_over_bboxonly_cells", ((over & slab_ok).float().sum(1) * lr).sum(), (oxy.float().sum(1) * lr).sum())
    add("cull/bbox_cells_per_proposal", (oxy.float().sum(1) * lr).sum(), lr.sum())
    add("cull/zpass_cells_per_proposal", (over.float().sum(1) * lr).sum(), lr.sum())
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
    # mask-term bound: ratio <= B / M when the hull's bounding box holds B < M pixels (M = mask area)
    bits = r.seg.bits[int(r.seg.off[b]):int(r.seg.off[b + 1])]
    M = float(sum(bin(int(x) & 0xffffffff).count("1") for x in bits.cpu().tolist()))
    pj = outs[True]["proj"][b]
    Bpix = ((torch.floor(pj[:, 2]) - torch.ceil(pj[:, 0]) + 3).clamp(min=0) * (torch.floor(pj[:, 3]) - torch.ceil(pj[:, 1]) + 3).clamp(min=0))
    ub_seg = torch.nan_to_num((Bpix / max(M, 1.0)).clamp(max=1.0), nan=1.0)
    Himg, Wimg = r.shard.depths[img].shape
    c0_, c1_ = float(h["clamp0"][b]), float(h["clamp1"][b])               # the reference clamps x with clamp0, y with clamp1
    qx = px.clamp(0, float(int(c0_ - 1))); qy = py.clamp(0, float(int(c1_ - 1)))   # (N, 8) clamped corners
    FACES = [[0, 1, 2, 3], [4, 5, 6, 7], [0, 1, 5, 4], [1, 2, 6, 5], [2, 3, 7, 6], [3, 0, 4, 7]]
    Af = torch.zeros(N, device="cuda")
    for f in FACES:
        x_, y_ = qx[:, f], qy[:, f]
        sa = 0.5 * (x_ * y_.roll(-1, 1) - x_.roll(-1, 1) * y_).sum(1)
        Af = torch.maximum(Af, sa.abs())
    bw = qx.max(1).values - qx.min(1).values + 2; bh = qy.max(1).values - qy.min(1).values + 2
    outside = bw * (qy.max(1).values - (Himg - 1)).clamp(min=0) + bh * (qx.max(1).values - (Wimg - 1)).clamp(min=0)
    Blow = (Af * 0.999 - outside * 1.001 - 2 * (bw + 16 + bh + 16) - 16).clamp(min=0)
    # exact hull area of the clamped corners (scipy), same margins
    from scipy.spatial import ConvexHull
    qxc, qyc = qx.cpu().numpy(), qy.cpu().numpy()
    Ah = np.zeros(N, dtype=np.float32)
    for n_ in range(N):
        try:
            Ah[n_] = ConvexHull(np.stack([qxc[n_], qyc[n_]], 1)).volume
        except Exception:
            Ah[n_] = 0.0
    Ah = torch.from_numpy(Ah).cuda()
    Blow_h = (Ah * 0.999 - outside * 1.001 - 2 * (bw + 16 + bh + 16) - 16).clamp(min=0)
    # mask pixels inside the hull's bounding box (integral image of the dense mask)
    dm = torch.as_tensor(all_masks[b]).cuda().float()
    integ = torch.zeros((Himg + 1, Wimg + 1), device="cuda"); integ[1:, 1:] = (dm[:Himg, :Wimg] != 0).float().cumsum(0).cumsum(1)
    # inter <= pixels of (hull bbox + margin) inside the ROI
    ix0 = torch.maximum(torch.floor(qx.min(1).values) - 1, torch.tensor(float(u0), device="cuda"))
    ix1 = torch.minimum(torch.ceil(qx.max(1).values) + 1, torch.tensor(float(u1 - 1), device="cuda"))
    iy0 = torch.maximum(torch.floor(qy.min(1).values) - 1, torch.tensor(float(v0), device="cuda"))
    iy1 = torch.minimum(torch.ceil(qy.max(1).values) + 1, torch.tensor(float(v1 - 1), device="cuda"))
    Iub = ((ix1 - ix0 + 1).clamp(min=0) * (iy1 - iy0 + 1).clamp(min=0))
    istar = torch.minimum(Iub, torch.tensor(M, device="cuda"))
    ub_seg2 = torch.nan_to_num(istar / (M + torch.maximum(Blow, istar) - istar).clamp(min=1), nan=1.0).clamp(max=1)
    ub_seg2 = torch.where(live, ub_seg2, torch.ones_like(ub_seg2))
    ub_seg3 = torch.nan_to_num(istar / (M + torch.maximum(Blow_h, istar) - istar).clamp(min=1), nan=1.0).clamp(max=1)
    ub_seg3 = torch.where(live, ub_seg3, torch.ones_like(ub_seg3))
    nn = lambda t: torch.nan_to_num(t, nan=0.0, posinf=0.0, neginf=0.0)
    jx0, jx1 = nn(ix0).clamp(0, Wimg - 1).long(), (nn(ix1).clamp(-1, Wimg - 1) + 1).long()
    jy0, jy1 = nn(iy0).clamp(0, Himg - 1).long(), (nn(iy1).clamp(-1, Himg - 1) + 1).long()
    Imask = (integ[jy1, jx1] - integ[jy0, jx1] - integ[jy1, jx0] + integ[jy0, jx0]).clamp(min=0)
    Imask = torch.where((ix1 >= ix0) & (iy1 >= iy0), Imask, torch.zeros_like(Imask))
    istar4 = torch.minimum(Imask, torch.tensor(M, device="cuda"))
    ub_seg4 = torch.nan_to_num(istar4 / (M + torch.maximum(Blow_h, istar4) - istar4).clamp(min=1), nan=0.0).clamp(max=1)
    ub_seg4 = torch.where(live, ub_seg4, torch.ones_like(ub_seg4))
    # cheap variants: extreme-point quadrilateral as the area bound; mask counts per cell of a 16 x 16 grid
    ar = torch.arange(N, device="cuda")
    iL, iR, iB, iT = qx.argmin(1), qx.argmax(1), qy.argmin(1), qy.argmax(1)
    ex = torch.stack([qx[ar, iL], qx[ar, iB], qx[ar, iR], qx[ar, iT]], 1); ey = torch.stack([qy[ar, iL], qy[ar, iB], qy[ar, iR], qy[ar, iT]], 1)
    Aq = (0.5 * (ex * ey.roll(-1, 1) - ex.roll(-1, 1) * ey).sum(1)).abs()
    Blow_q = (torch.maximum(Af, Aq) * 0.999 - outside * 1.001 - 2 * (bw + 16 + bh + 16) - 16).clamp(min=0)
    nr1, nc1 = v1 - v0, u1 - u0
    mcw, mch = max(1, (nc1 + 15) >> 4), max(1, (nr1 + 15) >> 4)
    roi_m = (dm[v0:v1, u0:u1] != 0).float()
    pad = torch.zeros((16 * mch, 16 * mcw), device="cuda"); pad[:nr1, :nc1] = roi_m
    cellm = pad.reshape(16, mch, 16, mcw).sum(3).sum(1)                      # (16,16) mask pixels per cell
    integc = torch.zeros((17, 17), device="cuda"); integc[1:, 1:] = cellm.cumsum(0).cumsum(1)
    cx0 = ((nn(ix0) - u0).clamp(0, nc1 - 1) // mcw).long().clamp(0, 15); cx1 = ((nn(ix1) - u0).clamp(0, nc1 - 1) // mcw).long().clamp(0, 15) + 1
    cy0 = ((nn(iy0) - v0).clamp(0, nr1 - 1) // mch).long().clamp(0, 15); cy1 = ((nn(iy1) - v0).clamp(0, nr1 - 1) // mch).long().clamp(0, 15) + 1
    Icell = integc[cy1, cx1] - integc[cy0, cx1] - integc[cy1, cx0] + integc[cy0, cx0]
    Icell = torch.where((ix1 >= ix0) & (iy1 >= iy0), Icell, torch.zeros_like(Icell))
    istar5 = torch.minimum(torch.minimum(Icell, Iub), torch.tensor(M, device="cuda"))
    ub_seg5 = torch.nan_to_num(istar5 / (M + torch.maximum(Blow_q, istar5) - istar5).clamp(min=1), nan=0.0).clamp(max=1)
    ub_seg5 = torch.where(live, ub_seg5, torch.ones_like(ub_seg5))
    ub_seg6 = torch.nan_to_num(istar5 / (M + torch.maximum(Blow, istar5) - istar5).clamp(min=1), nan=0.0).clamp(max=1)
    ub_seg6 = torch.where(live, ub_seg6, torch.ones_like(ub_seg6))
    extra = {}
    for G in (32, 64):
        gw, gh = max(1, (nc1 + G - 1) // G), max(1, (nr1 + G - 1) // G)
        padg = torch.zeros((G * gh, G * gw), device="cuda"); padg[:nr1, :nc1] = roi_m
        cm = padg.reshape(G, gh, G, gw).sum(3).sum(1)
        ig = torch.zeros((G + 1, G + 1), device="cuda"); ig[1:, 1:] = cm.cumsum(0).cumsum(1)
        gx0 = ((nn(ix0) - u0).clamp(0, nc1 - 1) // gw).long().clamp(0, G - 1); gx1 = ((nn(ix1) - u0).clamp(0, nc1 - 1) // gw).long().clamp(0, G - 1) + 1
        gy0 = ((nn(iy0) - v0).clamp(0, nr1 - 1) // gh).long().clamp(0, G - 1); gy1 = ((nn(iy1) - v0).clamp(0, nr1 - 1) // gh).long().clamp(0, G - 1) + 1
        Ig = ig[gy1, gx1] - ig[gy0, gx1] - ig[gy1, gx0] + ig[gy0, gx0]
        Ig = torch.where((ix1 >= ix0) & (iy1 >= iy0), Ig, torch.zeros_like(Ig))
        ist = torch.minimum(torch.minimum(Ig, Iub), torch.tensor(M, device="cuda"))
        u_ = torch.nan_to_num(ist / (M + torch.maximum(Blow, ist) - ist).clamp(min=1), nan=0.0).clamp(max=1)
        extra[f"G{G}"] = torch.where(live, u_, torch.ones_like(u_))
    u_ = torch.nan_to_num(istar4 / (M + torch.maximum(Blow, istar4) - istar4).clamp(min=1), nan=0.0).clamp(max=1)
    extra["exactI_maxface"] = torch.where(live, u_, torch.ones_like(u_))
    for nm_, u_ in (("3", ub_seg3), ("4", ub_seg4), ("5", ub_seg5), ("6", ub_seg6)):
        add("seg/viol_ubseg" + nm_, (outs[True]["seg"][b] > u_ * (1 + 1e-6)).sum(), N)
    vio = outs[True]["seg"][b] > ub_seg2 * (1 + 1e-6)
    if bool(vio.any()) and acc.get("dbg", [0])[0] < 6:
        acc.setdefault("dbg", [0.0, 1.0])[0] += 1
        k_ = int(vio.nonzero()[0])
        print("VIOL box", b, "prop", k_, "seg", float(outs[True]["seg"][b, k_]), "ub", float(ub_seg2[k_]), "Iub", float(Iub[k_]),
              "Blow", float(Blow[k_]), "Af", float(Af[k_]), "M", M, "bw,bh", float(bw[k_]), float(bh[k_]),
              "roi", (u0, v0, u1, v1), "qx", [round(float(t), 1) for t in qx[k_]], "qy", [round(float(t), 1) for t in qy[k_]],
              "clamp", float(h["clamp0"][b]), float(h["clamp1"][b]), "img", (Himg, Wimg), file=sys.stderr)
    add("seg/viol_ubseg2", vio.sum(), N)
    add("seg/mean_ubseg2", ub_seg2.sum(), N)
    add("seg/mean_seg", torch.nan_to_num(outs[True]["seg"][b]).sum(), N)
    add("seg/viol_ubseg", (outs[True]["seg"][b] > ub_seg * (1 + 1e-6)).sum(), N)
    add("seg/mean_ubseg", ub_seg.sum(), N)
    for use_seg in (True, False):
        o = outs[use_seg]
        sc = torch.nan_to_num(o["scores"][b], nan=0.0)
        iou = torch.nan_to_num(o["iou"][b], nan=1.0)
        cnt = o["count"][b].float()
        tag = "seg" if use_seg else "noseg"
        add(f"{tag}/viol_cell", (cnt > ub_cell + 0.5).sum(), N)
        add(f"{tag}/viol_slab", (cnt > ub_slab + 0.5).sum(), N)
        for name, ub in (("cell", ub_cell), ("slab", ub_slab), ("slab8", ub_slab8), ("slab8+ubseg", ub_slab8), ("slab8+ubseg2", ub_slab8), ("slab8+ubseg3", ub_slab8), ("slab8+ubseg4", ub_slab8), ("slab8+ubseg5", ub_slab8), ("slab8+ubseg6", ub_slab8), ("slab8+segoracle", ub_slab8)):
            ubs = iou * (ub / max(npts, 1.0)).clamp(max=1)
            if name.endswith("ubseg"):
                if not use_seg: continue
                ubs = ubs * ub_seg
            if name.endswith("ubseg2"):
                if not use_seg: continue
                ubs = ubs * ub_seg2
            if name.endswith("ubseg3"):
                if not use_seg: continue
                ubs = ubs * ub_seg3
            if name.endswith("ubseg4"):
                if not use_seg: continue
                ubs = ubs * ub_seg4
            if name.endswith("ubseg5"):
                if not use_seg: continue
                ubs = ubs * ub_seg5
            if name.endswith("ubseg6"):
                if not use_seg: continue
                ubs = ubs * ub_seg6
            if name.endswith("segoracle"):
                if not use_seg: continue
                ubs = ubs * torch.nan_to_num(o["seg"][b], nan=1.0)
            variants = [(name, ubs)]
            if name == "slab8" and use_seg:
                variants += [(f"slab8+{k_}", ubs * v_) for k_, v_ in extra.items()]
            for name, ubs in variants:
              for PA in (64,):
                order = ubs.argsort(descending=True)
                inA = torch.zeros(N, dtype=torch.bool, device="cuda"); inA[order[:PA]] = True
                TA = sc[inA].topk(K).values[-1]
                keepB = (~inA) & (ubs >= TA)
                add(f"{tag}/{name}/PA{PA}/count", inA.sum() + keepB.sum(), N)
                add(f"{tag}/{name}/PA{PA}/work", work[inA].sum() + work[keepB].sum(), work.sum())
                add(f"{tag}/{name}/PA{PA}/workA", work[inA].sum(), work.sum())
print(json.dumps({k: round(v[0] / max(v[1], 1e-9), 4) for k, v in acc.items()}, indent=0))
