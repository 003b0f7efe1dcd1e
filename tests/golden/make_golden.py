"""Generate tests/golden/*.npz by executing the REFERENCE's own code (read-only at
/root/reference, imported through oracle/refshim.py) on seeded synthetic inputs.

    python tests/golden/make_golden.py

Authoring container only. The fixtures are committed; the GPU box and the CPU test
suite read them without touching /root/reference. Inputs are stored in the files
so that the tests feed bit-identical data to the oracle and to the CUDA path.
"""
import os
import random
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "weak-cube-r-cnn_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from oracle import refshim, standins, loss as ol, loss_inputs as li, ref_loss  # noqa: E402
from tests.golden.golden_io import save_loss_case  # noqa: E402


def golden_ground():
    """normal_vector_from_maps + Plane_torch.fit_parallel (roi_heads.py:466-528,
    plane.py:79-134) with Python's random seeded: stores the drawn triples too."""
    from wcr import synth
    ns = refshim.load()
    head = refshim.roi_head_instance()
    for case, (H, W, seeds) in enumerate([(120, 160, (0, 1, 2)), (150, 210, (5, 6))]):
        ims = [synth.make_image(s, H, W, n_boxes=2, dummy_ground=(i == 1)) for i, s in enumerate(seeds)]
        for im in ims:
            im["K"][0, 0] *= 1.013
            im["K"][1, 1] *= 0.987
        depth = standins.ImageList.from_tensors([im["depth"] for im in ims])
        ground = standins.ImageList([None] * len(ims), [tuple(im["ground"].shape) for im in ims])
        gt = [im["ground"] for im in ims]
        ground.__class__ = type("IL", (standins.ImageList,), {
            "__getitem__": lambda self, i: gt[i], "device": property(lambda self: torch.device("cpu"))})
        Ks = [im["K"] for im in ims]
        # replay the draws: same seed, same order as plane.py:95
        random.seed(100 + case)
        normals = head.normal_vector_from_maps(ground, depth, Ks)
        random.seed(100 + case)
        ids, pts_all, eqs, inls = [], [], [], []
        from oracle import ground as og
        for im in ims:
            g = None if tuple(im["ground"].shape) == (1, 1) else im["ground"].numpy()
            pts, _ = og.unproject_ground_frame(im["depth"].numpy(), g, im["K"].numpy(), 5)
            state = random.getstate()
            eq, inl = ns.plane.Plane_torch().fit_parallel(torch.from_numpy(pts), thresh=0.05, maxIteration=1000)
            random.setstate(state)
            ids.append(og.draw_id_samples(len(pts), 1000, random))
            pts_all.append(pts); eqs.append(eq.numpy()); inls.append(inl.numpy())
        d = {"H": H, "W": W, "normals": normals.numpy()}
        for i, im in enumerate(ims):
            d[f"depth{i}"] = im["depth"].numpy(); d[f"ground{i}"] = im["ground"].numpy()
            d[f"K{i}"] = im["K"].numpy(); d[f"ids{i}"] = ids[i].astype(np.int32)
            d[f"pts{i}"] = pts_all[i]; d[f"eq{i}"] = eqs[i]; d[f"inl{i}"] = inls[i]
        d["n"] = len(ims)
        np.savez_compressed(os.path.join(HERE, f"ground_{case}.npz"), **d)
        print("ground", case, normals.numpy())


def golden_cubes():
    """Block-level outputs of the reference: orthobasis_from_normal_t, Cubes.get_all_corners,
    get_bube_corners, cubes_to_box, iou_2d (utils.py:119-187, spaces.py:192-245,
    conversions.py:25-48)."""
    ns = refshim.load()
    g = torch.Generator().manual_seed(11)
    I, Pn = 3, 40
    normal = torch.nn.functional.normalize(torch.tensor([0.03, 0.99, -0.05]), dim=0)
    yaw = torch.rand(Pn, generator=g) * np.pi
    R = ns.utils.orthobasis_from_normal_t(normal, yaw)
    R0 = ns.utils.orthobasis_from_normal_t(torch.tensor([0.0, 1.0, 0.0]), yaw)   # a == 0 branch
    cubes = torch.zeros(I, Pn, 15)
    cubes[:, :, 0:2] = torch.randn(I, Pn, 2, generator=g) * 0.8
    cubes[:, :, 2] = 1.0 + 4 * torch.rand(I, Pn, generator=g)
    cubes[:, :, 3:6] = 0.2 + 1.5 * torch.rand(I, Pn, 3, generator=g)
    cubes[:, :, 6:] = R.reshape(Pn, 9)
    cubes[0, 0, 2] = 0.3                       # close to the camera: clamp bounds hit
    K = torch.tensor([[525.0, 0, 320], [0, 525, 240], [0, 0, 1]])
    C = ns.spaces.Cubes(cubes)
    corners3d = C.get_all_corners()
    bube = C.get_bube_corners(K, [480, 640])
    boxes = torch.stack([b.tensor for b in ns.conversions.cubes_to_box(C, K, [480, 640])])
    gt = torch.tensor([[200.0, 150, 420, 380], [10, 10, 100, 90], [300, 200, 640, 480]])
    ious = torch.stack([ns.utils.iou_2d(standins.Boxes(gt[i:i + 1]), standins.Boxes(boxes[i])) for i in range(I)])
    np.savez_compressed(os.path.join(HERE, "cubes_0.npz"), normal=normal.numpy(), yaw=yaw.numpy(),
                        R=R.numpy(), R0=R0.numpy(), cubes=cubes.numpy(), K=K.numpy(),
                        corners3d=corners3d.numpy(), bube=bube.numpy(), boxes=boxes.numpy(),
                        gt=gt.numpy(), ious=ious.numpy())
    print("cubes", boxes.shape, float(ious.max()))


def golden_hull():
    """convex_hull = jarvis_march + fill_polygon, mask_iou, mod_mask_iou
    (utils.py:225-245, 410-493) on the reference's own five fixture point sets
    (utils.py:500-544, copied here as INPUT data) and on seeded random clamped
    corner sets. Stored: hull vertices, the binarised mask (bit-packed), every
    fractional soft value, and the two IoUs against a seeded random mask."""
    ns = refshim.load()
    U = ns.utils
    fixtures = [
        [[271.0, 356.0], [25.3744, 356.0], [0.0, 356.0], [0.0, 89.5266], [271.0, 159.3112],
         [95.5653, 201.7484], [0.0, 0.0], [271.0, 0.0]],
        [[150.3456, 0.0], [479.0, 0.0], [11.8427, 0.0], [0.0, 0.0], [121.4681, 232.5976],
         [375.6230, 383.9329], [12.8765, 630.0], [0.0, 344.7250]],
        [[290.9577, 171.1176], [197.7348, 483.7612], [383.0, 504.0], [383.0, 27.6211],
         [2.2419, 52.6505], [0.0, 399.6908], [0.0, 504.0], [0.0, 0.0]],
        [[271.0, 19.5241], [271.0, 356.0], [0.0, 0.0], [271.0, 0.0], [0.0, 0.0], [163.0264, 77.9408],
         [164.2467, 321.0222], [0.0, 356.0], [0.0, 0.0]],
        [[272.0, 1.0], [0.0, 173.5156], [74.8860, 141.3913], [253.8221, 0.0], [271.0, 0.0],
         [271.0, 356.0], [262.5294, 327.9978], [271.0, 120.8048]],
    ]
    rng = np.random.default_rng(17)

    def one_file(name, point_sets, H, W):
        d = {"H": H, "W": W, "n_sets": len(point_sets)}
        for i, pts in enumerate(point_sets):
            pts = np.asarray(pts, dtype=np.float32)
            hull = U.jarvis_march(torch.from_numpy(pts.copy()))
            soft = U.fill_polygon(torch.zeros(H, W), hull).numpy()
            seg = rng.uniform(size=(H, W)) < 0.35
            d[f"pts_{i}"] = pts
            d[f"hull_{i}"] = hull.numpy()
            d[f"maskbits_{i}"] = np.packbits((soft > 0).reshape(-1))
            frac = np.flatnonzero((soft > 0) & (soft < 1))
            d[f"frac_idx_{i}"] = frac
            d[f"frac_val_{i}"] = soft.reshape(-1)[frac]
            d[f"seg_{i}"] = seg
            d[f"miou_{i}"] = np.float32(U.mask_iou(torch.from_numpy(seg), soft))
            d[f"modiou_{i}"] = np.float32(U.mod_mask_iou(torch.from_numpy(seg).float(), soft))
        np.savez_compressed(os.path.join(HERE, name), **d)
        print(name, [d[f"hull_{i}"].shape[0] for i in range(len(point_sets))])

    one_file("hull_0.npz", fixtures, 640, 500)
    sets = []
    for t in range(12):
        pts = rng.uniform(-40, 200, size=(8, 2)).astype(np.float32)
        pts[:, 0] = np.clip(pts[:, 0], 0, 159)
        pts[:, 1] = np.clip(pts[:, 1], 0, 119)
        if t % 4 == 0:
            pts[3] = pts[6]
        sets.append(pts)
    one_file("hull_1.npz", sets, 120, 160)


def golden_zbox():
    """ROIHeads3DScore.pseudo_gt_z_box_loss (roi_heads.py:575-610) run as is: boxes partly
    and fully outside the image, even/odd patch sizes (lower median), 1-pixel patches."""
    head = refshim.roi_head_instance()
    g = torch.Generator().manual_seed(31)
    H, W = 90, 120
    depth = [(0.5 + 6 * torch.rand(H, W, generator=g)) for _ in range(3)]
    depth[1][::2] = depth[1][1::2]                          # repeated values
    boxes = []
    for i, nb in enumerate((6, 1, 9)):
        x1 = torch.rand(nb, generator=g) * W * 1.2 - 0.1 * W
        y1 = torch.rand(nb, generator=g) * H * 1.2 - 0.1 * H
        bw = 1.0 + torch.rand(nb, generator=g) * 50
        bh = 1.0 + torch.rand(nb, generator=g) * 40
        b = torch.stack((x1, y1, x1 + bw, y1 + bh), 1)
        boxes.append(b)
    boxes[0][1] = torch.tensor([-30.0, 10.0, -5.0, 40.0])   # fully outside: fallback + re-ordering
    boxes[2][4] = torch.tensor([50.0, 200.0, 80.0, 230.0])
    boxes[2][0] = torch.tensor([10.2, 10.3, 12.9, 11.8])    # 2 x 1 patch
    pred_z = (1 + 5 * torch.rand(16, generator=g)).requires_grad_(True)
    il = standins.ImageList.from_tensors(depth)
    loss = head.pseudo_gt_z_box_loss(il, tuple(boxes), pred_z)
    loss.sum().backward()
    np.savez_compressed(os.path.join(HERE, "zbox_0.npz"), depth=torch.stack(depth).numpy(),
                        boxes=torch.cat(boxes).numpy(), nb=np.array([6, 1, 9]), pred_z=pred_z.detach().numpy(),
                        loss=loss.detach().numpy(), grad=pred_z.grad.numpy())
    print("zbox", loss.detach().numpy().round(3))


def golden_replay():
    """Proposal attributes (u v z w h l yaw) drawn by the REFERENCE's sampler
    (sample_normal_in_range, utils.py:41-59, torch's global generator under manual_seed)
    for a few boxes of one small synthetic image: the replay stream of the parity tests
    (tests/test_oracle_compose.py on the CPU, tests/test_score_gpu.py on the GPU)."""
    from oracle import compose, cubes as oc
    from wcr import synth
    H, W, N = 120, 160, 192
    im = synth.make_image(77, H, W, n_boxes=4, dummy_ground=False)
    priors_all = synth.make_priors()
    depth, K = im["depth"].numpy(), im["K"].numpy()
    boxes = im["boxes"].numpy()
    normal = np.array([0.02, 0.999, -0.03], np.float32); normal /= np.linalg.norm(normal)
    z0 = np.array([oc.box_center_depth(depth, b) for b in boxes], np.float32)
    priors = priors_all[im["classes"]].numpy()
    seeds = np.arange(4) + 900
    attrs = np.stack([compose.draw_attributes(boxes[b], z0[b], priors[b], N, generator_seed=int(seeds[b]))
                      for b in range(4)])
    seg = synth.seg_masks(im["boxes"], H, W).numpy()
    np.savez_compressed(os.path.join(HERE, "replay_0.npz"), depth=depth, K=K, boxes=boxes, normal=normal,
                        z0=z0, priors=priors, classes=im["classes"].numpy(), seeds=seeds, attrs=attrs, seg=seg)
    print("replay", attrs.shape, attrs[:, :, 0])


def golden_loss():
    cases = [(0, [3, 1, 5, 2], False, {}),
             (1, [6, 2, 9], True, {"z_type": "sigmoid", "dims_priors_func": "sigmoid"}),
             (2, [1, 1], False, {}),
             (3, [12, 7, 20], False, {"z_type": "log", "use_confidence": 0.0}),
             (4, [5, 1, 8], False, {"loss_functions": ["iou", "z_pseudo_gt_patch", "pose_alignment",
                                                       "pose_ground", "dims"]})]
    for idx, (seed, nb, big, cfg_kw) in enumerate(cases):
        inp = li.make_loss_inputs(seed, nb, H=120, W=160, big=big)
        inp["depth_maps"] = [d.half().float() for d in inp["depth_maps"]]   # stored as fp16
        L, _, grads = li.run_with_grads(ref_loss.reference_weak_loss, ol.LossConfig(**cfg_kw), inp)
        save_loss_case(os.path.join(HERE, f"loss_{idx}.npz"), cfg_kw, inp, L, grads)
        print("loss", idx, {k: round(float(v), 5) for k, v in L.items()})


def golden_tier2():
    """Tier-2 terms ('z', 'depth', 'pose_ground2' on top of 'iou', 'dims') through the
    reference's own z_loss / depth_range_loss / normal_to_rotation. Inputs are re-derived
    from the seed by the tests (oracle/loss_inputs.py), only outputs are stored."""
    T2 = ("iou", "z", "depth", "pose_ground2", "dims")
    for idx, (seed, nb, conf) in enumerate([(0, [3, 1, 5, 2], 1.0), (4, [7, 12, 2], 0.0)]):
        H, W = 96, 128
        inp = li.add_masks(li.make_loss_inputs(seed, nb, H=H, W=W), seed, H=H, W=W)
        cfg = ol.LossConfig(loss_functions=T2, use_confidence=conf)
        L, pb, grads = li.run_with_grads(ref_loss.reference_weak_loss, cfg, inp)
        out = {"seed": seed, "nb": np.asarray(nb), "H": H, "W": W, "use_confidence": conf}
        out.update({"loss:" + k: v.numpy() for k, v in L.items()})
        out.update({"grad:" + k: v.numpy() for k, v in grads.items()})
        out.update({"box:" + k: pb[k].numpy() for k in ("raw_z", "raw_depth", "raw_normal_vec")})
        np.savez_compressed(os.path.join(HERE, f"tier2_{idx}.npz"), **out)
        print("tier2", idx, {k: round(float(v), 5) for k, v in L.items()})


def golden_tier2_seg():
    """'segmentation' (segment_loss through the reference's convex_hull raster, values and the
    autograd gradients that flow through the soft raster)."""
    for idx, (seed, nb) in enumerate([(0, [3, 1, 5, 2]), (6, [4, 7])]):
        H, W = 96, 128
        inp = li.add_masks(li.make_loss_inputs(seed, nb, H=H, W=W), seed, H=H, W=W)
        cfg = ol.LossConfig(loss_functions=("segmentation", "iou"), use_confidence=1.0)
        L, pb, grads = li.run_with_grads(ref_loss.reference_weak_loss, cfg, inp)
        out = {"seed": seed, "nb": np.asarray(nb), "H": H, "W": W, "use_confidence": 1.0}
        out.update({"loss:" + k: v.numpy() for k, v in L.items()})
        out.update({"grad:" + k: v.numpy() for k, v in grads.items()})
        out["box:raw_seg"] = pb["raw_seg"].numpy()
        np.savez_compressed(os.path.join(HERE, f"tier2seg_{idx}.npz"), **out)
        print("tier2seg", idx, {k: round(float(v), 6) for k, v in L.items()})


def golden_blocks2():
    """Round-2 blocks, straight from the reference: custom_mapping, randn_orthobasis_torch (the
    torch.randn draws are stored: replay), gt_in_norm_range, is_gt_included, the AUTOGRAD gradients
    of Cubes.get_all_corners / get_bube_corners / cubes_to_box (spaces.py:192-245,
    conversions.py:25-48), and the stand-alone z_loss / depth_range_loss / normal_to_rotation
    methods of ROIHeads3DScore (roi_heads.py:530-573, 658-696)."""
    import contextlib
    import io
    import types
    ns = refshim.load()
    U = ns.utils
    head = refshim.roi_head_instance()
    g = torch.Generator().manual_seed(31)
    out = {}
    xs = [0.0, -0.2, 1e-6, 0.05, 0.3, 0.5, 0.7, 0.95, 0.999999]
    out["cm_x"] = np.asarray(xs, dtype=np.float32)
    out["cm_betas"] = np.asarray([1.7, 3.0, 1.01])
    out["cm_out"] = np.asarray([U.custom_mapping([float(np.float32(v)) for v in xs], b) for b in out["cm_betas"]])
    torch.manual_seed(5)
    z = torch.randn(3, 7, 3, 3)
    torch.manual_seed(5)
    out["ro_z"], out["ro_out"] = z.numpy(), U.randn_orthobasis_torch(7, 3).numpy()
    lo, hi, gt = torch.randn(20, generator=g), torch.randn(20, generator=g), torch.randn(20, generator=g)
    out["gn_lo"], out["gn_hi"], out["gn_gt"] = lo.numpy(), hi.numpy(), gt.numpy()
    out["gn_out"] = np.stack([U.gt_in_norm_range((lo[i], hi[i]), gt[i]).numpy() for i in range(20)])
    # is_gt_included: 24 random cases around the decision boundaries
    n = 24
    cen, dim = torch.randn(n, 3, generator=g) * 2, 0.2 + 2 * torch.rand(n, 3, generator=g)
    rng = torch.sort(torch.randn(n, 3, 2, generator=g) * 2, dim=-1)[0]
    pri = torch.stack((0.2 + 2 * torch.rand(n, 3, generator=g), 0.1 + 0.5 * torch.rand(n, 3, generator=g)), -1)
    for i in range(0, n, 2):                   # every other case near / inside the acceptance region
        rng[i, :, 0] = cen[i] - 0.5 - torch.rand(3, generator=g)
        rng[i, :, 1] = cen[i] + 0.5 + torch.rand(3, generator=g)
        dim[i] = pri[i, :, 0] + pri[i, :, 1] * (torch.rand(3, generator=g) * 3.2 - 1.6) * (1.0 if i % 4 else 0.9)
    inc = []
    for i in range(n):
        cube = types.SimpleNamespace(center=cen[i], dimensions=dim[i])
        with contextlib.redirect_stdout(io.StringIO()):
            inc.append(bool(U.is_gt_included(cube, rng[i, 0], rng[i, 1], rng[i, 2], pri[i, 0], pri[i, 1], pri[i, 2])))
    out.update({"ig_center": cen.numpy(), "ig_dims": dim.numpy(), "ig_ranges": rng.numpy(),
                "ig_priors": pri.numpy(), "ig_included": np.asarray(inc)})
    # autograd through the geometry blocks
    I, Pn = 2, 9
    yaw = torch.rand(Pn, generator=g) * np.pi
    R = U.orthobasis_from_normal_t(torch.nn.functional.normalize(torch.tensor([0.03, 0.99, -0.05]), dim=0), yaw)
    cubes = torch.zeros(I, Pn, 15)
    cubes[:, :, 0:2] = torch.randn(I, Pn, 2, generator=g) * 0.8
    cubes[:, :, 2] = 1.0 + 4 * torch.rand(I, Pn, generator=g)
    cubes[:, :, 3:6] = 0.2 + 1.5 * torch.rand(I, Pn, 3, generator=g)
    cubes[:, :, 6:] = R.reshape(Pn, 9)
    cubes[0, 0, 2] = 0.3                       # close to the camera: clamp bounds hit, no gradient there
    K = torch.tensor([[525.0, 0, 320], [0, 525, 240], [0, 0, 1]])
    clamp = [480, 640]
    up3 = torch.randn(I, Pn, 8, 3, generator=g)
    up2 = torch.randn(I, Pn, 8, 2, generator=g)
    up4 = torch.randn(I, Pn, 4, generator=g)
    ct = cubes.clone().requires_grad_(True)
    (ns.spaces.Cubes(ct).get_all_corners() * up3).sum().backward()
    out["ag_cubes"], out["ag_K"], out["ag_clamp"] = cubes.numpy(), K.numpy(), np.asarray(clamp)
    out["ag_up3"], out["ag_up2"], out["ag_up4"] = up3.numpy(), up2.numpy(), up4.numpy()
    out["ag_g3"] = ct.grad.numpy().copy()
    ct = cubes.clone().requires_grad_(True)
    (ns.spaces.Cubes(ct).get_bube_corners(K, clamp) * up2).sum().backward()
    out["ag_g2"] = ct.grad.numpy().copy()
    ct = cubes.clone().requires_grad_(True)
    boxes = torch.stack([b.tensor for b in ns.conversions.cubes_to_box(ns.spaces.Cubes(ct), K, clamp)])
    (boxes * up4).sum().backward()
    out["ag_g4"] = ct.grad.numpy().copy()
    # stand-alone z_loss / depth_range_loss / normal_to_rotation on the tier-2 test inputs
    seed, nb, H, W = 2, [4, 1, 6], 96, 128
    inp = li.add_masks(li.make_loss_inputs(seed, nb, H=H, W=W), seed, H=H, W=W)
    dec = ol.decode(ol.LossConfig(), inp["deltas"], inp["z_raw"], inp["dims_raw"], inp["pose"], inp["src_boxes"],
                    inp["K4"], inp["virtual_to_real"], inp["prior_mean"], inp["prior_std"])
    nbox = inp["deltas"].shape[0]
    box6, Rr = dec[5].detach(), dec[4].detach()
    cubes_t = torch.cat((box6, Rr.reshape(nbox, 9)), 1).unsqueeze(1)
    Ks = torch.zeros(nbox, 3, 3)
    K4 = inp["K4"]
    Ks[:, 0, 0], Ks[:, 1, 1], Ks[:, 0, 2], Ks[:, 1, 2], Ks[:, 2, 2] = K4[:, 0], K4[:, 1], K4[:, 2], K4[:, 3], 1
    im_sizes = [[int(c[0]), int(c[1])] for c in inp["clamp"]]
    C = ns.spaces.Cubes(cubes_t)
    proj = torch.stack([ns.conversions.cubes_to_box(C[i], Ks[i], im_sizes[i])[0].tensor[0] for i in range(nbox)])
    out["sa_seed"], out["sa_nb"], out["sa_H"], out["sa_W"] = seed, np.asarray(nb), H, W
    out["sa_cubes"], out["sa_proj"] = cubes_t.numpy(), proj.numpy()
    out["sa_z_loss"] = head.z_loss(standins.Boxes(inp["gt_boxes"]), C, Ks, im_sizes, standins.Boxes(proj)).detach().numpy()
    for tag, pad in (("", (0, 0)), ("_pad", (11, 32))):
        masks = [torch.nn.functional.pad(m, (0, pad[1], 0, pad[0])) for m in inp["gt_masks"]]
        ct = cubes_t.clone().requires_grad_(True)
        d = head.depth_range_loss(masks, inp["at_which_mask_idx"], inp["depth_maps"], ns.spaces.Cubes(ct),
                                  standins.Boxes(inp["gt_boxes"]), nb)
        d.sum().backward()
        out["sa_depth" + tag], out["sa_depth_grad" + tag] = d.detach().numpy(), ct.grad.numpy().copy()
    out["sa_pad"] = np.asarray([11, 32])
    nrm = torch.nn.functional.normalize(torch.randn(5, 3, generator=g), dim=1)
    out["sa_normals"], out["sa_rot"] = nrm.numpy(), head.normal_to_rotation(nrm).numpy()
    np.savez_compressed(os.path.join(HERE, "blocks2_0.npz"), **out)
    print("blocks2", out["cm_out"][0][:4], out["ig_included"].sum(), out["sa_z_loss"][:4], out["sa_depth_pad"][:3])


def golden_tier2_padded():
    """'depth' with masks at a PADDED size (ImageList batch tensor: bottom / right padding), so
    depth_range_loss resizes every depth map with F.interpolate(bilinear, align_corners=True)
    (roi_heads.py:672): values and autograd gradients from the reference."""
    T2 = ("iou", "depth", "dims")
    seed, nb, conf, H, W, pad = 3, [5, 2, 4], 1.0, 96, 128, (16, 32)
    inp = li.add_masks(li.make_loss_inputs(seed, nb, H=H, W=W), seed, H=H, W=W)
    inp["gt_masks"] = [torch.nn.functional.pad(m, (0, pad[1], 0, pad[0])) for m in inp["gt_masks"]]
    cfg = ol.LossConfig(loss_functions=T2, use_confidence=conf)
    L, pb, grads = li.run_with_grads(ref_loss.reference_weak_loss, cfg, inp)
    out = {"seed": seed, "nb": np.asarray(nb), "H": H, "W": W, "use_confidence": conf, "pad": np.asarray(pad)}
    out.update({"loss:" + k: v.numpy() for k, v in L.items()})
    out.update({"grad:" + k: v.numpy() for k, v in grads.items()})
    out["box:raw_depth"] = pb["raw_depth"].numpy()
    np.savez_compressed(os.path.join(HERE, "tier2pad_0.npz"), **out)
    print("tier2pad", {k: round(float(v), 5) for k, v in L.items()})


FORWARD_CUBE_CASES = [
    dict(seed=0, nb=[3, 1, 4]),                                                         # config.py:186 default list
    dict(seed=5, nb=[6, 2, 5, 3], use_confidence=0.0,
         loss_functions=("iou", "z_pseudo_gt_center", "pose_alignment", "pose_ground2", "dims", "segmentation", "depth"),
         pad=(8, 16)),
]


def golden_forward_cube():
    """The reference's OWN ROIHeads3DScore._forward_cube training branch (roi_heads.py:698-1144),
    driven through oracle/ref_forward_cube.py with a stub pooler / cube head: the Cube/loss_*
    dict and the autograd gradients w.r.t. the five per-class head outputs."""
    from oracle import ref_forward_cube as rf
    for idx, kw in enumerate(FORWARD_CUBE_CASES):
        case = rf.make_case(**kw)
        L, grads = rf.run_reference_with_grads(case)
        out = {"kw": np.frombuffer(repr(kw).encode(), dtype=np.uint8)}
        out.update({"loss:" + k: v.numpy() for k, v in L.items()})
        out.update({"grad:" + k: v.numpy() for k, v in grads.items()})
        np.savez_compressed(os.path.join(HERE, f"forward_cube_{idx}.npz"), **out)
        print("forward_cube", idx, {k: round(float(v), 5) for k, v in L.items()})


if __name__ == "__main__":
    assert refshim.available(), "needs /root/reference"
    if "--round2" in sys.argv:
        golden_blocks2()
        golden_tier2_padded()
        golden_forward_cube()
        sys.exit(0)
    golden_ground()
    golden_cubes()
    golden_hull()
    golden_zbox()
    golden_replay()
    golden_loss()
    golden_tier2()
    golden_tier2_seg()
    golden_blocks2()
    golden_tier2_padded()
