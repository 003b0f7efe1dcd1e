"""Authoring container only: time the REFERENCE's own torch code on BASELINE configs[0] (C1: one
synthetic 640x480 SUN RGB-D-shaped image, 5 GT 2D boxes, 1000 proposals/box, mask term on) in
full -- SURVEY 8(d). The blocks are the reference's functions, imported read-only through
oracle/refshim.py and chained by oracle/compose.py (the snapshot ships no driver):

  ROIHeads3DScore.normal_vector_from_maps + Plane_torch.fit_parallel   (step 1)
  sample_normal_in_range, orthobasis_from_normal_t, lifting             (step 2)
  Cubes.get_bube_corners, cubes_to_box, iou_2d, convex_hull, mask_iou,
  point-in-cube in torch, top-k                                         (step 3)

    python tools/time_reference_c1.py > profiles/reference_torch_c1_r2.json
"""
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import numpy as np
import torch

from oracle import compose, cubes as oc, refshim, standins
from wcr import synth

H, W, NBOX, N, K_TOP = 480, 640, 5, 1000, 16


def one_pass(im, priors, seg):
    ns = refshim.load()
    head = refshim.roi_head_instance()
    t = {}
    t0 = time.perf_counter()
    depth = standins.ImageList.from_tensors([im["depth"]])
    gts = [im["ground"]]
    ground = standins.ImageList([None], [tuple(im["ground"].shape)])
    ground.__class__ = type("IL", (standins.ImageList,), {"__getitem__": lambda self, i: gts[i],
                                                          "device": property(lambda self: torch.device("cpu"))})
    normal = head.normal_vector_from_maps(ground, depth, [im["K"]])[0]
    t["ground_s"] = time.perf_counter() - t0
    t1 = time.perf_counter()
    cubes = []
    for b in range(NBOX):
        box = im["boxes"][b]
        z0 = float(oc.box_center_depth(im["depth"].numpy(), box.numpy()))
        a = compose.draw_attributes(box, z0, priors[int(im["classes"][b])], N, generator_seed=b)
        cubes.append(compose.assemble(im["K"], normal, a))
    t["sample_s"] = time.perf_counter() - t1
    t2 = time.perf_counter()
    for b in range(NBOX):
        r = compose.score(im["depth"], im["K"], im["boxes"][b].numpy(), (H, W), cubes[b], seg=seg[b].numpy(),
                          e_iou=1, e_pts=1, e_seg=1)
        oc.topk(r["score"], K_TOP)
    t["score_s"] = time.perf_counter() - t2
    t["total_s"] = time.perf_counter() - t0
    return t


def main():
    assert refshim.available(), "needs /root/reference"
    torch.set_num_threads(os.cpu_count())
    im = synth.make_image(0, H, W, n_boxes=NBOX, dummy_ground=False)
    seg = synth.seg_masks(im["boxes"], H, W)
    priors = synth.make_priors()
    one_pass(im, priors, seg)                          # warm-up
    runs = [one_pass(im, priors, seg) for _ in range(3)]
    med = statistics.median(r["total_s"] for r in runs)
    pick = min(runs, key=lambda r: abs(r["total_s"] - med))
    print(json.dumps({
        "impl": "reference (its own torch code through oracle/refshim.py + oracle/compose.py)",
        "config": f"C1: 1 image {W}x{H}, {NBOX} GT boxes, {N} proposals/box, IoU2D x point-in-cube x mask coverage, top-{K_TOP}",
        "metric": "cube proposals scored/sec", "value": NBOX * N / med, "unit": "proposals/s",
        "seconds_median_of_3": med, "seconds_all": [r["total_s"] for r in runs],
        "breakdown_s": pick, "cores": os.cpu_count(), "torch_threads": torch.get_num_threads(),
        "where": "authoring container (no GPU); the reference cannot travel to the GPU box",
    }))


if __name__ == "__main__":
    main()
