#!/usr/bin/env python
"""Timings of the rows next to the hot path, one JSON line each:

  iou3d    exact oriented-box IoU, 1264 GT cubes x 1000 proposals each (the C2 proposal set)
  resize   depth + ground maps 480x640 -> 512x683 for 64 images (the mapper's NEAREST resize)
  glue     class gather + 6-D pose + per-box constants for 1156 boxes x 50 classes, fwd + bwd

CUDA events, L2 not flushed (these are latency-bound helpers). CPU side: the float64 oracle on a
bounded sample (iou3d), PIL on the host (resize), torch on the host (glue).
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))


def timed(fn, steps=20, warmup=3):
    import torch
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    import numpy as np
    import torch
    from oracle import glue as og, iou3d as o3
    from wcr import _abi
    from wcr.glue import box_constants, class_gather, rotation_6d_to_matrix
    from wcr.io import resize_nearest
    dev = torch.device("cuda", 0)
    lib = _abi.load()
    _abi.check(lib.wcr_device_check(), "wcr_device_check")
    rng = np.random.default_rng(0)

    # ---- iou3d -------------------------------------------------------------------------------
    def rot():
        q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        return -q * np.sign(np.linalg.det(q))                    # det -1 like the reference's bases
    nb, P = 1264, 1000
    gts = [(rng.normal(size=3) * 0.5 + [0, 0, 4], rng.uniform(0.4, 2.0, 3), rot()) for _ in range(nb)]
    G = torch.tensor(np.stack([o3.cube_corners(*c) for c in gts]), dtype=torch.float32, device=dev)
    jit = torch.randn(nb, P, 1, 3, device=dev) * 0.3
    Pc = (G[:, None] + jit).contiguous()                          # proposals: the GT cube shifted
    vol = torch.empty(nb, P, device=dev)
    iou = torch.empty(nb, P, device=dev)

    def run_iou():   # one GT against its own P proposals, batched over boxes with one launch each... as one launch:
        for b in range(0, nb, nb):                                # (1 x P per box) == block-diagonal of (nb x nb*P)
            pass
        _abi.check(lib.wcr_box3d_overlap(G.data_ptr(), 1, Pc.data_ptr(), nb * P, vol.data_ptr(), iou.data_ptr(),
                                         _abi.stream_ptr(dev)), "wcr_box3d_overlap")
    ms = timed(run_iou, steps=10)
    k = 60
    t0 = time.perf_counter()
    for j in range(k):
        o3.box3d_overlap_pair(G[0].cpu().numpy(), Pc[0, j].cpu().numpy())
    cpu_pair = (time.perf_counter() - t0) / k
    print(json.dumps({"metric": "oriented-box IoU pairs/sec", "value": nb * P / (ms * 1e-3), "unit": "pairs/s",
                      "ms": ms, "config": {"workload": f"{nb * P} (box, proposal) pairs, one launch"},
                      "cpu_baseline": {"value": 1.0 / cpu_pair, "unit": "pairs/s", "cores": 1, "kind": "port",
                                       "sample": f"{k} pairs, oracle/iou3d.py (numpy float64)"}}), flush=True)

    # ---- resize ------------------------------------------------------------------------------
    from PIL import Image
    B = 64
    depth = [torch.rand(480, 640, device=dev) * 10 + 0.3 for _ in range(B)]
    ground = [(torch.rand(480, 640, device=dev) < 0.4) for _ in range(B)]

    def run_resize():
        for d, g in zip(depth, ground):
            resize_nearest(d, (512, 683)); resize_nearest(g, (512, 683))
    ms = timed(run_resize, steps=5)
    dn, gn = depth[0].cpu().numpy(), ground[0].cpu().numpy()
    t0 = time.perf_counter()
    for _ in range(B):
        np.array(Image.fromarray(dn).resize((683, 512), Image.NEAREST)); np.array(Image.fromarray(gn).resize((683, 512), Image.NEAREST))
    cpu_ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"metric": "map resize latency", "value": ms, "unit": "ms", "higher_is_better": False,
                      "config": {"workload": f"{B} depth (f32) + {B} ground (bool) maps 480x640 -> 512x683, device resident"},
                      "cpu_baseline": {"value": cpu_ms, "unit": "ms", "cores": 1, "kind": "reference",
                                       "sample": "PIL Image.resize(NEAREST) on the host, same maps"}}), flush=True)

    # ---- glue --------------------------------------------------------------------------------
    n, C = 1156, 50
    pose6 = torch.randn(n, C, 6, device=dev, requires_grad=True)
    dims = torch.randn(n, C, 3, device=dev, requires_grad=True)
    cls = torch.randint(0, C, (n,), device=dev)
    Ks = [torch.tensor([[500.0, 0, 320], [0, 500.0, 240], [0, 0, 1]])] * 16
    nbi = [n // 16] * 15 + [n - 15 * (n // 16)]

    def run_glue():
        pose6.grad = None; dims.grad = None
        R = rotation_6d_to_matrix(class_gather(pose6, cls))
        d = class_gather(dims, cls)
        box_constants(Ks, [1.0] * 16, [(480, 640)] * 16, nbi, device=dev)
        (R.sum() + d.sum()).backward()
    ms = timed(run_glue, steps=20)
    p6c, dc, clc = pose6.detach().cpu().requires_grad_(True), dims.detach().cpu().requires_grad_(True), cls.cpu()

    def cpu_glue():
        p6c.grad = None; dc.grad = None
        R = og.class_gather(og.rotation_6d_to_matrix(p6c.view(-1, 6)).view(n, C, 3, 3), clc)   # every class, as the reference
        d = og.class_gather(dc, clc)
        og.box_constants(Ks, [1.0] * 16, [(480, 640)] * 16, nbi)
        (R.sum() + d.sum()).backward()
    cpu_glue()
    t0 = time.perf_counter()
    for _ in range(5):
        cpu_glue()
    cpu_ms = (time.perf_counter() - t0) * 1e3 / 5
    print(json.dumps({"metric": "caller glue fwd+bwd latency", "value": ms, "unit": "ms", "higher_is_better": False,
                      "config": {"workload": f"class gather + 6-D pose + per-box constants, {n} boxes x {C} classes"},
                      "cpu_baseline": {"value": cpu_ms, "unit": "ms", "cores": torch.get_num_threads(), "kind": "port",
                                       "sample": "oracle/glue.py (torch, host), 6-D pose converted for every class"}}), flush=True)


if __name__ == "__main__":
    main()
