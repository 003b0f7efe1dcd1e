#!/usr/bin/env python
"""BASELINE.json configs[4]: proposal-count sweep 1k-64k per box at depth resolutions
480p-1080p (roofline study of the fused scoring kernel), one JSON line per point.

    python tools/sweep.py [--images 8] [--boxes 8] [--reps 5] [--seg]

Time = CUDA events around the fused scoring launch (sampling in-kernel), inputs resident,
256 MiB L2 flush before every timed launch. Rates: proposals/s and point tests/s;
`frac` = 18 FLOP x tests/s over the FFMA probe measured in the same process.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=8)
    ap.add_argument("--boxes", type=int, default=8)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--seg", action="store_true")
    ap.add_argument("--N", type=int, nargs="*", default=[1000, 2000, 4000, 8000, 16000, 32000, 64000])
    a = ap.parse_args()
    import torch
    from wcr import _abi, synth
    from wcr.pipeline import ShardRunner
    from wcr.proposals import ScoreConfig, score_topk, box_center_depth
    dev = torch.device("cuda", 0)
    lib = _abi.load()
    _abi.check(lib.wcr_device_check(), "wcr_device_check")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    probe_out = torch.empty(148 * 8 * 256 * 2, dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    peak = 0.0
    for _ in range(5):
        e0.record()
        flops = lib.wcr_probe_ffma(probe_out.data_ptr(), probe_out.numel(), 4096, _abi.stream_ptr(dev))
        e1.record(); torch.cuda.synchronize()
        peak = max(peak, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    priors = synth.make_priors()
    for (H, W) in ((480, 640), (720, 1280), (1080, 1920)):
        ims = synth.make_batch(a.images, H, W, n_boxes=a.boxes, first=500)
        if a.seg:
            for im in ims:
                im["seg"] = synth.seg_masks(im["boxes"], H, W)
        for N in a.N:
            runner = ShardRunner(ims, N, dev, priors=priors, score_cfg=ScoreConfig(k=16, e_seg=1 if a.seg else 0), seed=7)
            runner.upload()
            runner.step()
            z0 = box_center_depth(runner.shard, runner.boxes)
            full = score_topk(runner.shard, runner.boxes, None, N, runner.score_cfg, priors=runner.priors,
                              normals=runner.normals, z0=z0, seed=7, seg=runner.seg, want_all=True)
            tests = float(full["npts"].sum().item()) * N
            del full
            ms = []
            for _ in range(a.reps + 1):
                flush.zero_()
                e0.record(); runner._score(z0); e1.record()
                torch.cuda.synchronize()
                ms.append(e0.elapsed_time(e1))
            t = min(ms[1:]) * 1e-3
            print(json.dumps({"H": H, "W": W, "N": N, "boxes": runner.n_boxes, "mask_term": bool(a.seg),
                              "kernel_ms": t * 1e3, "proposals_per_s": runner.n_boxes * N / t,
                              "point_tests_per_s": tests / t, "alg_tflops_18": 18 * tests / t / 1e12,
                              "fp32_probe_tflops": peak, "frac": 18 * tests / t / 1e12 / peak}), flush=True)
            del runner
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
