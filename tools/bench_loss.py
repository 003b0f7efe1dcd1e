#!/usr/bin/env python
"""BASELINE.json configs[2]: weak-loss forward + backward at ResNet34-FPN-shaped proposal
counts (batch 16, n_i ~ U{1..128} foreground boxes per image, random-init head outputs).

    python tools/bench_loss.py [--steps 50] [--warmup 5] [--cpu-repeats 3]

Prints one JSON line: GPU time per fwd+bwd (CUDA events, median), kernel launches per
step, and the CPU time of the same block through the oracle's torch + autograd
restatement (`cpu_baseline`, kind "port": the reference's loss block needs detectron2
objects and /root/reference, neither of which exists on the GPU box). Latency-bound by
construction (a few hundred KB of data): no roofline claim.
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--cpu-repeats", type=int, default=3)
    ap.add_argument("--patch", action="store_true", help="use 'z_pseudo_gt_patch' (box median) instead of the centre lookup")
    ap.add_argument("--noise-depth", action="store_true",
                    help="white-noise depth maps (the parity tests' worst case) instead of the synthetic "
                         "floor + wall + objects scenes of bench.py")
    ap.add_argument("--default", action="store_true",
                    help="cfg.loss_functions of config.py:186: dims, pose_alignment, pose_ground, iou, z, "
                         "z_pseudo_gt_patch, depth (per-object masks)")
    ap.add_argument("--no-host-sync", action="store_true",
                    help="weak_loss(host_sync=False): keys decided on the host, flags checked one call late")
    a = ap.parse_args()
    import torch
    from oracle import loss as ol, loss_inputs as li
    from wcr import _abi
    from wcr.images import ImageShard
    from wcr.roi_heads import WeakLossConfig, weak_loss

    g = torch.Generator().manual_seed(0)
    nb = [int(x) for x in torch.randint(1, 129, (16,), generator=g)]
    inp = li.make_loss_inputs(7, nb, H=480, W=640)
    if not a.noise_depth:      # scene-like depth (smooth surfaces + 1 % noise), as in bench.py's workload
        from wcr import synth
        inp["depth_maps"] = [synth.make_image(1000 + i, 480, 640, n_boxes=8)["depth"].contiguous()
                             for i in range(len(nb))]
    lf = ("iou", "z_pseudo_gt_patch" if a.patch else "z_pseudo_gt_center", "pose_alignment", "pose_ground", "dims")
    if a.default:
        lf = ("dims", "pose_alignment", "pose_ground", "iou", "z", "z_pseudo_gt_patch", "depth")
        inp = li.add_masks(inp, 7, H=480, W=640)
    # published weights: IOU 4, NORMAL_VEC 40, Z 100, DIMS 0.1, POSE 4 (README.md:185-191)
    kw = dict(loss_functions=lf, loss_w_iou=4.0, loss_w_normal_vec=40.0, loss_w_z=100.0, loss_w_dims=0.1, loss_w_pose=4.0)
    dev = torch.device("cuda", 0)
    _abi.check(_abi.load().wcr_device_check(), "wcr_device_check")
    shard = ImageShard([d.to(dev) for d in inp["depth_maps"]], None, [torch.eye(3)] * len(nb), device=dev)
    leaves = {k: inp[k].to(dev).requires_grad_(True) for k in li.GRAD_KEYS}
    rest = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in inp.items()
            if k not in li.GRAD_KEYS and k != "depth_maps"}
    if a.default:      # masks resident on the device, as masks_all_images is in the reference
        rest["gt_masks"] = torch.stack([m[0] for m in inp["gt_masks"]]).to(dev)
    cfg = WeakLossConfig(**kw)

    def step():
        for v in leaves.values():
            v.grad = None
        losses = weak_loss(cfg, **leaves, **rest, depth_maps=shard, host_sync=not a.no_host_sync)
        sum(losses.values()).backward()
        return losses

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    ms = []
    _abi.reset_launch_count()
    for _ in range(a.steps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); step(); e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    launches = _abi.launch_count() / a.steps
    gpu_losses = {k: float(v) for k, v in step().items()}

    torch.set_num_threads(os.cpu_count() or 1)
    ocfg = ol.LossConfig(**kw)
    cpu, Lc = [], {}
    for _ in range(a.cpu_repeats):
        t0 = time.perf_counter()
        Lc, _, _ = li.run_with_grads(ol.weak_loss, ocfg, inp)
        cpu.append((time.perf_counter() - t0) * 1e3)
    worst = max(abs(gpu_losses[k] - float(Lc[k])) / max(abs(float(Lc[k])), 1e-6) for k in Lc) if cpu else None
    print(json.dumps({
        "metric": "weak-loss fwd+bwd latency", "value": statistics.median(ms), "unit": "ms",
        "higher_is_better": False, "n_gpus": 1, "steps": a.steps, "warmup": a.warmup, "dtype": "f32",
        "data": "synthetic", "config": {"workload": f"weak-loss fwd/bwd, batch 16, {sum(nb)} foreground boxes "
                                                    f"(n_i~U{{1..128}}), 480x640 {'white-noise' if a.noise_depth else 'scene-like'} depth, "
                                                    f"loss_functions={list(lf)}"},
        "gpu_launches_per_step": launches, "ms_min": min(ms), "ms_max": max(ms),
        "includes": "Python wrapper, autograd bookkeeping, " + ("NO host synchronisation (flags checked one call late)"
                                                                if a.no_host_sync else "one 52-byte D2H read of the 'present' flags"),
        "host_sync": not a.no_host_sync,
        "cpu_baseline": {"value": statistics.median(cpu) if cpu else None, "unit": "ms", "cores": torch.get_num_threads(),
                         "kind": "port", "sample": "the same inputs, oracle/loss.py (torch fp32 + autograd) on the host"},
        "max_rel_diff_vs_cpu": worst,
    }), flush=True)


if __name__ == "__main__":
    main()
