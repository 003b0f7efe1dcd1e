#!/usr/bin/env python
"""bench.py -- cube proposals scored/sec on synthetic SUN RGB-D-shaped inputs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl wcr|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the hot path over one batch (BASELINE.json configs[1]:
64 images 640x480, ~20 GT boxes/image, 1000 proposals/box) per GPU:
ground-plane RANSAC -> pseudo depth -> fused sample+score+top-k (+ the NCCL
all-gather of the top-k records when N > 1). Prints ONE JSON line on rank 0.

--impl reference times the CPU arm: the reference path's restatement in C
(oracle/, OpenMP over all host cores) on a bounded sample of the same workload.
The reference itself is Python/torch and cannot travel to the GPU box
(/root/reference does not exist there); DESIGN.md "Measurement" says so.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))

IMAGES_PER_GPU = 64
H, W, NPROP, TOPK = 480, 640, 1000, 16
ALG_FLOP_PER_TEST = 18.0      # 3 sub + 9 mul + 6 add per (proposal, point)
EXEC_FLOP_PER_TEST = 12.0     # what the kernel issues: 6 FFMA
FP32_NOMINAL_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.4, SURVEY §8d
METRIC = "cube proposals scored/sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="wcr", choices=["wcr", "reference"])
    ap.add_argument("--images", type=int, default=IMAGES_PER_GPU, help="images per GPU")
    ap.add_argument("--boxes", type=int, default=None, help="GT boxes per image (default: Poisson(20))")
    ap.add_argument("--cpu-sample-images", type=int, default=None,
                    help="images of the workload the CPU arm processes (default: 24 for the "
                         "cpu_baseline leg, 6 per step for --impl reference)")
    ap.add_argument("--no-seg", action="store_true", help="skip the mask-coverage term")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--c4-images", type=int, default=8192,
                    help="global image count of the strong-scaling block (BASELINE configs[3]); 0 = skip")
    ap.add_argument("--c4-passes", type=int, default=3)
    ap.add_argument("--e2e-reserve-ctas", type=int, default=None,
                    help="CTA slots the scoring kernel leaves free for the mask gather of the other pipeline slot "
                         "(default: ShardPipeline's, 18 with mapped dense masks; 0 = gather between the passes)")
    ap.add_argument("--dense-mode", default="mapped", choices=["roi", "rows", "mapped", "copy"],
                    help="how the e2e arm moves the dense per-box masks (ShardRunner dense_seg)")
    return ap.parse_args()


def workload_name(images, seg=True, boxes=None):
    return (f"SUN RGB-D val-shaped batch: {images} images {W}x{H} per GPU, "
            f"{'~20 GT boxes/image (Poisson)' if boxes is None else str(boxes) + ' GT boxes/image'}, "
            f"{NPROP} proposals/box, score = IoU2D x point-in-cube"
            f"{' x mask coverage' if seg else ''}, top-{TOPK}")


def make_images(n, first, seg, n_boxes=None, dense=True):
    """Synthetic images first..first+n-1. With ``dense=False`` the per-box masks are
    bit-packed per image right away (a 1024-image shard would hold 6 GB of dense masks);
    ``dense="both"`` keeps the dense masks beside the packed ones."""
    import numpy as np
    from wcr import synth
    from wcr.proposals import pack_box_masks
    ims = []
    for i in range(n):
        im = synth.make_image(first + i, H, W, n_boxes)
        if seg:
            m = synth.seg_masks(im["boxes"], H, W)
            if dense:
                im["seg"] = m
            if dense is not True:
                b = im["boxes"].numpy()
                rois = np.stack([np.minimum(np.maximum(b[:, 0], 0), W).astype(np.int32),
                                 np.minimum(np.maximum(b[:, 1], 0), H).astype(np.int32),
                                 np.minimum(np.maximum(b[:, 2], 0), W).astype(np.int32),
                                 np.minimum(np.maximum(b[:, 3], 0), H).astype(np.int32)], 1)
                im["seg_packed"] = pack_box_masks(rois, list(m))
        ims.append(im)
    return ims


# ------------------------------------------------------------------ CPU arm --
def cpu_arm(n_images: int, first: int = 0, repeats: int = 1, seg: bool = True, n_boxes=None):
    """Time the oracle's C restatement (OpenMP, all cores) on `n_images` images of
    the workload. Returns (proposals/s, cores, sample string, seconds)."""
    import numpy as np
    import torch
    from oracle import clib, cubes as oc, ground as og
    from wcr import synth
    import random
    clib.set_num_threads()          # torchrun exports OMP_NUM_THREADS=1: ask for every core
    ims = make_images(n_images, first, seg, n_boxes)
    priors = synth.make_priors()
    rng = np.random.default_rng(0)
    n_prop = 0
    t0 = time.perf_counter()
    for _ in range(repeats):
        for im in ims:
            depth = im["depth"].numpy()
            K = im["K"].numpy()
            g = None if tuple(im["ground"].shape) == (1, 1) else im["ground"].numpy()
            pts, _ = og.unproject_ground_frame(depth, g, K, 5)
            if len(pts) >= 3:
                ids = og.draw_id_samples(len(pts), 1000, random.Random(1))
                eq, _, _ = og.fit_parallel(pts, ids, 0.05)
                normal = og.fix_up_normal(eq[:3])
            else:
                normal = np.array([0, 1, 0], dtype=np.float32)
            for b in range(len(im["boxes"])):
                box = im["boxes"][b].numpy()
                z0 = float(oc.box_center_depth(depth, box))
                pm = priors[int(im["classes"][b])].numpy()
                a = np.empty((7, NPROP), dtype=np.float32)
                a[0] = np.clip(rng.normal((box[0] + box[2]) / 2, 0.25 * (box[2] - box[0]), NPROP), box[0], box[2])
                a[1] = np.clip(rng.normal((box[1] + box[3]) / 2, 0.25 * (box[3] - box[1]), NPROP), box[1], box[3])
                a[2] = np.clip(rng.normal(z0, 0.15 * z0, NPROP), 0.5 * z0, 1.5 * z0)
                for d in range(3):
                    a[3 + d] = np.clip(rng.normal(pm[0, d], pm[1, d], NPROP), 0.05, pm[0, d] + 3 * pm[1, d])
                a[6] = rng.uniform(0, np.pi, NPROP)
                cubes = oc.assemble_cubes(K, normal, a)
                r = oc.score_box(depth, K, box, (H, W), cubes, 1, 1, 1, 1 if seg else 0, 0,
                                 im["seg"][b].numpy() if seg else None, fast_seg=True)
                oc.topk(r["score"], TOPK)
                n_prop += NPROP
    dt = time.perf_counter() - t0
    cores = clib.num_threads()
    sample = (f"{n_images} images x{repeats} of the same synthetic workload "
              f"({n_prop // NPROP} boxes, {n_prop} proposals), C restatement with OpenMP")
    return n_prop / dt, cores, sample, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import clib
    clib.build()
    n = max(1, args.cpu_sample_images or 6)
    seg = not args.no_seg
    cpu_arm(1, seg=seg)  # warm
    vals, secs = [], 0.0
    for w in range(args.warmup):
        cpu_arm(1, first=w, seg=seg, n_boxes=args.boxes)
    for s in range(args.steps):
        v, cores, sample, dt = cpu_arm(n, first=s * n, seg=seg, n_boxes=args.boxes)
        vals.append(v); secs += dt
    value = sum(vals) / len(vals)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "proposals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.images, seg, args.boxes), "step_sample": sample},
        "cpu_baseline": {"value": value, "unit": "proposals/s", "cores": cores, "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": "proposals/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def ncu_traffic_bytes():
    """dram__bytes_read.sum + dram__bytes_write.sum of one scoring launch, from the
    newest committed `ncu --set full` summary (profiles/score_tiles_r*.txt); None when
    there is none. A profiler-derived figure, reported beside the event timing."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "score_tiles_r*.txt")))
    if not files:
        return None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, seen = 0.0, 0
    for line in open(files[-1]):
        m = re.match(r"(dram__bytes_(?:read|write)\.sum)\s+([0-9.]+)\s+(\w+)", line)
        if m and m.group(3) in unit:
            total += float(m.group(2)) * unit[m.group(3)]
            seen += 1
    return total if seen == 2 else None


# ------------------------------------------------- C4: strong scaling block --
C4_SEED = 4321


def run_c4(args, dev, rank, world, base_ims, priors, use_seg):
    """BASELINE.json configs[3]: `--c4-images` images (8192) over the N ranks of this run,
    image-sharded in blocks of 64 (wcr.job), one all-gather of the top-k records per pass.
    STRONG scaling: the global image list is the same at every N (tiled from the 64 base
    images of the headline workload through a fixed pseudo-random map, so host generation
    stays bounded while box counts per rank vary). Returns the `c4_strong` dict (rank 0)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from wcr import dist as wdist
    from wcr.job import BlockSet, ResidentBlocks, StreamedBlocks
    from wcr.proposals import ScoreConfig

    cfg = ScoreConfig(k=TOPK, e_seg=1 if use_seg else 0)
    bs = BlockSet(base_ims, args.c4_images, block=IMAGES_PER_GPU, map_seed=0, use_seg=use_seg)
    # blocks -> ranks by predicted work (longest first); contiguous image ranges
    # (wdist.shard_range) leave the slowest rank ~2 % above the mean at 8 GPUs
    rank_blocks = wdist.balanced_assignment(bs.block_weights(NPROP), world) if world > 1 else \
        [list(range(bs.n_blocks))]
    mine = rank_blocks[rank]
    order = torch.from_numpy(bs.global_order(rank_blocks)).to(dev)
    gather = wdist.TopkGather(bs.n_boxes_of(mine), TOPK, dev)
    res = ResidentBlocks(bs, mine, NPROP, dev, priors=priors, score_cfg=cfg, seed=C4_SEED,
                         records=gather.send)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    res.run_pass(); gather.gather_padded()          # warm-up pass
    sync_all()
    P = max(args.c4_passes, 3)
    ev = [[torch.cuda.Event(True) for _ in range(3)] for _ in range(P)]
    for p in range(P):
        sync_all()
        ev[p][0].record(); res.run_pass(); ev[p][1].record(); gather.gather_padded(); ev[p][2].record()
    sync_all()
    pass_ms = sum(e[0].elapsed_time(e[2]) for e in ev) / P
    work_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / P
    gather_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / P
    resident = gather.gather().index_select(0, order)   # records of every rank -> global box order

    # end to end: pinned host planes -> device -> records on the host, double-buffered
    st = StreamedBlocks(bs, mine, NPROP, dev, priors=priors, score_cfg=cfg, seed=C4_SEED,
                        records=gather.send)
    after = gather.gather_padded if world > 1 else None
    st.run_pass(after)
    e2e_s = 0.0
    for p in range(P):
        sync_all()
        t0 = time.perf_counter()
        st.run_pass(after)
        e2e_s += time.perf_counter() - t0
    e2e_s /= P
    same_local = bool(torch.equal(st.h_records[:st.n_boxes].view(torch.int32),
                                  res.records[:res.n_boxes].cpu().view(torch.int32)))

    t = torch.tensor([pass_ms, e2e_s, gather_ms], dtype=torch.float64, device=dev)
    per_rank = torch.tensor([work_ms, float(res.n_boxes), float(same_local)], dtype=torch.float64, device=dev)
    ranks = torch.empty((world, 3), dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_gather_into_tensor(ranks.view(-1), per_rank)
    else:
        ranks[0] = per_rank
    pass_ms, e2e_s, gather_ms = (float(x) for x in t.cpu())
    ranks = ranks.cpu().numpy()

    verified, how = None, None
    if rank == 0:
        if world > 1:
            # the same global list in ONE process (this GPU alone, block by block from the host planes)
            single = StreamedBlocks(bs, range(bs.n_blocks), NPROP, dev, priors=priors, score_cfg=cfg,
                                    seed=C4_SEED)
            want = single.run_pass()
            got = resident.cpu()
            verified = bool(got.shape == want.shape and torch.equal(got.view(torch.int32), want.view(torch.int32)))
            how = (f"TopkGather.gather() of {world} ranks == the single-process records of the same "
                   f"{args.c4_images} images (recomputed on rank 0), compared bit for bit")
            del single
        else:
            verified = bool(ranks[:, 2].all())
            how = "N=1: records of the host-streamed pass == records of the device-resident pass, bit for bit"
    if world > 1:
        dist.barrier()
    if rank != 0:
        return None
    total_boxes = float(ranks[:, 1].sum())
    total_props = total_boxes * NPROP
    return {
        "config": f"BASELINE configs[3]: {args.c4_images} images {W}x{H} over {world} GPU(s), strong scaling, "
                  f"blocks of {IMAGES_PER_GPU} images assigned to ranks by predicted work (longest first), "
                  f"{NPROP} proposals/box, mask term {'on' if use_seg else 'off'}, "
                  f"global list tiled from the 64 base images (fixed map)",
        "scaling": "strong", "images": args.c4_images, "boxes": int(total_boxes), "passes_timed": P,
        "value": total_props / (pass_ms * 1e-3), "unit": "proposals/s",
        "ms_per_pass": pass_ms, "l2": "inputs of a pass (>= 1.5 GB per rank) far exceed L2",
        "e2e": {"value": total_props / e2e_s, "unit": "proposals/s", "s_per_pass": e2e_s,
                "h2d_bytes_per_pass_rank0": st.h2d_bytes, "d2h_bytes_per_pass_rank0": st.d2h_bytes,
                "api": "wcr.job.StreamedBlocks (two slots, three streams)"},
        "allgather_ms": gather_ms, "allgather_bytes_per_rank": int(gather.send.numel() * 4),
        "rank_boxes": {"min": int(ranks[:, 1].min()), "max": int(ranks[:, 1].max())},
        "rank_work_ms": {"min": float(ranks[:, 0].min()), "max": float(ranks[:, 0].max()),
                         "all": [round(float(x), 3) for x in ranks[:, 0]]},
        "imbalance": float(ranks[:, 0].max() / max(ranks[:, 0].mean(), 1e-9)),
        "streamed_equals_resident": bool(ranks[:, 2].all()),
        "gather_verified": verified, "gather_check": how,
    }


# ------------------------------------------------------------------ GPU arm --
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-i", str(self.index), "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        mhz = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names)
                   if any(len(r) > 2 + i and r[2 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": mx[0] if mx else None,
                "samples": len(mhz), "reasons": reasons}


def run_wcr(args):
    import torch
    import torch.distributed as dist
    from wcr import _abi, synth, dist as wdist
    from wcr.pipeline import ShardRunner, ShardPipeline
    from wcr.proposals import ScoreConfig, score_topk, box_center_depth
    from wcr.ground import unproject_cloud

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = wdist.bind_to_gpu_numa(local)   # before any pinned allocation: first touch lands on the GPU's node
    if world > 1:
        # NCCL logs (its version banner included) go to stdout by default: keep stdout for the
        # one JSON line the driver parses
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    lib = _abi.load()                      # raises if the CUDA library is not built
    _abi.check(lib.wcr_device_check(), "wcr_device_check")

    # weak scaling: every rank owns `images` images of the global synthetic set
    use_seg = not args.no_seg
    ims = make_images(args.images, rank * args.images, use_seg, args.boxes, dense="both")
    priors = synth.make_priors()
    n_local_boxes = sum(len(im["boxes"]) for im in ims)
    gather = wdist.TopkGather(n_local_boxes, TOPK, dev)
    runner = ShardRunner(ims, NPROP, dev, priors=priors,
                         score_cfg=ScoreConfig(k=TOPK, e_seg=1 if use_seg else 0), seed=1234,
                         records=gather.send)
    runner.upload()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def one_step(ev=None):
        runner.step(ev)
        if world > 1:
            gather.gather_padded()

    # algorithmic work of one scoring launch: N * sum_b P_b point tests
    runner.step()
    z0 = box_center_depth(runner.shard, runner.boxes)
    full = score_topk(runner.shard, runner.boxes, None, NPROP, runner.score_cfg, priors=runner.priors,
                      normals=runner.normals, z0=z0, seed=1234, seg=runner.seg, want_all=True)
    tests_per_launch = float(full["npts"].sum().item()) * NPROP
    executed_unpruned = float(full["workspace"].executed_tests())     # after the projected-box cull, every proposal tested
    del full

    for _ in range(max(args.warmup, 3)):
        flush.zero_(); one_step()
    torch.cuda.synchronize()

    # ---- device-resident timing (value) + scoring-kernel timing (roofline) ----
    ev = [(torch.cuda.Event(True), torch.cuda.Event(True), torch.cuda.Event(True),
           torch.cuda.Event(True)) for _ in range(args.steps)]
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    if rank == 0:
        sampler.start()
    _abi.reset_launch_count()
    for s in range(args.steps):
        flush.zero_()                      # L2 flush between timed iterations (untimed)
        ev[s][0].record()
        one_step((ev[s][2], ev[s][3]))
        ev[s][1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches = _abi.launch_count()
    hulls = runner.ws.hulls_filled() if use_seg else 0
    # what the TIMED configuration issued (point term on demand: pruned proposals are never tested)
    executed_per_launch = float(runner.ws.executed_tests())
    pruned = runner.ws.proposals_pruned()
    step_ms = sum(e[0].elapsed_time(e[1]) for e in ev)
    score_ms = sum(e[2].elapsed_time(e[3]) for e in ev) / args.steps

    # ---- end-to-end through the public API with host buffers ------------------
    # (a) one shard at a time: upload -> pass -> download -> host sync (ShardRunner.run_host)
    for _ in range(2):
        runner.run_host()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(args.steps):
        runner.run_host()
        if world > 1:
            gather.gather_padded()
            torch.cuda.synchronize()
    e2e_sync_s = time.perf_counter() - t0
    # (b) the same work double-buffered (ShardPipeline): shard s + 1 is uploaded and the records
    # of shard s - 1 are downloaded while shard s is scored. Every step still copies all of its
    # inputs from pinned host memory and its records back.
    def pipelined(r_a, r_b, g_a, g_b):
        reserve = args.e2e_reserve_ctas
        pipe = ShardPipeline([r_a, r_b], after_step=[g_a.gather_padded, g_b.gather_padded] if world > 1 else None,
                             reserve_ctas=reserve)
        for s in range(4):
            pipe.wait(s - 2); pipe.submit(s)
        pipe.drain()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for s in range(args.steps):
            pipe.wait(s - 2)
            pipe.submit(s)
        pipe.wait(args.steps - 2); pipe.wait(args.steps - 1)
        pipe.drain()
        return time.perf_counter() - t0

    gather2 = wdist.TopkGather(n_local_boxes, TOPK, dev)
    runner2 = ShardRunner(ims, NPROP, dev, priors=priors, score_cfg=runner.score_cfg, seed=1234,
                          records=gather2.send)
    e2e_packed_s = pipelined(runner, runner2, gather, gather2)
    # (c) the headline e2e: the per-box masks arrive DENSE, as the reference's object_masks()
    # produces them (roi_heads.py:383-401), in pinned host memory; inside the timed region the ROI
    # bytes of every box's map cross PCIe (h2d_bytes_per_step counts them) and are bit-packed on the
    # device: by default the pack kernel gathers them straight from the mapped pinned maps (128-byte
    # reads, 32 GB/s; --dense-mode roi = one 2-D copy per box, 6 GB/s: 1264 small copies)
    e2e_s, dense_h2d, e2e_dense_sync_s = e2e_packed_s, None, None
    if use_seg:
        gather3 = wdist.TopkGather(n_local_boxes, TOPK, dev)
        rd = [ShardRunner(ims, NPROP, dev, priors=priors, score_cfg=runner.score_cfg, seed=1234,
                          records=g.send, dense_seg=args.dense_mode) for g in (gather2, gather3)]
        dense_h2d = rd[0].h2d_bytes
        e2e_s = pipelined(rd[0], rd[1], gather2, gather3)
        for _ in range(2):
            rd[0].run_host()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for s in range(args.steps):
            rd[0].run_host()
            if world > 1:
                gather2.gather_padded()
                torch.cuda.synchronize()
        e2e_dense_sync_s = time.perf_counter() - t0
        same = torch.equal(rd[0].h_records.view(torch.int32), runner.h_records.view(torch.int32))
        assert same, "dense-mask path and pre-packed path disagree"
        del rd, gather3
    del runner2, gather2
    clocks = sampler.stop() if rank == 0 else None

    # ---- FP32 probe + HBM-bound unprojection kernel (aux rooflines) -----------
    probe_out = torch.empty(148 * 8 * 256 * 2, dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    best = 0.0
    for _ in range(5):
        e0.record()
        flops = lib.wcr_probe_ffma(probe_out.data_ptr(), probe_out.numel(), 4096, _abi.stream_ptr(dev))
        e1.record(); torch.cuda.synchronize()
        best = max(best, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    fp32_probe_tflops = best
    cloud_ms = []
    for _ in range(5):
        flush.zero_(); e0.record(); unproject_cloud(runner.shard, 1); e1.record()
        torch.cuda.synchronize(); cloud_ms.append(e0.elapsed_time(e1))
    cloud_bytes = args.images * H * W * (4 + 16)
    cloud_gbs = cloud_bytes / (min(cloud_ms) * 1e-3) / 1e9

    t = torch.tensor([step_ms, e2e_s, score_ms, e2e_sync_s, e2e_packed_s, e2e_dense_sync_s or 0.0],
                     dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_ms, e2e_s, score_ms, e2e_sync_s, e2e_packed_s, e2e_dense_sync_s = (float(x) for x in t.cpu())
    props = torch.tensor([runner.n_proposals], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(props)
    total_props = float(props.item())

    c4 = None
    if args.c4_images > 0:
        del flush
        base = ims if (rank == 0 and args.images == IMAGES_PER_GPU and args.boxes is None) else \
            make_images(IMAGES_PER_GPU, 0, use_seg, None, dense=False)   # the same base set on every rank
        c4 = run_c4(args, dev, rank, world, base, priors, use_seg)

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        achieved = ALG_FLOP_PER_TEST * tests_per_launch / (score_ms * 1e-3) / 1e12
        line = {
            "metric": METRIC, "value": total_props * args.steps / (step_ms * 1e-3),
            "unit": "proposals/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": step_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.images, use_seg, args.boxes), "l2": "256 MiB flush between timed steps",
                       "parallelism": f"image-sharded x{world}, all-gather of top-k records",
                       "boxes_rank0": runner.n_boxes, "rng": "in-kernel Philox4x32-10"},
            "e2e": {"value": total_props * args.steps / e2e_s, "unit": "proposals/s",
                    "h2d_bytes_per_step": dense_h2d if dense_h2d is not None else runner.h2d_bytes,
                    "d2h_bytes_per_step": runner.d2h_bytes,
                    "api": "wcr.pipeline.ShardPipeline (double-buffered: copies of neighbouring steps overlap the pass)",
                    "masks": ("dense per-box masks (n_boxes x H x W u8, as object_masks() yields them) in pinned host memory; "
                              + {"mapped": "every step the pack kernel reads the ROI bytes of each box's map from the mapped "
                                           "host memory over PCIe (on the upload stream) and writes the bit-packed words",
                                 "roi": "every step the ROI rectangle of each box's map is copied H2D (one 2-D copy per box) and "
                                        "bit-packed on the device",
                                 "rows": "every step one contiguous copy per box (first to last ROI pixel) H2D, bit-packed on the device",
                                 "copy": "every step the whole maps are copied H2D and bit-packed on the device"}[args.dense_mode]
                              if dense_h2d is not None else "no mask term"),
                    "dense_mode": args.dense_mode,
                    "dense_mask_bytes_on_host": (int(runner.n_boxes) * H * W) if dense_h2d is not None else 0,
                    "one_shard_at_a_time": (total_props * args.steps / e2e_dense_sync_s) if e2e_dense_sync_s
                    else total_props * args.steps / e2e_sync_s,
                    "prepacked": {"value": total_props * args.steps / e2e_packed_s,
                                  "h2d_bytes_per_step": runner.h2d_bytes,
                                  "one_shard_at_a_time": total_props * args.steps / e2e_sync_s,
                                  "masks": "bit-packed on the host before the timed region (loader-side)"}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "numa_rank0": numa,
            "roofline": {"bound": "fp32", "kernel": "score_tiles_kernel", "achieved": achieved,
                         "peak": fp32_probe_tflops, "unit": "TFLOP/s",
                         "frac": achieved / fp32_probe_tflops,
                         "peak_source": "FFMA probe kernel timed in this run (MEASURED_PEAKS.json has no FP32 figure)",
                         "frac_of_nominal_74.4": achieved / FP32_NOMINAL_TFLOPS,
                         "point_tests_per_s": tests_per_launch / (score_ms * 1e-3),
                         "alg_flop_per_test": ALG_FLOP_PER_TEST,
                         "executed_flop_per_test": EXEC_FLOP_PER_TEST,
                         "tests_executed_frac": executed_per_launch / max(tests_per_launch, 1.0),
                         "tests_executed_frac_every_proposal": executed_unpruned / max(tests_per_launch, 1.0),
                         "proposals_pruned_frac": pruned / max(runner.n_proposals, 1),
                         "frac_executed": EXEC_FLOP_PER_TEST * executed_per_launch / (score_ms * 1e-3) / 1e12 / fp32_probe_tflops,
                         "hulls_filled_frac": hulls / max(runner.n_proposals, 1),
                         "kernel_ms": score_ms,
                         "kernel_share_of_step": score_ms * args.steps / step_ms,
                         "traffic": ncu_traffic_bytes()},
            "roofline_aux": {"bound": "hbm", "kernel": "unproject_cloud_kernel", "achieved": cloud_gbs,
                             "peak": hbm_peak, "unit": "GB/s", "frac": cloud_gbs / hbm_peak,
                             "bytes_per_launch": cloud_bytes},
        }
        if c4 is not None:
            line["c4_strong"] = c4
        if not args.no_cpu_baseline and world == 1:
            from oracle import clib
            clib.build()
            cpu_arm(1, seg=use_seg, n_boxes=args.boxes)
            v, cores, sample, dt = cpu_arm(min(args.cpu_sample_images or 24, max(args.images, 1)), seg=use_seg,
                                           n_boxes=args.boxes)
            line["cpu_baseline"] = {"value": v, "unit": "proposals/s", "cores": cores,
                                    "kind": "port", "sample": sample, "seconds": dt}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line: dict) -> None:
    """The ONE line of stdout. Everything else a run prints -- NCCL's version banner is a bare
    printf to fd 1 -- has been pointed at stderr by quiet_stdout()."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def quiet_stdout() -> None:
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


if __name__ == "__main__":
    quiet_stdout()
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_wcr(a)
