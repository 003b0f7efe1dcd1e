"""Full-size checks (BASELINE configs[1]: 64 images 640x480, ~1264 boxes, N = 1000).

1. The culled tile kernel and the first, unculled thread-per-proposal kernel must agree on every
   one of the 1.26 M point-in-cube counts and on every top-k record, bit for bit. The unculled
   kernel lives in its own library, libwcr_b200_baseline.so (the same sources compiled with
   -DWCR_KEEP_BASELINE_KERNEL, build.py), loaded by a child process: two implementations of the
   same definition.
2. 32 randomly chosen boxes of the same run are also scored by the C oracle
   (oracle/csrc/wcr_oracle.c, which never culls): counts, mask pixel ratios and the top-16 of
   the benchmarked configuration (mask term on) against the oracle itself, at the benchmarked
   size."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import cubes as oc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/weak-cube-r-cnn_b200")
from wcr import _abi
_abi.LIB_PATH = _abi.LIB_PATH.replace("libwcr_b200.so", "libwcr_b200_baseline.so")
from tests.test_score_fullsize_gpu import run_case
c, r = run_case()[:2]
np.savez({out!r}, count=c, records=r)
"""


def run_case(n_images=64, N=1000, seg=False, keep=False):
    from wcr import synth
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, ScoreConfig, SegMasks, score_topk, propose
    ims = [synth.make_image(i, 480, 640) for i in range(n_images)]
    shard = ImageShard([im["depth"].cuda() for im in ims], [im["ground"].cuda() for im in ims],
                       [im["K"] for im in ims])
    boxes = BoxTable(shard, [im["boxes"] for im in ims], [im["classes"] for im in ims])
    normals = torch.tensor([[0.0, 1.0, 0.0]] * n_images)
    priors = synth.make_priors()
    if not seg:
        res = score_topk(shard, boxes, None, N, ScoreConfig(k=16), priors=priors,
                         normals=normals, seed=77, want_all=True)
        torch.cuda.synchronize()
        return res["count"].cpu().numpy(), res["records"].cpu().numpy()
    masks = [m for im in ims for m in synth.seg_masks(im["boxes"], 480, 640)]
    segm = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, priors, normals, N, seed=77)
    cfg = ScoreConfig(k=16, e_seg=1)
    full = score_topk(shard, boxes, cubes, cfg=cfg, seg=segm, want_all=True)
    rec_full = full["records"].clone()
    lazy = score_topk(shard, boxes, cubes, cfg=cfg, seg=segm)      # the benchmarked path: mask term on demand
    torch.cuda.synchronize()
    return ims, masks, cubes.cpu().numpy(), full, rec_full.cpu().numpy(), lazy["records"].cpu().numpy()


def test_tile_kernel_equals_unculled_kernel_at_full_size(cuda_lib, tmp_path):
    out = str(tmp_path / "v4.npz")
    subprocess.run([sys.executable, "-c", CHILD.format(root=ROOT, out=out)], check=True,
                   cwd=ROOT, timeout=600)
    ref = np.load(out)
    count, records = run_case()
    assert count.shape[0] > 1000 and count.sum() > 10 ** 8
    assert np.array_equal(count, ref["count"])
    assert np.array_equal(records.view(np.uint32), ref["records"].view(np.uint32))


def test_full_size_sample_of_boxes_against_the_oracle(cuda_lib):
    ims, masks, cubes, full, rec_full, rec_lazy = run_case(seg=True)
    n_boxes = cubes.shape[0]
    assert n_boxes > 1000
    # on-demand mask term == full evaluation, every record of every box, bit for bit
    assert np.array_equal(rec_full.view(np.uint32), rec_lazy.view(np.uint32))
    owner = np.concatenate([[i] * len(im["boxes"]) for i, im in enumerate(ims)])
    first = np.concatenate([[0], np.cumsum([len(im["boxes"]) for im in ims])])
    count = full["count"].cpu().numpy(); segr = full["seg"].cpu().numpy()
    score = full["scores"].cpu().numpy(); npts = full["npts"].cpu().numpy()
    for gb in np.random.default_rng(5).choice(n_boxes, 32, replace=False):
        im = ims[owner[gb]]
        lb = gb - first[owner[gb]]
        ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), im["boxes"][lb].numpy(), (480, 640),
                           cubes[gb], 1, 1, 1, 1, 0, masks[gb].numpy(), fast_seg=True)
        assert np.array_equal(count[gb], ref["count"]), gb
        assert int(npts[gb]) == int(ref["npts"])
        assert np.array_equal(segr[gb], ref["seg"]), gb
        assert np.array_equal(score[gb], ref["score"]), gb
        want = oc.topk(ref["score"], 16)
        assert np.array_equal(rec_lazy[gb, :, 0].view(np.int32), want), gb
        assert np.array_equal(rec_lazy[gb, :, 1], ref["score"][want]), gb
