"""Full-size cross-check (BASELINE configs[1]: 64 images 640x480, ~1264 boxes, N = 1000): the
culled tile kernel and the first, unculled thread-per-proposal kernel (WCR_SCORE_VARIANT=4, run
in a child process because the variant is latched per process) must agree on every one of the
1.26 M point-in-cube counts and on every top-k record, bit for bit. Two independent
implementations of the same definition; the oracle pins that definition at small sizes
(test_score_gpu.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, {root!r}); sys.path.insert(0, {root!r} + "/weak-cube-r-cnn_b200")
from tests.test_score_fullsize_gpu import run_case
c, r = run_case()
np.savez({out!r}, count=c, records=r)
"""


def run_case(n_images=64, N=1000):
    from wcr import synth
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, ScoreConfig, score_topk
    ims = [synth.make_image(i, 480, 640) for i in range(n_images)]
    shard = ImageShard([im["depth"].cuda() for im in ims], [im["ground"].cuda() for im in ims],
                       [im["K"] for im in ims])
    boxes = BoxTable(shard, [im["boxes"] for im in ims], [im["classes"] for im in ims])
    normals = torch.tensor([[0.0, 1.0, 0.0]] * n_images)
    res = score_topk(shard, boxes, None, N, ScoreConfig(k=16), priors=synth.make_priors(),
                     normals=normals, seed=77, want_all=True)
    torch.cuda.synchronize()
    return res["count"].cpu().numpy(), res["records"].cpu().numpy()


def test_tile_kernel_equals_unculled_kernel_at_full_size(cuda_lib, tmp_path):
    assert os.environ.get("WCR_SCORE_VARIANT") in (None, ""), "run this test with the default kernel"
    out = str(tmp_path / "v4.npz")
    env = dict(os.environ, WCR_SCORE_VARIANT="4")
    subprocess.run([sys.executable, "-c", CHILD.format(root=ROOT, out=out)], check=True, env=env,
                   cwd=ROOT, timeout=600)
    ref = np.load(out)
    count, records = run_case()
    assert count.shape[0] > 1000 and count.sum() > 10 ** 8
    assert np.array_equal(count, ref["count"])
    assert np.array_equal(records.view(np.uint32), ref["records"].view(np.uint32))
