"""Multi-GPU result check (VERDICT r1 missing #1): the all-gathered top-k records of N ranks
equal the single-process records of the same global image list, bit for bit. The 2-rank case
needs two GPUs (skipped on a 1-GPU box); the 1-GPU cases check the same invariance by running
the two halves separately, and pin wcr.job to the existing ShardRunner path."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _base(n=5, H=240, W=320):
    from wcr import synth
    base = []
    for i in range(n):
        im = synth.make_image(60 + i, H + 8 * (i % 2), W, n_boxes=1 + i % 3, dummy_ground=(i == 1))
        im["seg"] = synth.seg_masks(im["boxes"], H + 8 * (i % 2), W)
        base.append(im)
    return base


def test_blocks_independent_of_sharding_and_match_shard_runner(cuda_lib):
    from wcr import synth
    from wcr.job import BlockSet, ResidentBlocks, StreamedBlocks
    from wcr.pipeline import ShardRunner
    from wcr.proposals import ScoreConfig
    base = _base()
    bs = BlockSet(base, 22, block=4, map_seed=1)         # 6 blocks, the last holds 2 images
    cfg = ScoreConfig(k=8, e_seg=1)
    priors = synth.make_priors()
    kw = dict(priors=priors, score_cfg=cfg, seed=9)
    whole = ResidentBlocks(bs, range(6), 250, "cuda", **kw).run_pass().clone()
    parts = [ResidentBlocks(bs, r, 250, "cuda", **kw).run_pass().clone() for r in (range(0, 2), range(2, 3), range(3, 6))]
    assert torch.equal(torch.cat(parts).view(torch.int32), whole.view(torch.int32))
    st = StreamedBlocks(bs, range(6), 250, "cuda", **kw)
    for _ in range(2):                                   # slots are reused across passes
        got = st.run_pass()
        assert torch.equal(got.view(torch.int32), whole.cpu().view(torch.int32))
    assert st.h2d_bytes > 22 * 240 * 320 * 4 and st.d2h_bytes == whole.numel() * 4
    # block b alone through the existing one-shard driver, seed + b
    for b in (0, 5):
        ims = [base[j] for j in bs.images_of(b)]
        r = ShardRunner(ims, 250, "cuda", priors=priors, score_cfg=cfg, seed=9 + b)
        want = r.run_host()
        lo, hi = int(bs.box_offset[b]), int(bs.box_offset[b + 1])
        assert torch.equal(want.view(torch.int32), whole[lo:hi].cpu().view(torch.int32))


def test_gathered_records_equal_single_process_2gpu(cuda_lib):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    port = 29600 + os.getpid() % 300
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port),
           os.path.join(ROOT, "tests", "dist_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    assert "DIST_OK world=2" in p.stdout, p.stdout[-2000:] + p.stderr[-2000:]
