"""ShardPipeline (double-buffered host -> device -> host driver) must hand back exactly the
records ShardRunner.run_host() produces for the same host buffers, also when the host buffers
of a slot are rewritten between its uses."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _images(first, n=3, H=240, W=320):
    from wcr import synth
    ims = [synth.make_image(first + i, H, W, n_boxes=3, dummy_ground=False) for i in range(n)]
    for im in ims:
        im["seg"] = synth.seg_masks(im["boxes"], H, W)
    return ims


def test_pipeline_matches_run_host(cuda_lib):
    from wcr import synth
    from wcr.pipeline import ShardRunner, ShardPipeline
    from wcr.proposals import ScoreConfig
    priors = synth.make_priors()
    cfg = ScoreConfig(k=8, e_seg=1)
    base = _images(100)
    batches = []
    for f in (1.0, 1.1, 0.9):                 # same boxes and masks, different depth maps
        ims = [dict(im) for im in base]
        for im in ims:
            im["depth"] = (im["depth"] * f).contiguous()
        batches.append(ims)
    # reference results: one shard at a time
    want = []
    for ims in batches:
        r = ShardRunner(ims, 400, "cuda", priors=priors, score_cfg=cfg, seed=5)
        want.append(r.run_host().clone())
    assert not torch.equal(want[0], want[1])
    slots = [ShardRunner(batches[0], 400, "cuda", priors=priors, score_cfg=cfg, seed=5) for _ in range(2)]
    pipe = ShardPipeline(slots)
    order = [0, 1, 2, 1, 0, 2, 2]
    got = []
    for i, bi in enumerate(order):
        prev = pipe.wait(i - 2)               # slot free: its previous records are on the host
        if prev is not None:
            got.append(prev.clone())
        s = pipe.slot(i)
        for k, im in enumerate(batches[bi]):  # refill the slot's pinned depth plane
            h, w = im["depth"].shape
            s.h_depth[k, :h, :w] = im["depth"]
        pipe.submit(i)
    got.append(pipe.wait(len(order) - 2).clone())
    got.append(pipe.wait(len(order) - 1).clone())
    pipe.drain()
    assert len(got) == len(order)
    for g, bi in zip(got, order):
        assert torch.equal(g, want[bi])
