"""Step 1 parity on the GPU: unprojection + RANSAC vs oracle/ground.py (which is
pinned bit-exactly to the reference, tests/test_oracle_ground.py)."""
import random

import numpy as np
import pytest
import torch

from oracle import ground as og

pytestmark = pytest.mark.gpu


def _scene(i, H=480, W=640, dummy=None):
    from wcr import synth
    return synth.make_image(i, H, W, n_boxes=3, dummy_ground=dummy)


@pytest.mark.parametrize("H,W,nth", [(480, 640, 5), (375, 1242, 5), (97, 133, 3), (1080, 1920, 5)])
def test_unproject_ground_bit_exact(cuda_lib, H, W, nth):
    from wcr.images import ImageShard
    from wcr.ground import unproject_ground
    ims = [_scene(i, H, W, dummy=(i == 1)) for i in range(3)]
    for im in ims:  # non-integer focal exercises the float floor-division quirk
        im["K"] = im["K"].clone(); im["K"][0, 0] *= 1.013; im["K"][1, 1] *= 0.987
    shard = ImageShard([im["depth"].cuda() for im in ims], [im["ground"].cuda() for im in ims],
                       [im["K"] for im in ims])
    pts, count = unproject_ground(shard, nth)
    pts, count = pts.cpu().numpy(), count.cpu().numpy()
    for i, im in enumerate(ims):
        g = None if tuple(im["ground"].shape) == (1, 1) else im["ground"].numpy()
        ref, _ = og.unproject_ground_frame(im["depth"].numpy(), g, im["K"].numpy(), nth)
        assert count[i] == len(ref)
        assert np.array_equal(pts[i, :count[i], :3], ref)  # bit-exact, same order


def test_unproject_cloud_bit_exact(cuda_lib):
    from wcr.images import ImageShard
    from wcr.ground import unproject_cloud
    ims = [_scene(i, 120 + 7 * i, 160 + 5 * i) for i in range(3)]
    ims[0]["depth"][5, 7] = 0.0
    ims[0]["depth"][6, 7] = float("nan")
    for stride in (1, 2):
        shard = ImageShard([im["depth"].cuda() for im in ims], None, [im["K"] for im in ims])
        out = unproject_cloud(shard, stride).cpu().numpy()
        for i, im in enumerate(ims):
            ref = og.unproject_K_frame(im["depth"].numpy(), im["K"].numpy(), stride)
            hs, ws = ref.shape[:2]
            got = out[i, :hs * ws, :3].reshape(hs, ws, 3)
            assert np.array_equal(got, ref, equal_nan=True)


@pytest.mark.parametrize("H,W", [(480, 640), (1080, 1920)])
def test_ransac_replay_bit_exact(cuda_lib, H, W):
    """Same points, same random.sample stream => same counts, same arg-max, same
    plane bits, same inlier set, same up vector."""
    from wcr.images import ImageShard
    from wcr.ground import unproject_ground, ransac_ground, draw_reference_samples
    ims = [_scene(10 + i, H, W, dummy=(i == 2)) for i in range(4)]
    shard = ImageShard([im["depth"].cuda() for im in ims], [im["ground"].cuda() for im in ims],
                       [im["K"] for im in ims])
    pts, count = unproject_ground(shard, 5)
    rng = random.Random(7)
    ids = draw_reference_samples(count.cpu().tolist(), 1000, rng)
    res = ransac_ground(pts, count, ids.cuda(), want_mask=True)
    P = pts.cpu().numpy(); C = count.cpu().numpy()
    assert (C >= 3).sum() >= 2
    for i in range(len(ims)):
        if C[i] < 3:
            assert torch.isnan(res["eq"][i]).all()
            continue
        p = P[i, :C[i], :3]
        eq, inl, counts = og.fit_parallel(p, ids[i].numpy().astype(np.int64))
        assert np.array_equal(res["counts"][i].cpu().numpy(), counts)
        assert int(res["best"][i]) == int(np.argmax(counts))
        assert np.array_equal(res["eq"][i].cpu().numpy(), eq)
        assert np.array_equal(np.flatnonzero(res["mask"][i, :C[i]].cpu().numpy()), inl)
        assert np.array_equal(res["normal"][i].cpu().numpy(), og.fix_up_normal(eq[:3]))


def test_plane_torch_dropin(cuda_lib):
    from wcr.plane import Plane_torch
    torch.manual_seed(0)
    pts = (torch.randn(3000, 3) * torch.tensor([2., 0.03, 3.]) + torch.tensor([0., 1.3, 4.]))
    random.seed(3)
    eq, inl = Plane_torch().fit_parallel(pts.cuda(), thresh=0.05, maxIteration=1000)
    random.seed(3)
    ids = og.draw_id_samples(3000, 1000, random)
    eq_o, inl_o, _ = og.fit_parallel(pts.numpy(), ids)
    assert np.array_equal(eq.cpu().numpy(), eq_o)
    assert np.array_equal(inl.cpu().numpy(), inl_o)
    assert inl.dtype == torch.int64


def test_normal_vector_from_maps_native_rng(cuda_lib):
    """Philox mode: not the reference's stream, so check the estimate, determinism
    and the fix-up invariants instead."""
    from wcr.ground import normal_vector_from_maps
    ims = [_scene(20 + i, dummy=False) for i in range(4)]
    args = ([im["ground"].cuda() for im in ims], [im["depth"].cuda() for im in ims],
            [im["K"] for im in ims])
    n1 = normal_vector_from_maps(*args, seed=5)
    n2 = normal_vector_from_maps(*args, seed=5)
    assert torch.equal(n1, n2)
    n = n1.cpu()
    assert (n[:, 1] > 0.99).all()  # synthetic floor is horizontal: up ~ +y
    assert torch.allclose(n.norm(dim=1), torch.ones(4), atol=1e-5)


def test_degenerate_inputs(cuda_lib):
    from wcr.images import ImageShard
    from wcr.ground import unproject_ground, ransac_ground
    im = _scene(1, 40, 40)
    im["ground"] = torch.zeros(40, 40, dtype=torch.uint8)  # no ground pixel at all
    shard = ImageShard([im["depth"].cuda()], [im["ground"].cuda()], [im["K"]])
    pts, count = unproject_ground(shard, 5)
    assert int(count[0]) == 0
    res = ransac_ground(pts, count, None, iters=100)
    assert torch.isnan(res["eq"]).all()
    with pytest.raises(Exception):
        unproject_ground(ImageShard([im["depth"]], None, [im["K"]]))  # CPU tensor: no fallback
