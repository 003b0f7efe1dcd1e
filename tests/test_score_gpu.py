"""Steps 2+3 parity on the GPU, through the C ABI, against oracle/cubes.py.

Bars (north_star): point-in-cube counts and top-k indices bit-exact against the
oracle's stated rounding order; IoU / scores within 1e-5 relative (here they are
bit-exact too, the assertion still states the tolerance)."""
import numpy as np
import pytest
import torch

from oracle import clib, cubes as oc

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def _setup(n_images, H=480, W=640, n_boxes=4, first=0, clamp=None):
    from wcr import synth
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable
    ims = [synth.make_image(first + i, H, W, n_boxes=n_boxes, dummy_ground=False)
           for i in range(n_images)]
    shard = ImageShard([im["depth"].cuda() for im in ims], [im["ground"].cuda() for im in ims],
                       [im["K"] for im in ims])
    boxes = BoxTable(shard, [im["boxes"] for im in ims], [im["classes"] for im in ims],
                     clamp=None if clamp is None else [clamp] * n_images)
    normals = torch.tensor([[0.01 * i, 1.0, -0.02 * i] for i in range(n_images)])
    normals = normals / normals.norm(dim=1, keepdim=True)
    return ims, shard, boxes, synth.make_priors(), normals


def test_philox_known_answers(cuda_lib):
    """Random123 kat_vectors for philox4x32-10 pin the device generator."""
    from wcr import _abi
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    out = torch.zeros(4, dtype=torch.int32, device="cuda")
    for ctr, key, want in kat:
        _abi.check(cuda_lib.wcr_philox_block(key[0] | (key[1] << 32), *ctr, out.data_ptr(),
                                             _abi.stream_ptr()), "wcr_philox_block")
        got = out.cpu().numpy().view(np.uint32)
        assert tuple(int(x) for x in got) == want
        assert tuple(int(x) for x in clib.philox4x32_10(ctr, key)) == want


def test_box_center_depth_bit_exact(cuda_lib):
    from wcr.proposals import box_center_depth
    ims, shard, boxes, _, _ = _setup(3, n_boxes=6)
    z0 = box_center_depth(shard, boxes).cpu().numpy()
    ref = [oc.box_center_depth(im["depth"].numpy(), im["boxes"][b].numpy())
           for im in ims for b in range(len(im["boxes"]))]
    assert np.array_equal(z0, np.array(ref, dtype=np.float32))


def _replay_attrs(ims, priors, N, seed):
    """Plausible (u v z w h l yaw) samples per box, host-drawn (stands in for the
    reference sampler here; tests/test_compose_ref.py uses the real one)."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for im in ims:
        for b in range(len(im["boxes"])):
            x1, y1, x2, y2 = im["boxes"][b].tolist()
            pm = priors[int(im["classes"][b])]
            z0 = float(oc.box_center_depth(im["depth"].numpy(), im["boxes"][b].numpy()))
            a = torch.empty(7, N)
            a[0] = (x1 + x2) / 2 + 0.25 * (x2 - x1) * torch.randn(N, generator=g)
            a[1] = (y1 + y2) / 2 + 0.25 * (y2 - y1) * torch.randn(N, generator=g)
            a[2] = z0 * (1 + 0.15 * torch.randn(N, generator=g)).clamp(0.5, 1.5)
            for d in range(3):
                a[3 + d] = (pm[0, d] + pm[1, d] * torch.randn(N, generator=g)).clamp(min=0.05)
            a[6] = np.pi * torch.rand(N, generator=g)
            out.append(a)
    return torch.stack(out)


def test_sample_replay_matches_oracle(cuda_lib):
    from wcr.proposals import propose
    N = 257
    ims, shard, boxes, priors, normals = _setup(2, n_boxes=3)
    replay = _replay_attrs(ims, priors, N, 5)
    cubes = propose(shard, boxes, priors, normals, N, replay=replay.cuda()).cpu().numpy()
    gb = 0
    for i, im in enumerate(ims):
        for b in range(len(im["boxes"])):
            ref = oc.assemble_cubes(im["K"].numpy(), normals[i].numpy(), replay[gb].numpy())
            # centres / dims bit-exact; R within 2 ulp of libm's sin/cos
            assert np.array_equal(cubes[gb][:, :6], ref[:, :6])
            np.testing.assert_allclose(cubes[gb][:, 6:], ref[:, 6:], rtol=0, atol=3e-7)
            gb += 1


def test_sample_native_deterministic_and_in_range(cuda_lib):
    from wcr.proposals import propose, ProposalConfig, box_center_depth
    N = 1000
    ims, shard, boxes, priors, normals = _setup(2, n_boxes=5)
    cfg = ProposalConfig()
    c1 = propose(shard, boxes, priors, normals, N, cfg, seed=11)
    c2 = propose(shard, boxes, priors, normals, N, cfg, seed=11)
    c3 = propose(shard, boxes, priors, normals, N, cfg, seed=12)
    assert torch.equal(c1, c2) and not torch.equal(c1, c3)
    c = c1.cpu()
    z0 = box_center_depth(shard, boxes).cpu()
    gb = 0
    for i, im in enumerate(ims):
        for b in range(len(im["boxes"])):
            z = c[gb, :, 2]
            assert (z >= z0[gb] * cfg.z_lo - 1e-6).all() and (z <= z0[gb] * cfg.z_hi + 1e-6).all()
            pm = priors[int(im["classes"][b])]
            d = c[gb, :, 3:6]
            assert (d >= cfg.dim_lo).all() and (d <= pm[0] + cfg.dim_hi_sigmas * pm[1] + 1e-6).all()
            # sample mean of the dims near the prior mean (truncation is mild)
            assert torch.allclose(d.mean(0), pm[0], rtol=0.08)
            R = c[gb, :, 6:].reshape(N, 3, 3)
            eye = torch.eye(3).expand(N, 3, 3)
            assert torch.allclose(R.transpose(1, 2) @ R, eye, atol=2e-6)       # orthonormal
            assert torch.allclose(torch.linalg.det(R), -torch.ones(N), atol=1e-5)  # det = -1 (A2.4)
            assert torch.allclose(R[:, :, 1], normals[i].expand(N, 3), atol=1e-7)  # column 1 = normal
            gb += 1


@pytest.mark.parametrize("N,H,W,stride", [(1000, 480, 640, 1), (333, 375, 1242, 2), (2500, 240, 320, 1)])
def test_score_topk_bit_exact(cuda_lib, N, H, W, stride):
    from wcr.proposals import propose, score_topk, ScoreConfig
    ims, shard, boxes, priors, normals = _setup(2, H, W, n_boxes=3, first=40)
    cubes = propose(shard, boxes, priors, normals, N, seed=3)
    cfg = ScoreConfig(point_stride=stride, k=16)
    res = score_topk(shard, boxes, cubes, cfg=cfg, want_all=True)
    cubes_h = cubes.cpu().numpy()
    gb = 0
    for i, im in enumerate(ims):
        for b in range(len(im["boxes"])):
            ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), im["boxes"][b].numpy(),
                               (H, W), cubes_h[gb], stride)
            assert int(res["npts"][gb]) == ref["npts"]
            assert np.array_equal(res["count"][gb].cpu().numpy(), ref["count"])      # bit-exact
            np.testing.assert_allclose(res["proj"][gb].cpu().numpy(), ref["proj"], rtol=RTOL)
            np.testing.assert_allclose(res["iou"][gb].cpu().numpy(), ref["iou"], rtol=RTOL)
            np.testing.assert_allclose(res["scores"][gb].cpu().numpy(), ref["score"], rtol=RTOL)
            want = oc.topk(res["scores"][gb].cpu().numpy(), 16)
            assert np.array_equal(res["idx"][gb].cpu().numpy(), want)                 # bit-exact
            assert np.array_equal(oc.topk(ref["score"], 16), want)
            rec = res["records"][gb].cpu().numpy()
            for r, j in enumerate(want):
                assert np.array_equal(rec[r, 2:17], cubes_h[gb][j])
                assert rec[r, 1] == res["scores"][gb, j].item()
            assert ref["count"].max() > 0   # the test actually exercises the counter
            gb += 1


def test_score_fused_sampling_equals_two_step(cuda_lib):
    """cubes == NULL (a proposal is regenerated from its Philox counter wherever the launch needs
    it -- prep kernel, hull fills, record writer -- and never stored) must select the same top-k,
    with the same records, as sampling to HBM first (wcr_sample_cubes) and scoring that array."""
    from wcr.proposals import propose, score_topk, ScoreConfig
    N = 1000
    ims, shard, boxes, priors, normals = _setup(3, n_boxes=4, first=7)
    cubes = propose(shard, boxes, priors, normals, N, seed=21)
    a = score_topk(shard, boxes, cubes, cfg=ScoreConfig())
    b = score_topk(shard, boxes, None, N, cfg=ScoreConfig(), priors=priors, normals=normals, seed=21)
    assert torch.equal(a["records"], b["records"])


def test_score_edge_cases(cuda_lib):
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, score_topk, ScoreConfig
    from wcr import synth
    im = synth.make_image(3, 120, 160, n_boxes=1, dummy_ground=False)
    im["depth"][:60] = 0.0          # invalid depth rows
    im["depth"][70, 5] = float("nan")
    boxes_t = torch.tensor([[-20.0, -5.0, 50.0, 90.0],    # partly outside the image
                            [30.2, 40.7, 31.1, 41.3],     # empty integer ROI
                            [10.0, 0.0, 150.0, 59.0]])    # only invalid depth -> P_b = 0
    shard = ImageShard([im["depth"].cuda()], None, [im["K"]])
    boxes = BoxTable(shard, [boxes_t], [torch.zeros(3, dtype=torch.long)])
    g = torch.Generator().manual_seed(0)
    N = 40
    cubes = torch.zeros(3, N, 15)
    cubes[:, :, 0:2] = torch.randn(3, N, 2, generator=g) * 0.5
    cubes[:, :, 2] = 1.0 + torch.rand(3, N, generator=g) * 3
    cubes[:, :, 3:6] = 0.2 + torch.rand(3, N, 3, generator=g)
    cubes[:, :, 6:] = torch.eye(3).reshape(9)
    cubes[0, 0, 2] = 0.0            # centre at z = 0: corners straddle the camera plane
    cubes[0, 1, 3:6] = 0.0          # zero-size cube -> 0/0 nowhere, but a point box
    res = score_topk(shard, boxes, cubes.cuda(), cfg=ScoreConfig(k=32), want_all=True)
    for b in range(3):
        ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), boxes_t[b].numpy(), (120, 160),
                           cubes[b].numpy())
        assert int(res["npts"][b]) == ref["npts"]
        assert np.array_equal(res["count"][b].cpu().numpy(), ref["count"])
        np.testing.assert_allclose(res["scores"][b].cpu().numpy(), ref["score"], rtol=RTOL)
        np.testing.assert_array_equal(np.isnan(res["proj"][b].cpu().numpy()), np.isnan(ref["proj"]))
        assert np.array_equal(res["idx"][b].cpu().numpy(), oc.topk(ref["score"], 32))
    assert int(res["npts"][1]) == 0 and int(res["npts"][2]) == 0
    # k > N: the tail is padded with idx = -1
    res2 = score_topk(shard, boxes, cubes[:, :5].contiguous().cuda(), cfg=ScoreConfig(k=8))
    assert (res2["idx"][:, 5:] == -1).all() and (res2["idx"][:, :5] >= 0).all()
    with pytest.raises(Exception):
        score_topk(shard, boxes, cubes, cfg=ScoreConfig())   # CPU tensor: no fallback


@pytest.mark.parametrize("n_images,n_boxes", [(3, 20), (6, 20)])
def test_score_mid_size_shards(cuda_lib, n_images, n_boxes):
    """The tile kernel sizes its items per launch (128 ... 1024 proposals per tile item, from the
    number of boxes): 60 and 120 boxes land on the two middle sizes (256, 512); the small and the
    full-size cases are covered by the other tests. Counts and top-k bit-exact vs the oracle."""
    from wcr.proposals import propose, score_topk, ScoreConfig
    H, W, N = 240, 320, 700
    ims, shard, boxes, priors, normals = _setup(n_images, H, W, n_boxes=n_boxes, first=60)
    cubes = propose(shard, boxes, priors, normals, N, seed=9)
    res = score_topk(shard, boxes, cubes, cfg=ScoreConfig(k=16), want_all=True)
    executed = res["workspace"].executed_tests()
    cubes_h = cubes.cpu().numpy()
    gb, total = 0, 0
    for im in ims:
        for b in range(len(im["boxes"])):
            ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), im["boxes"][b].numpy(), (H, W), cubes_h[gb])
            assert np.array_equal(res["count"][gb].cpu().numpy(), ref["count"]), gb
            assert np.array_equal(res["idx"][gb].cpu().numpy(), oc.topk(ref["score"], 16))
            total += ref["npts"] * N
            gb += 1
    assert gb == n_images * n_boxes
    assert 0 < executed < total          # the cull removed work, and the counter saw the rest


@pytest.mark.parametrize("stride", [1, 3])
def test_score_culling_adversarial(cuda_lib, stride):
    """The tile path skips every (proposal, pixel) pair the proposal's projected box cannot
    reach. The oracle never culls, so bit-exact counts on cubes chosen to break a careless
    cull (arbitrary rotations, corners behind / on / just in front of the camera plane,
    cubes that swallow the camera, sliver cubes, boxes of every tile shape) prove the cull
    is conservative."""
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, score_topk, ScoreConfig
    from wcr import synth
    H, W = 300, 500
    im = synth.make_image(11, H, W, n_boxes=1, dummy_ground=False)
    im["depth"][::7, ::5] *= 0.1                      # a lattice of very near points
    boxes_t = torch.tensor([[3.0, 2.0, 497.0, 41.0],      # wide and flat  -> 8-row tiles
                            [10.5, 5.0, 27.0, 299.0],     # narrow and tall -> 32-row tiles
                            [100.0, 60.0, 330.0, 250.0],  # large
                            [200.2, 100.1, 233.9, 117.5],  # small
                            [0.0, 0.0, 500.0, 300.0]])    # the whole image
    nb = len(boxes_t)
    shard = ImageShard([im["depth"].cuda()], None, [im["K"]])
    boxes = BoxTable(shard, [boxes_t], [torch.zeros(nb, dtype=torch.long)])
    g = torch.Generator().manual_seed(5)
    N = 700
    f = float(im["K"][0, 0])
    cubes = torch.zeros(nb, N, 15)
    # centres lifted from random pixels (inside and outside the box) at random depths
    u = torch.rand(nb, N, generator=g) * 1.4 * W - 0.2 * W
    v = torch.rand(nb, N, generator=g) * 1.4 * H - 0.2 * H
    zc = torch.exp(torch.rand(nb, N, generator=g) * 5.0 - 2.5)            # 0.08 .. 12 m
    cubes[:, :, 0] = zc * (u - W / 2) / f
    cubes[:, :, 1] = zc * (v - H / 2) / f
    cubes[:, :, 2] = zc
    cubes[:, :, 3:6] = torch.exp(torch.rand(nb, N, 3, generator=g) * 6.0 - 4.0)   # 0.02 .. 7.4 m
    q = torch.randn(nb, N, 4, generator=g)
    q = q / q.norm(dim=-1, keepdim=True)
    w_, x_, y_, z_ = q.unbind(-1)
    R = torch.stack([1 - 2 * (y_ * y_ + z_ * z_), 2 * (x_ * y_ - z_ * w_), 2 * (x_ * z_ + y_ * w_),
                     2 * (x_ * y_ + z_ * w_), 1 - 2 * (x_ * x_ + z_ * z_), 2 * (y_ * z_ - x_ * w_),
                     2 * (x_ * z_ - y_ * w_), 2 * (y_ * z_ + x_ * w_), 1 - 2 * (x_ * x_ + y_ * y_)], -1)
    cubes[:, :, 6:] = R
    cubes[:, 0, 0:3] = torch.tensor([0.0, 0.0, 0.5])      # swallows the camera
    cubes[:, 0, 3:6] = 3.0
    cubes[:, 1, 2] = -1.0                                  # entirely behind the camera
    cubes[:, 2, 2] = 0.0105                                # nearest corner within a hair of the 1 cm guard
    cubes[:, 2, 3:6] = 0.001
    cubes[:, 3, 3] = float("nan")
    cubes[:, 4, 0] = float("inf")
    # NON-orthonormal R (a raw head pose with drift): the accepted region {|R^T (p - c)| <= half} is
    # no longer the box spanned by the corners, so the cull must switch itself off
    cubes[:, 5:60, 6:] *= (0.6 + 0.8 * torch.rand(nb, 55, 9, generator=g))
    cubes[:, 60:80, 6:] += 0.3 * torch.randn(nb, 20, 9, generator=g)
    res = score_topk(shard, boxes, cubes.cuda(), cfg=ScoreConfig(k=16, point_stride=stride), want_all=True)
    total = 0
    for b in range(nb):
        ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), boxes_t[b].numpy(), (H, W),
                           cubes[b].numpy(), stride)
        assert int(res["npts"][b]) == ref["npts"]
        assert np.array_equal(res["count"][b].cpu().numpy(), ref["count"]), b
        assert np.array_equal(res["idx"][b].cpu().numpy(), oc.topk(ref["score"], 16))
        total += int(ref["count"].sum())
    assert total > 10000   # the cubes do hold points


@pytest.mark.parametrize("N,H,W,mode,eseg", [(600, 480, 640, 0, 1), (300, 375, 1242, 1, 1),
                                             (1500, 240, 320, 0, 2)])
def test_score_seg_term_bit_exact(cuda_lib, N, H, W, mode, eseg):
    """Mask coverage (A3.7/A3.8 inside the fused epilogue): hull march + interval
    fill + mask IoU must equal the oracle's per-pixel half-plane test bit for bit
    (integer pixel counts), and so must the top-k with the seg factor."""
    from wcr import synth
    from wcr.proposals import propose, score_topk, ScoreConfig, SegMasks
    # the ROI head hands [H, W] as the clamp tuple (x clamped by H!); on the wide image
    # use (W, H) so that the hulls are not all squeezed into x < H
    clamp = (H, W) if W < 2 * H else (W, H)
    ims, shard, boxes, priors, normals = _setup(2, H, W, n_boxes=3, first=60, clamp=clamp)
    masks = [m for im in ims for m in synth.seg_masks(im["boxes"], H, W)]
    seg = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, priors, normals, N, seed=5)
    # push some cubes far off-centre so that their hulls clamp onto the image border
    cubes[:, ::7, 0] += 2.5
    cubes[:, ::11, 1] -= 2.0
    cfg = ScoreConfig(e_seg=eseg, seg_mode=mode, k=16)
    res = score_topk(shard, boxes, cubes, cfg=cfg, seg=seg, want_all=True)
    cubes_h = cubes.cpu().numpy()
    gb = 0
    for im in ims:
        for b in range(len(im["boxes"])):
            ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), im["boxes"][b].numpy(), clamp,
                               cubes_h[gb], 1, 1, 1, eseg, mode, masks[gb].numpy())
            assert np.array_equal(res["count"][gb].cpu().numpy(), ref["count"])
            np.testing.assert_allclose(res["seg"][gb].cpu().numpy(), ref["seg"], rtol=RTOL)
            assert np.array_equal(res["seg"][gb].cpu().numpy(), ref["seg"])          # integers / integers
            np.testing.assert_allclose(res["scores"][gb].cpu().numpy(), ref["score"], rtol=RTOL)
            want = oc.topk(ref["score"], 16)
            assert np.array_equal(res["idx"][gb].cpu().numpy(), want)
            rec = res["records"][gb].cpu().numpy()
            for r, j in enumerate(want):
                assert rec[r, 19] == ref["seg"][j] and rec[r, 17] == ref["iou"][j]
            assert (ref["seg"] > 0).any() and (ref["seg"] == 0).any()
            gb += 1


def test_score_seg_lazy_mask_term(cuda_lib):
    """Without per-proposal outputs the kernel skips the hull fill of proposals that provably hold
    no point (their score is 0 whatever the mask term is) and evaluates it lazily for those that
    still reach the top-k. The records -- index, score, s_iou, s_pts, s_seg -- must equal the
    oracle's, which computes every term for every proposal. Box 0: ordinary proposals; box 1:
    half of them pushed far behind the scene (zero count, lazy); box 2: ALL of them (every
    winner's mask term comes from the lazy path)."""
    from wcr import synth
    from wcr.proposals import propose, score_topk, ScoreConfig, SegMasks
    H, W, N = 240, 320, 300
    ims, shard, boxes, priors, normals = _setup(1, H, W, n_boxes=3, first=33)
    masks = [m for im in ims for m in synth.seg_masks(im["boxes"], H, W)]
    seg = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, priors, normals, N, seed=13)
    scale = torch.ones(3, N, 1, device=cubes.device)
    scale[1, ::2] = 6.0          # same pixel footprint, six times the depth: no point inside
    scale[2, :] = 6.0
    cubes[:, :, 0:3] *= scale
    cfg = ScoreConfig(e_seg=1, k=16)
    res = score_topk(shard, boxes, cubes, cfg=cfg, seg=seg)          # want_all=False: lazy path on
    cubes_h = cubes.cpu().numpy()
    im = ims[0]
    for b in range(3):
        ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), im["boxes"][b].numpy(), (H, W),
                           cubes_h[b], 1, 1, 1, 1, 0, masks[b].numpy())
        want = oc.topk(ref["score"], 16)
        assert np.array_equal(res["idx"][b].cpu().numpy(), want), b
        rec = res["records"][b].cpu().numpy()
        for r, j in enumerate(want):
            assert rec[r, 1] == ref["score"][j]
            assert rec[r, 17] == ref["iou"][j]
            assert rec[r, 18] == (ref["count"][j] / ref["npts"] if ref["npts"] else 0.0) or \
                abs(rec[r, 18] - ref["count"][j] / max(ref["npts"], 1)) < 1e-7
            assert rec[r, 19] == ref["seg"][j], (b, r, j)
        if b == 2:
            assert (ref["count"] == 0).all() and (ref["seg"][want] > 0).any()
        if b == 1:
            assert (ref["count"][::2] == 0).all() and (ref["count"][1::2] > 0).any()


@pytest.mark.parametrize("N,k,e_pts,e_seg,e_iou", [(300, 16, 1, 1, 1), (1000, 16, 1, 1, 1), (1500, 32, 1, 2, 1),
                                                   (1000, 1, 0, 1, 1), (700, 8, 1, 1, 0), (2500, 16, 2, 1, 1)])
def test_score_seg_on_demand_equals_full_evaluation(cuda_lib, N, k, e_pts, e_seg, e_iou):
    """Without per-proposal outputs the mask term is evaluated only for proposals whose bound
    iou^a * pts^b can still reach the k-th best score (chunk_epilogue_warp). The records must be
    those of the full evaluation (want_all=True fills every hull through the seg items), bit
    for bit, for one and several 1024-proposal chunks, k = 1 .. 32 and every exponent mix."""
    from wcr import synth
    from wcr.proposals import propose, score_topk, ScoreConfig, SegMasks
    H, W = 240, 320
    ims, shard, boxes, priors, normals = _setup(3, H, W, n_boxes=4, first=51)
    masks = [m for im in ims for m in synth.seg_masks(im["boxes"], H, W)]
    seg = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, priors, normals, N, seed=21)
    cubes[5, ::3, 0:3] *= 6.0                  # a box where a third of the proposals hold no point
    cubes[7, :, 0:3] *= 6.0                    # ... and one where none does (all bounds are zero)
    cfg = ScoreConfig(e_iou=e_iou, e_pts=e_pts, e_seg=e_seg, k=k)
    full = score_topk(shard, boxes, cubes, cfg=cfg, seg=seg, want_all=True)
    rec_full = full["records"].clone()
    for mode in (2, 1, 0):        # on demand, every hull through the seg items, the library's own choice
        cfg_m = ScoreConfig(e_iou=e_iou, e_pts=e_pts, e_seg=e_seg, k=k, seg_eval=mode)
        lazy = score_topk(shard, boxes, cubes, cfg=cfg_m, seg=seg)
        assert torch.equal(lazy["records"].view(torch.int32), rec_full.view(torch.int32)), mode
        filled = lazy["workspace"].hulls_filled()
        assert filled >= cubes.shape[0] * N if mode == 1 else filled > 0
    # the winners' scores are the full evaluation's scores
    idx = rec_full[:, :, 0].contiguous().view(torch.int32).long()
    assert (idx >= 0).all()
    assert torch.equal(torch.gather(full["scores"], 1, idx), rec_full[:, :, 1])


@pytest.mark.parametrize("N,k,e_pts,e_seg,e_iou", [(300, 16, 1, 1, 1), (1000, 16, 1, 1, 1), (1000, 16, 1, 0, 1),
                                                   (1500, 32, 1, 2, 1), (50, 4, 1, 0, 1), (100, 4, 1, 1, 1), (700, 8, 2, 1, 0),
                                                   (2500, 1, 2, 0, 1), (1000, 16, 1, 1, 2), (1025, 4, 1, 1, 1), (33, 2, 1, 0, 1)])
def test_score_pts_on_demand_equals_every_proposal(cuda_lib, N, k, e_pts, e_seg, e_iou):
    """pts_eval=2: the tiles are walked twice; the second pass skips every proposal whose score bound
    (IoU factor x share of the ROI's points its cull box can reach) is below the k-th best score of
    the first pass (chunk_epilogue_first). The records must be those of pts_eval=1 (every proposal
    tested) and of the full evaluation with per-proposal outputs, bit for bit, with and without the
    mask term, for one and several chunks, chunks shorter than the first pass, k = 1 .. 32."""
    from wcr import synth
    from wcr.proposals import propose, score_topk, ScoreConfig, SegMasks
    H, W = 240, 320
    ims, shard, boxes, priors, normals = _setup(3, H, W, n_boxes=4, first=51)
    seg = None
    if e_seg:
        masks = [m for im in ims for m in synth.seg_masks(im["boxes"], H, W)]
        seg = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, priors, normals, N, seed=21)
    cubes[5, ::3, 0:3] *= 6.0                  # a box where a third of the proposals hold no point
    cubes[7, :, 0:3] *= 6.0                    # ... and one where none does (T = 0: nothing may be pruned)
    cubes[2, 5, 6:15] *= 1.01                  # a cube the cull does not apply to (bound = every point)
    cubes[2, 7, 3] = float("nan")
    cfg = dict(e_iou=e_iou, e_pts=e_pts, e_seg=e_seg, k=k)
    full = score_topk(shard, boxes, cubes, cfg=ScoreConfig(**cfg), seg=seg, want_all=True)
    rec_full = full["records"].clone()
    every = score_topk(shard, boxes, cubes, cfg=ScoreConfig(pts_eval=1, **cfg), seg=seg)
    assert every["workspace"].proposals_pruned() == 0
    assert torch.equal(every["records"].view(torch.int32), rec_full.view(torch.int32))
    lazy = score_topk(shard, boxes, cubes, cfg=ScoreConfig(pts_eval=2, seg_eval=2, **cfg), seg=seg)
    assert torch.equal(lazy["records"].view(torch.int32), rec_full.view(torch.int32))
    pruned = lazy["workspace"].proposals_pruned()
    if N > 64:
        assert 0 < pruned < cubes.shape[0] * N, pruned
    else:
        assert pruned == 0
    # sampled in-kernel (no cube array): same equality, and the library's own choice of mode
    a = score_topk(shard, boxes, None, N, ScoreConfig(pts_eval=1, **cfg), priors=priors, normals=normals, seed=9, seg=seg)
    ra = a["records"].clone()
    for mode in (2, 0):
        bq = score_topk(shard, boxes, None, N, ScoreConfig(pts_eval=mode, **cfg), priors=priors, normals=normals,
                        seed=9, seg=seg)
        assert torch.equal(bq["records"].view(torch.int32), ra.view(torch.int32)), mode


def test_score_many_boxes_unsorted_plan(cuda_lib):
    """More than 2048 boxes: the plan kernel keeps input order (no largest-first sort) and the
    scoring kernel searches the item prefix in global memory instead of its shared-memory copy;
    the library picks the on-demand paths by itself (>= 444 chunks). Records == full evaluation."""
    from wcr import synth
    from wcr.proposals import propose, score_topk, ScoreConfig, SegMasks
    H, W, N = 120, 160, 96
    ims, shard, boxes, priors, normals = _setup(3, H, W, n_boxes=700, first=91)
    assert len(boxes) > 2048
    masks = [m for im in ims for m in synth.seg_masks(im["boxes"], H, W)]
    seg = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, priors, normals, N, seed=5)
    cfg = dict(e_seg=1, k=4)
    full = score_topk(shard, boxes, cubes, cfg=ScoreConfig(**cfg), seg=seg, want_all=True)
    rec_full = full["records"].clone()
    auto = score_topk(shard, boxes, cubes, cfg=ScoreConfig(**cfg), seg=seg)
    assert auto["workspace"].proposals_pruned() > 0
    assert torch.equal(auto["records"].view(torch.int32), rec_full.view(torch.int32))
    every = score_topk(shard, boxes, cubes, cfg=ScoreConfig(pts_eval=1, seg_eval=1, **cfg), seg=seg)
    assert torch.equal(every["records"].view(torch.int32), rec_full.view(torch.int32))


def test_score_custom_mapping_shaping(cuda_lib):
    """ScoreConfig.map_beta > 0 passes every term through custom_mapping (utils.py:207-223) before
    its exponent. Scores within 1e-5 of the composition of the oracle's terms with
    oracle.blocks2.custom_mapping; the top-k may differ only among proposals whose scores are
    closer than 1e-6 relative (powf in double on both sides, one float rounding apart); the
    on-demand mask path gives the records of the full evaluation."""
    from oracle import blocks2 as ob
    from wcr import synth
    from wcr.proposals import propose, score_topk, ScoreConfig, SegMasks
    H, W, N = 240, 320, 500
    ims, shard, boxes, priors, normals = _setup(1, H, W, n_boxes=3, first=61)
    masks = [m for im in ims for m in synth.seg_masks(im["boxes"], H, W)]
    seg = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, priors, normals, N, seed=4)
    cfg = ScoreConfig(e_seg=1, k=16, map_beta=1.7)
    full = score_topk(shard, boxes, cubes, cfg=cfg, seg=seg, want_all=True)
    rec_full = full["records"].clone()
    lazy = score_topk(shard, boxes, cubes, cfg=cfg, seg=seg)
    assert torch.equal(lazy["records"].view(torch.int32), rec_full.view(torch.int32))
    ch = cubes.cpu().numpy()
    im = ims[0]
    for b in range(3):
        ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), im["boxes"][b].numpy(), (H, W), ch[b],
                           1, 1, 1, 1, 0, masks[b].numpy())
        pts = ref["count"].astype(np.float32) / np.float32(max(ref["npts"], 1))

        def cm(x):   # the reference divides by zero at x == 1 (Python floats); the limit is 1
            return np.asarray([1.0 if v >= 1 else ob.custom_mapping([v], 1.7)[0] for v in x], dtype=np.float32)
        want = (cm(ref["iou"]) * cm(pts)) * cm(ref["seg"])
        got = full["scores"][b].cpu().numpy()
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-12)
        idx = rec_full[b, :, 0].contiguous().view(torch.int32).cpu().numpy()
        order = np.argsort(-want.astype(np.float64), kind="stable")[:16]
        for r in range(16):
            assert idx[r] == order[r] or abs(want[idx[r]] - want[order[r]]) <= 1e-6 * abs(want[order[r]]), (b, r)
    plain = score_topk(shard, boxes, cubes, cfg=ScoreConfig(e_seg=1, k=16), seg=seg, want_all=True)
    assert not torch.equal(plain["scores"], full["scores"])


def test_pack_box_masks_device_equals_host(cuda_lib):
    """wcr_pack_box_masks (dense per-box masks -> seg_bits on the device) == the host packer, word
    for word: boxes cut by the image border, empty ROIs, widths that are / are not multiples of
    32, a mask_idx indirection, and a pitched (padded) mask tensor."""
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, SegMasks
    from wcr import synth
    H, W = 97, 133
    ims = [synth.make_image(70 + i, H, W, n_boxes=1, dummy_ground=True) for i in range(2)]
    shard = ImageShard([im["depth"].cuda() for im in ims], None, [im["K"] for im in ims])
    bx = [torch.tensor([[-5.0, -3.0, 40.7, 50.2], [10.0, 10.0, 74.0, 30.0], [100.0, 60.0, 200.0, 120.0],
                        [20.3, 20.3, 20.9, 40.0], [0.0, 0.0, 133.0, 97.0]]),
          torch.tensor([[5.5, 6.5, 69.5, 80.0], [130.0, 90.0, 140.0, 99.0]])]
    boxes = BoxTable(shard, bx, [torch.zeros(len(b), dtype=torch.int64) for b in bx])
    g = torch.Generator().manual_seed(3)
    n = len(boxes)
    padded = torch.zeros((n, H + 3, W + 11), dtype=torch.uint8)
    padded[:, :H, :W] = (torch.rand((n, H, W), generator=g) < 0.4).to(torch.uint8) * 7   # any non-zero == object
    dense = padded[:, :H, :W]
    host = SegMasks(shard, boxes, [dense[i] for i in range(n)])
    dev = SegMasks.from_dense(shard, boxes, padded.cuda()[:, :H, :W])
    assert np.array_equal(dev.host_off, host.host_off)
    assert np.array_equal(dev.bits.cpu().numpy().view(np.uint32)[:host.host_off[-1]], host.host_bits[:host.host_off[-1]])
    perm = torch.randperm(n, generator=g)
    inv = torch.empty_like(perm); inv[perm] = torch.arange(n)
    dev2 = SegMasks.from_dense(shard, boxes, dense[perm].contiguous().cuda(), mask_idx=inv)
    assert torch.equal(dev2.bits, dev.bits)
    # odd pitches and a misaligned base: every byte alignment of a row start (the kernel reads aligned
    # 32-bit words and funnel-shifts), and the same maps read from mapped pinned host memory
    for pad, skew in ((10, 0), (8, 1), (9, 3)):
        big = torch.zeros((n * (H + 3) * (W + pad) + 8,), dtype=torch.uint8)
        view = big[skew:skew + n * (H + 3) * (W + pad)].view(n, H + 3, W + pad)
        view[:, :H, :W] = dense
        dev3 = SegMasks.from_dense(shard, boxes, big.cuda()[skew:skew + n * (H + 3) * (W + pad)].view(n, H + 3, W + pad)[:, :H, :W])
        assert torch.equal(dev3.bits, dev.bits), (pad, skew)
    pinned = padded.pin_memory()
    dev4 = SegMasks.from_dense(shard, boxes, pinned[:, :H, :W])
    assert torch.equal(dev4.bits, dev.bits)


def test_shard_runner_dense_masks_equal_prepacked(cuda_lib):
    from wcr import synth
    from wcr.pipeline import ShardRunner
    from wcr.proposals import ScoreConfig
    H, W = 240, 320
    ims = [synth.make_image(80 + i, H, W, n_boxes=3, dummy_ground=(i == 1)) for i in range(3)]
    for im in ims:
        im["seg"] = synth.seg_masks(im["boxes"], H, W)
    kw = dict(priors=synth.make_priors(), score_cfg=ScoreConfig(k=8, e_seg=1), seed=3)
    a = ShardRunner(ims, 400, "cuda", **kw).run_host().clone()
    for mode in (True, "rows", "copy", "mapped"):   # ROI rectangles / contiguous row bands / whole maps copied, ROI bytes read from host
        r = ShardRunner(ims, 400, "cuda", dense_seg=mode, **kw)
        if mode == "copy":
            assert r.h2d_bytes > 9 * H * W
        else:
            assert 0 < r.seg.roi_bytes < 9 * H * W and (r.d_seg_dense is None) == (mode == "mapped")
        assert torch.equal(r.run_host().view(torch.int32), a.view(torch.int32)), mode
        r.h_seg_dense.zero_()               # new mask contents reach the kernel
        b = r.run_host()
        assert not torch.equal(b, a) and (b[:, :, 19] == 0).all()


def test_score_seg_degenerate(cuda_lib):
    """Empty masks, cubes behind the camera (non-finite corners), cubes entirely
    outside the image (hull collapses onto the border, duplicate corners)."""
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, score_topk, ScoreConfig, SegMasks
    from wcr import synth
    H, W = 120, 160
    im = synth.make_image(5, H, W, n_boxes=1, dummy_ground=False)
    boxes_t = torch.tensor([[20.0, 30.0, 120.0, 100.0], [0.0, 0.0, 160.0, 120.0],
                            [30.2, 40.7, 31.1, 41.3]])
    shard = ImageShard([im["depth"].cuda()], None, [im["K"]])
    boxes = BoxTable(shard, [boxes_t], [torch.zeros(3, dtype=torch.long)])
    masks = list(synth.seg_masks(boxes_t, H, W))
    masks[1] = torch.zeros(H, W, dtype=torch.uint8)           # empty mask
    seg = SegMasks(shard, boxes, masks)
    g = torch.Generator().manual_seed(1)
    N = 64
    cubes = torch.zeros(3, N, 15)
    cubes[:, :, 0:2] = torch.randn(3, N, 2, generator=g) * 1.5
    cubes[:, :, 2] = 0.5 + torch.rand(3, N, generator=g) * 4
    cubes[:, :, 3:6] = 0.2 + torch.rand(3, N, 3, generator=g)
    cubes[:, :, 6:] = torch.eye(3).reshape(9)
    cubes[:, 0, 2] = 0.0            # straddles the camera plane
    cubes[:, 1, 2] = -2.0           # behind the camera
    cubes[:, 2, 0] = 50.0           # far right: every corner clamps to the border
    cubes[:, 3, 0:2] = -50.0        # top-left: all corners clamp to (0, 0) -> 8 duplicates
    res = score_topk(shard, boxes, cubes.cuda(), cfg=ScoreConfig(e_seg=1, k=8), seg=seg, want_all=True)
    for b in range(3):
        ref = oc.score_box(im["depth"].numpy(), im["K"].numpy(), boxes_t[b].numpy(), (H, W),
                           cubes[b].numpy(), 1, 1, 1, 1, 0, masks[b].numpy())
        assert np.array_equal(res["seg"][b].cpu().numpy(), ref["seg"])
        np.testing.assert_allclose(res["scores"][b].cpu().numpy(), ref["score"], rtol=RTOL)
        assert np.array_equal(res["idx"][b].cpu().numpy(), oc.topk(ref["score"], 8))
    assert (res["seg"][1] == 0).all()


def test_replay_of_reference_sampler_stream(cuda_lib):
    """north_star: 'same inputs and the same RNG stream as the reference's own torch
    path'. tests/golden/replay_0.npz holds the (u v z w h l yaw) samples drawn by the
    reference's sample_normal_in_range under torch.manual_seed; replayed through
    wcr_sample_cubes + wcr_score_topk they must give the oracle's cubes, counts, scores
    and top-k bit for bit."""
    import os
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, propose, score_topk, ScoreConfig, SegMasks
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "replay_0.npz"))
    depth, K = torch.from_numpy(z["depth"]), torch.from_numpy(z["K"])
    H, W = depth.shape
    nb, _, N = z["attrs"].shape
    shard = ImageShard([depth.cuda()], None, [K])
    boxes = BoxTable(shard, [torch.from_numpy(z["boxes"])], [torch.from_numpy(z["classes"])])
    priors = torch.zeros(50, 2, 3)
    priors[torch.from_numpy(z["classes"])] = torch.from_numpy(z["priors"])
    normals = torch.from_numpy(z["normal"]).view(1, 3)
    cubes = propose(shard, boxes, priors, normals, N, replay=torch.from_numpy(z["attrs"]), seed=0)
    seg = SegMasks(shard, boxes, list(torch.from_numpy(z["seg"])))
    res = score_topk(shard, boxes, cubes, cfg=ScoreConfig(e_seg=1, k=16), seg=seg, want_all=True)
    for b in range(nb):
        # cos/sin of the yaw differ in the last ulp between libm and CUDA: feed the device's
        # cubes to the oracle for everything downstream, and bound the assembly separately
        want = oc.assemble_cubes(z["K"], z["normal"], z["attrs"][b])
        got = cubes[b].cpu().numpy()
        np.testing.assert_allclose(got, want, rtol=0, atol=3e-7)
        assert np.array_equal(got[:, :6], want[:, :6])                    # centre + dims: bit-exact
        ref = oc.score_box(z["depth"], z["K"], z["boxes"][b], (H, W), got, 1, 1, 1, 1, 0, z["seg"][b])
        assert np.array_equal(res["count"][b].cpu().numpy(), ref["count"])
        assert np.array_equal(res["seg"][b].cpu().numpy(), ref["seg"])
        np.testing.assert_allclose(res["scores"][b].cpu().numpy(), ref["score"], rtol=RTOL)
        assert np.array_equal(res["idx"][b].cpu().numpy(), oc.topk(ref["score"], 16))
