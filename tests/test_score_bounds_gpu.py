"""The bounds the on-demand paths of wcr_score_topk prune with (score_prep_kernel: upper bound of the
inside-count from the ROI's cell grid, bounds of the mask ratio from the hull's bounding box and
the largest projected face) must hold for EVERY proposal, not just leave the top-k intact: the
values the kernel stored (ScoreWorkspace.bounds) against the full evaluation's counts and ratios."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, H, W, n_images, n_boxes, N, clamp):
    from wcr import synth
    from wcr.images import ImageShard
    from wcr.proposals import BoxTable, SegMasks, propose
    ims = [synth.make_image(seed + i, H, W, n_boxes=n_boxes, dummy_ground=False) for i in range(n_images)]
    shard = ImageShard([im["depth"].cuda() for im in ims], [im["ground"].cuda() for im in ims], [im["K"] for im in ims])
    boxes = BoxTable(shard, [im["boxes"] for im in ims], [im["classes"] for im in ims],
                     clamp=None if clamp is None else [clamp] * n_images)
    normals = torch.tensor([[0.01 * i, 1.0, -0.02 * i] for i in range(n_images)])
    normals = normals / normals.norm(dim=1, keepdim=True)
    masks = [m for im in ims for m in synth.seg_masks(im["boxes"], H, W)]
    seg = SegMasks(shard, boxes, masks)
    cubes = propose(shard, boxes, synth.make_priors(), normals, N, seed=seed)
    g = torch.Generator().manual_seed(seed)
    nb = cubes.shape[0]
    # adversarial proposals: huge cubes (hulls leave the image, corners pile up on the clamp
    # bounds -> duplicate bumps), tiny ones, cubes pushed sideways / towards the camera
    cubes[:, 0:N:7, 3:6] *= 6.0
    cubes[:, 1:N:7, 3:6] *= 0.05
    cubes[:, 2:N:7, 0] += (torch.rand(nb, len(range(2, N, 7)), generator=g).cuda() - 0.5) * 6
    cubes[:, 3:N:7, 2] *= 0.3
    cubes[:, 4:N:7, 1] += 2.0
    return shard, boxes, seg, cubes


@pytest.mark.parametrize("seed,H,W,clamp", [(3, 240, 320, None), (11, 240, 320, (320.0, 240.0)),
                                            (17, 480, 640, None), (23, 120, 400, (150.0, 500.0))])
def test_bounds_hold_for_every_proposal(cuda_lib, seed, H, W, clamp):
    from wcr.proposals import score_topk, ScoreConfig, ScoreWorkspace
    N, k = 700, 8
    shard, boxes, seg, cubes = _case(seed, H, W, 3, 4, N, clamp)
    nb = cubes.shape[0]
    ws = ScoreWorkspace(nb, N, k, "cuda")
    lazy = score_topk(shard, boxes, cubes, cfg=ScoreConfig(e_seg=1, k=k, pts_eval=2, seg_eval=2), seg=seg, workspace=ws)
    ub, sb, sw, mi = [t.clone() for t in ws.bounds()]
    assert ws.proposals_pruned() > 0
    full = score_topk(shard, boxes, cubes, cfg=ScoreConfig(e_seg=1, k=k), seg=seg, want_all=True)
    assert torch.equal(lazy["records"].view(torch.int32), full["records"].view(torch.int32))
    count, ratio = full["count"], full["seg"]
    # inside-count <= its bound (0x7fffffff: no bound)
    assert bool((count <= ub).all()), int((count > ub).sum())
    assert int((ub < 0x7fffffff).sum()) > nb * N // 2           # ... and the bound exists for most proposals
    # mask ratio <= seg_ratio_bound(iub, blow, area), evaluated like the kernel does (float32)
    off = seg.off.cpu().numpy()
    bits = seg.bits.cpu().numpy().view(np.uint32)
    area = torch.tensor([int(np.unpackbits(bits[off[b]:off[b + 1]].view(np.uint8)).sum()) for b in range(nb)],
                        dtype=torch.int64, device="cuda")[:, None]
    iub, blow = sb[..., 0].long(), sb[..., 1].long()
    # ... |mask & hull| also bounded by the mask pixels of the grid cells the scan window touches
    gx0, gx1, gy0, gy1 = (sw & 63).long(), ((sw >> 6) & 63).long(), ((sw >> 12) & 63).long(), ((sw >> 18) & 63).long()
    bi = torch.arange(nb, device="cuda")[:, None].expand(nb, N)
    mil = mi.long()
    in_cells = (mil[bi, gy1, gx1] - mil[bi, gy0, gx1]) - (mil[bi, gy1, gx0] - mil[bi, gy0, gx0])
    assert bool((mil[:, 32, 32] == area[:, 0]).all())            # the table's total is the mask area
    assert int((in_cells < torch.minimum(iub, area)).sum()) > 0   # ... and it does cut
    iub = torch.minimum(iub, in_cells)
    i = torch.minimum(iub, area)
    h = torch.maximum(blow, i)
    bound = torch.where(i > 0, i.float() / (area + h - i).clamp(min=1).float(), torch.zeros_like(i, dtype=torch.float32))
    r = torch.nan_to_num(ratio, nan=0.0)
    bad = r > bound
    assert not bool(bad.any()), (int(bad.sum()), r[bad][:5].tolist(), bound[bad][:5].tolist())
    # the bounds are not vacuous: they cut a fair share of the ratios' head room
    assert float(bound.mean()) < 0.9
    assert int((blow > 0).sum()) > 0 and int((iub < area).sum()) > 0
