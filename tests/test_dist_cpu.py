"""World-size-2 gloo test of the only collective on the path: image
sharding + the all-gather of fixed-stride top-k records."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from wcr import dist as wd
    lo, hi = wd.shard_range(7, rank, world)
    boxes_per_image = [3, 1, 4, 1, 5, 9, 2]
    n_local = sum(boxes_per_image[lo:hi])
    g = wd.TopkGather(n_local, 4, torch.device("cpu"))
    first_box = sum(boxes_per_image[:lo])
    for b in range(n_local):            # record content = global box id, so order is checkable
        g.send[b] = float(first_box + b)
    allrec = g.gather()
    if rank == 0:
        torch.save({"rec": allrec, "counts": g.counts}, out)
    dist.destroy_process_group()


def test_shard_range_partitions():
    sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
    from wcr.dist import shard_range
    for n in (0, 1, 7, 64, 8192):
        for w in (1, 2, 3, 8):
            r = [shard_range(n, i, w) for i in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def test_topk_allgather_world2(tmp_path):
    out = str(tmp_path / "r0.pt")
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    d = torch.load(out)
    assert d["counts"] == [9, 16]       # images [0,4) and [4,7)
    rec = d["rec"]
    assert rec.shape == (25, 4, 20)
    assert torch.equal(rec[:, 0, 0], torch.arange(25, dtype=torch.float32))   # global box order kept


def test_blockset_arithmetic():
    """wcr.job.BlockSet: blocks cut at global image positions, box offsets, ragged last block."""
    sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
    import numpy as np
    from wcr import synth
    from wcr.job import BlockSet
    from wcr.dist import shard_range
    base = [synth.make_image(i, 48, 64, n_boxes=1 + i) for i in range(3)]
    bs = BlockSet(base, 11, block=4, map_seed=0, use_seg=False)
    assert bs.n_blocks == 3 and len(bs.images_of(2)) == 3
    per_image = np.array([1 + j for j in bs.image_map])
    assert bs.block_boxes.tolist() == [per_image[0:4].sum(), per_image[4:8].sum(), per_image[8:].sum()]
    assert bs.box_offset[-1] == per_image.sum()
    img, boxes, bits, off = bs.tables(1, np.arange(4, dtype=np.uint64) * 100 + 1000, np.arange(4, dtype=np.uint64) + 7)
    assert bits is None and off is None
    assert img["depth"].tolist() == [1000, 1100, 1200, 1300] and len(boxes) == bs.block_boxes[1]
    assert boxes["image"].tolist() == sum(([i] * (1 + j) for i, j in enumerate(bs.images_of(1))), [])
    # ranks own whole blocks; with 8192 images / 64 per block that is shard_range over images
    for world in (1, 2, 4, 8):
        for r in range(world):
            lo, hi = shard_range(8192 // 64, r, world)
            assert (lo * 64, hi * 64) == shard_range(8192, r, world)


def test_balanced_assignment_and_global_order():
    sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
    import numpy as np
    from wcr import synth
    from wcr.dist import balanced_assignment
    from wcr.job import BlockSet
    rng = np.random.default_rng(0)
    w = rng.uniform(0.8, 1.2, 128)
    a = balanced_assignment(w, 8)
    assert sorted(b for r in a for b in r) == list(range(128)) and all(r == sorted(r) for r in a)
    loads = [w[r].sum() for r in a]
    assert max(loads) / np.mean(loads) < 1.005          # contiguous ranges: ~1.02
    assert balanced_assignment(w, 8) == a               # deterministic
    base = [synth.make_image(i, 48, 64, n_boxes=1 + i) for i in range(3)]
    bs = BlockSet(base, 23, block=4, map_seed=0, use_seg=False)
    rb = balanced_assignment(bs.block_weights(100), 2)
    idx = bs.global_order(rb)
    # tag every box with its global id in rank-major order, then undo
    tags = np.concatenate([np.arange(bs.box_offset[b], bs.box_offset[b + 1]) for r in rb for b in r])
    assert np.array_equal(tags[idx], np.arange(bs.box_offset[-1]))
