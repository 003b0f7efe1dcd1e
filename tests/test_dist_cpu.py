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
