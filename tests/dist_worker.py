"""Worker of tests/test_dist_gpu.py, launched with torch.distributed.run (one rank per GPU):
a small strong-scaling job (wcr.job) sharded by block, all-gathered with TopkGather, and
checked on rank 0 against the single-process result for the same global image list."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))


def main():
    from wcr import synth, dist as wd
    from wcr.job import BlockSet, ResidentBlocks, StreamedBlocks
    from wcr.proposals import ScoreConfig
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=dev)
    H, W, N, K = 240, 320, 300, 8
    base = []
    for i in range(6):
        im = synth.make_image(40 + i, H, W, n_boxes=1 + i % 4, dummy_ground=(i == 2))
        im["seg"] = synth.seg_masks(im["boxes"], H, W)
        base.append(im)
    bs = BlockSet(base, 37, block=4, map_seed=3)          # 10 blocks, the last one ragged
    cfg = ScoreConfig(k=K, e_seg=1)
    priors = synth.make_priors()
    # blocks -> ranks by predicted work (not contiguous): the gathered records are put back into
    # global box order before the comparison
    rank_blocks = wd.balanced_assignment(bs.block_weights(N), world)
    gather = wd.TopkGather(bs.n_boxes_of(rank_blocks[rank]), K, dev)
    mine = ResidentBlocks(bs, rank_blocks[rank], N, dev, priors=priors, score_cfg=cfg, seed=77,
                          records=gather.send)
    mine.run_pass()
    allrec = gather.gather().index_select(0, torch.from_numpy(bs.global_order(rank_blocks)).to(dev))
    torch.cuda.synchronize()
    if rank == 0:
        single = ResidentBlocks(bs, range(bs.n_blocks), N, dev, priors=priors, score_cfg=cfg, seed=77)
        want = single.run_pass().clone()
        torch.cuda.synchronize()
        assert allrec.shape == want.shape, (allrec.shape, want.shape)
        assert torch.equal(allrec.view(torch.int32), want.view(torch.int32)), "gathered records != single-process records"
        assert (want[:, 0, 0].view(torch.int32) >= 0).all()
        streamed = StreamedBlocks(bs, range(bs.n_blocks), N, dev, priors=priors, score_cfg=cfg, seed=77)
        got = streamed.run_pass()
        assert torch.equal(got.view(torch.int32), want.cpu().view(torch.int32)), "streamed != resident"
        print(f"DIST_OK world={world} boxes={want.shape[0]} counts={gather.counts}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
