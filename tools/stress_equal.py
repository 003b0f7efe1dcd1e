"""Stress (GPU box): the on-demand paths against the every-proposal evaluation on C2-sized shards,
many seeds and repetitions -- the lock-free queues and flags of score_tiles_kernel must give the
same records on every run. Prints one JSON line; exit code 1 on any mismatch."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import torch
import bench
from wcr import synth
from wcr.pipeline import ShardRunner
from wcr.proposals import ScoreConfig, score_topk, box_center_depth

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
ims = bench.make_images(64, 0, True, None, dense=False)
priors = synth.make_priors()
r = ShardRunner(ims, 1000, "cuda", priors=priors, score_cfg=ScoreConfig(k=16, e_seg=1), seed=1234)
r.upload(); r.step(); torch.cuda.synchronize()
z0 = box_center_depth(r.shard, r.boxes)
bad, runs = 0, 0
for seed in range(reps):
    for e_seg in (1, 0):
        kw = dict(priors=r.priors, normals=r.normals, z0=z0, seed=seed, seg=r.seg if e_seg else None, workspace=r.ws)
        ref = score_topk(r.shard, r.boxes, None, 1000, ScoreConfig(k=16, e_seg=e_seg, pts_eval=1, seg_eval=1), **kw)["records"].clone()
        for _ in range(3):
            out = score_topk(r.shard, r.boxes, None, 1000, ScoreConfig(k=16, e_seg=e_seg), **kw)["records"]
            runs += 1
            if not torch.equal(out.view(torch.int32), ref.view(torch.int32)):
                bad += 1
print(json.dumps({"runs": runs, "mismatches": bad}))
sys.exit(1 if bad else 0)
