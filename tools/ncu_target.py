"""Smallest program that launches the benchmarked scoring chain (C2: 64 images 640x480, ~1264
boxes, N = 1000, mask term on demand) a few times -- the target of the ncu captures under
profiles/ (B200_PROFILING.md recipe). `--no-seg` drops the mask term; `--small` = C1-like."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import torch
from wcr import _abi
for a_ in sys.argv[1:]:
    if a_.startswith("--lib="):
        _abi.LIB_PATH = _abi.LIB_PATH.replace("libwcr_b200.so", a_[6:])
import bench
from wcr import synth
from wcr.pipeline import ShardRunner
from wcr.proposals import ScoreConfig
from wcr.ground import unproject_cloud

seg = "--no-seg" not in sys.argv
pts_eval = next((int(a_[6:]) for a_ in sys.argv[1:] if a_.startswith("--pts=")), 0)
n_img = 1 if "--small" in sys.argv else 64
ims = bench.make_images(n_img, 0, seg, 5 if n_img == 1 else None, dense=False)
r = ShardRunner(ims, 1000, "cuda", priors=synth.make_priors(),
                score_cfg=ScoreConfig(k=16, e_seg=1 if seg else 0, pts_eval=pts_eval), seed=1234)
r.upload()
for _ in range(3):
    r.step()
    unproject_cloud(r.shard, 1)
torch.cuda.synchronize()
print("ok", r.n_boxes)
