"""Diagnostic (GPU box): device and host time of ShardRunner.upload() per dense-mask mode."""
import os, sys, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import torch
import bench
from wcr import synth
from wcr.pipeline import ShardRunner
from wcr.proposals import ScoreConfig

ims = bench.make_images(64, 0, True, None, dense="both")
res = {}
for mode in (False, "roi", "rows", "mapped"):
    r = ShardRunner(ims, 1000, "cuda", priors=synth.make_priors(), score_cfg=ScoreConfig(k=16, e_seg=1), seed=1234,
                    dense_seg=mode)
    for _ in range(3):
        r.upload(); r.step()
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(True) for _ in range(3))
    host, up, st = [], [], []
    for _ in range(10):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); e0.record(); r.upload(); e1.record(); t1 = time.perf_counter()
        r.step(); e2.record(); torch.cuda.synchronize()
        host.append((t1 - t0) * 1e3); up.append(e0.elapsed_time(e1)); st.append(e1.elapsed_time(e2))
    res[str(mode)] = {"upload_host_ms": min(host), "upload_dev_ms": min(up), "step_dev_ms": min(st), "h2d_bytes": r.h2d_bytes,
                      "GBps": r.h2d_bytes / (min(up) * 1e-3) / 1e9}
    del r
print(json.dumps(res))
