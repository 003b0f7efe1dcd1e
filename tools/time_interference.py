"""Diagnostic (GPU box): does a mask gather running on another stream slow the scoring pass?
Times device-resident steps alone, beside a gather from mapped HOST memory, and beside the same
gather from DEVICE memory, with and without CTA slots reserved for it."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import torch
import bench
from wcr import synth, _abi
from wcr.pipeline import ShardRunner
from wcr.proposals import ScoreConfig

ims = bench.make_images(64, 0, True, None, dense="both")
kw = dict(priors=synth.make_priors(), score_cfg=ScoreConfig(k=16, e_seg=1), seed=1234)
r = ShardRunner(ims, 1000, "cuda", **kw)                       # pre-packed masks: the pass under test
g = ShardRunner(ims, 1000, "cuda", dense_seg="mapped", **kw)   # source of gathers
dmask = g.h_seg_dense.cuda()
r.upload(); g.upload(); r.step(); torch.cuda.synchronize()
s_run, s_g = torch.cuda.Stream(), torch.cuda.Stream()
lib = _abi.load()
res = {}
s_c = torch.cuda.Stream()
for name, src, reserve, copies in (("alone", None, 0, False), ("copies", None, 0, True),
                                   ("host_gather_reserved9", g.h_seg_dense, 9, False),
                                   ("host_gather_reserved9+copies", g.h_seg_dense, 9, True),
                                   ("device_gather_reserved9+copies", dmask, 9, True)):
    lib.wcr_score_set_reserved_ctas(reserve)
    ts = []
    for it in range(8):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        if src is not None:
            with torch.cuda.stream(s_g):
                for _ in range(3):
                    g.seg.repack(src)
        if copies:
            with torch.cuda.stream(s_c):
                g.d_depth.copy_(g.h_depth, non_blocking=True); g.d_ground.copy_(g.h_ground, non_blocking=True)
        with torch.cuda.stream(s_run):
            e0.record(s_run); r.step(); e1.record(s_run)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    res[name] = round(sorted(ts)[len(ts) // 2], 3)
lib.wcr_score_set_reserved_ctas(0)
print(json.dumps(res))
