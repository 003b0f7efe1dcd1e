"""Diagnostic (GPU box): scoring launch chain at C2 in three modes + hull counter."""
import os, sys, json, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import torch
from wcr import _abi
for a_ in sys.argv[1:]:
    if a_.startswith("--lib="):
        _abi.LIB_PATH = _abi.LIB_PATH.replace("libwcr_b200.so", a_[6:])
PTS_ONLY = "--pts-only" in sys.argv
sys.argv = [sys.argv[0]] + [a_ for a_ in sys.argv[1:] if not a_.startswith("--")]
import bench
from wcr import synth
from wcr.pipeline import ShardRunner
from wcr.proposals import ScoreConfig, score_topk, box_center_depth

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n_box = int(sys.argv[2]) if len(sys.argv) > 2 else None
ims = bench.make_images(n_img, 0, True, n_box, dense=False)
priors = synth.make_priors()
r = ShardRunner(ims, 1000, "cuda", priors=priors, score_cfg=ScoreConfig(k=16, e_seg=1), seed=1234)
r.upload(); r.step(); torch.cuda.synchronize()
z0 = box_center_depth(r.shard, r.boxes)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

def timed(fn, n=5):
    ts = []
    for _ in range(n):
        flush.zero_(); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); out = fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return min(ts), out

def run(cfg, want_all=False, seg=True):
    return score_topk(r.shard, r.boxes, None, 1000, cfg, priors=r.priors, normals=r.normals, z0=z0, seed=1234,
                      seg=r.seg if seg else None, workspace=r.ws, want_all=want_all)

res = {}
for pe in (1, 2):
    t, out = timed(lambda: run(ScoreConfig(k=16, e_seg=1, pts_eval=pe)))
    res[f"seg_pts_eval{pe}_ms"] = t; res[f"seg_pts_eval{pe}_pruned"] = out["workspace"].proposals_pruned()
    res[f"seg_pts_eval{pe}_hulls"] = out["workspace"].hulls_filled(); res[f"seg_pts_eval{pe}_tests"] = out["workspace"].executed_tests()
    t, out = timed(lambda: run(ScoreConfig(k=16, e_seg=0, pts_eval=pe), seg=False))
    res[f"noseg_pts_eval{pe}_ms"] = t; res[f"noseg_pts_eval{pe}_pruned"] = out["workspace"].proposals_pruned()
    res[f"noseg_pts_eval{pe}_tests"] = out["workspace"].executed_tests()
a_ = run(ScoreConfig(k=16, e_seg=1, pts_eval=1))["records"].clone()
b_ = run(ScoreConfig(k=16, e_seg=1, pts_eval=2))["records"].clone()
res["seg_records_equal"] = bool(torch.equal(a_.view(torch.int32), b_.view(torch.int32)))
a_ = run(ScoreConfig(k=16, e_seg=0, pts_eval=1), seg=False)["records"].clone()
b_ = run(ScoreConfig(k=16, e_seg=0, pts_eval=2), seg=False)["records"].clone()
res["noseg_records_equal"] = bool(torch.equal(a_.view(torch.int32), b_.view(torch.int32)))
def phases(ws):
    v = ws.ws[64:64 + 8 * 13].view(torch.int64).tolist()
    d = dict(zip(["tiles_pass1", "tiles_pass2", "epilogue_first", "epilogue_final", "seg_batches", "wait_ready",
                  "claim_decode", "retire", "tail", "lost_race"], v))
    d = {k: round(x / (444 * 4 * 1.92e6), 3) for k, x in d.items()}   # ms per warp at 1.92 GHz, 1776 warps
    d["kernel_ms(last exit - first start)"] = round((v[10] - v[12]) / 1e6, 3)
    d["mean_warp_lifetime_ms"] = round(v[11] / 1776 / 1e6, 3)
    return d
if "prof" in _abi.LIB_PATH:
    for pe in (1, 2):
        out = run(ScoreConfig(k=16, e_seg=1, pts_eval=pe)); torch.cuda.synchronize()
        res[f"seg_pts_eval{pe}_cycles"] = phases(out["workspace"])
        out = run(ScoreConfig(k=16, e_seg=0, pts_eval=pe), seg=False); torch.cuda.synchronize()
        res[f"noseg_pts_eval{pe}_cycles"] = phases(out["workspace"])
if PTS_ONLY:
    print(json.dumps(res)); sys.exit(0)
t, out = timed(lambda: run(ScoreConfig(k=16, e_seg=1, seg_eval=2)))
res["seg_on_demand_ms"] = t; res["hulls_on_demand"] = out["workspace"].hulls_filled()
t, out = timed(lambda: run(ScoreConfig(k=16, e_seg=1, seg_eval=1)))
res["seg_every_hull_ms"] = t
t, out = timed(lambda: run(ScoreConfig(k=16, e_seg=1)))
res["seg_auto_ms"] = t; res["hulls_auto"] = out["workspace"].hulls_filled()
t, out = timed(lambda: run(ScoreConfig(k=16, e_seg=1), want_all=True))
res["seg_all_ms(want_all)"] = t; res["hulls_all"] = out["workspace"].hulls_filled()
t, out = timed(lambda: run(ScoreConfig(k=16, e_seg=0), seg=False))
res["noseg_ms"] = t; res["executed_tests"] = out["workspace"].executed_tests()
res["n_proposals"] = r.n_proposals
out = run(ScoreConfig(k=16, e_seg=1))
res["records_checksum"] = int(out["records"].view(torch.int32).to(torch.int64).sum().item())
res["lib"] = os.path.basename(_abi.LIB_PATH)
print(json.dumps(res))
