"""Diagnostic (GPU box): timeline of one steady-state ShardPipeline step with dense (mapped) masks --
when the input copies, the mask gather, the pre-kernels and the scoring chain of a shard start and
end, relative to the end of the previous shard's pass (CUDA events on the three streams)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "weak-cube-r-cnn_b200"))
import torch
import bench
from wcr import synth, ground, _abi
from wcr.pipeline import ShardRunner, ShardPipeline
from wcr.proposals import ScoreConfig, box_center_depth

mode = sys.argv[1] if len(sys.argv) > 1 else "mapped"
ims = bench.make_images(64, 0, True, None, dense="both")
kw = dict(priors=synth.make_priors(), score_cfg=ScoreConfig(k=16, e_seg=1), seed=1234,
          dense_seg=False if mode == "packed" else mode)
rs = [ShardRunner(ims, 1000, "cuda", **kw) for _ in range(2)]
pipe = ShardPipeline(rs, reserve_ctas=int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else None)
E = lambda: torch.cuda.Event(enable_timing=True)
log = []

def submit(i):
    k = i % 2; r = rs[k]
    ev = {n: E() for n in ("up0", "in", "g0", "up1", "run0", "pre1", "run1")}
    gather = r.h_seg_dense is not None and r.d_seg_dense is None
    with torch.cuda.stream(pipe.s_up):
        if pipe.pending[k]:
            pipe.s_up.wait_event(pipe.ev_run[k])
        ev["up0"].record(pipe.s_up)
        if gather:
            r.upload_inputs(); ev["in"].record(pipe.s_up); pipe.ev_in[k].record(pipe.s_up)
        else:
            r.upload_inputs(); ev["in"].record(pipe.s_up); pipe.ev_in[k].record(pipe.s_up)
            ev["g0"].record(pipe.s_up); r.upload_masks(); ev["up1"].record(pipe.s_up); pipe.ev_up[k].record(pipe.s_up)
    if pipe.reserve_ctas:
        _abi.load().wcr_score_set_reserved_ctas(pipe.reserve_ctas)
    if gather:
        with torch.cuda.stream(pipe.s_mask):
            pipe.s_mask.wait_event(pipe.ev_in[k])
            ev["g0"].record(pipe.s_mask); r.upload_masks(); ev["up1"].record(pipe.s_mask); pipe.ev_up[k].record(pipe.s_mask)
    with torch.cuda.stream(pipe.s_run):
        pipe.s_run.wait_event(pipe.ev_in[k])
        if pipe.pending[k]:
            pipe.s_run.wait_event(pipe.ev_down[k])
        ev["run0"].record(pipe.s_run)
        r.normals = ground.normal_vector_from_maps(None, None, None, shard=r.shard, seed=r.seed)
        z0 = box_center_depth(r.shard, r.boxes)
        ev["pre1"].record(pipe.s_run)
        r._seg_ready = pipe.ev_up[k] if r.d_seg_dense is None else None
        if r.d_seg_dense is not None:
            pipe.s_run.wait_event(pipe.ev_up[k]); r.seg.repack(r.d_seg_dense)
        r._score(z0)
        ev["run1"].record(pipe.s_run); pipe.ev_run[k].record(pipe.s_run)
    _abi.load().wcr_score_set_reserved_ctas(0)
    with torch.cuda.stream(pipe.s_down):
        pipe.s_down.wait_event(pipe.ev_run[k]); r.download(); pipe.ev_down[k].record(pipe.s_down)
    pipe.pending[k] = True
    log.append(ev)

for i in range(12):
    pipe.wait(i - 2); submit(i)
pipe.drain(); torch.cuda.synchronize()
base = log[6]["run1"]          # end of shard 6's pass
out = {"mode": mode}
for i in (7, 8):
    out[f"shard{i}"] = {n: round(base.elapsed_time(e), 3) for n, e in log[i].items()}
out["step_ms"] = round(log[7]["run1"].elapsed_time(log[9]["run1"]) / 2, 3)
print(json.dumps(out))
