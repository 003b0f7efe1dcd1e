"""wcr_ransac_ground_static: the allocation-free variant of wcr_ransac_ground (hypothesis
workspace in a file-scope __device__ array, include/wcr.h). No reference counterpart; the
check is equality with the existing entry point on the same inputs -- bit for bit, the bar
tests/test_ground_gpu.py uses for this kernel -- on two calls in a row (the scratch keeps the
previous call's contents), called straight through the C ABI.

Collected last on purpose (file name).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_HYP = 65536   # WCR_RANSAC_STATIC_MAX_HYP


def _inputs(seed, n_images, H, W):
    from wcr import synth
    from wcr.images import ImageShard
    from wcr.ground import unproject_ground
    ims = [synth.make_image(seed + i, H, W, n_boxes=2, dummy_ground=(i == 1)) for i in range(n_images)]
    shard = ImageShard([im["depth"].cuda() for im in ims], [im["ground"].cuda() for im in ims],
                       [im["K"] for im in ims])
    pts, count = unproject_ground(shard, 5)
    return shard, pts, count


def _outputs(B, iters, cap, dev):
    return dict(eq=torch.full((B, 4), 7.0, device=dev), normal=torch.full((B, 3), 7.0, device=dev),
                best=torch.full((B,), -5, dtype=torch.int32, device=dev),
                counts=torch.full((B, iters), -5, dtype=torch.int32, device=dev),
                mask=torch.full((B, cap), 9, dtype=torch.uint8, device=dev))


def _call_existing(lib, pts, count, iters, seed, o):
    from wcr import _abi
    B, cap = pts.shape[0], pts.shape[1]
    ws = torch.empty((B, iters, 4), dtype=torch.float32, device=pts.device)
    _abi.check(lib.wcr_ransac_ground(pts.data_ptr(), count.data_ptr(), cap, B, None, iters,
                                     ctypes.c_float(0.05), seed, o["eq"].data_ptr(),
                                     o["normal"].data_ptr(), o["best"].data_ptr(),
                                     o["counts"].data_ptr(), o["mask"].data_ptr(), ws.data_ptr(),
                                     _abi.stream_ptr(pts.device)), "wcr_ransac_ground")


def _call_static(lib, pts, count, iters, seed, o):
    from wcr import _abi
    B, cap = pts.shape[0], pts.shape[1]
    _abi.check(lib.wcr_ransac_ground_static(pts.data_ptr(), count.data_ptr(), cap, B, None, iters,
                                            ctypes.c_float(0.05), seed, o["eq"].data_ptr(),
                                            o["normal"].data_ptr(), o["best"].data_ptr(),
                                            o["counts"].data_ptr(), o["mask"].data_ptr(),
                                            _abi.stream_ptr(pts.device)), "wcr_ransac_ground_static")


def _same(a, b, count):
    torch.cuda.synchronize()
    assert torch.equal(a["counts"], b["counts"])
    assert torch.equal(a["best"], b["best"])
    # bit patterns: NaN planes (images with < 3 points) must match too
    assert torch.equal(a["eq"].view(torch.int32), b["eq"].view(torch.int32))
    assert torch.equal(a["normal"].view(torch.int32), b["normal"].view(torch.int32))
    for i, n in enumerate(count.cpu().tolist()):
        assert torch.equal(a["mask"][i, :n], b["mask"][i, :n])


def test_ransac_static_matches_existing_entry_point_twice(cuda_lib):
    lib = cuda_lib
    # call 1: 3 images x 1000 hypotheses; call 2: different shard, fewer images and hypotheses,
    # so stale scratch beyond the second call's extent (and stale planes inside it) are in play
    shard1, pts1, count1 = _inputs(100, 3, 240, 320)
    shard2, pts2, count2 = _inputs(200, 2, 120, 200)
    before = torch.cuda.memory_allocated()
    for (pts, count, iters, seed) in ((pts1, count1, 1000, 11), (pts2, count2, 700, 12)):
        B, cap = pts.shape[0], pts.shape[1]
        ref = _outputs(B, iters, cap, pts.device)
        got = _outputs(B, iters, cap, pts.device)
        _call_existing(lib, pts, count, iters, seed, ref)
        mark = torch.cuda.memory_allocated()
        _call_static(lib, pts, count, iters, seed, got)
        assert torch.cuda.memory_allocated() == mark   # torch's allocator was not touched by the call
        _same(ref, got, count)
        assert (ref["counts"] >= 0).all() and int(ref["counts"].max()) > 0
    assert before >= 0


def test_ransac_static_rejects_oversize(cuda_lib):
    from wcr import _abi
    lib = cuda_lib
    shard, pts, count = _inputs(300, 2, 120, 160)
    B, cap = pts.shape[0], pts.shape[1]
    iters = MAX_HYP // B + 1
    o = _outputs(B, iters, cap, pts.device)
    rc = lib.wcr_ransac_ground_static(pts.data_ptr(), count.data_ptr(), cap, B, None, iters,
                                      ctypes.c_float(0.05), 1, o["eq"].data_ptr(),
                                      o["normal"].data_ptr(), o["best"].data_ptr(),
                                      o["counts"].data_ptr(), None, _abi.stream_ptr(pts.device))
    assert rc == -1   # WCR_EINVAL
    msg = lib.wcr_last_error().decode()
    assert "WCR_RANSAC_STATIC_MAX_HYP" in msg and str(B * iters) in msg
    torch.cuda.synchronize()
    assert (o["counts"] == -5).all()   # nothing was enqueued
    # exactly at the limit is accepted
    iters = MAX_HYP // B
    o = _outputs(B, iters, cap, pts.device)
    ref = _outputs(B, iters, cap, pts.device)
    _call_static(lib, pts, count, iters, 5, o)
    _call_existing(lib, pts, count, iters, 5, ref)
    _same(ref, o, count)
