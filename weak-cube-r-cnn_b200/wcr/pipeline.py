"""The whole hot path for one image shard, as one object with resident buffers.

    ground normal (unproject + RANSAC)  ->  pseudo depth per box
    ->  fused [sample -> corners -> project -> IoU -> point-in-cube -> score -> top-k]

NEW orchestration (the reference's driver for this is absent from the snapshot,
SURVEY.md §0). ``ShardRunner.run_host()`` is the end-to-end call bench.py times as
``e2e``: pinned host buffers in, top-k records out on the host.
"""
from __future__ import annotations

import torch

from . import _abi, ground
from .images import ImageShard
from .proposals import (BoxTable, ProposalConfig, ScoreConfig, ScoreWorkspace, SegMasks,
                        box_center_depth, score_topk, RECORD_FLOATS)


class ShardRunner:
    def __init__(self, images: list[dict], N: int = 1000, device="cuda", *,
                 sample_cfg: ProposalConfig | None = None, score_cfg: ScoreConfig | None = None,
                 priors: torch.Tensor, seed: int = 0, records: torch.Tensor | None = None,
                 dense_seg: bool = False):
        """``images``: host dicts with depth (H,W) f32, ground (H,W) u8 or (1,1),
        K (3,3), boxes (b,4), classes (b,) and -- when ``score_cfg.e_seg > 0`` -- seg
        (b,H,W) per-box masks or seg_packed (``pack_box_masks`` output). All images are uploaded into two padded device planes
        so a step's H2D is a handful of large copies. ``dense_seg``: the per-box masks stay DENSE
        on the host side (what the reference's ``object_masks`` produces, roi_heads.py:383-401), in
        pinned memory; every ``upload`` copies the ROI rectangle of each box's map H2D (one 2-D copy
        per box, ``wcr_h2d_box_mask_rois``: only what the packer reads crosses PCIe) and every
        ``step`` bit-packs them on the device (``wcr_pack_box_masks``). ``dense_seg="copy"``: whole
        maps, n_boxes x H x W bytes per step; ``dense_seg="rows"``: one CONTIGUOUS copy per box, from
        the ROI's first pixel to its last (``wcr_h2d_box_mask_rows``; ~3x the rectangles' bytes but at
        the link's full rate and on a copy engine, so it hides behind the previous shard's pass: the
        fastest mode inside a ``ShardPipeline``); ``dense_seg="mapped"``: no copy at all, the pack
        kernel gathers the ROI bytes from the pinned maps over PCIe during ``upload`` (fewest bytes;
        best for one shard at a time -- it needs SMs, so it cannot overlap the persistent scoring
        kernel). Default (False): packed once on the host.

        The geometry of a runner is FROZEN at construction: image sizes, intrinsics, boxes and
        mask offsets are part of the device tables. Re-uploading (``upload``, ``ShardPipeline``)
        may change the depth / ground / mask CONTENTS only; a batch with other boxes needs its
        own runner (or ``wcr.job.StreamedBlocks``, which re-uploads the tables per block)."""
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _abi.WcrError("wcr: CUDA device required (there is no CPU fallback on this path)")
        self.N, self.seed = N, seed
        self.sample_cfg = sample_cfg or ProposalConfig()
        self.score_cfg = score_cfg or ScoreConfig()
        B = len(images)
        Hm = max(im["depth"].shape[0] for im in images)
        Wm = max(im["depth"].shape[1] for im in images)
        self.h_depth = torch.zeros((B, Hm, Wm), dtype=torch.float32).pin_memory()
        self.h_ground = torch.zeros((B, Hm, Wm), dtype=torch.uint8).pin_memory()
        self.has_ground = []
        for i, im in enumerate(images):
            h, w = im["depth"].shape
            self.h_depth[i, :h, :w] = im["depth"]
            g = im["ground"]
            self.has_ground.append(tuple(g.shape) != (1, 1))
            if self.has_ground[-1]:
                self.h_ground[i, :h, :w] = (g > 0).to(torch.uint8)
        self.d_depth = torch.empty_like(self.h_depth, device=self.device)
        self.d_ground = torch.empty_like(self.h_ground, device=self.device)
        depth_views = [self.d_depth[i, :im["depth"].shape[0], :im["depth"].shape[1]]
                       for i, im in enumerate(images)]
        ground_views = [self.d_ground[i, :im["depth"].shape[0], :im["depth"].shape[1]]
                        if self.has_ground[i] else torch.ones((1, 1), dtype=torch.uint8)
                        for i, im in enumerate(images)]
        self.shard = ImageShard(depth_views, ground_views, [im["K"] for im in images],
                                device=self.device)
        self.boxes = BoxTable(self.shard, [im["boxes"] for im in images],
                              [im["classes"] for im in images])
        self.h_boxes = torch.from_numpy(self.boxes.host.view("uint8").copy()).pin_memory()
        self.seg = self.h_seg = self.h_seg_dense = self.d_seg_dense = None
        if self.score_cfg.e_seg > 0 and dense_seg:
            nb_ = len(self.boxes)
            self.h_seg_dense = torch.zeros((nb_, Hm, Wm), dtype=torch.uint8).pin_memory()
            g = 0
            for im in images:
                h, w = im["depth"].shape
                m = im["seg"]
                self.h_seg_dense[g:g + len(m), :h, :w] = (torch.as_tensor(m) != 0).to(torch.uint8)
                g += len(m)
            self.dense_mode = dense_seg if isinstance(dense_seg, str) else "roi"
            if self.dense_mode == "mapped":
                self.seg = SegMasks.from_dense(self.shard, self.boxes, self.h_seg_dense)
            else:
                self.d_seg_dense = self.h_seg_dense.to(self.device)
                self.seg = SegMasks.from_dense(self.shard, self.boxes, self.d_seg_dense)
                from .proposals import box_rois
                import numpy as _np
                self._h_rois = _np.ascontiguousarray(box_rois(self.boxes, self.shard), dtype=_np.int32)
                rr = self._h_rois.astype(_np.int64)
                ok = (rr[:, 2] > rr[:, 0]) & (rr[:, 3] > rr[:, 1])
                self.row_band_bytes = int((((rr[:, 3] - rr[:, 1] - 1) * self.h_seg_dense.stride(1) + rr[:, 2] - rr[:, 0]) * ok).sum())
        elif self.score_cfg.e_seg > 0:
            if all("seg_packed" in im for im in images):
                self.seg = SegMasks(self.shard, self.boxes, packed=[w for im in images for w in im["seg_packed"]])
            else:
                self.seg = SegMasks(self.shard, self.boxes, [m for im in images for m in im["seg"]])
            self.h_seg = torch.from_numpy(self.seg.host_bits.view("int32").copy()).pin_memory()
        self.priors = priors.to(self.device, torch.float32).contiguous()
        nb = len(self.boxes)
        self.ws = ScoreWorkspace(nb, N, self.score_cfg.k, self.device, records=records)
        self.h_records = torch.empty((nb, self.score_cfg.k, RECORD_FLOATS),
                                     dtype=torch.float32).pin_memory()
        self.n_boxes = nb
        self.n_proposals = nb * N
        self.normals = None

    # bytes a step moves over PCIe in run_host()
    @property
    def h2d_bytes(self) -> int:
        seg = self.h_seg.numel() * 4 if self.h_seg is not None else 0
        if self.h_seg_dense is not None:      # bytes that cross PCIe for the masks
            seg = self.h_seg_dense.numel() if self.dense_mode == "copy" else \
                self.row_band_bytes if self.dense_mode == "rows" else self.seg.roi_bytes
        return self.h_depth.numel() * 4 + self.h_ground.numel() + self.h_boxes.numel() + seg

    @property
    def d2h_bytes(self) -> int:
        return self.h_records.numel() * 4

    def upload(self):
        self.upload_inputs()
        self.upload_masks()

    def upload_inputs(self):
        """Depth, ground maps and boxes: everything the kernels ahead of the scoring kernel read."""
        self.upload_boxes()
        self.d_depth.copy_(self.h_depth, non_blocking=True)
        self.d_ground.copy_(self.h_ground, non_blocking=True)

    def upload_boxes(self):
        self.boxes.table.copy_(self.h_boxes, non_blocking=True)

    def upload_masks(self):
        """The per-box masks (packed words, or dense maps packed on the device): only the scoring
        kernel reads them (``step(seg_ready=event)``)."""
        if self.h_seg is not None:
            self.seg.bits.copy_(self.h_seg, non_blocking=True)
        if self.d_seg_dense is not None and self.dense_mode == "copy":
            self.d_seg_dense.copy_(self.h_seg_dense, non_blocking=True)
        elif self.h_seg_dense is not None and self.d_seg_dense is None:
            # "mapped": the pack kernel gathers the ROI bytes straight from the pinned maps (128-byte
            # reads over PCIe) -- here, on the stream that uploads, so that in a ShardPipeline it
            # overlaps the previous shard's pass
            self.seg.repack(self.h_seg_dense)
        elif self.d_seg_dense is not None and self.dense_mode == "rows":
            m = self.h_seg_dense
            _abi.check(_abi.load().wcr_h2d_box_mask_rows(
                m.data_ptr(), self.d_seg_dense.data_ptr(), m.stride(0), m.stride(1),
                self._h_rois.ctypes.data, None, self.n_boxes, None, _abi.stream_ptr(self.device)),
                "wcr_h2d_box_mask_rows")
        elif self.d_seg_dense is not None:
            m = self.h_seg_dense
            _abi.check(_abi.load().wcr_h2d_box_mask_rois(
                m.data_ptr(), self.d_seg_dense.data_ptr(), m.stride(0), m.stride(1),
                self._h_rois.ctypes.data, None, self.n_boxes, _abi.stream_ptr(self.device)),
                "wcr_h2d_box_mask_rois")

    def step(self, score_events=None, seg_ready=None):
        """Device-resident pass: 10 kernel launches (ground unprojection, RANSAC count + select, box depth,
        plan, depth grid, prep, order, mask grid, scoring), no host sync, no D2H.
        ``score_events`` = (start, end) torch.cuda.Events recorded around the fused
        scoring launch (bench.py's roofline timing). ``seg_ready``: event after which the masks
        of this step are on the device (``upload_masks`` on another stream); None: already there."""
        self._seg_ready = seg_ready
        self.normals = ground.normal_vector_from_maps(None, None, None, shard=self.shard,
                                                      seed=self.seed)
        z0 = box_center_depth(self.shard, self.boxes)
        if self.d_seg_dense is not None:         # dense masks of this step -> seg_bits, on the device
            self.seg.repack(self.d_seg_dense)
        if score_events is not None:
            score_events[0].record()
            out = self._score(z0)
            score_events[1].record()
            return out
        return self._score(z0)

    def _score(self, z0):
        return score_topk(self.shard, self.boxes, None, self.N, self.score_cfg,
                          priors=self.priors, normals=self.normals, z0=z0,
                          sample_cfg=self.sample_cfg, seed=self.seed, seg=self.seg,
                          workspace=self.ws, seg_ready=getattr(self, "_seg_ready", None))

    def download(self):
        self.h_records.copy_(self.ws.records[:self.n_boxes], non_blocking=True)

    def run_host(self) -> torch.Tensor:
        """End to end: pinned host inputs -> device -> hot path -> records on host."""
        self.upload()
        self.step()
        self.download()
        torch.cuda.current_stream(self.device).synchronize()
        return self.h_records


class ShardPipeline:
    """Double-buffered end-to-end driver: while shard i is being scored, shard i + 1 is
    uploaded and the records of shard i - 1 are downloaded (three CUDA streams, events
    between them; no host synchronisation inside ``submit``).

        pipe = ShardPipeline([runner_a, runner_b])
        for i, batch in enumerate(batches):
            pipe.wait(i - 2)            # slot free again (its records were consumed)
            fill(pipe.slot(i), batch)   # write the pinned host buffers of the slot
            pipe.submit(i)
        records = pipe.wait(i)          # pinned (n_boxes, k, 20) of shard i

    NEW orchestration (no reference counterpart)."""

    def __init__(self, runners: list[ShardRunner], after_step=None, reserve_ctas: int | None = None):
        """``after_step[k]()``, if given, is called on the compute stream right after slot k's pass
        (bench.py hangs the NCCL all-gather of the top-k records there)."""
        self.after_step = after_step
        assert len(runners) >= 2 and len({r.device for r in runners}) == 1
        self.runners = runners
        dev = runners[0].device
        self.device = dev
        # mapped-memory mask gathers run beside the pass in CTA slots the scoring kernel leaves free
        self.reserve_ctas = reserve_ctas if reserve_ctas is not None else \
            (18 if any(r.h_seg_dense is not None and r.d_seg_dense is None for r in runners) else 0)
        self.s_mask = torch.cuda.Stream(dev)
        self.s_up = torch.cuda.Stream(dev)
        self.s_run = torch.cuda.Stream(dev)
        self.s_down = torch.cuda.Stream(dev)
        n = len(runners)
        self.ev_in = [torch.cuda.Event() for _ in range(n)]
        self.ev_up = [torch.cuda.Event() for _ in range(n)]
        self.ev_run = [torch.cuda.Event() for _ in range(n)]
        self.ev_down = [torch.cuda.Event() for _ in range(n)]
        self.pending = [False] * n
        # work enqueued before the pipeline existed (allocations, warm-up) is on the current stream
        for s in (self.s_up, self.s_run, self.s_down, self.s_mask):
            s.wait_stream(torch.cuda.current_stream(dev))

    def slot(self, i: int) -> ShardRunner:
        return self.runners[i % len(self.runners)]

    def submit(self, i: int) -> None:
        k = i % len(self.runners)
        r = self.runners[k]
        gather = r.h_seg_dense is not None and r.d_seg_dense is None   # masks gathered from mapped host memory
        with torch.cuda.stream(self.s_up):
            if self.pending[k]:
                self.s_up.wait_event(self.ev_run[k])    # the slot's previous pass has read its inputs
            r.upload_inputs()
            self.ev_in[k].record(self.s_up)
            if not gather:
                r.upload_masks()                        # only the scoring kernel waits for these
                self.ev_up[k].record(self.s_up)
        if self.reserve_ctas:
            _abi.check(_abi.load().wcr_score_set_reserved_ctas(self.reserve_ctas), "wcr_score_set_reserved_ctas")
        if gather:
            # The gather is a KERNEL reading host memory: it runs in the CTA slots the scoring kernel of
            # the other slot's pass leaves free (reserve_ctas), AFTER the copies of this slot -- SM loads
            # over PCIe beside copy-engine traffic slow the pass itself (5.6 vs 5.0 ms per step,
            # tools/time_interference.py); one behind the other both hide inside the pass.
            with torch.cuda.stream(self.s_mask):
                self.s_mask.wait_event(self.ev_in[k])
                r.upload_masks()
                self.ev_up[k].record(self.s_mask)
        with torch.cuda.stream(self.s_run):
            self.s_run.wait_event(self.ev_in[k])
            if self.pending[k]:
                self.s_run.wait_event(self.ev_down[k])  # the slot's previous records are on the host
            dense_on_device = r.d_seg_dense is not None # packed by step() itself on this stream: needs the maps
            if dense_on_device:
                self.s_run.wait_event(self.ev_up[k])
            try:
                r.step(seg_ready=None if dense_on_device else self.ev_up[k])
            finally:
                if self.reserve_ctas:
                    _abi.load().wcr_score_set_reserved_ctas(0)
            if self.after_step is not None:
                self.after_step[k]()
            self.ev_run[k].record(self.s_run)
        with torch.cuda.stream(self.s_down):
            self.s_down.wait_event(self.ev_run[k])
            r.download()
            self.ev_down[k].record(self.s_down)
        self.pending[k] = True

    def wait(self, i: int):
        """Block until shard i's records are on the host; returns them (None if i < 0)."""
        if i < 0:
            return None
        k = i % len(self.runners)
        if self.pending[k]:
            self.ev_down[k].synchronize()
        return self.runners[k].h_records

    def drain(self) -> None:
        for s in (self.s_up, self.s_mask, self.s_run, self.s_down):
            s.synchronize()
