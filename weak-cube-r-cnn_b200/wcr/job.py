"""Strong-scaling job over a long image list (BASELINE.json configs[3]: 8192 images at
1/2/4/8 GPUs), cut into BLOCKS of ``block`` images.

A block is the unit every kernel of the path is launched on (what `ShardRunner` calls a
shard); a rank owns a contiguous range of blocks (``wcr.dist.shard_range`` over blocks, which
for 8192 images / 64 per block / 1, 2, 4, 8 ranks is exactly ``shard_range(8192, r, N)``).
Because blocks are cut at GLOBAL image positions and every generator is keyed by
(seed + global block id; index inside the block), the records of a box do not depend on how
many ranks there are: the all-gathered records of N ranks must equal the single-process
records bit for bit (``bench.py: gather_verified``, tests/test_dist_gpu.py).

NEW orchestration (no reference counterpart): per-image independence is
`roi_heads.py:472`; the only collective is the final all-gather of top-k records.

Two drivers over the same `BlockSet`:
  * `ResidentBlocks`  -- every image of the rank resident in HBM (device-resident `value`);
  * `StreamedBlocks`  -- host buffers in, records out, double-buffered over two slots (`e2e`).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _abi, ground
from .images import IMAGE_DTYPE, ImageShard
from .proposals import (BOX_DTYPE, RECORD_FLOATS, BoxTable, ProposalConfig, ScoreConfig,
                        ScoreWorkspace, SegMasks, box_center_depth, pack_box_masks, score_topk)


class BlockSet:
    """Host-side description of the global image list, tiled from a fixed BASE set.

    ``base``: host image dicts (depth (H,W) f32, ground (H,W) u8 or the (1,1) dummy, K (3,3),
    boxes (b,4), classes (b,), and -- for the mask term -- seg (b,H,W) or seg_packed).
    Global image g shows base image ``image_map[g]`` (a fixed pseudo-random map, so box counts
    per block vary like a real data set's); nothing else about g is shared between repeats:
    the proposal and RANSAC streams are keyed by the block and the position in it."""

    def __init__(self, base: list[dict], n_images: int, block: int = 64, map_seed: int = 0,
                 use_seg: bool = True):
        assert n_images >= 1 and block >= 1
        self.base, self.n_images, self.block, self.use_seg = base, n_images, block, use_seg
        self.n_blocks = -(-n_images // block)
        self.image_map = np.random.default_rng(map_seed).integers(0, len(base), n_images)
        self.Hm = max(im["depth"].shape[0] for im in base)
        self.Wm = max(im["depth"].shape[1] for im in base)
        nb = len(base)
        # pinned base planes: the host source of every H2D copy of StreamedBlocks
        self.h_depth = torch.zeros((nb, self.Hm, self.Wm), dtype=torch.float32)
        self.h_ground = torch.zeros((nb, self.Hm, self.Wm), dtype=torch.uint8)
        if torch.cuda.is_available():       # (host-only uses, e.g. the CPU tests of the block arithmetic, skip pinning)
            self.h_depth, self.h_ground = self.h_depth.pin_memory(), self.h_ground.pin_memory()
        self.has_ground = np.zeros(nb, dtype=bool)
        self._img = np.zeros(nb, dtype=IMAGE_DTYPE)       # pointers filled per use
        self._boxes, self._seg = [], []
        for j, im in enumerate(base):
            h, w = im["depth"].shape
            self.h_depth[j, :h, :w] = im["depth"]
            g = im["ground"]
            self.has_ground[j] = tuple(g.shape) != (1, 1)
            if self.has_ground[j]:
                self.h_ground[j, :h, :w] = (g > 0).to(torch.uint8)
            K = np.asarray(im["K"], dtype=np.float32)
            self._img[j] = (0, 0, h, w, self.Wm, self.Wm, K[0, 0], K[1, 1], K[0, 2], K[1, 2])
            bx = torch.as_tensor(im["boxes"], dtype=torch.float32).numpy().reshape(-1, 4)
            r = np.zeros(len(bx), dtype=BOX_DTYPE)
            r["x1"], r["y1"], r["x2"], r["y2"] = bx[:, 0], bx[:, 1], bx[:, 2], bx[:, 3]
            r["cls"] = torch.as_tensor(im["classes"]).numpy().reshape(-1)
            r["clamp0"], r["clamp1"] = h, w           # the ROI head's [H, W] (roi_heads.py:801-806)
            self._boxes.append(r)
            if use_seg:
                if "seg_packed" in im:
                    words = list(im["seg_packed"])
                else:
                    rois = np.stack([np.minimum(np.maximum(bx[:, 0], 0), w).astype(np.int32),
                                     np.minimum(np.maximum(bx[:, 1], 0), h).astype(np.int32),
                                     np.minimum(np.maximum(bx[:, 2], 0), w).astype(np.int32),
                                     np.minimum(np.maximum(bx[:, 3], 0), h).astype(np.int32)], 1)
                    words = pack_box_masks(rois, list(im["seg"]))
                self._seg.append(words)
        self.base_boxes = np.array([len(r) for r in self._boxes], dtype=np.int64)
        per_image = self.base_boxes[self.image_map]
        pad = self.n_blocks * block - n_images
        self.block_boxes = np.concatenate([per_image, np.zeros(pad, np.int64)]).reshape(self.n_blocks, block).sum(1)
        self.box_offset = np.concatenate([[0], np.cumsum(self.block_boxes)])   # global box index of each block
        self.max_block_boxes = int(self.block_boxes.max())

    def images_of(self, b: int) -> np.ndarray:
        """base indices of the images of global block b"""
        return self.image_map[b * self.block:min((b + 1) * self.block, self.n_images)]

    def tables(self, b: int, depth_ptrs, ground_ptrs):
        """Host tables of block b: (IMAGE_DTYPE records with the given device addresses,
        BOX_DTYPE records, seg words u32 or None, seg offsets i64 or None)."""
        idx = self.images_of(b)
        img = self._img[idx].copy()
        img["depth"] = depth_ptrs
        img["ground"] = np.where(self.has_ground[idx], ground_ptrs, 0)
        recs = []
        for i, j in enumerate(idx):
            r = self._boxes[j].copy()
            r["image"] = i
            recs.append(r)
        boxes = np.concatenate(recs) if recs else np.zeros(0, dtype=BOX_DTYPE)
        bits = off = None
        if self.use_seg:
            words = [w for j in idx for w in self._seg[j]]
            off = np.zeros(len(words) + 1, dtype=np.int64)
            off[1:] = np.cumsum([len(w) for w in words])
            bits = np.concatenate(words).astype(np.uint32) if off[-1] else np.zeros(1, dtype=np.uint32)
        return img, boxes, bits, off

    def n_boxes_of(self, blocks) -> int:
        return int(sum(self.block_boxes[b] for b in blocks))

    def block_weights(self, N: int = 1000):
        """Predicted work per block: point tests (N x ROI pixels) plus a per-proposal term for
        the epilogue and the mask hulls -- what ``wcr.dist.balanced_assignment`` balances."""
        w_img = np.zeros(len(self.base), dtype=np.float64)
        for j, r in enumerate(self._boxes):
            h, w = int(self._img["H"][j]), int(self._img["W"][j])
            u0, u1 = np.clip(r["x1"], 0, w).astype(np.int64), np.clip(r["x2"], 0, w).astype(np.int64)
            v0, v1 = np.clip(r["y1"], 0, h).astype(np.int64), np.clip(r["y2"], 0, h).astype(np.int64)
            area = np.maximum(u1 - u0, 0) * np.maximum(v1 - v0, 0)
            w_img[j] = float(area.sum()) * 0.45 + 6000.0 * len(r)      # 45 % of the pairs are tested
        per_image = w_img[self.image_map]
        pad = self.n_blocks * self.block - self.n_images
        return np.concatenate([per_image, np.zeros(pad)]).reshape(self.n_blocks, self.block).sum(1) * N

    def global_order(self, rank_blocks):
        """Index that puts the rank-major concatenation of the ranks' records (rank r holds its
        blocks in ascending id) back into global box order: ``records_global = cat[index]``."""
        pos, at = {}, 0
        for blocks in rank_blocks:
            for b in blocks:
                pos[b] = at
                at += int(self.block_boxes[b])
        idx = np.concatenate([np.arange(pos[b], pos[b] + int(self.block_boxes[b]), dtype=np.int64)
                              for b in range(self.n_blocks)]) if self.n_blocks else np.zeros(0, np.int64)
        return idx


class _Block:
    """Device descriptors of one block (image table, box table, masks, workspace view)."""
    __slots__ = ("b", "shard", "boxes", "seg", "ws", "lo", "hi", "pins")


def _run_block(blk: _Block, N, score_cfg, sample_cfg, priors, seed):
    """The whole hot path on one block: ground normals -> pseudo depth -> fused
    sample + score + top-k into the block's slice of the record buffer."""
    normals = ground.normal_vector_from_maps(None, None, None, shard=blk.shard, seed=seed + blk.b)
    z0 = box_center_depth(blk.shard, blk.boxes)
    score_topk(blk.shard, blk.boxes, None, N, score_cfg, priors=priors, normals=normals, z0=z0,
               sample_cfg=sample_cfg, seed=seed + blk.b, seg=blk.seg, workspace=blk.ws)


class ResidentBlocks:
    """Blocks ``blocks`` (global ids) of a `BlockSet` with every image resident in HBM.

    ``records`` (n_boxes_of(blocks), k, 20): where the top-k records go, box-major in block
    order -- normally ``TopkGather.send`` so the all-gather needs no staging copy."""

    def __init__(self, bs: BlockSet, blocks, N: int, device, *, priors, score_cfg: ScoreConfig,
                 sample_cfg: ProposalConfig | None = None, seed: int = 0, records=None):
        self.bs, self.N, self.seed = bs, N, seed
        self.device = dev = torch.device(device)
        self.score_cfg, self.sample_cfg = score_cfg, sample_cfg or ProposalConfig()
        self.priors = priors.to(dev, torch.float32).contiguous()
        self.block_ids = list(blocks)
        idx = np.concatenate([bs.images_of(b) for b in self.block_ids]) if self.block_ids else np.zeros(0, np.int64)
        n_img = len(idx)
        base_d, base_g = bs.h_depth.to(dev), bs.h_ground.to(dev)
        self.d_depth = torch.empty((max(n_img, 1), bs.Hm, bs.Wm), dtype=torch.float32, device=dev)
        self.d_ground = torch.empty((max(n_img, 1), bs.Hm, bs.Wm), dtype=torch.uint8, device=dev)
        t_idx = torch.from_numpy(idx).to(dev)
        for s in range(0, n_img, 256):      # bounded temporaries
            torch.index_select(base_d, 0, t_idx[s:s + 256], out=self.d_depth[s:s + 256])
            torch.index_select(base_g, 0, t_idx[s:s + 256], out=self.d_ground[s:s + 256])
        del base_d, base_g
        self.n_boxes = bs.n_boxes_of(self.block_ids)
        self.n_proposals = self.n_boxes * N
        k = score_cfg.k
        self.records = records if records is not None else torch.empty(
            (max(self.n_boxes, 1), k, RECORD_FLOATS), dtype=torch.float32, device=dev)
        assert self.records.shape[0] >= self.n_boxes
        ws0 = ScoreWorkspace(max(bs.max_block_boxes, 1), N, k, dev, records=self.records)
        plane_d, plane_g = bs.Hm * bs.Wm * 4, bs.Hm * bs.Wm
        self.blocks, lo, img0 = [], 0, 0
        for b in self.block_ids:
            n_b = len(bs.images_of(b))
            dp = self.d_depth.data_ptr() + (img0 + np.arange(n_b, dtype=np.uint64)) * plane_d
            gp = self.d_ground.data_ptr() + (img0 + np.arange(n_b, dtype=np.uint64)) * plane_g
            img, boxes, bits, off = bs.tables(b, dp, gp)
            blk = _Block()
            blk.b, blk.lo, blk.hi = b, lo, lo + len(boxes)
            blk.shard = ImageShard.from_host_table(img, dev)
            blk.boxes = BoxTable.from_host(boxes, dev)
            blk.seg = SegMasks.from_arrays(bits, off, dev) if bs.use_seg else None
            blk.ws = ScoreWorkspace(len(boxes), N, k, dev, records=self.records[blk.lo:blk.hi], ws=ws0.ws)
            blk.pins = None
            self.blocks.append(blk)
            lo, img0 = blk.hi, img0 + n_b
        assert lo == self.n_boxes

    def run_pass(self):
        """One device-resident pass over every block of the rank: no host sync, no copies."""
        for blk in self.blocks:
            if len(blk.boxes):
                _run_block(blk, self.N, self.score_cfg, self.sample_cfg, self.priors, self.seed)
        return self.records[:self.n_boxes]


class StreamedBlocks:
    """End-to-end driver: per block, the images' planes are copied host -> device from pinned
    memory, the block's tables with them, the hot path runs, and the block's records are
    copied device -> host; two slots and three streams, so block i + 1 uploads and block i - 1
    downloads while block i is scored (every block still moves all of its bytes).

    The host planes are the BlockSet's pinned base planes (the global list is tiled from
    them); tables and packed masks per block are prebuilt pinned arrays (set-up, like a data
    loader's collate output)."""

    def __init__(self, bs: BlockSet, blocks, N: int, device, *, priors, score_cfg: ScoreConfig,
                 sample_cfg: ProposalConfig | None = None, seed: int = 0, records=None,
                 n_slots: int = 2):
        self.bs, self.N, self.seed = bs, N, seed
        self.device = dev = torch.device(device)
        self.score_cfg, self.sample_cfg = score_cfg, sample_cfg or ProposalConfig()
        self.priors = priors.to(dev, torch.float32).contiguous()
        self.block_ids = list(blocks)
        self.n_boxes = bs.n_boxes_of(self.block_ids)
        self.n_proposals = self.n_boxes * N
        k = score_cfg.k
        self.records = records if records is not None else torch.empty(
            (max(self.n_boxes, 1), k, RECORD_FLOATS), dtype=torch.float32, device=dev)
        self.h_records = torch.empty((max(self.n_boxes, 1), k, RECORD_FLOATS), dtype=torch.float32).pin_memory()
        cap_boxes = max(bs.max_block_boxes, 1)
        ws0 = ScoreWorkspace(cap_boxes, N, k, dev, records=self.records)
        tabs, max_words = [], 1
        self.slots = []
        for s in range(n_slots):
            self.slots.append({
                "depth": torch.empty((bs.block, bs.Hm, bs.Wm), dtype=torch.float32, device=dev),
                "ground": torch.empty((bs.block, bs.Hm, bs.Wm), dtype=torch.uint8, device=dev),
                "img": torch.empty((bs.block * IMAGE_DTYPE.itemsize,), dtype=torch.uint8, device=dev),
                "box": torch.empty((cap_boxes * BOX_DTYPE.itemsize,), dtype=torch.uint8, device=dev),
                "off": torch.empty((cap_boxes + 1,), dtype=torch.int64, device=dev)})
        plane_d, plane_g = bs.Hm * bs.Wm * 4, bs.Hm * bs.Wm
        for i, b in enumerate(self.block_ids):
            sl = self.slots[i % n_slots]
            n_b = len(bs.images_of(b))
            dp = sl["depth"].data_ptr() + np.arange(n_b, dtype=np.uint64) * plane_d
            gp = sl["ground"].data_ptr() + np.arange(n_b, dtype=np.uint64) * plane_g
            t = bs.tables(b, dp, gp)
            tabs.append(t)
            if t[2] is not None:
                max_words = max(max_words, len(t[2]))
        for sl in self.slots:
            sl["bits"] = torch.empty((max_words,), dtype=torch.int32, device=dev)
        self.blocks, lo = [], 0
        self.h2d_bytes = self.d2h_bytes = 0
        for i, (b, (img, boxes, bits, off)) in enumerate(zip(self.block_ids, tabs)):
            sl = self.slots[i % n_slots]
            blk = _Block()
            blk.b, blk.lo, blk.hi = b, lo, lo + len(boxes)
            blk.shard = ImageShard.from_host_table(img, dev, table=sl["img"])
            blk.boxes = BoxTable.from_host(boxes, dev, table=sl["box"])
            blk.seg = SegMasks.from_arrays(bits, off, dev, bits=sl["bits"], off=sl["off"]) if bs.use_seg else None
            blk.ws = ScoreWorkspace(len(boxes), N, k, dev, records=self.records[blk.lo:blk.hi], ws=ws0.ws)
            blk.pins = {"img": torch.from_numpy(img.view(np.uint8).copy()).pin_memory(),
                        "box": torch.from_numpy(boxes.view(np.uint8).copy()).pin_memory() if len(boxes) else None,
                        "bits": torch.from_numpy(bits.view(np.int32).copy()).pin_memory() if bits is not None else None,
                        "off": torch.from_numpy(off.copy()).pin_memory() if off is not None else None}
            idx = bs.images_of(b)
            self.h2d_bytes += len(idx) * plane_d + int(bs.has_ground[idx].sum()) * plane_g
            self.h2d_bytes += sum(v.numel() * v.element_size() for v in blk.pins.values() if v is not None)
            self.d2h_bytes += len(boxes) * k * RECORD_FLOATS * 4
            self.blocks.append(blk)
            lo = blk.hi
        self.s_up, self.s_run, self.s_down = (torch.cuda.Stream(dev) for _ in range(3))
        self.ev_up = [torch.cuda.Event() for _ in range(n_slots)]
        self.ev_run = [torch.cuda.Event() for _ in range(n_slots)]
        self.used = [False] * n_slots

    def _upload(self, i: int):
        blk, sl = self.blocks[i], self.slots[i % len(self.slots)]
        idx = self.bs.images_of(blk.b)
        for n, j in enumerate(idx):
            sl["depth"][n].copy_(self.bs.h_depth[j], non_blocking=True)
            if self.bs.has_ground[j]:
                sl["ground"][n].copy_(self.bs.h_ground[j], non_blocking=True)
        p = blk.pins
        sl["img"][:p["img"].numel()].copy_(p["img"], non_blocking=True)
        if p["box"] is not None:
            sl["box"][:p["box"].numel()].copy_(p["box"], non_blocking=True)
        if p["bits"] is not None:
            sl["bits"][:p["bits"].numel()].copy_(p["bits"], non_blocking=True)
            sl["off"][:p["off"].numel()].copy_(p["off"], non_blocking=True)

    def run_pass(self, after=None):
        """Every block of the rank, host buffers -> records on the host (``self.h_records``).
        ``after()``, if given, is enqueued on the compute stream after the last block (the
        all-gather of the record buffer). Returns after a host synchronisation."""
        cur = torch.cuda.current_stream(self.device)
        for s in (self.s_up, self.s_run, self.s_down):
            s.wait_stream(cur)
        n_slots = len(self.slots)
        self.used = [False] * n_slots
        for i, blk in enumerate(self.blocks):
            k = i % n_slots
            with torch.cuda.stream(self.s_up):
                if self.used[k]:
                    self.s_up.wait_event(self.ev_run[k])      # the slot's previous block has been scored
                self._upload(i)
                self.ev_up[k].record(self.s_up)
            with torch.cuda.stream(self.s_run):
                self.s_run.wait_event(self.ev_up[k])
                if len(blk.boxes):
                    _run_block(blk, self.N, self.score_cfg, self.sample_cfg, self.priors, self.seed)
                self.ev_run[k].record(self.s_run)
            with torch.cuda.stream(self.s_down):
                self.s_down.wait_event(self.ev_run[k])
                if blk.hi > blk.lo:
                    self.h_records[blk.lo:blk.hi].copy_(self.records[blk.lo:blk.hi], non_blocking=True)
            self.used[k] = True
        if after is not None:
            with torch.cuda.stream(self.s_run):
                after()
        for s in (self.s_up, self.s_run, self.s_down):
            s.synchronize()
        return self.h_records[:self.n_boxes]
