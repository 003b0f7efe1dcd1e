// score.cu -- fused proposal scoring + per-box top-k (north_star step 3, SURVEY
// A3.2-A3.6, A3.11, A3.12).
//
// Work decomposition. A box's ROI can hold 1.6 k ... 110 k depth pixels, so one
// CTA per box leaves a long tail. The ROI is cut into row blocks of ~SC_ITEM_PX
// pixels; an ITEM is (box, chunk of 256*T proposals, row block). A persistent grid
// (2 CTAs per SM) pulls items from an atomic queue. Per item: proposal state in
// registers, the item's depth pixels stream through shared memory as (a_u, z)
// pairs, per-proposal inside-counts are added to a global counter array. The CTA
// that retires the last item of a (box, chunk) runs the epilogue for that chunk
// (corners -> projection -> clamp -> min/max -> IoU -> score -> top-k); the CTA that
// retires the last chunk of a box merges the chunk candidates. One launch, no
// host round trip, records written straight into the caller's (NCCL send) buffer.
//
// Point test. Every tested point is an unprojected pixel p = z * (a_u, b_v, 1),
// a_u = (u - cx)/fx, b_v = (v - cy)/fy, so the local-frame coordinate factors per
// image row:
//     q_k = R[:,k] . (p - c) = fma(z, fma(A_k, a_u, G_k), -O_k),   (rows pre-scaled by 1/half_k)
//     G_k = fma(B_k, b_v, C_k) per row,  O_k = R[:,k] . c per proposal
// 6 FFMA + 3 compares per (proposal, point) instead of 3 FADD + 9 FFMA. This
// rounding order IS the definition oracle/csrc/wcr_oracle.c restates; counts are
// bit-exact against it (DESIGN.md "Point-in-cube test").
//
// FP32 CUDA-core kernel by construction: nothing here is a dense contraction.
#include <stdlib.h>
#include <string.h>
#include "sample.cuh"
#include "hull.cuh"

namespace wcr {

constexpr int SC_CHUNK = 1024;                 // proposals per item = threads x proposals per thread
constexpr int SC_TILE = 2048;                  // (a,a,z,z) pixels per smem tile = 32 KB
constexpr int SC_ITEM_PX = 4096;               // target depth pixels per item (large shards)
constexpr int SC_ITEM_PX_MIN = 512;            // ... shrunk for small shards so every SM gets items
constexpr int SC_MAX_CHUNKS = 64;

// Read-once global data (depth pixels, cube parameters): keep it out of L1 so that the
// box mask the hull fill re-reads stays resident.
__device__ __forceinline__ float ld_stream(const float *p)
{
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
// 1.0f if max(|q0|,|q1|,|q2|) <= 1 else 0.0f. One FMNMX3 + one FSET; NaN -> 0.
__device__ __forceinline__ float inside_unit(float q0, float q1, float q2)
{
    float m, r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(fabsf(q0)), "f"(fabsf(q1)), "f"(fabsf(q2)));
    asm("set.le.f32.f32 %0, %1, 0f3F800000;" : "=f"(r) : "f"(m));
    return r;
}

// Same decision on coordinates pre-scaled by 2^11 (exact), with the compare moved from the
// half-rate ALU pipe (FSET) to the FMA pipe: C - m^2 with C = 2^22 + 1 is >= 1 exactly when
// m <= 2^11 and <= 0 otherwise (m^2 - 2^22 is a multiple of 2^-24 * 2^22 away from 0, never
// in (0, 1)), one rounding, so .sat yields exactly 1.0f / 0.0f; NaN saturates to 0.
__device__ __forceinline__ float inside_sat(float q0, float q1, float q2)
{
    float m, r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(m) : "f"(fabsf(q0)), "f"(fabsf(q1)), "f"(fabsf(q2)));
    asm("fma.rn.sat.f32 %0, %1, %2, 0f4A800002;" : "=f"(r) : "f"(-m), "f"(m));
    return r;
}

// Integer ROI of a GT box: Boxes.clip((h, w)) then .long() slicing, as the
// reference slices depth patches (roi_heads.py:590-597), sub-sampled by `s`.
struct Roi { int u0, v0, ncols, nrows, rows_per_item, n_rowblocks; };

__device__ __forceinline__ Roi box_roi(const WcrBox &bx, int H, int W, int s, bool use_pts,
                                       int item_px = SC_ITEM_PX)
{
    Roi r;
    r.u0 = (int)fminf(fmaxf(bx.x1, 0.f), (float)W);
    const int u1 = (int)fminf(fmaxf(bx.x2, 0.f), (float)W);
    r.v0 = (int)fminf(fmaxf(bx.y1, 0.f), (float)H);
    const int v1 = (int)fminf(fmaxf(bx.y2, 0.f), (float)H);
    r.ncols = max(0, (u1 - r.u0 + s - 1) / s);
    r.nrows = max(0, (v1 - r.v0 + s - 1) / s);
    if (!use_pts || r.ncols == 0) r.nrows = 0;
    r.rows_per_item = max(1, item_px / max(r.ncols, 1));
    r.n_rowblocks = max(1, (r.nrows + r.rows_per_item - 1) / r.rows_per_item);
    return r;
}

struct PParam;   // score_tiles.cuh

// Where a proposal's 15 floats come from (fetch_cube, score_tiles_kernel.cuh).
struct CubeSrc {
    const float *cubes;                    // caller's [n_boxes][N][15], or NULL: regenerated in-kernel
    const float *priors, *normals, *z0;    // sampler inputs (cubes == NULL)
    WcrSampleParams sp;
    uint64_t seed;
    int N;
};

// Device workspace layout (caller-owned, zeroed by cudaMemsetAsync up to `zero_bytes`).
struct ScoreWs {
    int *queue;          // [1]   next item
    unsigned long long *executed;  // [1] at byte 8: point tests issued (tile path, after culling)
    unsigned long long *hulls;     // [1] at byte 16: hulls filled for the mask term
    int *dynq;           // [3] at byte 24: dynamic hull-batch queue head, tail, chunks through phase A
    int *queue_b;        // [1] at byte 36: next second-pass tile item (point term on demand)
    unsigned long long *pruned;    // [1] at byte 40: proposals never tested against the depth points
    int *segq;           // [n_boxes * n_chunks * 128] published batches (code + 1; 0 = not yet written)
    int *seg_pending;    // [n_boxes][n_chunks] retired batches per chunk
    int *ncand;          // [n_boxes][n_chunks] candidates per chunk
    uint16_t *candlist;  // [n_boxes][n_chunks][SC_CHUNK] candidate proposals (index in the chunk)
    float *ub;           // [n_boxes][N] score bounds, overwritten by actual scores as hulls are filled
    int *item_prefix;    // [n_boxes + 1] exclusive prefix of items per box, in serving order
    int *box_order;      // [n_boxes] box served at each position (largest first)
    int *done_chunk;     // [n_boxes][n_chunks] retired items per (box, chunk)
    int *done_box;       // [n_boxes] retired chunks per box
    int *npts;           // [n_boxes][n_chunks] valid depth points (P_b), one copy per chunk
    int *count;          // [n_boxes][N] inside-counts
    float *cand;         // [n_boxes][n_chunks][k][20] chunk candidates (n_chunks > 1)
    float *cubes;        // [n_boxes][N][15]: baseline build only (the tile path regenerates sampled cubes)
    int2 *segcnt;        // [n_boxes][N] (|seg & hull|, |hull|) pixel counts of the mask term
    PParam *params;      // [n_boxes][N] point-test constants (tile path)
    int4 *cullbox;       // [n_boxes][N] cull boxes (tile path)
    float2 *zgrid;       // [n_boxes][256] depth range of the ROI's 16 x 16 cells (tile path)
    uint8_t *mayhit;     // [n_boxes][N] 0: the proposal provably holds no point (tile path)
    float *siou;         // [n_boxes][N] IoU term, written by score_prep_kernel
    // point term on demand (a.pts_lazy)
    int *zcnt;           // [n_boxes][256] valid depth points of the ROI's 16 x 16 cells
    float *cellab;       // [n_boxes][64] pixel-ray slopes of the cell columns' / rows' first and last pixel
    int *ubcnt;          // [n_boxes][N] upper bound of the inside-count (points of the cells the cube can reach)
    float *okey;         // [n_boxes][N] ordering key: IoU factor x bound factor
    float2 *zblk;        // [n_boxes][64] depth range of the 2 x 2 blocks of cells
    int2 *segb;          // [n_boxes][N] mask-term bounds: (max |mask & hull|, min |hull|) pixels (a.seg_tight)
    int *mint;           // [n_boxes][33 * 33] integral table of the mask pixels per cell of a 32 x 32 ROI grid
    unsigned *segwin;    // [n_boxes][N] cell range of the hull's scan window in that grid (gx0 | gx1 << 6 | gy0 << 12 | gy1 << 18)
    uint16_t *order;     // [n_boxes][n_chunks][SC_CHUNK] chunk proposals by descending key; the first
                         //   pass tests [0, pa), its epilogue compacts the survivors of the rest behind them
    int *done_b;         // [n_boxes][n_chunks] retired second-pass items
    int *bready;         // [n_boxes][n_chunks] 0: first pass running; else 1 + survivors of the second pass
    unsigned *evald;     // [n_boxes][n_chunks][32] proposals whose hull the first pass's epilogue filled (ws.segcnt is valid)
};

struct ScoreArgs {
    const WcrImage *images;
    const WcrBox *boxes;
    int n_boxes;
    const float *cubes;
    CubeSrc src;
    int N;
    WcrScoreParams sc;
    int k;
    int n_chunks;
    int item_px;         // depth pixels per item (plan and item kernel must agree)
    int grp;             // tile path: registers per column group (plan and kernel must agree)
    int ps;              // tile path: proposals per tile item (power of two, 128 .. SC_CHUNK)
    int seg_lazy;        // tile path: mask term evaluated on demand by the epilogues (no seg items)
    int siou_ready;      // tile path: ws.siou holds every proposal's IoU term (written by score_prep_kernel)
    int pts_lazy;        // tile path: point term on demand (two passes over the tiles, see chunk_epilogue_first)
    int pa;              // ... proposals per chunk tested in the first pass (multiple of 32)
    int seg_tight;       // tile path: bounds carry the mask-term bound seg_ratio_bound (ws.segb is filled)
    ScoreWs ws;
    float *records;
    float *out_scores, *out_iou;
    int32_t *out_count;
    float *out_proj;
    int32_t *out_npts;
    const uint32_t *seg_bits;   // bit-packed per-box masks (e_seg > 0)
    const int64_t *seg_off;     // [n_boxes + 1] word offsets
    float *out_seg;
};

}  // namespace wcr
#include "score_tiles.cuh"
namespace wcr {

// Merge n_chunks*k candidate records of one box (warp 0 of the retiring CTA).
__device__ __forceinline__ void warp_merge(const float *cand, int n, int k, float *out)
{
    const int lane = threadIdx.x & 31;
    unsigned long long taken = 0ull;  // candidate lane + 32*i  <->  bit i (n <= 2048)
    for (int r = 0; r < k; ++r) {
        float bv = -INFINITY; int bj = 0x7fffffff, bc = -1;
        for (int i = 0; lane + 32 * i < n; ++i) {
            if (taken >> i & 1ull) continue;
            const int c = lane + 32 * i;
            const int j = __float_as_int(__ldcg(cand + (size_t)c * WCR_RECORD_FLOATS));
            if (j < 0) continue;
            const float v = __ldcg(cand + (size_t)c * WCR_RECORD_FLOATS + 1);
            if (v > bv || (v == bv && j < bj)) { bv = v; bj = j; bc = c; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (ov > bv || (ov == bv && oj < bj)) { bv = ov; bj = oj; bc = oc; }
        }
        float *dst = out + (size_t)r * WCR_RECORD_FLOATS;
        if (bc < 0) {
            if (lane < WCR_RECORD_FLOATS) dst[lane] = lane == 0 ? __int_as_float(-1) : 0.f;
        } else {
            if (lane < WCR_RECORD_FLOATS) dst[lane] = __ldcg(cand + (size_t)bc * WCR_RECORD_FLOATS + lane);
            if ((bc & 31) == lane) taken |= 1ull << (bc >> 5);
        }
    }
}

}  // namespace wcr
#include "score_tiles_kernel.cuh"
#ifdef WCR_KEEP_BASELINE_KERNEL
#include "score_items_baseline.cuh"
#endif

namespace wcr {
cudaError_t launch_sample_cubes(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                                const float *priors, const float *normals, const float *z0,
                                const WcrSampleParams &sp, int N, const float *replay,
                                uint64_t seed, float *out_cubes, cudaStream_t st);
}

namespace {
inline long long sms_resident_warps() { return (long long)wcr::sm_count() * 12; }   // 3 CTAs x 4 warps
inline int n_chunks_for(int N) { return (N + wcr::SC_CHUNK - 1) / wcr::SC_CHUNK; }
inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Carve the caller's workspace. Returns total bytes; *zero_bytes = leading region
// that must be zero at launch.
size_t carve(void *base, int n_boxes, int N, int k, bool need_cubes, bool need_seg,
             wcr::ScoreWs *ws, size_t *zero_bytes)
{
    const int nc = n_chunks_for(N);
    char *p = (char *)base;
    size_t off = 0;
    auto take = [&](size_t bytes) { char *q = p ? p + off : nullptr; off += align256(bytes); return q; };
    int *queue = (int *)take(256);  // header: queues and counters, see ScoreWs (+ phase cycles at byte 64, profiling build)
    int *done_chunk = (int *)take((size_t)n_boxes * nc * sizeof(int));
    int *done_box = (int *)take((size_t)n_boxes * sizeof(int));
    int *npts = (int *)take((size_t)n_boxes * nc * sizeof(int));
    int *count = (int *)take((size_t)n_boxes * N * sizeof(int));
    int *segq = (int *)take(need_seg ? (size_t)n_boxes * nc * 128 * sizeof(int) : 0);
    int *seg_pending = (int *)take(need_seg ? (size_t)n_boxes * nc * sizeof(int) : 0);
    int *done_b = (int *)take((size_t)n_boxes * nc * sizeof(int));
    int *bready = (int *)take((size_t)n_boxes * nc * sizeof(int));
    unsigned *evald = (unsigned *)take(need_seg ? (size_t)n_boxes * nc * 32 * sizeof(unsigned) : 0);
    if (zero_bytes) *zero_bytes = off;
    int *ncand = (int *)take(need_seg ? (size_t)n_boxes * nc * sizeof(int) : 0);
    uint16_t *candlist = (uint16_t *)take(need_seg ? (size_t)n_boxes * nc * wcr::SC_CHUNK * sizeof(uint16_t) : 0);
    float *ub = (float *)take(need_seg ? (size_t)n_boxes * N * sizeof(float) : 0);
    int *prefix = (int *)take((size_t)(n_boxes + 1) * sizeof(int));
    int *box_order = (int *)take((size_t)n_boxes * sizeof(int));
    float *cand = (float *)take(nc > 1 ? (size_t)n_boxes * nc * k * WCR_RECORD_FLOATS * sizeof(float) : 0);
    float *cubes = (float *)take(need_cubes ? (size_t)n_boxes * N * 15 * sizeof(float) : 0);
    int2 *segcnt = (int2 *)take(need_seg ? (size_t)n_boxes * N * sizeof(int2) : 0);
    wcr::PParam *params = (wcr::PParam *)take((size_t)n_boxes * N * sizeof(wcr::PParam));
    int4 *cullbox = (int4 *)take((size_t)n_boxes * N * sizeof(int4));
    float2 *zgrid = (float2 *)take((size_t)n_boxes * 256 * sizeof(float2));
    uint8_t *mayhit = (uint8_t *)take((size_t)n_boxes * N);
    float *siou = (float *)take((size_t)n_boxes * N * sizeof(float));
    int *zcnt = (int *)take((size_t)n_boxes * 256 * sizeof(int));
    float *cellab = (float *)take((size_t)n_boxes * 64 * sizeof(float));
    int *ubcnt = (int *)take((size_t)n_boxes * N * sizeof(int));
    float *okey = (float *)take((size_t)n_boxes * N * sizeof(float));
    float2 *zblk = (float2 *)take((size_t)n_boxes * 64 * sizeof(float2));
    int2 *segb = (int2 *)take(need_seg ? (size_t)n_boxes * N * sizeof(int2) : 0);
    int *mint = (int *)take(need_seg ? (size_t)n_boxes * 33 * 33 * sizeof(int) : 0);
    unsigned *segwin = (unsigned *)take(need_seg ? (size_t)n_boxes * N * sizeof(unsigned) : 0);
    uint16_t *order = (uint16_t *)take((size_t)n_boxes * nc * wcr::SC_CHUNK * sizeof(uint16_t));
    if (ws) {
        ws->zcnt = zcnt; ws->cellab = cellab; ws->ubcnt = ubcnt; ws->okey = okey; ws->zblk = zblk; ws->segb = segb; ws->mint = mint; ws->segwin = segwin; ws->order = order; ws->done_b = done_b; ws->bready = bready; ws->evald = evald;
        ws->queue_b = (int *)(queue ? (char *)queue + 36 : nullptr);
        ws->pruned = (unsigned long long *)(queue ? (char *)queue + 40 : nullptr);
        ws->segcnt = segcnt; ws->params = params; ws->cullbox = cullbox; ws->zgrid = zgrid; ws->mayhit = mayhit; ws->siou = siou;
        ws->queue = queue; ws->executed = (unsigned long long *)(queue ? (char *)queue + 8 : nullptr);
        ws->hulls = (unsigned long long *)(queue ? (char *)queue + 16 : nullptr);
        ws->dynq = (int *)(queue ? (char *)queue + 24 : nullptr);
        ws->segq = segq; ws->seg_pending = seg_pending; ws->ncand = ncand; ws->candlist = candlist; ws->ub = ub; ws->done_chunk = done_chunk; ws->done_box = done_box; ws->npts = npts;
        ws->count = count; ws->item_prefix = prefix; ws->box_order = box_order; ws->cand = cand; ws->cubes = cubes;
    }
    return off;
}
}  // namespace

extern "C" int64_t wcr_score_workspace_bytes(int n_boxes, int N, int k)
{
    if (n_boxes < 0 || N < 1 || k < 1) return -1;
#ifdef WCR_KEEP_BASELINE_KERNEL
    return (int64_t)carve(nullptr, n_boxes, N, k, true, true, nullptr, nullptr);
#else
    // no cube buffer: a sampled proposal is regenerated from its Philox counter where it is needed
    return (int64_t)carve(nullptr, n_boxes, N, k, false, true, nullptr, nullptr);
#endif
}

// Byte offsets, inside a workspace of wcr_score_workspace_bytes(n_boxes, N, k), of the per-proposal
// bounds the on-demand paths prune with (tests check them against the full evaluation).
extern "C" int wcr_score_bounds_offsets(int n_boxes, int N, int k, int64_t *off_ubcnt, int64_t *off_segb,
                                        int64_t *off_segwin, int64_t *off_mint)
{
    if (n_boxes < 0 || N < 1 || k < 1 || !off_ubcnt || !off_segb || !off_segwin || !off_mint) return WCR_EINVAL;
    wcr::ScoreWs w;
    char *base = reinterpret_cast<char *>(4096);
#ifdef WCR_KEEP_BASELINE_KERNEL
    carve(base, n_boxes, N, k, true, true, &w, nullptr);
#else
    carve(base, n_boxes, N, k, false, true, &w, nullptr);
#endif
    *off_ubcnt = (int64_t)((char *)w.ubcnt - base);
    *off_segb = (int64_t)((char *)w.segb - base);
    *off_segwin = (int64_t)((char *)w.segwin - base);
    *off_mint = (int64_t)((char *)w.mint - base);
    return WCR_OK;
}

// CTAs the persistent scoring kernel leaves unlaunched (of its 3 per SM): the hardware then keeps
// one CTA's worth of registers and shared memory free on that many SMs, where a kernel of ANOTHER
// stream -- the next shard's mask gather from host memory -- runs underneath the pass instead of
// after it (wcr_pack_box_masks then launches exactly that many 16-warp CTAs, one per freed slot). Process-wide; 0 (the default) = the whole GPU.
static int g_reserved_ctas = 0;
extern "C" int wcr_score_set_reserved_ctas(int n)
{
    if (n < 0) { wcr::set_error("wcr_score_set_reserved_ctas: n must be >= 0"); return WCR_EINVAL; }
    g_reserved_ctas = n;
    return WCR_OK;
}

extern "C" int wcr_score_topk(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                              const float *cubes, int N, const float *priors,
                              const float *normals, const float *z0,
                              const WcrSampleParams *h_sample, uint64_t seed,
                              const uint32_t *seg_bits, const int64_t *seg_off,
                              const WcrScoreParams *h_score, int k, float *out_records,
                              float *out_scores, float *out_iou, float *out_seg,
                              int32_t *out_count, float *out_proj, int32_t *out_npts, void *ws,
                              void *stream)
{
    return wcr_score_topk_after(images, boxes, n_boxes, cubes, N, priors, normals, z0, h_sample, seed, seg_bits,
                                seg_off, h_score, k, out_records, out_scores, out_iou, out_seg, out_count, out_proj,
                                out_npts, ws, nullptr, stream);
}

// wcr_score_topk whose scoring kernel -- the first launch that reads seg_bits -- waits for
// `seg_ready` (a cudaEvent_t, or NULL): the plan, depth-grid, prep and order kernels of the chain do
// not need the masks and run while another stream is still packing them.
extern "C" int wcr_score_topk_after(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                                    const float *cubes, int N, const float *priors,
                                    const float *normals, const float *z0,
                                    const WcrSampleParams *h_sample, uint64_t seed,
                                    const uint32_t *seg_bits, const int64_t *seg_off,
                                    const WcrScoreParams *h_score, int k, float *out_records,
                                    float *out_scores, float *out_iou, float *out_seg,
                                    int32_t *out_count, float *out_proj, int32_t *out_npts, void *ws,
                                    void *seg_ready, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && boxes && h_score && out_records && ws, "wcr_score_topk: null pointer");
    WCR_REQUIRE(n_boxes >= 0 && N >= 1, "wcr_score_topk: bad sizes");
    WCR_REQUIRE(k >= 1 && k <= WCR_MAX_TOPK, "wcr_score_topk: k must be in [1, %d]", WCR_MAX_TOPK);
    WCR_REQUIRE(cubes || (priors && normals && z0 && h_sample),
                "wcr_score_topk: cubes == NULL needs priors, normals, z0 and h_sample");
    WCR_REQUIRE(h_score->point_stride >= 1, "wcr_score_topk: point_stride must be >= 1");
    WCR_REQUIRE(h_score->e_iou >= 0 && h_score->e_pts >= 0 && h_score->e_seg >= 0,
                "wcr_score_topk: negative exponent");
    WCR_REQUIRE(h_score->seg_eval >= 0 && h_score->seg_eval <= 2, "wcr_score_topk: seg_eval must be 0, 1 or 2");
    WCR_REQUIRE(h_score->pts_eval >= 0 && h_score->pts_eval <= 2, "wcr_score_topk: pts_eval must be 0, 1 or 2");
    WCR_REQUIRE(h_score->map_beta >= 0.f, "wcr_score_topk: map_beta must be >= 0 (0 = no shaping)");
    WCR_REQUIRE(h_score->e_seg == 0 || (seg_bits && seg_off),
                "wcr_score_topk: e_seg > 0 needs seg_bits and seg_off");
    WCR_REQUIRE(((uintptr_t)ws & 255) == 0, "wcr_score_topk: workspace must be 256-byte aligned");
    const int nc = n_chunks_for(N);
    WCR_REQUIRE(nc <= wcr::SC_MAX_CHUNKS, "wcr_score_topk: N too large (max %d)",
                wcr::SC_MAX_CHUNKS * wcr::SC_CHUNK);
    // item indices are 32-bit: seg items alone are n_boxes * chunks * 32, a box has a few hundred
    // tile items per chunk at most
    WCR_REQUIRE((long long)n_boxes * nc * 2048 < 0x7fffffffLL, "wcr_score_topk: n_boxes * ceil(N / 1024) too large");
    if (n_boxes == 0) return WCR_OK;
    cudaStream_t st = (cudaStream_t)stream;

    wcr::ScoreArgs a;
    size_t zero_bytes = 0;
#ifdef WCR_KEEP_BASELINE_KERNEL
    carve(ws, n_boxes, N, k, cubes == nullptr, h_score->e_seg > 0, &a.ws, &zero_bytes);
#else
    carve(ws, n_boxes, N, k, false, h_score->e_seg > 0, &a.ws, &zero_bytes);
#endif
    cudaError_t e = cudaMemsetAsync(ws, 0, zero_bytes, st);
    if (e != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
#ifdef WCR_PROFILE_PHASES
    cudaMemsetAsync((char *)ws + 64 + 12 * 8, 0xff, 8, st);   // earliest start: atomicMin
#endif
    a.images = images; a.boxes = boxes; a.n_boxes = n_boxes; a.cubes = cubes; a.N = N;
    a.sc = *h_score; a.k = k;
    a.n_chunks = nc; a.records = out_records;
    a.out_scores = out_scores; a.out_iou = out_iou; a.out_count = out_count;
    a.out_proj = out_proj; a.out_npts = out_npts;
    a.seg_bits = seg_bits; a.seg_off = seg_off; a.out_seg = out_seg;
    a.item_px = 0; a.grp = 8; a.ps = wcr::SC_CHUNK; a.seg_lazy = 0; a.siou_ready = 0; a.pts_lazy = 0; a.pa = 64; a.seg_tight = 0;
    const int sms = wcr::sm_count();
    WCR_REQUIRE((N + 255) / 256 <= 65535, "wcr_score_topk: N too large");
    a.src.cubes = cubes; a.src.priors = priors; a.src.normals = normals; a.src.z0 = z0;
    a.src.seed = seed; a.src.N = N;
    if (h_sample) a.src.sp = *h_sample; else memset(&a.src.sp, 0, sizeof(a.src.sp));
#ifdef WCR_KEEP_BASELINE_KERNEL
    if (!cubes) {
        wcr::launch_sample_cubes(images, boxes, n_boxes, priors, normals, z0, *h_sample, N, nullptr,
                                 seed, a.ws.cubes, st);
        WCR_CHECK_LAUNCH("sample_cubes_kernel");
    }
    {   // libwcr_b200_baseline.so: the first, unculled kernel (score_items_baseline.cuh)
        // A 64-image shard has ~10^4 items of 4096 px; a single image with a handful of boxes
        // would leave most of the 296 resident CTAs idle, so its items are cut finer.
        const long long bc = (long long)n_boxes * nc;
        long long px = wcr::SC_ITEM_PX * bc / 296;
        int item_px = wcr::SC_ITEM_PX_MIN;
        while (item_px < wcr::SC_ITEM_PX && item_px < px) item_px <<= 1;
        a.item_px = item_px;
        wcr::score_plan_kernel<<<1, 1024, 0, st>>>(images, boxes, n_boxes, nc, h_score->point_stride,
                                                   h_score->e_pts > 0, a.item_px, a.ws.item_prefix);
        WCR_CHECK_LAUNCH("score_plan_kernel");
        if (seg_ready) cudaStreamWaitEvent(st, (cudaEvent_t)seg_ready, 0);
        wcr::score_items_kernel<4, 2, 256><<<sms * 2, 256, 0, st>>>(a);
        WCR_CHECK_LAUNCH("score_items_kernel");
        return WCR_OK;
    }
#else
    // Mask term on demand (chunk_epilogue_warp): valid when the ratio is bounded by 1 (mask_iou;
    // mod_mask_iou = inter^5 / union is not) and nobody asked for per-proposal scores or ratios.
    a.seg_lazy = h_score->e_seg > 0 && h_score->seg_mode == 0 && out_seg == nullptr && out_scores == nullptr;
    // Mask term on demand needs enough (box, chunk) pairs to hide the two dependent phases of a
    // chunk behind other chunks' work; a small shard fills every hull through the parallel seg
    // items instead. seg_eval: 0 = decide here, 1 = every hull, 2 = on demand.
    const bool many_chunks = (long long)n_boxes * nc >= sms_resident_warps() / 4;   // measured: 4.23 vs 4.51 ms at 486 chunks, 1.37 vs 0.90 ms at 64
    if (a.seg_lazy && h_score->seg_eval != 2) a.seg_lazy = h_score->seg_eval == 0 && many_chunks;
    // per-proposal IoU term in the prep kernel unless per-proposal outputs want the projected box
    a.siou_ready = !(out_scores || out_iou || out_seg || out_count || out_proj);
    // Point term on demand (chunk_epilogue_first): the score is bounded by the IoU factor times the
    // share of the ROI's points the cube's cull box can reach, every other factor being <= 1. Valid
    // when nobody wants per-proposal outputs, the terms are not reshaped (custom_mapping goes
    // through pow(): monotone only up to an ulp) and the mask ratio, if any, is bounded by 1.
    a.pts_lazy = h_score->e_pts > 0 && a.siou_ready && h_score->map_beta == 0.f &&
                 (h_score->e_seg == 0 || a.seg_lazy) && h_score->pts_eval != 1 &&
                 (h_score->pts_eval == 2 || many_chunks);
    // Bounds sharpened by a bound of the mask ratio (score_prep_kernel: hull bounding box against
    // the ROI, largest face against the mask area): needs the per-thread IoU pass and plain terms.
    a.seg_tight = a.seg_lazy && a.siou_ready && h_score->map_beta == 0.f;
    {   // proposals per tile item: the whole chunk when the shard has enough tiles for every
        // resident warp (a box has ~30 tiles), finer for small shards
        const long long bc = (long long)n_boxes * nc, want = 4LL * sms_resident_warps();
        int ps = wcr::SC_CHUNK;
        while (!a.pts_lazy && ps > 128 && bc * 32 * (wcr::SC_CHUNK / ps) < want) ps >>= 1;
        a.ps = ps;
    }
    wcr::score_plan_tiles_kernel<<<1, 1024, 0, st>>>(images, boxes, n_boxes, nc, h_score->point_stride,
                                                     h_score->e_pts > 0, 0, a.grp, wcr::SC_CHUNK / a.ps,
                                                     a.ws.item_prefix, a.ws.box_order);
    WCR_CHECK_LAUNCH("score_plan_tiles_kernel");
    if (h_score->e_pts > 0) {
        wcr::score_zgrid_kernel<<<dim3(n_boxes, 8), 256, 0, st>>>(images, boxes, h_score->point_stride, 1, a.ws.zgrid,
                                                                   a.pts_lazy ? a.ws.zcnt : nullptr, a.pts_lazy ? a.ws.cellab : nullptr,
                                                                   a.pts_lazy ? a.ws.zblk : nullptr);
        WCR_CHECK_LAUNCH("score_zgrid_kernel");
    }
    if (h_score->e_pts > 0 || a.siou_ready) {
        const dim3 grid(n_boxes, (N + 255) / 256);
        wcr::score_prep_kernel<<<grid, 256, 0, st>>>(images, boxes, n_boxes, a.src, N,
                                                      h_score->point_stride, h_score->e_pts > 0, a.ws.params, a.ws.cullbox,
                                                      a.ws.zgrid, a.ws.mayhit, a.siou_ready ? a.ws.siou : nullptr,
                                                      a.pts_lazy ? a.ws.zcnt : nullptr, a.ws.cellab, a.ws.ubcnt, a.ws.okey, a.ws.zblk, a.seg_tight ? a.ws.segb : nullptr,
                                                      a.ws.segwin, h_score->e_iou, h_score->e_pts);
        WCR_CHECK_LAUNCH("score_prep_kernel");
    }
    if (a.pts_lazy) {
        wcr::score_order_kernel<<<n_boxes * nc, 1024, 0, st>>>(a.ws.okey, N, nc, a.ws.order);
        WCR_CHECK_LAUNCH("score_order_kernel");
    }
    if (seg_ready) {
        cudaError_t we = cudaStreamWaitEvent(st, (cudaEvent_t)seg_ready, 0);
        if (we != cudaSuccess) { wcr::set_error("cudaStreamWaitEvent: %s", cudaGetErrorString(we)); return WCR_ELAUNCH; }
    }
    if (a.seg_tight) {     // mask pixels per cell of the ROI grid, for the epilogues' bound of |mask & hull|
        wcr::score_maskgrid_kernel<<<n_boxes, wcr::MG * wcr::MG, 0, st>>>(images, boxes, seg_bits, seg_off, a.ws.mint);
        WCR_CHECK_LAUNCH("score_maskgrid_kernel");
    }
    // FFMA2 point test, inside indicator on the integer pipe, 3 CTAs x 4 warps per SM
    wcr::score_tiles_kernel<4, 3, 8, true, true><<<max(sms, sms * 3 - g_reserved_ctas), 128, 0, st>>>(a);
    WCR_CHECK_LAUNCH("score_tiles_kernel");
    return WCR_OK;
#endif
}

// ---- dense per-box masks -> the bit-packed seg_bits layout --------------------------------
// The reference's object masks arrive dense, one H x W map per box (roi_heads.py:383-401,
// masks_all_images); wcr_score_topk wants them bit-packed over each box's integer ROI. A warp
// packs 128 pixels of a row per trip: every lane reads ONE aligned 32-bit word (128 contiguous
// bytes per warp request -- the maps may live in mapped pinned HOST memory, where the request size
// is what PCIe efficiency hangs on), a funnel shift with the neighbour's word undoes the row's
// byte misalignment, the lane's four "non-zero" bits are OR-ed across groups of eight lanes into
// the four output words.
// At most 32 registers, two independent requests in flight per warp, 4 CTAs x 8 warps per box: PCIe
// reads want tens of thousands of bytes in flight (a 148-warp "background" shape meant to sit beside
// the persistent scoring kernel reached 5 GB/s against 32 GB/s and was dropped; one-warp CTAs take
// all 32 CTA slots of an SM and lock the next shard's prep kernels out).
namespace wcr {
// one 128-pixel unit of a row: issue the loads (pack_load), later turn them into 4 words (pack_store)
__device__ __forceinline__ void pack_geom(const uint8_t *m, const Roi &r, int mask_pitch, int upr, int unit,
                                          int &row, int &q, int &npx, int &sh, const uint32_t *&wp)
{
    row = unit / upr; q = unit - row * upr;
    const int c0 = q << 7;
    npx = min(128, r.ncols - c0);                         // pixels of this unit
    const uintptr_t addr = (uintptr_t)(m + (size_t)(r.v0 + row) * mask_pitch + r.u0 + c0);
    sh = (int)(addr & 3);
    wp = reinterpret_cast<const uint32_t *>(addr - sh);   // the pixels occupy bytes [sh, sh + npx) of wp[0 ..]
}

__device__ __forceinline__ void pack_load(const uint8_t *m, const Roi &r, int mask_pitch, int upr, int unit, int lane,
                                          uint32_t &w, uint32_t &wx)
{
    int row, q, npx, sh;
    const uint32_t *wp;
    pack_geom(m, r, mask_pitch, upr, unit, row, q, npx, sh, wp);
    w = 0u; wx = 0u;
    if (4 * lane < sh + npx) w = __ldcs(wp + lane);
    if (lane == 31 && sh + npx > 128) wx = __ldcs(wp + 32);
}

__device__ __forceinline__ void pack_store(const uint8_t *m, const Roi &r, int mask_pitch, int upr, int unit, int lane,
                                           uint32_t w, uint32_t wx, uint32_t *out, int wpr)
{
    int row, q, npx, sh;
    const uint32_t *wp;
    pack_geom(m, r, mask_pitch, upr, unit, row, q, npx, sh, wp);
    uint32_t nxt = __shfl_down_sync(0xffffffffu, w, 1);
    if (lane == 31) nxt = wx;
    const uint32_t by = __funnelshift_r(w, nxt, 8 * sh);  // the lane's pixels 4 lane .. 4 lane + 3
    const int left = npx - 4 * lane;                      // valid pixels from this lane's first on
    uint32_t nib = 0u;
    if (left > 0 && (by & 0x000000ffu)) nib |= 1u;
    if (left > 1 && (by & 0x0000ff00u)) nib |= 2u;
    if (left > 2 && (by & 0x00ff0000u)) nib |= 4u;
    if (left > 3 && (by & 0xff000000u)) nib |= 8u;
    uint32_t v = nib << (4 * (lane & 7));
    v |= __shfl_xor_sync(0xffffffffu, v, 1);
    v |= __shfl_xor_sync(0xffffffffu, v, 2);
    v |= __shfl_xor_sync(0xffffffffu, v, 4);
    const int k = lane >> 3;                              // output word of this lane's group
    if ((lane & 7) == 0 && 32 * k < npx) out[(size_t)row * wpr + q * 4 + k] = v;
}

// grid (x, y): CTA (bx, by) takes boxes bx, bx + gridDim.x, ... and of each the units by, by + gridDim.y,
// ..., U of them at a time (U independent requests in flight per warp)
template <int U>
__global__ void __maxnreg__(32)
pack_box_masks_kernel(const uint8_t *__restrict__ masks, int64_t mask_stride, int mask_pitch,
                      const int32_t *__restrict__ mask_idx, const WcrImage *__restrict__ images,
                      const WcrBox *__restrict__ boxes, int n_boxes, const int64_t *__restrict__ seg_off,
                      uint32_t *__restrict__ out_bits)
{
    const int lane = threadIdx.x & 31, wpc = blockDim.x >> 5, step = gridDim.y * wpc;
    const int first = blockIdx.y * wpc + (threadIdx.x >> 5);
    for (int b = blockIdx.x; b < n_boxes; b += gridDim.x) {
        const WcrBox bx = boxes[b];
        const WcrImage im = images[bx.image];
        const Roi r = box_roi(bx, im.H, im.W, 1, true);
        const int nrows = r.ncols > 0 ? r.nrows : 0, wpr = (r.ncols + 31) >> 5;
        const int upr = (r.ncols + 127) >> 7;             // 128-pixel units per row
        const int n_units = nrows * upr;
        const uint8_t *m = masks + (size_t)(mask_idx ? mask_idx[b] : b) * mask_stride;
        uint32_t *out = out_bits + seg_off[b];
        for (int unit = first; unit < n_units; unit += U * step) {
            uint32_t w[U], wx[U];
#pragma unroll
            for (int i = 0; i < U; ++i)
                if (unit + i * step < n_units) pack_load(m, r, mask_pitch, upr, unit + i * step, lane, w[i], wx[i]);
#pragma unroll
            for (int i = 0; i < U; ++i)
                if (unit + i * step < n_units) pack_store(m, r, mask_pitch, upr, unit + i * step, lane, w[i], wx[i], out, wpr);
        }
    }
}
}  // namespace wcr

static int pack_box_masks_launch(const char *who, const uint8_t *masks, int64_t mask_stride, int mask_pitch,
                                 const int32_t *mask_idx, const WcrImage *images, const WcrBox *boxes,
                                 int n_boxes, const int64_t *seg_off, uint32_t *out_bits, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(masks && images && boxes && seg_off && out_bits, "%s: null pointer", who);
    WCR_REQUIRE(n_boxes >= 0 && mask_pitch >= 1 && mask_stride >= 0, "%s: bad sizes", who);
    if (n_boxes == 0) return WCR_OK;
    WCR_REQUIRE(n_boxes <= 0x7fffffff / 4, "%s: too many boxes", who);
    // a 640 x 480 ROI is 2400 units of 128 pixels: 4 CTAs x 8 warps per box keep a small shard busy too
    // At most 3 CTAs x 8 warps per SM's worth of CTAs, each walking boxes in a grid-stride loop: plenty
    // of requests in flight for PCIe (148 x 24 x 256 B), while two thirds of every SM's thread slots,
    // its registers and all of its shared memory stay free for the next shard's unprojection, RANSAC
    // and prep kernels that run beside a gather from host memory.
    if (g_reserved_ctas > 0) {
        // background shape (wcr_score_set_reserved_ctas): one 16-warp CTA per slot the scoring kernel
        // leaves free -- 16 K registers, beside two scoring CTAs -- each walking whole boxes
        wcr::pack_box_masks_kernel<4><<<dim3(g_reserved_ctas, 1), 512, 0, (cudaStream_t)stream>>>(
            masks, mask_stride, mask_pitch, mask_idx, images, boxes, n_boxes, seg_off, out_bits);
    } else {
        const int ctas = wcr::sm_count() * 3;
        const int gx = n_boxes * 4 <= ctas ? n_boxes : max(1, ctas / 4);
        wcr::pack_box_masks_kernel<2><<<dim3(gx, 4), 256, 0, (cudaStream_t)stream>>>(
            masks, mask_stride, mask_pitch, mask_idx, images, boxes, n_boxes, seg_off, out_bits);
    }
    WCR_CHECK_LAUNCH("pack_box_masks_kernel");
    return WCR_OK;
}

extern "C" int wcr_pack_box_masks(const uint8_t *masks, int64_t mask_stride, int mask_pitch,
                                  const int32_t *mask_idx, const WcrImage *images, const WcrBox *boxes,
                                  int n_boxes, const int64_t *seg_off, uint32_t *out_bits, void *stream)
{
    return pack_box_masks_launch("wcr_pack_box_masks", masks, mask_stride, mask_pitch, mask_idx, images, boxes,
                                 n_boxes, seg_off, out_bits, stream);
}

// Host -> device copy of just the ROI rectangle of every box's dense mask: one cudaMemcpy2DAsync
// per box on `stream` (copy engine; overlaps kernels of other streams). The dense maps are the
// reference's masks_all_images in PINNED host memory; the device buffer has the same layout, only
// the rectangles wcr_pack_box_masks reads are refreshed. h_rois [n][4] = u0 v0 u1 v1 (HOST array,
// the integer ROIs: Boxes.clip + .long(), as wcr_score_topk slices them).
extern "C" int wcr_h2d_box_mask_rois(const uint8_t *h_masks, uint8_t *d_masks, int64_t mask_stride,
                                     int mask_pitch, const int32_t *h_rois, const int32_t *h_mask_idx,
                                     int n_boxes, void *stream)
{
    WCR_REQUIRE(h_masks && d_masks && h_rois, "wcr_h2d_box_mask_rois: null pointer");
    WCR_REQUIRE(n_boxes >= 0 && mask_pitch >= 1 && mask_stride >= 0, "wcr_h2d_box_mask_rois: bad sizes");
    for (int b = 0; b < n_boxes; ++b) {
        const int u0 = h_rois[4 * b], v0 = h_rois[4 * b + 1], u1 = h_rois[4 * b + 2], v1 = h_rois[4 * b + 3];
        if (u1 <= u0 || v1 <= v0) continue;
        const size_t off = (size_t)(h_mask_idx ? h_mask_idx[b] : b) * (size_t)mask_stride +
                           (size_t)v0 * (size_t)mask_pitch + (size_t)u0;
        cudaError_t e = cudaMemcpy2DAsync(d_masks + off, (size_t)mask_pitch, h_masks + off, (size_t)mask_pitch,
                                          (size_t)(u1 - u0), (size_t)(v1 - v0), cudaMemcpyHostToDevice,
                                          (cudaStream_t)stream);
        if (e != cudaSuccess) {
            wcr::set_error("wcr_h2d_box_mask_rois: cudaMemcpy2DAsync (box %d): %s", b, cudaGetErrorString(e));
            return WCR_ELAUNCH;
        }
    }
    return WCR_OK;
}

// The same refresh as ONE contiguous copy per box: from the ROI's first pixel to its last, whole
// rows in between ((nrows - 1) * pitch + ncols bytes). About three times the bytes of the
// rectangles on SUN RGB-D-sized boxes, but a 2-D copy of short rows runs at a fraction of the
// link's rate (6 GB/s measured for ~200-byte rows against 55 GB/s for contiguous blocks), and
// copy engines need no SM: the transfer hides behind the previous shard's pass.
extern "C" int wcr_h2d_box_mask_rows(const uint8_t *h_masks, uint8_t *d_masks, int64_t mask_stride,
                                     int mask_pitch, const int32_t *h_rois, const int32_t *h_mask_idx,
                                     int n_boxes, int64_t *out_bytes, void *stream)
{
    WCR_REQUIRE(h_masks && d_masks && h_rois, "wcr_h2d_box_mask_rows: null pointer");
    WCR_REQUIRE(n_boxes >= 0 && mask_pitch >= 1 && mask_stride >= 0, "wcr_h2d_box_mask_rows: bad sizes");
    int64_t total = 0;
    for (int b = 0; b < n_boxes; ++b) {
        const int u0 = h_rois[4 * b], v0 = h_rois[4 * b + 1], u1 = h_rois[4 * b + 2], v1 = h_rois[4 * b + 3];
        if (u1 <= u0 || v1 <= v0) continue;
        const size_t off = (size_t)(h_mask_idx ? h_mask_idx[b] : b) * (size_t)mask_stride +
                           (size_t)v0 * (size_t)mask_pitch + (size_t)u0;
        const size_t bytes = (size_t)(v1 - v0 - 1) * (size_t)mask_pitch + (size_t)(u1 - u0);
        cudaError_t e = cudaMemcpyAsync(d_masks + off, h_masks + off, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream);
        if (e != cudaSuccess) {
            wcr::set_error("wcr_h2d_box_mask_rows: cudaMemcpyAsync (box %d): %s", b, cudaGetErrorString(e));
            return WCR_ELAUNCH;
        }
        total += (int64_t)bytes;
    }
    if (out_bytes) *out_bytes = total;
    return WCR_OK;
}
