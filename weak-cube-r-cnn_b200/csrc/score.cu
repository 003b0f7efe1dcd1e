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

typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi)
{
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(u64 v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
// Blackwell packed FP32: two independent IEEE fma.rn per instruction (FFMA2).
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c)
{
    u64 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ u64 fadd2(u64 a, u64 b)
{
    u64 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
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

// Device workspace layout (caller-owned, zeroed by cudaMemsetAsync up to `zero_bytes`).
struct ScoreWs {
    int *queue;          // [1]   next item
    unsigned long long *executed;  // [1] at byte 8: point tests issued (tile path, after culling)
    int *item_prefix;    // [n_boxes + 1] exclusive prefix of items per box
    int *done_chunk;     // [n_boxes][n_chunks] retired items per (box, chunk)
    int *done_box;       // [n_boxes] retired chunks per box
    int *npts;           // [n_boxes][n_chunks] valid depth points (P_b), one copy per chunk
    int *count;          // [n_boxes][N] inside-counts
    float *cand;         // [n_boxes][n_chunks][k][20] chunk candidates (n_chunks > 1)
    float *cubes;        // [n_boxes][N][15] when sampled by the library
    int2 *segcnt;        // [n_boxes][N] (|seg & hull|, |hull|) pixel counts of the mask term
    PParam *params;      // [n_boxes][N] point-test constants (tile path)
    int4 *cullbox;       // [n_boxes][N] cull boxes (tile path)
    float2 *zgrid;       // [n_boxes][256] depth range of the ROI's 16 x 16 cells (tile path)
    uint8_t *mayhit;     // [n_boxes][N] 0: the proposal provably holds no point (tile path)
};

struct ScoreArgs {
    const WcrImage *images;
    const WcrBox *boxes;
    int n_boxes;
    const float *cubes;
    int N;
    WcrScoreParams sc;
    int k;
    int n_chunks;
    int item_px;         // depth pixels per item (plan and item kernel must agree)
    int grp;             // tile path: registers per column group (plan and kernel must agree)
    int split_queues;    // tile path: separate queues for tile and seg items
    int ps;              // tile path: proposals per tile item (power of two, 128 .. SC_CHUNK)
    int seg_lazy;        // tile path: mask term skipped for proposals that provably hold no point
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

// One CTA: items per box -> exclusive prefix. n_boxes is at most a few 100 k.
__global__ void __launch_bounds__(1024)
score_plan_kernel(const WcrImage *__restrict__ images, const WcrBox *__restrict__ boxes,
                  int n_boxes, int n_chunks, int stride, int use_pts, int item_px,
                  int *__restrict__ prefix)
{
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n_boxes; base += 1024) {
        const int b = base + tid;
        int n = 0;
        if (b < n_boxes) {
            const WcrBox bx = boxes[b];
            const WcrImage im = images[bx.image];
            n = box_roi(bx, im.H, im.W, stride, use_pts != 0, item_px).n_rowblocks * n_chunks;
        }
        int incl = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            const int t = s_warp[lane];
            int w = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += x;
            }
            s_warp[lane] = w - t;  // exclusive over warps
        }
        __syncthreads();
        const int carry = s_carry;
        if (b < n_boxes) prefix[b] = carry + s_warp[wid] + incl - n;
        __syncthreads();
        if (tid == 1023) s_carry = carry + s_warp[wid] + incl;
        __syncthreads();
    }
    if (tid == 0) prefix[n_boxes] = s_carry;
}

// k rounds of (max score, lowest index) over the CTA's SC_CHUNK candidates in smem.
// Winner r's local index (or -1) lands in s_win[r], its score in s_wsc[r].
template <int SC_THREADS>
__device__ __forceinline__ void cta_topk(float *s_score, int k, float *s_red_v, int *s_red_i,
                                         int *s_win, float *s_wsc)
{
    constexpr int SC_T = SC_CHUNK / SC_THREADS;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int r = 0; r < k; ++r) {
        float bv = -INFINITY; int bi = 0x7fffffff;
#pragma unroll
        for (int t = 0; t < SC_T; ++t) {
            const int loc = t * SC_THREADS + tid;
            const float v = s_score[loc];
            if (v > bv) { bv = v; bi = loc; }  // ascending loc: keeps the lowest index
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_red_v[wid] = bv; s_red_i[wid] = bi; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < SC_THREADS / 32; ++w)
                if (s_red_v[w] > bv || (s_red_v[w] == bv && s_red_i[w] < bi)) { bv = s_red_v[w]; bi = s_red_i[w]; }
            if (bv == -INFINITY) bi = -1;
            s_win[r] = bi; s_wsc[r] = bv;
            if (bi >= 0) s_score[bi] = -INFINITY;  // taken
        }
        __syncthreads();
    }
}

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
namespace wcr {

// The FIRST scoring kernel: thread = T proposals in registers, every ROI pixel broadcast to them
// from shared memory, no culling. Superseded by score_tiles_kernel; kept behind
// WCR_SCORE_VARIANT=4 (or 1) as the measured baseline and as an independent implementation
// the tile kernel is cross-checked against at full size (tests/test_score_fullsize_gpu.py).
// V selects the inside indicator: 1: FMNMX3 + FSET, 4: FMNMX3 + FFMA.SAT on 2^11-scaled rows.
template <int V, int MINB, int SC_THREADS>
__global__ void __launch_bounds__(SC_THREADS, MINB) score_items_kernel(const ScoreArgs a)
{
    constexpr int T = SC_CHUNK / SC_THREADS;   // proposals per thread
    __shared__ __align__(16) float4 tile[SC_TILE];
    __shared__ float s_score[SC_CHUNK];
    __shared__ float s_iou[SC_CHUNK];
    __shared__ float s_seg[SC_CHUNK];
    __shared__ float s_red_v[SC_THREADS / 32];
    __shared__ int s_red_i[SC_THREADS / 32];
    __shared__ int s_win[WCR_MAX_TOPK];
    __shared__ float s_wsc[WCR_MAX_TOPK];
    __shared__ int s_item, s_flag, s_area;

    const int tid = threadIdx.x, lane = tid & 31;
    const int *__restrict__ prefix = a.ws.item_prefix;
    const int total = prefix[a.n_boxes];
    const float *__restrict__ cubes = a.cubes ? a.cubes : a.ws.cubes;
    const bool use_pts = a.sc.e_pts > 0;
    const int s = a.sc.point_stride;

    for (;;) {
        __syncthreads();
        if (tid == 0) s_item = atomicAdd(a.ws.queue, 1);
        __syncthreads();
        const int item = s_item;
        if (item >= total) break;
        int b = 0, hi = a.n_boxes;  // prefix[b] <= item < prefix[hi]
        while (hi - b > 1) {
            const int mid = (b + hi) >> 1;
            if (prefix[mid] <= item) b = mid; else hi = mid;
        }
        const WcrBox bx = a.boxes[b];
        const WcrImage im = a.images[bx.image];
        const Intr K = {im.fx, im.fy, im.cx, im.cy};
        const Roi roi = box_roi(bx, im.H, im.W, s, use_pts, a.item_px);
        const int local = item - prefix[b];
        const int chunk = local / roi.n_rowblocks, rb = local - chunk * roi.n_rowblocks;
        const int row_begin = rb * roi.rows_per_item;
        const int row_end = min(roi.nrows, row_begin + roi.rows_per_item);

        // ---- prologue: proposal state -> registers, two proposals per FFMA2 lane pair
        // unit-cube form: row k of the test is scaled by 1/half_k so that
        // inside <=> max_k |q'_k| <= 1 (one FMNMX3 + one FSET instead of 3 FSETP).
        u64 A2[T / 2][3], B2[T / 2][3], C2[T / 2][3], NO2[T / 2][3];
#pragma unroll
        for (int pr = 0; pr < T / 2; ++pr) {
            float Af[2][3], Bf[2][3], Cf[2][3], NOf[2][3];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int j = chunk * SC_CHUNK + (2 * pr + e) * SC_THREADS + tid;
                bool live = j < a.N && row_end > row_begin;
                float cube[15];
                if (live) {
                    const float *src = cubes + ((size_t)b * a.N + j) * 15;
#pragma unroll
                    for (int i = 0; i < 15; ++i) cube[i] = ld_stream(src + i);
                    live = cube[3] > 0.f && cube[4] > 0.f && cube[5] > 0.f;  // empty cube holds no point
                }
                if (live) {
                    const float half[3] = {cube[5] / 2.f, cube[4] / 2.f, cube[3] / 2.f};
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        Af[e][k] = cube[6 + k] / half[k];
                        Bf[e][k] = cube[9 + k] / half[k];
                        Cf[e][k] = cube[12 + k] / half[k];
                        NOf[e][k] = -fmaf(Cf[e][k], cube[2], fmaf(Bf[e][k], cube[1], Af[e][k] * cube[0]));
                        if (V == 4) {   // exact power-of-two scaling for inside_sat
                            Af[e][k] *= 2048.f; Bf[e][k] *= 2048.f; Cf[e][k] *= 2048.f; NOf[e][k] *= 2048.f;
                        }
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 3; ++k) { Af[e][k] = Bf[e][k] = Cf[e][k] = 0.f; NOf[e][k] = nanf_(); }
                }
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                A2[pr][k] = pack2(Af[0][k], Af[1][k]); B2[pr][k] = pack2(Bf[0][k], Bf[1][k]);
                C2[pr][k] = pack2(Cf[0][k], Cf[1][k]); NO2[pr][k] = pack2(NOf[0][k], NOf[1][k]);
            }
        }
        float cntf[T];
#pragma unroll
        for (int t = 0; t < T; ++t) cntf[t] = 0.f;

        // ---- main loop: this item's rows, tile by tile --------------------------
        int my_valid = 0;
        if (row_end > row_begin) {
            const int ncols = roi.ncols;
            const int pieces = (ncols + SC_TILE - 1) / SC_TILE;
            const int segpad = min(ncols, SC_TILE);
            const int spt = SC_TILE / segpad;  // row segments per tile, >= 1
            const int total_segs = (row_end - row_begin) * pieces;
            for (int seg0 = 0; seg0 < total_segs; seg0 += spt) {
                const int nseg = min(spt, total_segs - seg0);
                if (seg0) __syncthreads();
                for (int idx = tid; idx < nseg * segpad; idx += SC_THREADS) {
                    const int ss = idx / segpad, pos = idx - ss * segpad;
                    const int seg = seg0 + ss;
                    const int row = seg / pieces, piece = seg - row * pieces;
                    const int col = piece * SC_TILE + pos;
                    float av = 0.f, z = nanf_();
                    if (col < ncols) {
                        const int u = roi.u0 + col * s, v = roi.v0 + (row_begin + row) * s;
                        const float d = ld_stream(im.depth + (size_t)v * im.depth_pitch + u);
                        av = ((float)u - K.cx) / K.fx;
                        if (d > 0.f && d <= 3.402823466e38f) { z = d; ++my_valid; }
                    }
                    tile[idx] = make_float4(av, z, av, z);
                }
                __syncthreads();
                for (int ss = 0; ss < nseg; ++ss) {
                    const int row = (seg0 + ss) / pieces;
                    const float bv = ((float)(roi.v0 + (row_begin + row) * s) - K.cy) / K.fy;
                    const u64 bvbv = pack2(bv, bv);
                    u64 G2[T / 2][3];
#pragma unroll
                    for (int pr = 0; pr < T / 2; ++pr)
#pragma unroll
                        for (int k = 0; k < 3; ++k) G2[pr][k] = ffma2(B2[pr][k], bvbv, C2[pr][k]);
                    {
                        float Af[T][3], Gf[T][3], NOf[T][3];
#pragma unroll
                        for (int pr = 0; pr < T / 2; ++pr)
#pragma unroll
                            for (int k = 0; k < 3; ++k) {
                                unpack2(A2[pr][k], Af[2 * pr][k], Af[2 * pr + 1][k]);
                                unpack2(G2[pr][k], Gf[2 * pr][k], Gf[2 * pr + 1][k]);
                                unpack2(NO2[pr][k], NOf[2 * pr][k], NOf[2 * pr + 1][k]);
                            }
                        const float4 *src = tile + ss * segpad;
#pragma unroll 4
                        for (int p = 0; p < segpad; ++p) {
                            const float2 pp = *reinterpret_cast<const float2 *>(src + p);   // (a_u, z): one LDS.64
#pragma unroll
                            for (int t = 0; t < T; ++t) {
                                const float q0 = fmaf(pp.y, fmaf(Af[t][0], pp.x, Gf[t][0]), NOf[t][0]);
                                const float q1 = fmaf(pp.y, fmaf(Af[t][1], pp.x, Gf[t][1]), NOf[t][1]);
                                const float q2 = fmaf(pp.y, fmaf(Af[t][2], pp.x, Gf[t][2]), NOf[t][2]);
                                cntf[t] += V == 1 ? inside_unit(q0, q1, q2) : inside_sat(q0, q1, q2);
                            }
                        }
                    }
                }
            }
        }
        int cnt[T];
#pragma unroll
        for (int t = 0; t < T; ++t) cnt[t] = (int)cntf[t];

        // ---- mask term of this item's share of the chunk's proposals ---------------
        // The hull rasterisation does not depend on the points, so the chunk's proposals
        // are dealt out over the box's row-block items (no serial epilogue, no tail):
        // item rb takes proposals [rb*per, rb*per + per) of the chunk. Thread = proposal
        // for projection + Jarvis march (hull left in the thread's scratch column of the
        // now idle point tile), then each warp walks its 32 hulls with lanes = pixel rows.
        const bool use_seg = a.sc.e_seg > 0;
        SegBox sb = {nullptr, 0, 0, 0, 0, 0, 0};
        float lo0, hi0, lo1, hi1;
        clamp_bounds(bx.clamp0, lo0, hi0);
        clamp_bounds(bx.clamp1, lo1, hi1);
        if (use_seg) {
            const Roi r1 = box_roi(bx, im.H, im.W, 1, true);
            sb.u0 = r1.u0; sb.v0 = r1.v0; sb.ncols = r1.ncols; sb.nrows = r1.ncols > 0 ? r1.nrows : 0;
            sb.wpr = (sb.ncols + 31) >> 5;
            sb.bits = a.seg_bits + a.seg_off[b];
            const int chunk_n = min(SC_CHUNK, a.N - chunk * SC_CHUNK);
            const int per = (chunk_n + roi.n_rowblocks - 1) / roi.n_rowblocks;
            const int q0 = rb * per, q1 = min(q0 + per, chunk_n);
            float *scratch = reinterpret_cast<float *>(tile) + tid;
            const Strided spx = {scratch, SC_THREADS}, spy = {scratch + 8 * SC_THREADS, SC_THREADS};
            const Strided shx = {scratch + 16 * SC_THREADS, SC_THREADS}, shy = {scratch + 24 * SC_THREADS, SC_THREADS};
            float *wbase = reinterpret_cast<float *>(tile) + (tid & ~31);
            const float hx2 = (float)(int)(bx.clamp0 - 1.f), hy2 = (float)(int)(bx.clamp1 - 1.f);
            __syncthreads();  // every thread is done reading the point tile
#pragma unroll 1
            for (int base = q0; base < q1; base += SC_THREADS) {
                // hull h of this round -> (lane h / 8, warp h % 8): spreads a short round over all warps
                const int q = base + lane * (SC_THREADS / 32) + (tid >> 5);
                int m_own = -1;
                if (q < q1) {
                    const float *src = cubes + ((size_t)b * a.N + chunk * SC_CHUNK + q) * 15;
                    float cube[15], cxy[16];
#pragma unroll
                    for (int i = 0; i < 15; ++i) cube[i] = ld_stream(src + i);
                    project_cube(cube, K, lo0, hi0, lo1, hi1, cxy);
                    // second clamp of the corners (roi_heads.py:950-952)
                    bool finite = true;
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        finite = finite && fabsf(cxy[2 * i]) <= 3.402823466e38f && fabsf(cxy[2 * i + 1]) <= 3.402823466e38f;
                        spx(i) = fminf(fmaxf(cxy[2 * i], 0.f), hx2);
                        spy(i) = fminf(fmaxf(cxy[2 * i + 1], 0.f), hy2);
                    }
                    if (finite) {
                        m_own = jarvis_march(spx, spy, 8, shx, shy);
                        if (m_own > 0) hull_edges(shx, shy, m_own, spx, spy);  // corners are dead: reuse
                    }
                }
                __syncwarp();
                int my_inter = 0, my_bube = 0;
#pragma unroll 1
                for (int p = 0; p < 32; ++p) {
                    const int mp = __shfl_sync(0xffffffffu, m_own, p);
                    if (mp < 1) continue;
                    const Strided exp_ = {wbase + p, SC_THREADS}, eyp = {wbase + p + 8 * SC_THREADS, SC_THREADS};
                    const Strided hxp = {wbase + p + 16 * SC_THREADS, SC_THREADS};
                    const Strided hyp = {wbase + p + 24 * SC_THREADS, SC_THREADS};
                    int it, bu;
                    seg_counts_warp(hxp, hyp, exp_, eyp, mp, im.H, im.W, sb, it, bu);
                    if (lane == p) { my_inter = it; my_bube = bu; }
                }
                __syncwarp();
                if (q < q1) a.ws.segcnt[(size_t)b * a.N + chunk * SC_CHUNK + q] = make_int2(my_inter, my_bube);
            }
        }

        // ---- retire the item: counts -> global, ticket ---------------------------
#pragma unroll
        for (int t = 0; t < T; ++t) {
            const int j = chunk * SC_CHUNK + t * SC_THREADS + tid;
            if (cnt[t] > 0) atomicAdd(a.ws.count + (size_t)b * a.N + j, cnt[t]);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) my_valid += __shfl_xor_sync(0xffffffffu, my_valid, o);
        if (lane == 0 && my_valid) atomicAdd(a.ws.npts + (size_t)b * a.n_chunks + chunk, my_valid);
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const int ticket = atomicAdd(a.ws.done_chunk + (size_t)b * a.n_chunks + chunk, 1);
            s_flag = (ticket == roi.n_rowblocks - 1);
        }
        __syncthreads();
        if (!s_flag) continue;
        __threadfence();

        // ---- epilogue of (box, chunk): corners -> projection -> IoU -> score ------
        const int npts = __ldcg(a.ws.npts + (size_t)b * a.n_chunks + chunk);
        if (a.out_npts && chunk == 0 && tid == 0) a.out_npts[b] = npts;
        // |seg| of the box mask: CTA-wide popcount (the per-proposal pixel counts were
        // written by the items)
        if (use_seg) {
            if (tid == 0) s_area = 0;
            __syncthreads();
            int part = 0;
            for (int w = tid; w < sb.nrows * sb.wpr; w += SC_THREADS) part += __popc(__ldg(sb.bits + w));
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            if (lane == 0 && part) atomicAdd(&s_area, part);
            __syncthreads();
            sb.area = s_area;
        }
#pragma unroll 1
        for (int t = 0; t < T; ++t) {
            const int loc = t * SC_THREADS + tid;
            const int j = chunk * SC_CHUNK + loc;
            const size_t o = (size_t)b * a.N + j;
            float score = -INFINITY, siou = 0.f, spts = 0.f, sseg = 1.f;
            int c = 0;
            ProjBox pb = {0.f, 0.f, 0.f, 0.f};
            if (j < a.N) {
                const float *src = cubes + o * 15;
                float cube[15];
#pragma unroll
                for (int i = 0; i < 15; ++i) cube[i] = ld_stream(src + i);
                c = __ldcg(a.ws.count + o);
                pb = project_cube(cube, K, lo0, hi0, lo1, hi1, nullptr);
                siou = iou_2d(bx.x1, bx.y1, bx.x2, bx.y2, pb);
                spts = npts > 0 ? (float)c / (float)npts : 0.f;
                if (use_seg) {
                    const int2 sc2 = __ldcg(a.ws.segcnt + o);
                    sseg = seg_ratio(sc2.x, sc2.y, sb.area, a.sc.seg_mode);
                }
            }
            if (j < a.N) {
                score = (ipow(siou, a.sc.e_iou) * ipow(spts, a.sc.e_pts)) * ipow(sseg, a.sc.e_seg);
                if (a.out_scores) a.out_scores[o] = score;
                if (a.out_iou) a.out_iou[o] = siou;
                if (a.out_seg) a.out_seg[o] = sseg;
                if (a.out_count) a.out_count[o] = c;
                if (a.out_proj) {
                    float *d = a.out_proj + o * 4;
                    d[0] = pb.x1; d[1] = pb.y1; d[2] = pb.x2; d[3] = pb.y2;
                }
            }
            s_score[loc] = score; s_iou[loc] = siou; s_seg[loc] = sseg;
        }
        __syncthreads();
        cta_topk<SC_THREADS>(s_score, a.k, s_red_v, s_red_i, s_win, s_wsc);
        float *rec_base = a.n_chunks == 1
                              ? a.records + (size_t)b * a.k * WCR_RECORD_FLOATS
                              : a.ws.cand + ((size_t)b * a.n_chunks + chunk) * a.k * WCR_RECORD_FLOATS;
        if (tid < a.k) {
            const int loc = s_win[tid];
            float *rec = rec_base + (size_t)tid * WCR_RECORD_FLOATS;
            if (loc < 0) {
                rec[0] = __int_as_float(-1);
                for (int i = 1; i < WCR_RECORD_FLOATS; ++i) rec[i] = 0.f;
            } else {
                const int j = chunk * SC_CHUNK + loc;
                const float *src = cubes + ((size_t)b * a.N + j) * 15;
                rec[0] = __int_as_float(j);
                rec[1] = s_wsc[tid];
                for (int i = 0; i < 15; ++i) rec[2 + i] = __ldg(src + i);
                const int c = __ldcg(a.ws.count + (size_t)b * a.N + j);
                rec[17] = s_iou[loc]; rec[18] = npts > 0 ? (float)c / (float)npts : 0.f;
                rec[19] = s_seg[loc];
            }
        }
        if (a.n_chunks > 1) {
            __threadfence();
            __syncthreads();
            if (tid == 0) s_flag = (atomicAdd(a.ws.done_box + b, 1) == a.n_chunks - 1);
            __syncthreads();
            if (s_flag && tid < 32) {
                __threadfence();
                warp_merge(a.ws.cand + (size_t)b * a.n_chunks * a.k * WCR_RECORD_FLOATS,
                           a.n_chunks * a.k, a.k, a.records + (size_t)b * a.k * WCR_RECORD_FLOATS);
            }
        }
    }
}

}  // namespace wcr

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
    int *queue = (int *)take(sizeof(int));
    int *done_chunk = (int *)take((size_t)n_boxes * nc * sizeof(int));
    int *done_box = (int *)take((size_t)n_boxes * sizeof(int));
    int *npts = (int *)take((size_t)n_boxes * nc * sizeof(int));
    int *count = (int *)take((size_t)n_boxes * N * sizeof(int));
    if (zero_bytes) *zero_bytes = off;
    int *prefix = (int *)take((size_t)(n_boxes + 1) * sizeof(int));
    float *cand = (float *)take(nc > 1 ? (size_t)n_boxes * nc * k * WCR_RECORD_FLOATS * sizeof(float) : 0);
    float *cubes = (float *)take(need_cubes ? (size_t)n_boxes * N * 15 * sizeof(float) : 0);
    int2 *segcnt = (int2 *)take(need_seg ? (size_t)n_boxes * N * sizeof(int2) : 0);
    wcr::PParam *params = (wcr::PParam *)take((size_t)n_boxes * N * sizeof(wcr::PParam));
    int4 *cullbox = (int4 *)take((size_t)n_boxes * N * sizeof(int4));
    float2 *zgrid = (float2 *)take((size_t)n_boxes * 256 * sizeof(float2));
    uint8_t *mayhit = (uint8_t *)take((size_t)n_boxes * N);
    if (ws) {
        ws->segcnt = segcnt; ws->params = params; ws->cullbox = cullbox; ws->zgrid = zgrid; ws->mayhit = mayhit;
        ws->queue = queue; ws->executed = (unsigned long long *)(queue ? (char *)queue + 8 : nullptr); ws->done_chunk = done_chunk; ws->done_box = done_box; ws->npts = npts;
        ws->count = count; ws->item_prefix = prefix; ws->cand = cand; ws->cubes = cubes;
    }
    return off;
}
}  // namespace

extern "C" int64_t wcr_score_workspace_bytes(int n_boxes, int N, int k)
{
    if (n_boxes < 0 || N < 1 || k < 1) return -1;
    // sized for the larger case (library-side sampling needs the cube buffer)
    return (int64_t)carve(nullptr, n_boxes, N, k, true, true, nullptr, nullptr);
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
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && boxes && h_score && out_records && ws, "wcr_score_topk: null pointer");
    WCR_REQUIRE(n_boxes >= 0 && N >= 1, "wcr_score_topk: bad sizes");
    WCR_REQUIRE(k >= 1 && k <= WCR_MAX_TOPK, "wcr_score_topk: k must be in [1, %d]", WCR_MAX_TOPK);
    WCR_REQUIRE(cubes || (priors && normals && z0 && h_sample),
                "wcr_score_topk: cubes == NULL needs priors, normals, z0 and h_sample");
    WCR_REQUIRE(h_score->point_stride >= 1, "wcr_score_topk: point_stride must be >= 1");
    WCR_REQUIRE(h_score->e_iou >= 0 && h_score->e_pts >= 0 && h_score->e_seg >= 0,
                "wcr_score_topk: negative exponent");
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
    carve(ws, n_boxes, N, k, cubes == nullptr, h_score->e_seg > 0, &a.ws, &zero_bytes);
    cudaError_t e = cudaMemsetAsync(ws, 0, zero_bytes, st);
    if (e != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    static int forced = -2;
    if (forced == -2) {  // development knob; the default is what DESIGN.md measures
        const char *ev = getenv("WCR_SCORE_VARIANT");
        forced = ev ? atoi(ev) : -1;
    }
    // 411: FFMA2 point test with the inside indicator on the integer pipe, 12 warps per SM
    const int variant = forced >= 0 ? forced : 411;
    const bool tiles = variant >= 100;
    if (tiles) {
        { static int sq = -1; if (sq < 0) { const char *e2 = getenv("WCR_SPLIT_QUEUES"); sq = e2 ? atoi(e2) : 1; } a.split_queues = sq; }
        {   // proposals per tile item: the whole chunk when the shard has enough tiles for every
            // resident warp (a box has ~30 tiles), finer for small shards
            const long long bc = (long long)n_boxes * nc, want = 4LL * sms_resident_warps();
            int ps = wcr::SC_CHUNK;
            while (ps > 128 && bc * 32 * (wcr::SC_CHUNK / ps) < want) ps >>= 1;
            a.ps = ps;
        }
        a.grp = 4 << ((variant / 10) % 10);   // x0y: 4, x1y: 8, x2y: 16, x3y: 32
        wcr::score_plan_tiles_kernel<<<1, 1024, 0, st>>>(images, boxes, n_boxes, nc, h_score->point_stride,
                                                         h_score->e_pts > 0, a.split_queues ? 0 : h_score->e_seg > 0, a.grp, wcr::SC_CHUNK / a.ps,
                                                         a.ws.item_prefix);
        WCR_CHECK_LAUNCH("score_plan_tiles_kernel");
    } else {
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
    }
    if (!cubes) {
        WCR_REQUIRE((N + 255) / 256 <= 65535, "wcr_score_topk: N too large");
        wcr::launch_sample_cubes(images, boxes, n_boxes, priors, normals, z0, *h_sample, N, nullptr,
                                 seed, a.ws.cubes, st);
        WCR_CHECK_LAUNCH("sample_cubes_kernel");
    }
    a.images = images; a.boxes = boxes; a.n_boxes = n_boxes; a.cubes = cubes; a.N = N;
    a.sc = *h_score; a.k = k;
    a.n_chunks = nc; a.records = out_records;
    a.out_scores = out_scores; a.out_iou = out_iou; a.out_count = out_count;
    a.out_proj = out_proj; a.out_npts = out_npts;
    a.seg_bits = seg_bits; a.seg_off = seg_off; a.out_seg = out_seg;
    const int sms = wcr::sm_count();
    if (tiles) {
        // mask term lazily: only when the point term is part of the score (a zero count then
        // forces a zero score) and the caller did not ask for the per-proposal mask ratios
        static int lazy_knob = -1;
        if (lazy_knob < 0) { const char *e3 = getenv("WCR_SEG_LAZY"); lazy_knob = e3 ? atoi(e3) : 1; }
        a.seg_lazy = lazy_knob && h_score->e_seg > 0 && h_score->e_pts > 0 && out_seg == nullptr;
        if (h_score->e_pts > 0) {
            wcr::score_zgrid_kernel<<<dim3(n_boxes, 8), 1024, 0, st>>>(images, boxes, h_score->point_stride, 1, a.ws.zgrid);
            WCR_CHECK_LAUNCH("score_zgrid_kernel");
            const dim3 grid(n_boxes, (N + 255) / 256);
            wcr::score_prep_kernel<<<grid, 256, 0, st>>>(images, boxes, n_boxes, cubes ? cubes : a.ws.cubes, N,
                                                          h_score->point_stride, 1, a.ws.params, a.ws.cullbox,
                                                          a.ws.zgrid, a.ws.mayhit);
            WCR_CHECK_LAUNCH("score_prep_kernel");
        }
        switch (variant) {
        case 100: wcr::score_tiles_kernel<8, 2, 4, false><<<sms * 2, 256, 0, st>>>(a); break;   // scalar FFMA
        case 120: wcr::score_tiles_kernel<8, 2, 16, false><<<sms * 2, 256, 0, st>>>(a); break;
        case 210: wcr::score_tiles_kernel<8, 2, 8, true><<<sms * 2, 256, 0, st>>>(a); break;    // FFMA2, 16 warps/SM
        case 221: wcr::score_tiles_kernel<4, 3, 16, true><<<sms * 3, 128, 0, st>>>(a); break;
        case 421: wcr::score_tiles_kernel<4, 3, 16, true, true><<<sms * 3, 128, 0, st>>>(a); break;
        case 410: wcr::score_tiles_kernel<8, 2, 8, true, true><<<sms * 2, 256, 0, st>>>(a); break;
        case 231: wcr::score_tiles_kernel<4, 3, 32, true><<<sms * 3, 128, 0, st>>>(a); break;   // no in-tile culling
        case 211: wcr::score_tiles_kernel<4, 3, 8, true><<<sms * 3, 128, 0, st>>>(a); break;    // FFMA.SAT indicator
        default: wcr::score_tiles_kernel<4, 3, 8, true, true><<<sms * 3, 128, 0, st>>>(a); break;   // 411
        }
        WCR_CHECK_LAUNCH("score_tiles_kernel");
        return WCR_OK;
    }
    switch (variant) {
    case 1: wcr::score_items_kernel<1, 2, 256><<<sms * 2, 256, 0, st>>>(a); break;
    default: wcr::score_items_kernel<4, 2, 256><<<sms * 2, 256, 0, st>>>(a); break;
    }
    WCR_CHECK_LAUNCH("score_items_kernel");
    return WCR_OK;
}
