// score.cu -- fused proposal scoring + per-box top-k (north_star step 3, SURVEY
// A3.2-A3.6, A3.11, A3.12).
//
// One CTA owns one (box, chunk of 256*T proposals). Proposal state lives in
// registers; the depth pixels of the GT-box ROI stream once through shared memory
// as (a_u, z) pairs, a_u = (u - cx)/fx. Because every tested point is an
// unprojected pixel p = z * (a_u, b_v, 1), the local-frame coordinate factors per
// image row:
//     q_k = R[:,k] . (p - c) = z * (R0k*a_u + (R1k*b_v + R2k)) - R[:,k].c
//         = fma(z, fma(A_k, a_u, G_k), -O_k),   G_k = fma(B_k, b_v, C_k) per row
// i.e. 6 FFMA + 3 compares per (proposal, point) instead of 3 FADD + 9 FFMA. The
// rounding order above IS the definition oracle/csrc/wcr_oracle.c restates; counts
// are bit-exact against it (DESIGN.md "Point-in-cube test").
//
// FP32 CUDA-core kernel by construction: nothing here is a dense contraction.
#include <string.h>
#include "sample.cuh"

namespace wcr {

constexpr int SC_THREADS = 256;
constexpr int SC_TILE = 2048;  // (a_u, z) pairs per shared-memory tile = 16 KB

__device__ __forceinline__ void count_inside(int &cnt, float q0, float q1, float q2, float h0,
                                             float h1, float h2)
{
    asm("{\n\t.reg .pred p;\n\t.reg .f32 a0, a1, a2;\n\t"
        "abs.f32 a0, %1;\n\tabs.f32 a1, %2;\n\tabs.f32 a2, %3;\n\t"
        "setp.le.f32 p, a0, %4;\n\t"
        "setp.le.and.f32 p, a1, %5, p;\n\t"
        "setp.le.and.f32 p, a2, %6, p;\n\t"
        "@p add.s32 %0, %0, 1;\n\t}"
        : "+r"(cnt)
        : "f"(q0), "f"(q1), "f"(q2), "f"(h0), "f"(h1), "f"(h2));
}

struct ScoreArgs {
    const WcrImage *images;
    const WcrBox *boxes;
    const float *cubes;
    int N;
    const float *priors, *normals, *z0;
    WcrSampleParams sp;
    uint64_t seed;
    const uint32_t *seg_bits;
    const int64_t *seg_off;
    WcrScoreParams sc;
    int k;
    int n_chunks;
    float *records;  // out_records if n_chunks == 1 else workspace
    float *out_scores, *out_iou;
    int32_t *out_count;
    float *out_proj;
    int32_t *out_npts;
};

// Integer ROI of a GT box: Boxes.clip((h, w)) then .long() slicing, as the
// reference slices depth patches (roi_heads.py:590-597).
__device__ __forceinline__ void box_roi(const WcrBox &bx, int H, int W, int &u0, int &v0, int &u1,
                                        int &v1)
{
    u0 = (int)fminf(fmaxf(bx.x1, 0.f), (float)W);
    u1 = (int)fminf(fmaxf(bx.x2, 0.f), (float)W);
    v0 = (int)fminf(fmaxf(bx.y1, 0.f), (float)H);
    v1 = (int)fminf(fmaxf(bx.y2, 0.f), (float)H);
}

template <int T>
__global__ void __launch_bounds__(SC_THREADS, 2) score_topk_kernel(const ScoreArgs a)
{
    constexpr int CHUNK = SC_THREADS * T;
    __shared__ __align__(16) float2 tile[SC_TILE];
    __shared__ float s_score[CHUNK];
    __shared__ float s_iou[CHUNK];
    __shared__ float s_pts[CHUNK];
    __shared__ float s_red_v[SC_THREADS / 32];
    __shared__ int s_red_i[SC_THREADS / 32];
    __shared__ int s_win[WCR_MAX_TOPK];
    __shared__ int s_npts;

    const int b = blockIdx.x, chunk = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const WcrBox bx = a.boxes[b];
    const WcrImage im = a.images[bx.image];
    const Intr K = {im.fx, im.fy, im.cx, im.cy};
    const bool fused_sampling = (a.cubes == nullptr);
    BoxCtx ctx;
    if (fused_sampling) ctx = make_box_ctx(im, bx, a.priors, a.normals, a.z0[b], a.sp);
    if (tid == 0) s_npts = 0;

    // ---- prologue: proposal state -> registers ---------------------------------
    float A[T][3], B[T][3], C[T][3], NO[T][3], Hh[T][3];
    int cnt[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const int j = chunk * CHUNK + t * SC_THREADS + tid;
        cnt[t] = 0;
        if (j < a.N) {
            float cube[15];
            if (fused_sampling) {
                sample_cube(ctx, nullptr, a.N, a.seed, (uint32_t)b, (uint32_t)j, cube);
            } else {
                const float *src = a.cubes + ((size_t)b * a.N + j) * 15;
#pragma unroll
                for (int i = 0; i < 15; ++i) cube[i] = __ldg(src + i);
            }
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                A[t][k] = cube[6 + k]; B[t][k] = cube[9 + k]; C[t][k] = cube[12 + k];
                NO[t][k] = -fmaf(C[t][k], cube[2], fmaf(B[t][k], cube[1], A[t][k] * cube[0]));
            }
            Hh[t][0] = cube[5] / 2.f; Hh[t][1] = cube[4] / 2.f; Hh[t][2] = cube[3] / 2.f;
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) { A[t][k] = B[t][k] = C[t][k] = NO[t][k] = 0.f; Hh[t][k] = -1.f; }
        }
    }

    // ---- main loop: depth ROI tiles ---------------------------------------------
    int u0, v0, u1, v1;
    box_roi(bx, im.H, im.W, u0, v0, u1, v1);
    const int s = a.sc.point_stride;
    const int ncols = (u1 - u0 + s - 1) / s, nrows = (v1 - v0 + s - 1) / s;
    const bool use_pts = (a.sc.e_pts > 0) && ncols > 0 && nrows > 0;
    int my_valid = 0;
    if (use_pts) {
        const int pieces = (ncols + SC_TILE - 1) / SC_TILE;
        const int segpad = (min(ncols, SC_TILE) + 1) & ~1;
        const int spt = SC_TILE / segpad;  // segments per tile, >= 1
        const int total_segs = nrows * pieces;
        for (int seg0 = 0; seg0 < total_segs; seg0 += spt) {
            const int nseg = min(spt, total_segs - seg0);
            __syncthreads();
            for (int idx = tid; idx < nseg * segpad; idx += SC_THREADS) {
                const int ss = idx / segpad, pos = idx - ss * segpad;
                const int seg = seg0 + ss;
                const int row = seg / pieces, piece = seg - row * pieces;
                const int col = piece * SC_TILE + pos;
                float av = 0.f, z = nanf_();
                if (col < ncols && pos < SC_TILE) {
                    const int u = u0 + col * s, v = v0 + row * s;
                    const float d = __ldg(im.depth + (size_t)v * im.depth_pitch + u);
                    av = ((float)u - K.cx) / K.fx;
                    if (d > 0.f && d <= 3.402823466e38f) { z = d; ++my_valid; }
                }
                tile[idx] = make_float2(av, z);
            }
            __syncthreads();
            for (int ss = 0; ss < nseg; ++ss) {
                const int seg = seg0 + ss;
                const int row = seg / pieces;
                const float bv = ((float)(v0 + row * s) - K.cy) / K.fy;
                float G[T][3];
#pragma unroll
                for (int t = 0; t < T; ++t)
#pragma unroll
                    for (int k = 0; k < 3; ++k) G[t][k] = fmaf(B[t][k], bv, C[t][k]);
                const float4 *src = reinterpret_cast<const float4 *>(tile + ss * segpad);
#pragma unroll 2
                for (int p = 0; p < segpad / 2; ++p) {
                    const float4 pp = src[p];
#pragma unroll
                    for (int t = 0; t < T; ++t) {
                        const float q0 = fmaf(pp.y, fmaf(A[t][0], pp.x, G[t][0]), NO[t][0]);
                        const float q1 = fmaf(pp.y, fmaf(A[t][1], pp.x, G[t][1]), NO[t][1]);
                        const float q2 = fmaf(pp.y, fmaf(A[t][2], pp.x, G[t][2]), NO[t][2]);
                        count_inside(cnt[t], q0, q1, q2, Hh[t][0], Hh[t][1], Hh[t][2]);
                    }
#pragma unroll
                    for (int t = 0; t < T; ++t) {
                        const float q0 = fmaf(pp.w, fmaf(A[t][0], pp.z, G[t][0]), NO[t][0]);
                        const float q1 = fmaf(pp.w, fmaf(A[t][1], pp.z, G[t][1]), NO[t][1]);
                        const float q2 = fmaf(pp.w, fmaf(A[t][2], pp.z, G[t][2]), NO[t][2]);
                        count_inside(cnt[t], q0, q1, q2, Hh[t][0], Hh[t][1], Hh[t][2]);
                    }
                }
            }
        }
    }
    // valid-point total P_b
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) my_valid += __shfl_xor_sync(0xffffffffu, my_valid, o);
    __syncthreads();
    if (lane == 0 && my_valid) atomicAdd(&s_npts, my_valid);
    __syncthreads();
    const int npts = s_npts;
    if (a.out_npts && chunk == 0 && tid == 0) a.out_npts[b] = npts;

    // ---- epilogue: corners -> projection -> box -> IoU -> score -----------------
    float lo0, hi0, lo1, hi1;
    clamp_bounds(bx.clamp0, lo0, hi0);
    clamp_bounds(bx.clamp1, lo1, hi1);
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const int loc = t * SC_THREADS + tid;
        const int j = chunk * CHUNK + loc;
        float score = -INFINITY, siou = 0.f, spts = 0.f;
        if (j < a.N) {
            float cube[15];
            if (fused_sampling) {
                sample_cube(ctx, nullptr, a.N, a.seed, (uint32_t)b, (uint32_t)j, cube);
            } else {
                const float *src = a.cubes + ((size_t)b * a.N + j) * 15;
#pragma unroll
                for (int i = 0; i < 15; ++i) cube[i] = __ldg(src + i);
            }
            const ProjBox pb = project_cube(cube, K, lo0, hi0, lo1, hi1, nullptr);
            siou = iou_2d(bx.x1, bx.y1, bx.x2, bx.y2, pb);
            spts = npts > 0 ? (float)cnt[t] / (float)npts : 0.f;
            score = (ipow(siou, a.sc.e_iou) * ipow(spts, a.sc.e_pts)) * 1.f;  // s_seg: seg.cuh
            const size_t o = (size_t)b * a.N + j;
            if (a.out_scores) a.out_scores[o] = score;
            if (a.out_iou) a.out_iou[o] = siou;
            if (a.out_count) a.out_count[o] = cnt[t];
            if (a.out_proj) {
                float *d = a.out_proj + o * 4;
                d[0] = pb.x1; d[1] = pb.y1; d[2] = pb.x2; d[3] = pb.y2;
            }
        }
        s_score[loc] = score; s_iou[loc] = siou; s_pts[loc] = spts;
    }
    __syncthreads();

    // ---- top-k of this chunk: k rounds of (max score, lowest index) -------------
    for (int r = 0; r < a.k; ++r) {
        float bv = -INFINITY; int bi = 0x7fffffff;
#pragma unroll
        for (int t = 0; t < T; ++t) {
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
            s_win[r] = bi;
            if (bi >= 0) s_score[bi] = -INFINITY;  // taken; its value is kept in s_red_v[0]
            s_red_v[0] = bv;
        }
        __syncthreads();
        // winner's record
        if (tid == 0) {
            const int loc = s_win[r];
            float *rec = a.records + (((size_t)b * a.n_chunks + chunk) * a.k + r) * WCR_RECORD_FLOATS;
            if (loc < 0) {
                rec[0] = __int_as_float(-1);
                for (int i = 1; i < WCR_RECORD_FLOATS; ++i) rec[i] = 0.f;
            } else {
                const int j = chunk * CHUNK + loc;
                float cube[15];
                if (fused_sampling) sample_cube(ctx, nullptr, a.N, a.seed, (uint32_t)b, (uint32_t)j, cube);
                else {
                    const float *src = a.cubes + ((size_t)b * a.N + j) * 15;
                    for (int i = 0; i < 15; ++i) cube[i] = src[i];
                }
                rec[0] = __int_as_float(j);
                rec[1] = s_red_v[0];
                for (int i = 0; i < 15; ++i) rec[2 + i] = cube[i];
                rec[17] = s_iou[loc]; rec[18] = s_pts[loc]; rec[19] = 1.f;
            }
        }
        __syncthreads();
    }
}

// Merge the per-chunk candidates of one box: one warp per box, k rounds.
__global__ void __launch_bounds__(32)
topk_merge_kernel(const float *__restrict__ cand, int n_chunks, int k, float *__restrict__ out)
{
    const int b = blockIdx.x, lane = threadIdx.x;
    const int n = n_chunks * k;
    const float *C = cand + (size_t)b * n * WCR_RECORD_FLOATS;
    // candidates are few (<= 64 chunks * 32): mark taken ones in a per-lane bitmask
    unsigned long long taken = 0ull;  // candidate lane + 32*i  <->  bit i
    for (int r = 0; r < k; ++r) {
        float bv = -INFINITY; int bj = 0x7fffffff, bc = -1;
        for (int i = 0; lane + 32 * i < n; ++i) {
            if (taken >> i & 1ull) continue;
            const int c = lane + 32 * i;
            const int j = __float_as_int(C[(size_t)c * WCR_RECORD_FLOATS]);
            if (j < 0) continue;
            const float v = C[(size_t)c * WCR_RECORD_FLOATS + 1];
            if (v > bv || (v == bv && j < bj)) { bv = v; bj = j; bc = c; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
            const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
            if (ov > bv || (ov == bv && oj < bj)) { bv = ov; bj = oj; bc = oc; }
        }
        float *dst = out + ((size_t)b * k + r) * WCR_RECORD_FLOATS;
        if (bc < 0) {
            if (lane < WCR_RECORD_FLOATS) dst[lane] = lane == 0 ? __int_as_float(-1) : 0.f;
        } else {
            if (lane < WCR_RECORD_FLOATS) dst[lane] = C[(size_t)bc * WCR_RECORD_FLOATS + lane];
            if ((bc & 31) == lane) taken |= 1ull << (bc >> 5);
        }
    }
}

}  // namespace wcr

namespace {
constexpr int kT = 4;
constexpr int kChunk = wcr::SC_THREADS * kT;
inline int n_chunks_for(int N) { return (N + kChunk - 1) / kChunk; }
}  // namespace

extern "C" int64_t wcr_score_workspace_bytes(int n_boxes, int N, int k)
{
    if (n_boxes < 0 || N < 1 || k < 1) return -1;
    const int nc = n_chunks_for(N);
    if (nc <= 1) return 16;  // never dereferenced
    return (int64_t)n_boxes * nc * k * WCR_RECORD_FLOATS * (int64_t)sizeof(float);
}

extern "C" int wcr_score_topk(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                              const float *cubes, int N, const float *priors,
                              const float *normals, const float *z0,
                              const WcrSampleParams *h_sample, uint64_t seed,
                              const uint32_t *seg_bits, const int64_t *seg_off,
                              const WcrScoreParams *h_score, int k, float *out_records,
                              float *out_scores, float *out_iou, int32_t *out_count,
                              float *out_proj, int32_t *out_npts, void *ws, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && boxes && h_score && out_records, "wcr_score_topk: null pointer");
    WCR_REQUIRE(n_boxes >= 0 && N >= 1, "wcr_score_topk: bad sizes");
    WCR_REQUIRE(k >= 1 && k <= WCR_MAX_TOPK, "wcr_score_topk: k must be in [1, %d]", WCR_MAX_TOPK);
    WCR_REQUIRE(cubes || (priors && normals && z0 && h_sample),
                "wcr_score_topk: cubes == NULL needs priors, normals, z0 and h_sample");
    WCR_REQUIRE(h_score->point_stride >= 1, "wcr_score_topk: point_stride must be >= 1");
    WCR_REQUIRE(h_score->e_iou >= 0 && h_score->e_pts >= 0 && h_score->e_seg >= 0,
                "wcr_score_topk: negative exponent");
    WCR_REQUIRE(h_score->e_seg == 0 || (seg_bits && seg_off),
                "wcr_score_topk: e_seg > 0 needs seg_bits and seg_off");
    const int nc = n_chunks_for(N);
    WCR_REQUIRE(nc <= 64, "wcr_score_topk: N too large (max %d)", 64 * kChunk);
    WCR_REQUIRE(nc == 1 || ws, "wcr_score_topk: workspace required when N > %d", kChunk);
    if (n_boxes == 0) return WCR_OK;

    wcr::ScoreArgs a;
    a.images = images; a.boxes = boxes; a.cubes = cubes; a.N = N;
    a.priors = priors; a.normals = normals; a.z0 = z0;
    if (h_sample) a.sp = *h_sample; else memset(&a.sp, 0, sizeof(a.sp));
    a.seed = seed; a.seg_bits = seg_bits; a.seg_off = seg_off; a.sc = *h_score; a.k = k;
    a.n_chunks = nc;
    a.records = nc == 1 ? out_records : (float *)ws;
    a.out_scores = out_scores; a.out_iou = out_iou; a.out_count = out_count;
    a.out_proj = out_proj; a.out_npts = out_npts;
    dim3 grid(n_boxes, nc);
    wcr::score_topk_kernel<kT><<<grid, wcr::SC_THREADS, 0, (cudaStream_t)stream>>>(a);
    WCR_CHECK_LAUNCH("score_topk_kernel");
    if (nc > 1) {
        wcr::topk_merge_kernel<<<n_boxes, 32, 0, (cudaStream_t)stream>>>((const float *)ws, nc, k,
                                                                        out_records);
        WCR_CHECK_LAUNCH("topk_merge_kernel");
    }
    return WCR_OK;
}
