// score_items_baseline.cuh -- the FIRST scoring kernel (thread = T proposals in registers, every
// ROI pixel broadcast to them from shared memory, no culling), superseded by score_tiles_kernel.
// Compiled only with -DWCR_KEEP_BASELINE_KERNEL into libwcr_b200_baseline.so (build.py), where
// wcr_score_topk runs THIS kernel: an independent implementation of the same definition that
// tests/test_score_fullsize_gpu.py cross-checks the tile kernel against at the full C2 size, and
// the measured "unculled" baseline of DESIGN.md 3.4. Not part of the product library.
#pragma once

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



// The FIRST scoring kernel: thread = T proposals in registers, every ROI pixel broadcast to them
// from shared memory, no culling. Superseded by score_tiles_kernel; kept behind
// -DWCR_KEEP_BASELINE_KERNEL (libwcr_b200_baseline.so) as the measured baseline and as an independent implementation
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
