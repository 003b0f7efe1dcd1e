// hull.cuh -- convex hull of projected cube corners, polygon fill and mask IoU
// (SURVEY A3.7 / A3.8), shared by the fused scoring epilogue (score.cu) and the
// block-level drop-ins (blocks.cu).
//
// Reference: ProposalNetwork/utils/utils.py:373-408 (direction, distance_sq,
// findDuplicates), :410-454 (jarvis_march), :456-488 (fill_polygon), :225-245
// (mask_iou, mod_mask_iou); call site roi_heads.py:428, 950-952.
// Mirrored operation by operation in oracle/csrc/wcr_oracle.c (orc_jarvis_march,
// orc_fill_polygon, orc_seg_term). Built with -fmad=false: every product and
// difference below is rounded separately, like the reference's 0-d torch ops.
#pragma once
#include "common.cuh"

namespace wcr {

// Strided scratch accessor: element i of a per-thread array interleaved across
// the CTA (bank-conflict free) or a plain array (stride 1).
struct Strided {
    float *p;
    int stride;
    __device__ __forceinline__ float &operator()(int i) const { return p[i * stride]; }
};

__device__ __forceinline__ float hull_direction(float p1x, float p1y, float p2x, float p2y,
                                                float p3x, float p3y)
{
    const float a = (p2x - p1x) * (p3y - p1y);
    const float b = (p2y - p1y) * (p3x - p1x);
    return a - b;
}
__device__ __forceinline__ float hull_dist_sq(float p1x, float p1y, float p2x, float p2y)
{
    const float dx = p2x - p1x, dy = p2y - p1y;
    return dx * dx + dy * dy;
}

// jarvis_march (utils.py:410-454) on n <= 32 points held in (px, py), which are
// modified in place by the duplicate bump (:413-420). The hull lands in (hx, hy)
// in the reference's flipped order. Returns the vertex count m, or -1 when the
// march does not return to its start within n steps (the reference would never
// terminate; defined as "empty mask" here and in the oracle).
template <class A>
__device__ int jarvis_march(A px, A py, int n, A hx, A hy)
{
    unsigned dup = 0u;
    for (int i = 0; i < n - 1; ++i) {
        const float xi = px(i), yi = py(i);
        for (int j = i + 1; j < n; ++j)
            if (xi == px(j) && yi == py(j)) dup |= 1u << i;
    }
    int k = 0;
    for (int i = 0; i < n; ++i)
        if (dup >> i & 1u) {
            ++k;  // k-th duplicated index in ascending order gets +k on x and y
            px(i) = px(i) + (float)k;
            py(i) = py(i) + (float)k;
        }
    float min_x = px(0);
    for (int i = 1; i < n; ++i) min_x = fminf(min_x, px(i));
    int index = -1;
    float best_y = 0.f;
    for (int i = 0; i < n; ++i) {
        const float y = py(i);
        if (px(i) == min_x && (index < 0 || y > best_y)) { index = i; best_y = y; }
    }
    if (index < 0) return -1;
    int m = 0;
    hx(m) = px(index); hy(m) = py(index); ++m;
    int l = index;
    for (;;) {
        int q = (l + 1) % n;
        const float lx = px(l), ly = py(l);
        float qx = px(q), qy = py(q);
        for (int i = 0; i < n; ++i) {
            if (i == l) continue;
            const float ix = px(i), iy = py(i);
            const float d = hull_direction(lx, ly, ix, iy, qx, qy);
            if (d > 0.f || (d == 0.f && hull_dist_sq(ix, iy, lx, ly) > hull_dist_sq(qx, qy, lx, ly))) {
                q = i; qx = ix; qy = iy;
            }
        }
        l = q;
        if (l == index) break;
        if (m == n) return -1;
        hx(m) = qx; hy(m) = qy; ++m;
    }
    for (int i = 0; i < m / 2; ++i) {  // torch.flip(result, [0])
        const float tx = hx(i), ty = hy(i);
        hx(i) = hx(m - 1 - i); hy(i) = hy(m - 1 - i);
        hx(m - 1 - i) = tx; hy(m - 1 - i) = ty;
    }
    return m;
}

// fill_polygon value of one pixel (utils.py:469-486): prod_i clamp(val_i, 0, 1).
template <class A>
__device__ __forceinline__ float fill_value(A hx, A hy, int m, float X, float Y)
{
    float prod = 1.f;
    for (int i = 0; i < m; ++i) {
        const int i2 = (i + 1 == m) ? 0 : i + 1;
        const float v1x = hx(i), v1y = hy(i);
        const float ex = hx(i2) - v1x, ey = hy(i2) - v1y;
        const float t1 = (X - v1x) * ey, t2 = (Y - v1y) * ex;
        float val = t1 - t2;
        if (val == val) val = fminf(fmaxf(val, 0.f), 1.f);  // torch.max/min keep NaN
        prod = prod * val;
    }
    return prod;
}

// Bit-packed mask of one GT box over its integer ROI [u0, u0+ncols) x [v0, v0+nrows):
// bit (c & 31) of word r*wpr + (c >> 5); padding bits are zero.
struct SegBox {
    const uint32_t *bits;
    int u0, v0, ncols, nrows, wpr;
    int area;  // popcount of the whole mask
};

__device__ __forceinline__ int seg_popc_range(const SegBox &sb, int row, int a, int b)
{
    // columns a..b (inclusive, ROI-relative, 0 <= a <= b < ncols) of ROI row `row`
    const uint32_t *w = sb.bits + (size_t)row * sb.wpr;
    const int wa = a >> 5, wb = b >> 5;
    const uint32_t ma = 0xffffffffu << (a & 31);
    const uint32_t mb = 0xffffffffu >> (31 - (b & 31));
    if (wa == wb) return __popc(__ldg(w + wa) & ma & mb);
    // all loads issued before the first popcount (a word-by-word loop pays one L1 latency
    // per word): the two end words + up to 6 interior words cover intervals of 256 pixels
    const uint32_t first = __ldg(w + wa), last = __ldg(w + wb);
    uint32_t mid[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) mid[i] = (wa + 1 + i < wb) ? __ldg(w + wa + 1 + i) : 0u;
    int c = __popc(first & ma) + __popc(last & mb);
#pragma unroll
    for (int i = 0; i < 6; ++i) c += __popc(mid[i]);
    for (int i = wa + 7; i < wb; ++i) c += __popc(__ldg(w + i));
    return c;
}

// Composite mask term of one cube (project spec, see orc_seg_term): pixels (X, Y)
// of the H x W image with val_i > 0 for every hull edge, restricted to the hull's
// bounding box grown by one pixel. val_i(X) is monotone in X under IEEE rounding,
// so every edge cuts an integer half-line out of a pixel row: the row's interior is
// an interval [lo, hi] found by one (approximate) division + an exact fix-up instead
// of a per-pixel test. inter = |seg & bube|, uni = |seg| + |bube| - inter.
//
// Smallest X in [c0, c1 + 1] with g(X) > t2 (ey > 0, g non-decreasing), resp. largest
// X in [c0 - 1, c1] (ey < 0, g non-increasing), g(X) = fl(fl(X - v1x) * ey).
// Slow, always-right version (walks from the candidate).
static __device__ __forceinline__ int edge_cut_walk(float v1x, float ey, float t2, int c0, int c1, int X)
{
    if (ey > 0.f) {
        X = min(max(X, c0), c1 + 1);
        while (X > c0 && ((float)(X - 1) - v1x) * ey > t2) --X;
        while (X <= c1 && !(((float)X - v1x) * ey > t2)) ++X;
        return X;
    }
    X = min(max(X, c0 - 1), c1);
    while (X < c1 && ((float)(X + 1) - v1x) * ey > t2) ++X;
    while (X >= c0 && !(((float)X - v1x) * ey > t2)) --X;
    return X;
}

// Edge vectors of a closed hull, e_i = v_{i+1} - v_i (utils.py:465-467), written once
// per hull so that the per-row loop reads them instead of recomputing.
template <class A>
__device__ __forceinline__ void hull_edges(A hx, A hy, int m, A ex, A ey)
{
    for (int i = 0; i < m; ++i) {
        const int i2 = (i + 1 == m) ? 0 : i + 1;
        ex(i) = hx(i2) - hx(i);
        ey(i) = hy(i2) - hy(i);
    }
}

// One hull edge held in registers for the row loop.
struct EdgeRegs { float v1x, v1y, ex, ey, rcp; };

// Exact row interval, one edge at a time (always right, slow): every edge cuts an
// integer half-line out of the row; walk from an approximate cut to the exact one.
static __device__ __noinline__ void row_interval_walk(const float *hx, const float *hy, const float *ex,
                                               const float *ey, int stride, int m, float Y, int c0,
                                               int c1, int *lo_out, int *hi_out)
{
    int lo = c0, hi = c1;
    for (int i = 0; i < m && lo <= hi; ++i) {
        const float v1x = hx[i * stride], eyi = ey[i * stride];
        const float t2 = (Y - hy[i * stride]) * ex[i * stride];
        if (eyi == 0.f) {
            if (!(((float)c0 - v1x) * eyi > t2)) hi = -1;
        } else {
            const float xc = v1x + __fdividef(t2, eyi);
            const float g = fminf(fmaxf(floorf(xc), (float)(c0 - 1)), (float)(c1 + 1));  // NaN -> c0 - 1
            const int X = edge_cut_walk(v1x, eyi, t2, c0, c1, (int)g);
            if (eyi > 0.f) lo = max(lo, X); else hi = min(hi, X);
        }
    }
    *lo_out = lo; *hi_out = hi;
}

struct HullBox { int r0, r1, c0, c1; };
template <class A>
__device__ __forceinline__ HullBox hull_bbox(A hx, A hy, int m, int H, int W)
{
    float mny = hy(0), mxy = mny, mnx = hx(0), mxx = mnx;
    for (int i = 1; i < m; ++i) {
        const float y = hy(i), x = hx(i);
        mny = fminf(mny, y); mxy = fmaxf(mxy, y);
        mnx = fminf(mnx, x); mxx = fmaxf(mxx, x);
    }
    HullBox b;
    b.r0 = max((int)floorf(mny) - 1, 0); b.r1 = min((int)ceilf(mxy) + 1, H - 1);
    b.c0 = max((int)floorf(mnx) - 1, 0); b.c1 = min((int)ceilf(mxx) + 1, W - 1);
    return b;
}

__device__ __forceinline__ float seg_ratio(int inter, int bube, int area, int mode)
{
    if (inter == 0) return 0.f;
    const float fi = (float)inter, fu = (float)(area + bube - inter);
    if (mode == 1) return ((((fi * fi) * fi) * fi) * fi) / fu;  // utils.py:245
    return fi / fu;
}

// Upper bound of seg_ratio(inter, bube, area, 0) from inter <= iub and bube >= max(blow, inter):
// i / (area + h - i) grows with i and falls with h, so the bound is its value at i = min(iub, area),
// h = max(blow, i). Integers below 2^24 and one correctly rounded division on both sides: the
// float ratio is <= the float bound.
__device__ __forceinline__ float seg_ratio_bound(int iub, int blow, int area)
{
    const int i = min(iub, area);
    if (i <= 0) return 0.f;
    const int h = max(blow, i);
    return (float)i / (float)(area + h - i);
}

// Rows r0 + lane, r0 + 32 + lane, ... of one hull with M edges (compile-time, so the
// edge loops are straight-line code with no per-edge branches).
//
// Per row, the reference's test for pixel (X, Y) and edge i is
//     g_i(X) = fl(fl(X - v1x) * ey) > t2_i,   t2_i = fl(fl(Y - v1y) * ex)      (a - b > 0 <=> a > b)
// g_i is a monotone step in X (IEEE rounding is monotone): 0..0 1..1 for a rising edge
// (ey > 0), 1..1 0..0 for a falling one, constant for a horizontal one. The row's
// interior is therefore [L, R], L = first column where ALL rising edges pass, R = last
// column where ALL falling edges pass. L and R are guessed from approximate cuts
// (one FFMA per edge) and then PROVED with the exact expression: all rising pass at L
// and not all at L - 1, all falling pass at R and not all at R + 1 (two exact
// evaluations per edge). A row whose guess cannot be proved (cut within rounding of a
// pixel centre, ill-conditioned edge) is redone by row_interval_walk.
template <int M, class A>
__device__ __forceinline__ void hull_rows(const EdgeRegs (&e)[8], A hx, A hy, A ex, A ey, int m_real, int r0,
                                          int r1, int c0, int c1, const SegBox &sb, int &it, int &bu)
{
    const int lane = threadIdx.x & 31;
    const float fc0 = (float)c0, fc1 = (float)c1;
    for (int Y0 = r0; Y0 <= r1; Y0 += 32) {
        const int Yi = Y0 + lane;
        if (Yi <= r1) {
            const float Y = (float)Yi;
            // ---- guess: approximate cuts
            float t2[M];
            float lo_c = -INFINITY, hi_c = INFINITY, dacc = -1.f;
#pragma unroll
            for (int i = 0; i < M; ++i) {
                t2[i] = (Y - e[i].v1y) * e[i].ex;
                const float xc = __fmaf_rn(t2[i], e[i].rcp, e[i].v1x);
                lo_c = fmaxf(lo_c, e[i].ey > 0.f ? xc : -INFINITY);
                hi_c = fminf(hi_c, e[i].ey < 0.f ? xc : INFINITY);
                dacc = fmaxf(dacc, e[i].ey == 0.f ? t2[i] : -1.f);  // horizontal: passes iff (+-0) > t2
            }
            const float fL = fminf(fmaxf(floorf(lo_c) + 1.f, fc0), fc1 + 1.f);   // in [c0, c1 + 1]
            const float fR = fmaxf(fminf(ceilf(hi_c) - 1.f, fc1), fc0 - 1.f);    // in [c0 - 1, c1]
            // ---- proof: exact evaluations at L, L - 1 (rising edges) and R, R + 1 (falling)
            unsigned pL = 1u, pLm = 1u, pR = 1u, pRp = 1u;
#pragma unroll
            for (int i = 0; i < M; ++i) {
                const unsigned up = e[i].ey > 0.f, dn = e[i].ey < 0.f;
                const float xa = up ? fL : fR, xb = up ? fL - 1.f : fR + 1.f;
                const unsigned ga = (xa - e[i].v1x) * e[i].ey > t2[i];
                const unsigned gb = (xb - e[i].v1x) * e[i].ey > t2[i];
                pL &= ga | (up ^ 1u); pLm &= gb | (up ^ 1u);
                pR &= ga | (dn ^ 1u); pRp &= gb | (dn ^ 1u);
            }
            const bool left_ok = (fL == fc0 || !pLm) && (fL > fc1 || pL);
            const bool right_ok = (fR == fc1 || !pRp) && (fR < fc0 || pR);
            int lo = (int)fL, hi = (int)fR;
            if (!(left_ok && right_ok))
                row_interval_walk(&hx(0), &hy(0), &ex(0), &ey(0), hx.stride, m_real, Y, c0, c1, &lo, &hi);
            else if (!(dacc < 0.f))
                hi = -1;
            if (lo <= hi) {
                bu += hi - lo + 1;
                const int row = Yi - sb.v0;
                if (row >= 0 && row < sb.nrows) {
                    const int a = max(lo, sb.u0) - sb.u0, b = min(hi, sb.u0 + sb.ncols - 1) - sb.u0;
                    if (a <= b) it += seg_popc_range(sb, row, a, b);
                }
            }
        }
    }
}

// Warp-cooperative pixel counts of ONE hull (1 <= m <= 8, hull and m uniform across
// the warp); every lane returns the totals. The hull's edges are read once (broadcast
// loads) into registers.
template <class A>
__device__ __forceinline__ void seg_counts_warp(A hx, A hy, A ex, A ey, int m, int H, int W,
                                                const SegBox &sb, int &inter, int &bube)
{
    const int lane = threadIdx.x & 31;
    EdgeRegs e[8];
    float mny = 0.f, mxy = 0.f, mnx = 0.f, mxx = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        if (i < m) {
            e[i].v1x = hx(i); e[i].v1y = hy(i); e[i].ex = ex(i); e[i].ey = ey(i);
            e[i].rcp = __fdividef(1.f, e[i].ey);   // approximate is enough: the cut is only a guess, proved below
            if (i == 0) { mnx = mxx = e[i].v1x; mny = mxy = e[i].v1y; }
            else {
                mnx = fminf(mnx, e[i].v1x); mxx = fmaxf(mxx, e[i].v1x);
                mny = fminf(mny, e[i].v1y); mxy = fmaxf(mxy, e[i].v1y);
            }
        } else {
            // padding edge: horizontal (ey = 0) with t2 = (Y + 1e30) * -1 < 0 on every row, i.e. an
            // edge every pixel passes. It takes no part in the cuts (ey == 0) nor in the proof, so
            // hulls with 3..5 vertices run the 6-edge row kernel and 7 the 8-edge one: two row kernels
            // instead of six in the instruction caches (one 8-edge kernel for all was slower).
            e[i].v1x = 0.f; e[i].v1y = -1.0e30f; e[i].ex = -1.f; e[i].ey = 0.f; e[i].rcp = 0.f;
        }
    }
    const int r0 = max((int)floorf(mny) - 1, 0), r1 = min((int)ceilf(mxy) + 1, H - 1);
    const int c0 = max((int)floorf(mnx) - 1, 0), c1 = min((int)ceilf(mxx) + 1, W - 1);
    int it = 0, bu = 0;
    switch (m) {
    case 3: case 4: case 5: case 6: hull_rows<6>(e, hx, hy, ex, ey, m, r0, r1, c0, c1, sb, it, bu); break;
    case 7: case 8: hull_rows<8>(e, hx, hy, ex, ey, m, r0, r1, c0, c1, sb, it, bu); break;
    default:  // m = 1, 2: no interior in exact arithmetic; evaluated the slow, literal way
        for (int Yi = r0 + lane; Yi <= r1; Yi += 32) {
            int lo, hi;
            row_interval_walk(&hx(0), &hy(0), &ex(0), &ey(0), hx.stride, m, (float)Yi, c0, c1, &lo, &hi);
            if (lo <= hi) {
                bu += hi - lo + 1;
                const int row = Yi - sb.v0;
                if (row >= 0 && row < sb.nrows) {
                    const int a = max(lo, sb.u0) - sb.u0, b = min(hi, sb.u0 + sb.ncols - 1) - sb.u0;
                    if (a <= b) it += seg_popc_range(sb, row, a, b);
                }
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        it += __shfl_xor_sync(0xffffffffu, it, o);
        bu += __shfl_xor_sync(0xffffffffu, bu, o);
    }
    inter = it; bube = bu;
}

}  // namespace wcr
