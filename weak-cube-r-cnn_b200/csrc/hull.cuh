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
    int c = __popc(__ldg(w + wa) & ma) + __popc(__ldg(w + wb) & mb);
    for (int i = wa + 1; i < wb; ++i) c += __popc(__ldg(w + i));
    return c;
}

// Composite mask term of one cube (project spec, see orc_seg_term): pixels (X, Y)
// of the H x W image with val_i > 0 for every hull edge, restricted to the hull's
// bounding box grown by one pixel. val_i(X) is monotone in X under IEEE
// rounding, so every edge cuts an integer half-line out of the row: the row's
// interior is an interval [lo, hi] found by one division + an exact fix-up
// instead of a per-pixel test. inter = |seg & bube|, uni = |seg| + |bube| - inter.
template <class A>
__device__ float seg_term(A hx, A hy, int m, int H, int W, const SegBox &sb, int mode)
{
    if (m < 1) return 0.f;
    float mny = hy(0), mxy = mny, mnx = hx(0), mxx = mnx;
    for (int i = 1; i < m; ++i) {
        const float y = hy(i), x = hx(i);
        mny = fminf(mny, y); mxy = fmaxf(mxy, y);
        mnx = fminf(mnx, x); mxx = fmaxf(mxx, x);
    }
    const int r0 = max((int)floorf(mny) - 1, 0), r1 = min((int)ceilf(mxy) + 1, H - 1);
    const int c0 = max((int)floorf(mnx) - 1, 0), c1 = min((int)ceilf(mxx) + 1, W - 1);
    int inter = 0, bube = 0;
    for (int Yi = r0; Yi <= r1; ++Yi) {
        const float Y = (float)Yi;
        int lo = c0, hi = c1;
        for (int i = 0; i < m && lo <= hi; ++i) {
            const int i2 = (i + 1 == m) ? 0 : i + 1;
            const float v1x = hx(i), v1y = hy(i);
            const float ex = hx(i2) - v1x, ey = hy(i2) - v1y;
            const float t2 = (Y - v1y) * ex;
            // inside this edge  <=>  g(X) = fl(fl(X - v1x) * ey) > t2   (a - b > 0 <=> a > b)
            if (ey == 0.f) {
                if (!(((float)lo - v1x) * ey > t2)) hi = -1;  // +-0 > t2, same for every X
            } else if (ey > 0.f) {  // g non-decreasing: smallest X with g(X) > t2
                float xs = floorf(v1x + t2 / ey) + 1.f;
                int X = (int)fminf(fmaxf(xs, (float)lo), (float)(hi + 1));
                while (X > lo && ((float)(X - 1) - v1x) * ey > t2) --X;
                while (X <= hi && !(((float)X - v1x) * ey > t2)) ++X;
                lo = X;
            } else {  // g non-increasing: largest X with g(X) > t2
                float xs = ceilf(v1x + t2 / ey) - 1.f;
                int X = (int)fminf(fmaxf(xs, (float)(lo - 1)), (float)hi);
                while (X < hi && ((float)(X + 1) - v1x) * ey > t2) ++X;
                while (X >= lo && !(((float)X - v1x) * ey > t2)) --X;
                hi = X;
            }
        }
        if (lo > hi) continue;
        bube += hi - lo + 1;
        const int row = Yi - sb.v0;
        if (row >= 0 && row < sb.nrows) {
            const int a = max(lo, sb.u0) - sb.u0, b = min(hi, sb.u0 + sb.ncols - 1) - sb.u0;
            if (a <= b) inter += seg_popc_range(sb, row, a, b);
        }
    }
    if (inter == 0) return 0.f;
    const float fi = (float)inter, fu = (float)(sb.area + bube - inter);
    if (mode == 1) return ((((fi * fi) * fi) * fi) * fi) / fu;  // utils.py:245
    return fi / fu;
}

}  // namespace wcr
