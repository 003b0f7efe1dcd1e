/* TEST INFRASTRUCTURE ONLY (oracle). Never linked or loaded by the product path.
 *
 * Plain-C restatement of the pieces of the hot path whose bit pattern depends
 * on where a fused multiply-add sits. Compiled with -ffp-contract=off so the
 * only fused operations are the explicit fmaf() calls below.
 *
 * Each function cites the reference lines it follows (paths relative to
 * /root/reference). Where the reference is torch, the rounding order stated
 * here was established by running the reference in the authoring container
 * (torch 2.11 CPU) -- see tests/golden/make_golden.py and DESIGN.md "Oracle".
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ProposalNetwork/utils/plane.py:96-113 -- hypothesis plane through 3 sampled
 * points. torch.cross(A,B) on CPU evaluates a1*b2 - a2*b1 as
 * fma(a1, b2, -(a2*b1)); torch.norm(p=2) over 3 values is
 * sqrt(fma(z,z,fma(y,y,x*x))); k = -((c0*q0 + c1*q1) + c2*q2) with q = sample
 * point 1 (plane.py:112), products rounded separately. */
static inline float cross_c(float a1, float b2, float a2, float b1)
{
    float t = a2 * b1;
    return fmaf(a1, b2, -t);
}

void orc_plane_hypotheses(const float *pts, int64_t n, const int64_t *ids,
                          int64_t iters, float *eqs)
{
    (void)n;
    for (int64_t i = 0; i < iters; ++i) {
        const float *p0 = pts + 3 * ids[3 * i + 0];
        const float *p1 = pts + 3 * ids[3 * i + 1];
        const float *p2 = pts + 3 * ids[3 * i + 2];
        float ax = p1[0] - p0[0], ay = p1[1] - p0[1], az = p1[2] - p0[2];
        float bx = p2[0] - p0[0], by = p2[1] - p0[1], bz = p2[2] - p0[2];
        float cx = cross_c(ay, bz, az, by);
        float cy = cross_c(az, bx, ax, bz);
        float cz = cross_c(ax, by, ay, bx);
        float nrm = sqrtf(fmaf(cz, cz, fmaf(cy, cy, cx * cx)));
        cx = cx / nrm; cy = cy / nrm; cz = cz / nrm;
        float k = -((cx * p1[0] + cy * p1[1]) + cz * p1[2]);
        eqs[4 * i + 0] = cx; eqs[4 * i + 1] = cy; eqs[4 * i + 2] = cz; eqs[4 * i + 3] = k;
    }
}

/* plane.py:117-126 -- dist = (((a*x + b*y) + c*z) + d) / sqrt((a*a + b*b) + c*c),
 * inlier iff |dist| <= thresh. Elementwise torch kernels: every op rounded. */
void orc_plane_counts(const float *pts, int64_t n, const float *eqs, int64_t iters,
                      float thresh, int64_t *counts)
{
    for (int64_t i = 0; i < iters; ++i) {
        float a = eqs[4 * i], b = eqs[4 * i + 1], c = eqs[4 * i + 2], d = eqs[4 * i + 3];
        float den = sqrtf((a * a + b * b) + c * c);
        int64_t cnt = 0;
        for (int64_t j = 0; j < n; ++j) {
            float x = pts[3 * j], y = pts[3 * j + 1], z = pts[3 * j + 2];
            float dist = (((a * x + b * y) + c * z) + d) / den;
            cnt += (fabsf(dist) <= thresh);
        }
        counts[i] = cnt;
    }
}

/* ======================================================================== */
/* steps 2+3: cube assembly, projection, IoU, point-in-cube, top-k           */
/* ======================================================================== */

typedef struct { float fx, fy, cx, cy; } orc_intr;

static inline float clamp_keep_nan(float x, float lo, float hi)
{
    if (x != x) return x;
    float t = x < lo ? lo : x;
    return t > hi ? hi : t;
}

/* roi_heads.py:645-652 applied to the GT box centre (Boxes.get_centers). */
float orc_box_center_depth(const float *depth, int H, int W, int pitch, const float *box)
{
    float x = (box[0] + box[2]) / 2.f, y = (box[1] + box[3]) / 2.f;
    x = clamp_keep_nan(x, 10.f, (float)(W - 11));
    y = clamp_keep_nan(y, 10.f, (float)(H - 11));
    return depth[(size_t)(int)y * pitch + (int)x];
}

/* utils.py:111-131 vec_perp_t / rotate_vector_t / orthobasis_from_normal_t and the
 * lifting of roi_heads.py:927-928. attrs = (u v z w h l yaw). cos/sin are passed
 * in so that a caller can inject the device's values (libm and CUDA differ in the
 * last ulp); orc_assemble_cube() below uses libm. */
void orc_assemble_cube_cs(const float *K4, const float *n, const float *attrs, float c, float s,
                          float *cube)
{
    const float n0 = n[0], n1 = n[1], n2 = n[2];
    float vp[3], kxv[3], nk[3], x[3];
    if (n0 == 0.f) { vp[0] = 0.f; vp[1] = n2; vp[2] = -n1; }
    else {
        float m = sqrtf((n1 * n1 + n0 * n0) + 0.f);
        if (!(m > 1e-8f)) m = 1e-8f;
        vp[0] = n1 / m; vp[1] = (-n0) / m; vp[2] = 0.f / m;
    }
    kxv[0] = cross_c(n1, vp[2], n2, vp[1]);
    kxv[1] = cross_c(n2, vp[0], n0, vp[2]);
    kxv[2] = cross_c(n0, vp[1], n1, vp[0]);
    const float kv = (n0 * vp[0] + n1 * vp[1]) + n2 * vp[2];
    nk[0] = n0 * kv; nk[1] = n1 * kv; nk[2] = n2 * kv;
    const float omc = 1.f - c;
    for (int r = 0; r < 3; ++r) x[r] = (vp[r] * c + kxv[r] * s) + nk[r] * omc;
    const float y0 = cross_c(n1, x[2], n2, x[1]);
    const float y1 = cross_c(n2, x[0], n0, x[2]);
    const float y2 = cross_c(n0, x[1], n1, x[0]);
    const float u = attrs[0], v = attrs[1], z = attrs[2];
    cube[0] = (z * (u - K4[2])) / K4[0];
    cube[1] = (z * (v - K4[3])) / K4[1];
    cube[2] = z;
    cube[3] = attrs[3]; cube[4] = attrs[4]; cube[5] = attrs[5];
    cube[6] = x[0]; cube[7] = n0; cube[8] = y0;
    cube[9] = x[1]; cube[10] = n1; cube[11] = y1;
    cube[12] = x[2]; cube[13] = n2; cube[14] = y2;
}

void orc_assemble_cubes(const float *K4, const float *n, const float *attrs /* [7][N] */, int N,
                        float *cubes /* [N][15] */)
{
    for (int j = 0; j < N; ++j) {
        float a[7];
        for (int t = 0; t < 7; ++t) a[t] = attrs[(size_t)t * N + j];
        orc_assemble_cube_cs(K4, n, a, cosf(a[6]), sinf(a[6]), cubes + (size_t)j * 15);
    }
}

/* math_util.py:198-217 + spaces.py:224-245 + conversions.py:25-48.
 * out_box = x1 y1 x2 y2 ; corners (optional) 8 x (x, y) clamped. */
void orc_project_cube(const float *cube, const float *K4, float clamp0, float clamp1,
                      float *out_box, float *corners)
{
    static const int NX[8] = {1, 0, 0, 1, 1, 0, 0, 1};  /* vertices with -l/2 */
    static const int NY[8] = {1, 1, 0, 0, 1, 1, 0, 0};  /* vertices with -h/2 */
    const float hx = cube[5] / 2.f, hy = cube[4] / 2.f, hz = cube[3] / 2.f;
    const float *R = cube + 6;
    const float lo0 = (float)(int)(-clamp0 / 2.f + 1.f), hi0 = (float)(int)(clamp0 - 1.f + clamp0);
    const float lo1 = (float)(int)(-clamp1 / 2.f + 1.f), hi1 = (float)(int)(clamp1 - 1.f + clamp1);
    float mnx = 0, mxx = 0, mny = 0, mxy = 0;
    int nanx = 0, nany = 0;
    for (int i = 0; i < 8; ++i) {
        const float lx = (NX[i] ? -1.f : 1.f) * hx, ly = (NY[i] ? -1.f : 1.f) * hy,
                    lz = (i < 4 ? -1.f : 1.f) * hz;
        const float X = fmaf(R[2], lz, fmaf(R[1], ly, R[0] * lx)) + cube[0];
        const float Y = fmaf(R[5], lz, fmaf(R[4], ly, R[3] * lx)) + cube[1];
        const float Z = fmaf(R[8], lz, fmaf(R[7], ly, R[6] * lx)) + cube[2];
        float px = fmaf(K4[2], Z, K4[0] * X) / Z;
        float py = fmaf(K4[3], Z, K4[1] * Y) / Z;
        px = clamp_keep_nan(px, lo0, hi0);
        py = clamp_keep_nan(py, lo1, hi1);
        if (corners) { corners[2 * i] = px; corners[2 * i + 1] = py; }
        nanx |= (px != px); nany |= (py != py);
        if (i == 0) { mnx = mxx = px; mny = mxy = py; }
        else {
            mnx = fminf(mnx, px); mxx = fmaxf(mxx, px);
            mny = fminf(mny, py); mxy = fmaxf(mxy, py);
        }
    }
    out_box[0] = nanx ? NAN : mnx; out_box[2] = nanx ? NAN : mxx;
    out_box[1] = nany ? NAN : mny; out_box[3] = nany ? NAN : mxy;
}

/* detectron2 pairwise_iou as called from utils.py:181-187. */
float orc_iou_2d(const float *g, const float *p)
{
    const float a1 = (g[2] - g[0]) * (g[3] - g[1]);
    const float a2 = (p[2] - p[0]) * (p[3] - p[1]);
    float iw = fminf(g[2], p[2]) - fmaxf(g[0], p[0]);
    float ih = fminf(g[3], p[3]) - fmaxf(g[1], p[1]);
    if (p[0] != p[0] || p[1] != p[1]) return 0.f;
    if (iw < 0.f) iw = 0.f;
    if (ih < 0.f) ih = 0.f;
    const float inter = iw * ih;
    return inter > 0.f ? inter / ((a1 + a2) - inter) : 0.f;
}

/* ------------------------------------------------------------------------ */
/* Convex hull of projected corners + polygon fill + mask IoU (A3.7, A3.8)     */
/* ------------------------------------------------------------------------ */

/* ProposalNetwork/utils/utils.py:373-377. torch 0-d fp32 arithmetic: every
 * product and difference rounded separately (no fused multiply-add). */
static inline float hull_direction(const float *p1, const float *p2, const float *p3)
{
    const float a = (p2[0] - p1[0]) * (p3[1] - p1[1]);
    const float b = (p2[1] - p1[1]) * (p3[0] - p1[0]);
    return a - b;
}
static inline float hull_distance_sq(const float *p1, const float *p2)
{
    const float dx = p2[0] - p1[0], dy = p2[1] - p1[1];
    return dx * dx + dy * dy;
}

/* utils.py:410-454 jarvis_march on n 2-D points (n <= 32).
 *  - duplicates (utils.py:379-408, 413-420): index i is "duplicated" when some
 *    j > i holds the same point; the k-th such index in ascending order (the
 *    iteration order of a Python set of small ints) gets (k+1) added to x and y;
 *  - start: min x, ties -> first arg-max of y among them (:423-430);
 *  - march (:437-452): q = (l+1)%n; for i != l: d = direction(p_l, p_i, p_q);
 *    take i when d > 0, or d == 0 and i is farther from l than q;
 *  - result flipped (:454).
 * Returns the number of hull vertices m (hull[m][2], flipped order), or -1 when
 * the march does not return to its start within n steps (the reference would
 * loop forever; this project defines the mask as empty then). */
int orc_jarvis_march(const float *pts_in, int n, float *hull)
{
    float p[32][2];
    int dup[32], ndup = 0;
    if (n < 1 || n > 32) return -1;
    for (int i = 0; i < n; ++i) { p[i][0] = pts_in[2 * i]; p[i][1] = pts_in[2 * i + 1]; }
    for (int i = 0; i < n - 1; ++i) {
        int is_dup = 0;
        for (int j = i + 1; j < n; ++j)
            if (pts_in[2 * i] == pts_in[2 * j] && pts_in[2 * i + 1] == pts_in[2 * j + 1]) is_dup = 1;
        if (is_dup) dup[ndup++] = i;
    }
    for (int k = 0; k < ndup; ++k) {
        p[dup[k]][0] = p[dup[k]][0] + (float)(k + 1);
        p[dup[k]][1] = p[dup[k]][1] + (float)(k + 1);
    }
    float min_x = p[0][0];
    for (int i = 1; i < n; ++i) if (p[i][0] < min_x) min_x = p[i][0];   /* torch.min: NaN excluded by the caller */
    int index = -1;
    for (int i = 0; i < n; ++i)
        if (p[i][0] == min_x && (index < 0 || p[i][1] > p[index][1])) index = i;
    if (index < 0) return -1;
    float res[32][2];
    int m = 0;
    res[m][0] = p[index][0]; res[m][1] = p[index][1]; ++m;
    int l = index;
    for (;;) {
        int q = (l + 1) % n;
        for (int i = 0; i < n; ++i) {
            if (i == l) continue;
            const float d = hull_direction(p[l], p[i], p[q]);
            if (d > 0.f || (d == 0.f && hull_distance_sq(p[i], p[l]) > hull_distance_sq(p[q], p[l])))
                q = i;
        }
        l = q;
        if (l == index) break;
        if (m == n) return -1;
        res[m][0] = p[q][0]; res[m][1] = p[q][1]; ++m;
    }
    for (int i = 0; i < m; ++i) { hull[2 * i] = res[m - 1 - i][0]; hull[2 * i + 1] = res[m - 1 - i][1]; }
    return m;
}

/* utils.py:456-488 fill_polygon: soft mask = prod_i clamp(val_i, 0, 1),
 * val_i = (X - v1x)*e_y - (Y - v1y)*e_x over integer pixel coordinates. */
void orc_fill_polygon(const float *hull, int m, int H, int W, float *mask)
{
#pragma omp parallel for
    for (int Y = 0; Y < H; ++Y)
        for (int X = 0; X < W; ++X) {
            float prod = 1.f;
            for (int i = 0; i < m; ++i) {
                const float *v1 = hull + 2 * i, *v2 = hull + 2 * ((i + 1) % m);
                const float ex = v2[0] - v1[0], ey = v2[1] - v1[1];
                const float t1 = ((float)X - v1[0]) * ey, t2 = ((float)Y - v1[1]) * ex;
                float val = t1 - t2;
                val = val != val ? val : (val < 0.f ? 0.f : val);   /* torch.max(val, 0) keeps NaN */
                val = val != val ? val : (val > 1.f ? 1.f : val);
                prod = prod * val;
            }
            mask[(size_t)Y * W + X] = prod;
        }
}

/* Second clamp of the projected corners before the hull, roi_heads.py:950-952. */
static inline void seg_clamp_corners(const float *corners, float clamp0, float clamp1, float *out)
{
    const float hx = (float)(int)(clamp0 - 1.f), hy = (float)(int)(clamp1 - 1.f);
    for (int i = 0; i < 8; ++i) {
        out[2 * i] = clamp_keep_nan(corners[2 * i], 0.f, hx);
        out[2 * i + 1] = clamp_keep_nan(corners[2 * i + 1], 0.f, hy);
    }
}

/* Composite mask term (project spec built from utils.py:225-245,
 * 410-493 and roi_heads.py:428, 950-952):
 *   coords = second-clamped projected corners; hull = jarvis_march(coords);
 *   bube   = pixels (X, Y) of the H x W image with val_i > 0 for every hull edge,
 *            restricted to the hull's bounding box grown by one pixel,
 *            [floor(min) - 1, ceil(max) + 1] in x and y (identical to
 *            fill_polygon(...) > 0 for non-degenerate hulls);
 *   seg    = the box's mask restricted to its integer ROI [u0,u1) x [v0,v1);
 *   inter  = |seg & bube|, uni = |seg| + |bube| - inter;
 *   s_seg  = inter == 0 ? 0 : inter / uni   (mode 1: inter^5 / uni, utils.py:236-245)
 * Non-finite corners or an unclosed march give 0. */
float orc_seg_term(const float *corners, float clamp0, float clamp1, const uint8_t *seg, int H, int W,
                   int pitch, const int *roi /* u0 v0 u1 v1 */, int seg_area /* < 0: count here */,
                   int mode, int *out_inter, int *out_bube)
{
    float c2[16], hull[16];
    if (out_inter) *out_inter = 0;
    if (out_bube) *out_bube = 0;
    for (int i = 0; i < 16; ++i) if (!(fabsf(corners[i]) <= 3.402823466e38f)) return 0.f;
    seg_clamp_corners(corners, clamp0, clamp1, c2);
    const int m = orc_jarvis_march(c2, 8, hull);
    if (m < 1) return 0.f;
    float mny = hull[1], mxy = hull[1], mnx = hull[0], mxx = hull[0];
    for (int i = 1; i < m; ++i) {
        mny = fminf(mny, hull[2 * i + 1]); mxy = fmaxf(mxy, hull[2 * i + 1]);
        mnx = fminf(mnx, hull[2 * i]); mxx = fmaxf(mxx, hull[2 * i]);
    }
    int r0 = (int)floorf(mny) - 1, r1 = (int)ceilf(mxy) + 1;
    int c0 = (int)floorf(mnx) - 1, c1 = (int)ceilf(mxx) + 1;
    if (r0 < 0) r0 = 0;
    if (r1 > H - 1) r1 = H - 1;
    if (c0 < 0) c0 = 0;
    if (c1 > W - 1) c1 = W - 1;
    if (seg_area < 0) {
        seg_area = 0;
        for (int v = roi[1]; v < roi[3]; ++v)
            for (int u = roi[0]; u < roi[2]; ++u) seg_area += seg[(size_t)v * pitch + u] > 0;
    }
    int inter = 0, bube = 0;
    for (int Y = r0; Y <= r1; ++Y)
        for (int X = c0; X <= c1; ++X) {
            int in = 1;
            for (int i = 0; i < m && in; ++i) {
                const float *v1 = hull + 2 * i, *v2 = hull + 2 * ((i + 1) % m);
                const float ex = v2[0] - v1[0], ey = v2[1] - v1[1];
                const float t1 = ((float)X - v1[0]) * ey, t2 = ((float)Y - v1[1]) * ex;
                in = (t1 - t2) > 0.f;
            }
            if (!in) continue;
            ++bube;
            if (X >= roi[0] && X < roi[2] && Y >= roi[1] && Y < roi[3] && seg[(size_t)Y * pitch + X] > 0)
                ++inter;
        }
    if (out_inter) *out_inter = inter;
    if (out_bube) *out_bube = bube;
    if (inter == 0) return 0.f;
    const float fi = (float)inter, fu = (float)(seg_area + bube - inter);
    if (mode == 1) return ((((fi * fi) * fi) * fi) * fi) / fu;
    return fi / fu;
}

/* Same composite term, evaluated the way a tuned CPU port would (used by the timed
 * CPU baseline, checked equal to orc_seg_term in tests/test_oracle_hull.py):
 * val_i(X) is monotone in X under IEEE rounding, so each hull edge cuts an integer
 * half-line out of a pixel row; the row's interior is an interval found by one
 * division plus an exact fix-up, and |seg & row interval| comes from per-row prefix
 * sums of the box mask (rowpre [nrows][ncols + 1]). */
float orc_seg_term_fast(const float *corners, float clamp0, float clamp1, int H, int W,
                        const int *roi, const int *rowpre, int seg_area, int mode)
{
    float c2[16], hull[16];
    for (int i = 0; i < 16; ++i) if (!(fabsf(corners[i]) <= 3.402823466e38f)) return 0.f;
    seg_clamp_corners(corners, clamp0, clamp1, c2);
    const int m = orc_jarvis_march(c2, 8, hull);
    if (m < 1) return 0.f;
    float mny = hull[1], mxy = hull[1], mnx = hull[0], mxx = hull[0];
    for (int i = 1; i < m; ++i) {
        mny = fminf(mny, hull[2 * i + 1]); mxy = fmaxf(mxy, hull[2 * i + 1]);
        mnx = fminf(mnx, hull[2 * i]); mxx = fmaxf(mxx, hull[2 * i]);
    }
    int r0 = (int)floorf(mny) - 1, r1 = (int)ceilf(mxy) + 1;
    int c0 = (int)floorf(mnx) - 1, c1 = (int)ceilf(mxx) + 1;
    if (r0 < 0) r0 = 0;
    if (r1 > H - 1) r1 = H - 1;
    if (c0 < 0) c0 = 0;
    if (c1 > W - 1) c1 = W - 1;
    const int ncols = roi[2] - roi[0], nrows = roi[3] - roi[1];
    int inter = 0, bube = 0;
    for (int Yi = r0; Yi <= r1; ++Yi) {
        const float Y = (float)Yi;
        int lo = c0, hi = c1;
        for (int i = 0; i < m && lo <= hi; ++i) {
            const float *v1 = hull + 2 * i, *v2 = hull + 2 * ((i + 1) % m);
            const float ex = v2[0] - v1[0], ey = v2[1] - v1[1];
            const float t2 = (Y - v1[1]) * ex;
            if (ey == 0.f) {
                if (!(((float)lo - v1[0]) * ey > t2)) hi = -1;
            } else if (ey > 0.f) {
                float xs = floorf(v1[0] + t2 / ey) + 1.f;
                xs = xs != xs ? (float)lo : xs;
                int X = (int)fminf(fmaxf(xs, (float)lo), (float)(hi + 1));
                while (X > lo && ((float)(X - 1) - v1[0]) * ey > t2) --X;
                while (X <= hi && !(((float)X - v1[0]) * ey > t2)) ++X;
                lo = X;
            } else {
                float xs = ceilf(v1[0] + t2 / ey) - 1.f;
                xs = xs != xs ? (float)hi : xs;
                int X = (int)fminf(fmaxf(xs, (float)(lo - 1)), (float)hi);
                while (X < hi && ((float)(X + 1) - v1[0]) * ey > t2) ++X;
                while (X >= lo && !(((float)X - v1[0]) * ey > t2)) --X;
                hi = X;
            }
        }
        if (lo > hi) continue;
        bube += hi - lo + 1;
        const int row = Yi - roi[1];
        if (row >= 0 && row < nrows) {
            const int a = (lo > roi[0] ? lo : roi[0]) - roi[0];
            const int b = (hi < roi[2] - 1 ? hi : roi[2] - 1) - roi[0];
            if (a <= b) inter += rowpre[(size_t)row * (ncols + 1) + b + 1] - rowpre[(size_t)row * (ncols + 1) + a];
        }
    }
    if (inter == 0) return 0.f;
    const float fi = (float)inter, fu = (float)(seg_area + bube - inter);
    if (mode == 1) return ((((fi * fi) * fi) * fi) * fi) / fu;
    return fi / fu;
}

static inline float ipow_(float x, int e) { float r = 1.f; for (int i = 0; i < e; ++i) r = r * x; return r; }

/* SURVEY A3.11/A3.12 for ONE box: N cubes against the depth pixels of the box's
 * integer ROI (Boxes.clip + .long(), roi_heads.py:590-597), sub-sampled by
 * `stride`. Point p = z*(a_u, b_v, 1), a_u = (u-cx)/fx, b_v = (v-cy)/fy. The test
 * "|R[:,k].(p-c)| <= half_k for k = x,y,z" with half = (l/2, h/2, w/2)
 * (math_util.py:198-207 axis mapping) is evaluated in unit-cube form:
 *   A'_k = R0k/half_k, B'_k = R1k/half_k, C'_k = R2k/half_k      (IEEE division)
 *   O'_k = fma(C'_k, cz, fma(B'_k, cy, A'_k*cx))
 *   q'_k = fma(z, fma(A'_k, a_u, fma(B'_k, b_v, C'_k)), -O'_k)
 *   inside iff max(|q'_0|, |q'_1|, |q'_2|) <= 1      (max ignores NaN operands)
 * A cube with a non-positive dimension contains no point. Returns P_b. */
int orc_score_box_seg(const float *depth, int H, int W, int pitch, const float *K4, const float *gbox,
                      float clamp0, float clamp1, const float *cubes, int N, int stride, int e_iou,
                      int e_pts, int e_seg, int seg_mode, const uint8_t *seg /* H x W or NULL */,
                      int seg_pitch, float *out_score, float *out_iou, int *out_count,
                      float *out_proj, float *out_seg)
{
    float t;
    t = gbox[0] < 0.f ? 0.f : gbox[0]; const int u0 = (int)(t > W ? (float)W : t);
    t = gbox[2] < 0.f ? 0.f : gbox[2]; const int u1 = (int)(t > W ? (float)W : t);
    t = gbox[1] < 0.f ? 0.f : gbox[1]; const int v0 = (int)(t > H ? (float)H : t);
    t = gbox[3] < 0.f ? 0.f : gbox[3]; const int v1 = (int)(t > H ? (float)H : t);
    int npts = 0;
    /* the mask term always uses the stride-1 integer ROI */
    const int roi[4] = {u0, v0, u1, v1};
    int seg_area = 0;
    /* seg_mode bit 8: use orc_seg_term_fast (timed CPU baseline) instead of the
     * per-pixel definition */
    const int fast = (seg_mode >> 8) & 1;
    seg_mode &= 0xff;
    int *rowpre = NULL;
    if (e_seg > 0 && seg) {
        for (int v = v0; v < v1; ++v)
            for (int u = u0; u < u1; ++u) seg_area += seg[(size_t)v * seg_pitch + u] > 0;
        if (fast && u1 > u0 && v1 > v0) {
            rowpre = (int *)malloc(sizeof(int) * (size_t)(v1 - v0) * (u1 - u0 + 1));
            for (int v = v0; v < v1; ++v) {
                int *rp = rowpre + (size_t)(v - v0) * (u1 - u0 + 1);
                rp[0] = 0;
                for (int u = u0; u < u1; ++u) rp[u - u0 + 1] = rp[u - u0] + (seg[(size_t)v * seg_pitch + u] > 0);
            }
        }
    }
    if (e_pts > 0)
        for (int v = v0; v < v1; v += stride)
            for (int u = u0; u < u1; u += stride) {
                const float d = depth[(size_t)v * pitch + u];
                npts += (d > 0.f && d <= 3.402823466e38f);
            }
#pragma omp parallel for schedule(dynamic, 8)
    for (int j = 0; j < N; ++j) {
        const float *cube = cubes + (size_t)j * 15;
        float A[3], B[3], C[3], NO[3];
        const float half[3] = {cube[5] / 2.f, cube[4] / 2.f, cube[3] / 2.f};
        const int live = cube[3] > 0.f && cube[4] > 0.f && cube[5] > 0.f;
        for (int k = 0; k < 3; ++k) {
            A[k] = cube[6 + k] / half[k]; B[k] = cube[9 + k] / half[k]; C[k] = cube[12 + k] / half[k];
            NO[k] = -fmaf(C[k], cube[2], fmaf(B[k], cube[1], A[k] * cube[0]));
        }
        int cnt = 0;
        if (e_pts > 0 && live)
            for (int v = v0; v < v1; v += stride) {
                const float bv = ((float)v - K4[3]) / K4[1];
                const float G0 = fmaf(B[0], bv, C[0]), G1 = fmaf(B[1], bv, C[1]),
                            G2 = fmaf(B[2], bv, C[2]);
                for (int u = u0; u < u1; u += stride) {
                    const float d = depth[(size_t)v * pitch + u];
                    if (!(d > 0.f && d <= 3.402823466e38f)) continue;
                    const float au = ((float)u - K4[2]) / K4[0];
                    const float q0 = fmaf(d, fmaf(A[0], au, G0), NO[0]);
                    const float q1 = fmaf(d, fmaf(A[1], au, G1), NO[1]);
                    const float q2 = fmaf(d, fmaf(A[2], au, G2), NO[2]);
                    cnt += fmaxf(fmaxf(fabsf(q0), fabsf(q1)), fabsf(q2)) <= 1.f;
                }
            }
        float pb[4], corners[16];
        orc_project_cube(cube, K4, clamp0, clamp1, pb, corners);
        const float siou = orc_iou_2d(gbox, pb);
        const float spts = npts > 0 ? (float)cnt / (float)npts : 0.f;
        float sseg = 1.f;
        if (e_seg > 0 && seg)
            sseg = rowpre ? orc_seg_term_fast(corners, clamp0, clamp1, H, W, roi, rowpre, seg_area, seg_mode)
                 : (fast ? 0.f
                         : orc_seg_term(corners, clamp0, clamp1, seg, H, W, seg_pitch, roi, seg_area,
                                        seg_mode, NULL, NULL));
        out_score[j] = (ipow_(siou, e_iou) * ipow_(spts, e_pts)) * ipow_(sseg, e_seg);
        if (out_seg) out_seg[j] = sseg;
        if (out_iou) out_iou[j] = siou;
        if (out_count) out_count[j] = cnt;
        if (out_proj) { out_proj[4 * j] = pb[0]; out_proj[4 * j + 1] = pb[1]; out_proj[4 * j + 2] = pb[2]; out_proj[4 * j + 3] = pb[3]; }
    }
    free(rowpre);
    return npts;
}

int orc_score_box(const float *depth, int H, int W, int pitch, const float *K4, const float *gbox,
                  float clamp0, float clamp1, const float *cubes, int N, int stride, int e_iou,
                  int e_pts, float *out_score, float *out_iou, int *out_count, float *out_proj)
{
    return orc_score_box_seg(depth, H, W, pitch, K4, gbox, clamp0, clamp1, cubes, N, stride, e_iou,
                             e_pts, 0, 0, NULL, 0, out_score, out_iou, out_count, out_proj, NULL);
}

/* Philox4x32-10 (Salmon et al., SC'11), for the known-answer test of the
 * in-kernel generator. */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4])
{
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int i = 0; i < 10; ++i) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU arm asks for
 * every host core explicitly. */
void orc_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int orc_num_threads(void)
{
    int n = 1;
#ifdef _OPENMP
#pragma omp parallel
    {
#pragma omp single
        n = omp_get_num_threads();
    }
#endif
    return n;
}
