// cube_math.cuh -- per-cube arithmetic shared by the sampler (sample.cu), the fused
// scoring kernel (score.cu) and the weak-loss kernels (loss.cu).
//
// Every function states the reference lines it restates and fixes ONE rounding
// order, mirrored operation by operation in oracle/csrc/wcr_oracle.c. The library
// is compiled with -fmad=false: the only fused multiply-adds are the fmaf() calls
// written here.
#pragma once
#include "common.cuh"

namespace wcr {

struct Intr { float fx, fy, cx, cy; };

// torch.cross on CPU: a1*b2 - a2*b1 evaluated as fma(a1, b2, -(a2*b1)).
__device__ __forceinline__ float cross_term(float a1, float b2, float a2, float b1)
{
    const float t = a2 * b1;
    return fmaf(a1, b2, -t);
}

// Yaw-independent part of orthobasis_from_normal_t (utils.py:111-131):
//   vp  = vec_perp_t(n)            (:111-117; NOT normalised when n.x == 0)
//   kxv = cross(n, vp)             (rotate_vector_t term2, :143)
//   nk  = n * (n . vp)             (term3 without (1-cos), :144)
struct YawBasis {
    float n[3], vp[3], kxv[3], nk[3];
};

__device__ __forceinline__ YawBasis make_yaw_basis(float n0, float n1, float n2)
{
    YawBasis b;
    b.n[0] = n0; b.n[1] = n1; b.n[2] = n2;
    if (n0 == 0.f) {
        b.vp[0] = 0.f; b.vp[1] = n2; b.vp[2] = -n1;
    } else {
        // normalize_vector (utils.py:9-15): v / max(sqrt(sum v^2), 1e-8), v = [b, -a, 0]
        float m = sqrtf((n1 * n1 + n0 * n0) + 0.f);
        m = fmaxf(m, 1e-8f);
        b.vp[0] = n1 / m; b.vp[1] = (-n0) / m; b.vp[2] = 0.f / m;
    }
    b.kxv[0] = cross_term(n1, b.vp[2], n2, b.vp[1]);
    b.kxv[1] = cross_term(n2, b.vp[0], n0, b.vp[2]);
    b.kxv[2] = cross_term(n0, b.vp[1], n1, b.vp[0]);
    const float kv = (n0 * b.vp[0] + n1 * b.vp[1]) + n2 * b.vp[2];
    b.nk[0] = n0 * kv; b.nk[1] = n1 * kv; b.nk[2] = n2 * kv;
    return b;
}

// orthobasis_from_normal_t (utils.py:119-131): columns [x, n, y], det = -1.
// R is row-major: R[3*r + c].
__device__ __forceinline__ void yaw_to_R(const YawBasis &b, float yaw, float R[9])
{
    const float c = cosf(yaw), s = sinf(yaw);
    const float omc = 1.f - c;
    float x[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) x[r] = (b.vp[r] * c + b.kxv[r] * s) + b.nk[r] * omc;
    const float y0 = cross_term(b.n[1], x[2], b.n[2], x[1]);
    const float y1 = cross_term(b.n[2], x[0], b.n[0], x[2]);
    const float y2 = cross_term(b.n[0], x[1], b.n[1], x[0]);
    R[0] = x[0]; R[1] = b.n[0]; R[2] = y0;
    R[3] = x[1]; R[4] = b.n[1]; R[5] = y1;
    R[6] = x[2]; R[7] = b.n[2]; R[8] = y2;
}

// Lift a sampled (u, v, z) through K (roi_heads.py:927-928 order) and assemble the
// 15-float cube [c(3), w, h, l, R(9)] (spaces.py:130-150).
__device__ __forceinline__ void assemble_cube(const Intr &K, const YawBasis &yb, float u, float v,
                                              float z, float w, float h, float l, float yaw,
                                              float cube[15])
{
    cube[0] = (z * (u - K.cx)) / K.fx;
    cube[1] = (z * (v - K.cy)) / K.fy;
    cube[2] = z;
    cube[3] = w; cube[4] = h; cube[5] = l;
    yaw_to_R(yb, yaw, cube + 6);
}

// torch.clamp(x, lo, hi): NaN stays NaN.
__device__ __forceinline__ float clamp_keep_nan(float x, float lo, float hi)
{
    return (x != x) ? x : fminf(fmaxf(x, lo), hi);
}

struct ProjBox { float x1, y1, x2, y2; };

// Corner i of get_cuboid_verts_faces (math_util.py:198-217): local x = -+l/2 on
// vertices {0,3,4,7} / {1,2,5,6}; local y = -+h/2 on {0,1,4,5} / {2,3,6,7};
// local z = -+w/2 on {0,1,2,3} / {4,5,6,7}.
__device__ __forceinline__ void corner_signs(int i, float &sx, float &sy, float &sz)
{
    sx = (i == 0 || i == 3 || i == 4 || i == 7) ? -1.f : 1.f;
    sy = (i == 0 || i == 1 || i == 4 || i == 5) ? -1.f : 1.f;
    sz = (i < 4) ? -1.f : 1.f;
}

// 8 corners -> K projection -> clamp -> min/max (math_util.py:198-217,
// spaces.py:224-245, conversions.py:25-48). lo/hi are the integer clamp bounds
// int(-c/2+1), int(c-1+c) (spaces.py:241-242). NaN coordinates propagate to the
// box like torch.min/max do. Optionally returns the clamped corners.
__device__ __forceinline__ ProjBox project_cube(const float cube[15], const Intr &K, float lo0,
                                                float hi0, float lo1, float hi1,
                                                float *corners_xy /* 16 or nullptr */)
{
    const float hx = cube[5] / 2.f, hy = cube[4] / 2.f, hz = cube[3] / 2.f;
    const float *R = cube + 6;
    float mnx = 0.f, mxx = 0.f, mny = 0.f, mxy = 0.f;
    bool nanx = false, nany = false;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float sx, sy, sz;
        corner_signs(i, sx, sy, sz);
        const float lx = sx * hx, ly = sy * hy, lz = sz * hz;
        const float X = fmaf(R[2], lz, fmaf(R[1], ly, R[0] * lx)) + cube[0];
        const float Y = fmaf(R[5], lz, fmaf(R[4], ly, R[3] * lx)) + cube[1];
        const float Z = fmaf(R[8], lz, fmaf(R[7], ly, R[6] * lx)) + cube[2];
        float px = fmaf(K.cx, Z, K.fx * X) / Z;
        float py = fmaf(K.cy, Z, K.fy * Y) / Z;
        px = clamp_keep_nan(px, lo0, hi0);
        py = clamp_keep_nan(py, lo1, hi1);
        if (corners_xy) { corners_xy[2 * i] = px; corners_xy[2 * i + 1] = py; }
        nanx |= (px != px); nany |= (py != py);
        if (i == 0) { mnx = mxx = px; mny = mxy = py; }
        else {
            mnx = fminf(mnx, px); mxx = fmaxf(mxx, px);
            mny = fminf(mny, py); mxy = fmaxf(mxy, py);
        }
    }
    ProjBox b;
    const float q = nanf_();
    b.x1 = nanx ? q : mnx; b.x2 = nanx ? q : mxx;
    b.y1 = nany ? q : mny; b.y2 = nany ? q : mxy;
    return b;
}

__device__ __forceinline__ void clamp_bounds(float c, float &lo, float &hi)
{
    lo = (float)(int)(-c / 2.f + 1.f);   // int(-clamp/2+1), spaces.py:241
    hi = (float)(int)(c - 1.f + c);      // int(clamp-1+clamp)
}

// detectron2 pairwise_iou (utils.py:181-187): inter>0 ? inter/(a1+a2-inter) : 0.
__device__ __forceinline__ float iou_2d(float gx1, float gy1, float gx2, float gy2, const ProjBox &p)
{
    const float a1 = (gx2 - gx1) * (gy2 - gy1);
    const float a2 = (p.x2 - p.x1) * (p.y2 - p.y1);
    float iw = fminf(gx2, p.x2) - fmaxf(gx1, p.x1);
    float ih = fminf(gy2, p.y2) - fmaxf(gy1, p.y1);
    if (p.x1 != p.x1 || p.y1 != p.y1) return 0.f;  // NaN box: inter is NaN, (NaN > 0) false
    iw = fmaxf(iw, 0.f); ih = fmaxf(ih, 0.f);
    const float inter = iw * ih;
    return inter > 0.f ? inter / ((a1 + a2) - inter) : 0.f;
}

// custom_mapping (utils.py:207-223) of one score term, in double like the reference's Python
// floats: 0 for x <= 0, 1 / (1 + (x / (1 - x)) ** -beta) else (x == 1 -> 1). beta <= 0: identity.
__device__ __forceinline__ float shape_term(float x, float beta)
{
    if (!(beta > 0.f)) return x;
    if (!(x > 0.f)) return x != x ? x : 0.f;
    const double v = (double)x;
    return (float)(1.0 / (1.0 + pow(v / (1.0 - v), -(double)beta)));
}

__device__ __forceinline__ float ipow(float x, int e)
{
    float r = 1.f;
    for (int i = 0; i < e; ++i) r = r * x;
    return r;
}

}  // namespace wcr
