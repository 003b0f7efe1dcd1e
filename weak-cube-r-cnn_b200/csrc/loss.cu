// loss.cu -- fused weak-loss forward / backward (north_star step 4, SURVEY A4.1-A4.9,
// A4.11). Reference: cubercnn/modeling/roi_heads/roi_heads.py:786-1144 and the
// helpers cited in include/wcr.h.
//
// The reference issues O(n) tiny kernels from Python loops (one cubes_to_box call per
// box, an O(n_i^2) CPU pair loop for the pose term); here:
//   forward : box kernel (decode -> cube -> 8 corners -> projection -> clamp -> min/max
//             -> GIoU, pseudo-depth L1, ground-normal cosine, dims hinge)
//             + pair kernel (one CTA per image) + one reduction CTA
//   backward: pair kernel + box kernel with the hand-derived chain rule, following
//             autograd's sub-gradient choices (first arg-extremum over corners,
//             closed-interval clamp, tie-splitting binary min/max, sign(0) = 0).
// Latency-bound by nature (n <= a few thousand boxes): no roofline claim.
#include <math.h>
#include <string.h>
#include "cube_math.cuh"
#include "hull.cuh"

namespace wcr {

// decoded state of one box (roi_heads.py:786-905, 927-931)
struct Decoded {
    float cube_x, cube_y, sw, sh;
    float zf;          // z after z_type, before virtual_to_real
    float z;           // cube_z
    float dims[3];     // w h l
    float M[9];        // allocentric -> egocentric rotation (constant in backward)
    bool  validM;      // angle > 0
    float R[9];        // egocentric pose
    float cube[15];
    Intr K;
};

// pytorch3d axis_angle_to_matrix (quaternion route), math_util.py:827
__device__ __forceinline__ void axis_angle_to_matrix(float ax, float ay, float az, float M[9])
{
    const float angles = sqrtf((ax * ax + ay * ay) + az * az);
    const float half = angles * 0.5f;
    const float s = (fabsf(angles) < 1e-6f) ? 0.5f - (angles * angles) / 48.f : sinf(half) / angles;
    const float r = cosf(half), i = ax * s, j = ay * s, k = az * s;
    const float two_s = 2.0f / (((r * r + i * i) + j * j) + k * k);
    M[0] = 1.f - two_s * (j * j + k * k); M[1] = two_s * (i * j - k * r); M[2] = two_s * (i * k + j * r);
    M[3] = two_s * (i * j + k * r); M[4] = 1.f - two_s * (i * i + k * k); M[5] = two_s * (j * k - i * r);
    M[6] = two_s * (i * k - j * r); M[7] = two_s * (j * k + i * r); M[8] = 1.f - two_s * (i * i + j * j);
}

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

__device__ __forceinline__ Decoded decode_box(const WcrLossIO &io, const WcrLossParams &P, int i)
{
    Decoded d;
    const float *sb = io.src_boxes + 4 * i;
    const float *k4 = io.K4 + 4 * i;
    d.K.fx = k4[0]; d.K.fy = k4[1]; d.K.cx = k4[2]; d.K.cy = k4[3];
    d.sw = sb[2] - sb[0]; d.sh = sb[3] - sb[1];
    const float scx = sb[0] + 0.5f * d.sw, scy = sb[1] + 0.5f * d.sh;
    d.cube_x = scx + d.sw * io.deltas[2 * i];        // roi_heads.py:843
    d.cube_y = scy + d.sh * io.deltas[2 * i + 1];    // :844
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float raw = io.dims_raw[3 * i + k], pm = io.prior_mean[3 * i + k];
        if (P.dims_func == 1) {                      // scaled_sigmoid, :856-859
            const float ps = io.prior_std[3 * i + k];
            const float dmin = fmaxf(pm - 3.f * ps, 0.f), dmax = pm + 3.f * ps;
            d.dims[k] = dmin + (dmax - dmin) * sigmoidf_(raw);
        } else {
            d.dims[k] = expf(fminf(raw, 5.f)) * pm;  // :861
        }
    }
    // R_from_allocentric (math_util.py:802-830)
    {
        float ox = (d.cube_x - d.K.cx) / d.K.fx, oy = (d.cube_y - d.K.cy) / d.K.fy, oz = 1.f;
        const float nrm = sqrtf((ox * ox + oy * oy) + oz * oz);
        ox /= nrm; oy /= nrm; oz /= nrm;
        const float angle = acosf(oz);
        const float axx = 0.f - oy, axy = 0.f + ox;
        const float norms = sqrtf((axx * axx + axy * axy) + 0.f);
        d.validM = angle > 0.f;
        const float *Rv = io.pose + 9 * i;
        if (d.validM) {
            axis_angle_to_matrix(angle * axx / norms, angle * axy / norms, angle * 0.f / norms, d.M);
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    d.R[3 * r + c] = fmaf(d.M[3 * r + 2], Rv[6 + c],
                                          fmaf(d.M[3 * r + 1], Rv[3 + c], d.M[3 * r] * Rv[c]));
        } else {
#pragma unroll
            for (int t = 0; t < 9; ++t) { d.R[t] = Rv[t]; d.M[t] = (t % 4 == 0) ? 1.f : 0.f; }
        }
    }
    const float zr = io.z_raw[i];
    d.zf = P.z_type == 1 ? sigmoidf_(zr) * 100.f : (P.z_type == 2 ? expf(zr) : zr);  // :874-880
    d.z = d.zf * io.v2r[i];                                                          // :904-905
    d.cube[0] = d.z * (d.cube_x - d.K.cx) / d.K.fx;   // :927
    d.cube[1] = d.z * (d.cube_y - d.K.cy) / d.K.fy;   // :928
    d.cube[2] = d.z;
    d.cube[3] = d.dims[0]; d.cube[4] = d.dims[1]; d.cube[5] = d.dims[2];
#pragma unroll
    for (int t = 0; t < 9; ++t) d.cube[6 + t] = d.R[t];
    return d;
}

// torchvision generalized_box_iou_loss(boxes1 = gt, boxes2 = proj); optionally the
// gradient w.r.t. proj (x1g y1g x2g y2g).
__device__ __forceinline__ float giou_loss(const float *g, const float *p, float *dp)
{
    const float eps = 1e-7f;
    const float x1 = g[0], y1 = g[1], x2 = g[2], y2 = g[3];
    const float x1g = p[0], y1g = p[1], x2g = p[2], y2g = p[3];
    const float xk1 = fmaxf(x1, x1g), yk1 = fmaxf(y1, y1g), xk2 = fminf(x2, x2g), yk2 = fminf(y2, y2g);
    const bool nanb = (x1g != x1g) || (y1g != y1g) || (x2g != x2g) || (y2g != y2g);
    const bool mask = !nanb && (yk2 > yk1) && (xk2 > xk1);
    const float I = mask ? (xk2 - xk1) * (yk2 - yk1) : 0.f;
    const float A2 = (x2g - x1g) * (y2g - y1g);
    const float U = ((x2 - x1) * (y2 - y1) + A2) - I;
    const float iou = I / (U + eps);
    const float xc1 = fminf(x1, x1g), yc1 = fminf(y1, y1g), xc2 = fmaxf(x2, x2g), yc2 = fmaxf(y2, y2g);
    const float C = (xc2 - xc1) * (yc2 - yc1);
    float loss = 1.f - (iou - ((C - U) / (C + eps)));
    if (nanb) loss = nanf_();
    if (dp) {
        const float Ue = U + eps, Ce = C + eps;
        const float gI = -(1.f / Ue + I / (Ue * Ue)) + 1.f / Ce;
        const float gA2 = I / (Ue * Ue) - 1.f / Ce;
        const float gC = Ue / (Ce * Ce);
        // binary max/min backward: the larger (smaller) operand takes it, ties split in half
        auto w_max = [](float other, float self) { return self > other ? 1.f : (self == other ? 0.5f : 0.f); };
        auto w_min = [](float other, float self) { return self < other ? 1.f : (self == other ? 0.5f : 0.f); };
        const float iw = xk2 - xk1, ih = yk2 - yk1;
        const float gI_ = mask ? gI : 0.f;
        dp[0] = gI_ * (-ih) * w_max(x1, x1g) + gA2 * (-(y2g - y1g)) + gC * (-(yc2 - yc1)) * w_min(x1, x1g);
        dp[1] = gI_ * (-iw) * w_max(y1, y1g) + gA2 * (-(x2g - x1g)) + gC * (-(xc2 - xc1)) * w_min(y1, y1g);
        dp[2] = gI_ * (ih) * w_min(x2, x2g) + gA2 * (y2g - y1g) + gC * (yc2 - yc1) * w_max(x2, x2g);
        dp[3] = gI_ * (iw) * w_min(y2, y2g) + gA2 * (x2g - x1g) + gC * (xc2 - xc1) * w_max(y2, y2g);
    }
    return loss;
}

// depth[(long) clamp(y,10,h-11)][(long) clamp(x,10,w-11)]  (roi_heads.py:645-652)
__device__ __forceinline__ float depth_at(const WcrImage &im, float x, float y)
{
    x = clamp_keep_nan(x, 10.f, (float)(im.W - 11));
    y = clamp_keep_nan(y, 10.f, (float)(im.H - 11));
    return im.depth[(size_t)(int)y * im.depth_pitch + (int)x];
}

__device__ __forceinline__ int image_of(const int32_t *off, int n_images, int i)
{
    int lo = 0, hi = n_images;  // off[lo] <= i < off[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (off[mid] <= i) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ float sgn(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }

// F.cosine_similarity(n, p): x / max(||x||, eps) each, then dot.
__device__ __forceinline__ float cos_sim3(const float *a, const float *b, float *an, float *bn,
                                          float *bnorm)
{
    const float na = fmaxf(sqrtf((a[0] * a[0] + a[1] * a[1]) + a[2] * a[2]), 1e-8f);
    const float nb = fmaxf(sqrtf((b[0] * b[0] + b[1] * b[1]) + b[2] * b[2]), 1e-8f);
#pragma unroll
    for (int k = 0; k < 3; ++k) { an[k] = a[k] / na; bn[k] = b[k] / nb; }
    *bnorm = nb;
    return (an[0] * bn[0] + an[1] * bn[1]) + an[2] * bn[2];
}

constexpr int LT = WCR_LOSS_TERMS - 1;  // per-box term rows (uncert is not one)

// ---------------------------------------------------------------- forward --
struct ZSteps { float v[WCR_Z_LOSS_STEPS]; };

// depth extent of the 8 cube corners (get_all_corners()[..., 2], roi_heads.py:665-667) with
// the FIRST arg-max / arg-min corner (torch.max / torch.min over a dim)
__device__ __forceinline__ float corner_depth_range(const Decoded &d, int &cmax, int &cmin)
{
    const float hx = d.dims[2] / 2.f, hy = d.dims[1] / 2.f, hz = d.dims[0] / 2.f;
    float zmax = 0.f, zmin = 0.f;
    cmax = 0; cmin = 0;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        float sx, sy, sz;
        corner_signs(c, sx, sy, sz);
        const float Z = fmaf(d.R[8], sz * hz, fmaf(d.R[7], sy * hy, d.R[6] * (sx * hx))) + d.cube[2];
        if (c == 0) { zmax = Z; zmin = Z; }
        else {
            if (Z > zmax || (Z != Z && zmax == zmax)) { zmax = Z; cmax = c; }
            if (Z < zmin || (Z != Z && zmin == zmin)) { zmin = Z; cmin = c; }
        }
    }
    return zmax - zmin;
}

// 1 - cos of the relative angle between the predicted pose and the pseudo-GT rotation built
// from the ground normal (roi_heads.py:1003-1007); also reports pytorch3d's trace check
__device__ __forceinline__ float ground2_term(const float *R, const float *G, float conf, bool *bad)
{
    const float d0 = fmaf(R[2], G[2], fmaf(R[1], G[1], R[0] * G[0]));
    const float d1 = fmaf(R[5], G[5], fmaf(R[4], G[4], R[3] * G[3]));
    const float d2 = fmaf(R[8], G[8], fmaf(R[7], G[7], R[6] * G[6]));
    const float tr = (d0 + d1) + d2;
    if (bad) *bad = (tr < -1.0f - 1e-4f) || (tr > 3.0f + 1e-4f);
    return (1.f - (tr - 1.0f) * 0.5f) * conf;
}

__global__ void __launch_bounds__(128)
loss_box_fwd_kernel(const WcrLossIO io, const WcrLossParams P, const ZSteps zs)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= io.n) return;
    const Decoded d = decode_box(io, P, i);
    float lo0, hi0, lo1, hi1;
    clamp_bounds(io.clamp[2 * i], lo0, hi0);
    clamp_bounds(io.clamp[2 * i + 1], lo1, hi1);
    const ProjBox pb = project_cube(d.cube, d.K, lo0, hi0, lo1, hi1, nullptr);
    const float proj[4] = {pb.x1, pb.y1, pb.x2, pb.y2};
#pragma unroll
    for (int t = 0; t < 4; ++t) io.proj[4 * i + t] = proj[t];
#pragma unroll
    for (int t = 0; t < 9; ++t) io.R_ego[9 * i + t] = d.R[t];
    if (io.z_dec) io.z_dec[i] = d.z;
    const float q = nanf_();
    float term[LT];
#pragma unroll
    for (int t = 0; t < LT; ++t) term[t] = q;
    if (P.loss_flags & WCR_LOSS_IOU) term[WCR_T_IOU - 1] = giou_loss(io.gt_boxes + 4 * i, proj, nullptr);
    const int img = image_of(io.img_off, io.n_images, i);
    if (P.loss_flags & WCR_LOSS_POSE_GROUND) {
        float an[3], bn[3], nb;
        const float c = cos_sim3(io.normals + 3 * img, d.R + 3, an, bn, &nb);   // row 1 of R (:992)
        term[WCR_T_NORMAL - 1] = (1.f - fabsf(c)) * io.ground_conf[img];
    }
    if (P.loss_flags & WCR_LOSS_POSE_GROUND2) {      // overrides 'pose_ground' (:997-1007)
        bool bad;
        term[WCR_T_NORMAL - 1] = ground2_term(d.R, io.gt_rot + 9 * img, io.ground_conf[img], &bad);
        if (bad) io.stats[WCR_STAT_TRACE_ERR] = 1.f;
    }
    if ((P.loss_flags & WCR_LOSS_Z_CENTER) && !(P.loss_flags & WCR_LOSS_Z_PATCH))
        term[WCR_T_PSEUDO_Z - 1] = fabsf(d.z - depth_at(io.images[img], d.cube_x, d.cube_y));
    if (P.loss_flags & WCR_LOSS_Z_SEARCH) {           // z_loss, roi_heads.py:530-573
        const float *g = io.gt_boxes + 4 * i;
        const float gt_area = (g[2] - g[0]) * (g[3] - g[1]);
        const float pcx = (proj[0] + proj[2]) / 2.f, pcy = (proj[1] + proj[3]) / 2.f;
        const float pred_area = (proj[2] - proj[0]) * (proj[3] - proj[1]);
        // :545-546 compares the BOOL (lower <= centre) with the upper bound
        const float bx = (g[0] - 50.f <= pcx) ? 1.f : 0.f, by = (g[1] - 50.f <= pcy) ? 1.f : 0.f;
        const bool within = (bx <= g[2] + 50.f) && (by <= g[3] + 50.f);
        float score = 5.0f;                           // 0.1 * max_count
        if (within) {
            const bool smaller = gt_area < pred_area;
            float cube[15];
#pragma unroll
            for (int t = 0; t < 15; ++t) cube[t] = d.cube[t];
            float best = 0.f, best_z = d.z;
            for (int k = 0; k < WCR_Z_LOSS_STEPS; ++k) {
                const float zk = smaller ? d.z + zs.v[k] : d.z - zs.v[k];
                cube[2] = zk;
                const ProjBox mb = project_cube(cube, d.K, lo0, hi0, lo1, hi1, nullptr);
                float area = (mb.x2 - mb.x1) * (mb.y2 - mb.y1);
                area = area + (area == 0.f ? 10000000.f : 0.f);
                const float diff = fabsf(gt_area - area);
                // torch.argmin: first minimum, a NaN counts as the minimum
                if (k == 0 || (best == best && (diff != diff || diff < best))) { best = diff; best_z = zk; }
            }
            score = fabsf(d.z - best_z);
        }
        term[WCR_T_Z - 1] = score / 2.f;
    }
    if (P.loss_flags & WCR_LOSS_DEPTH) {              // depth_range_loss, :658-684
        int cmax, cmin;
        term[WCR_T_DEPTH - 1] = fabsf(io.depth_gt[i] - corner_depth_range(d, cmax, cmin));
    }
    if (P.loss_flags & WCR_LOSS_DIMS) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float s = fabsf(d.dims[k] - io.prior_mean[3 * i + k]) / io.prior_std[3 * i + k];
            term[WCR_T_DIMS_W - 1 + k] = fmaxf(s - 1.0f, 0.f);
            if (s != s) term[WCR_T_DIMS_W - 1 + k] = s;  // torch.max propagates NaN
        }
    }
#pragma unroll
    for (int t = 0; t < LT; ++t) io.terms[(size_t)t * io.n + i] = term[t];
}

__global__ void box_z_loss_kernel(const int32_t *__restrict__ img_off, int n_images,
                                  const float *__restrict__ gt, const int *__restrict__ inside,
                                  const float *__restrict__ pred_z, float *__restrict__ out_loss,
                                  float *__restrict__ out_sign);

// block-wide sum of (a, b) over 256 threads, result valid in every thread
__device__ __forceinline__ void block_sum2(float &a, float &b, float *s_a, float *s_b)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    __syncthreads();
    if (lane == 0) { s_a[wid] = a; s_b[wid] = b; }
    __syncthreads();
    a = 0.f; b = 0.f;
    for (int w = 0; w < nw; ++w) { a += s_a[w]; b += s_b[w]; }
}

// cos of the relative angle of two poses: (tr(Ri Rj^T) - 1) / 2  (math_util.py:1179, 1222-1226)
__device__ __forceinline__ float pair_cos(const float *Ri, const float *Rj)
{
    const float d0 = fmaf(Ri[2], Rj[2], fmaf(Ri[1], Rj[1], Ri[0] * Rj[0]));
    const float d1 = fmaf(Ri[5], Rj[5], fmaf(Ri[4], Rj[4], Ri[3] * Rj[3]));
    const float d2 = fmaf(Ri[8], Rj[8], fmaf(Ri[7], Rj[7], Ri[6] * Rj[6]));
    return (((d0 + d1) + d2) - 1.0f) * 0.5f;
}

// One CTA per image: mean over pairs (i > j) of 1 - |cos|  (roi_heads.py:453-460).
__global__ void __launch_bounds__(256) loss_pose_fwd_kernel(const WcrLossIO io)
{
    __shared__ float s_a[8], s_b[8];
    const int im = blockIdx.x;
    const int b0 = io.img_off[im], ni = io.img_off[im + 1] - b0;
    float sum = 0.f, dummy = 0.f;
    for (int i = 1 + threadIdx.x; i < ni; i += blockDim.x) {
        float Ri[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) Ri[t] = io.R_ego[9 * (size_t)(b0 + i) + t];
        for (int j = 0; j < i; ++j) sum += 1.f - fabsf(pair_cos(Ri, io.R_ego + 9 * (size_t)(b0 + j)));
    }
    block_sum2(sum, dummy, s_a, s_b);
    if (threadIdx.x == 0) {
        const float npairs = 0.5f * (float)ni * (float)(ni - 1);
        io.img_pose[im] = ni == 0 ? nanf_() : (ni == 1 ? 0.f : sum / npairs);
    }
}

__device__ __forceinline__ float term_weight(const WcrLossParams &P, int t)
{
    switch (t) {
    case WCR_T_UNCERT: return P.use_confidence;
    case WCR_T_IOU: return P.w_iou * P.w_3d;
    case WCR_T_POSE: return P.w_pose * P.w_3d;
    case WCR_T_NORMAL: return P.w_normal * P.w_3d;
    case WCR_T_PSEUDO_Z: return P.w_z * P.w_3d;
    case WCR_T_Z: return P.w_z * P.w_3d;              // roi_heads.py:1125
    case WCR_T_DEPTH: return P.w_depth * P.w_3d;      // :1143
    case WCR_T_SEG: return P.w_seg * P.w_3d;          // :1121
    default: return P.w_dims * P.w_3d;
    }
}

__device__ __forceinline__ bool term_enabled(const WcrLossParams &P, int t)
{
    switch (t) {
    case WCR_T_UNCERT: return P.use_confidence > 0.f;
    case WCR_T_IOU: return P.loss_flags & WCR_LOSS_IOU;
    case WCR_T_POSE: return P.loss_flags & WCR_LOSS_POSE_ALIGN;
    case WCR_T_NORMAL: return P.loss_flags & (WCR_LOSS_POSE_GROUND | WCR_LOSS_POSE_GROUND2);
    case WCR_T_PSEUDO_Z: return P.loss_flags & (WCR_LOSS_Z_CENTER | WCR_LOSS_Z_PATCH);
    case WCR_T_Z: return P.loss_flags & WCR_LOSS_Z_SEARCH;
    case WCR_T_DEPTH: return P.loss_flags & WCR_LOSS_DEPTH;
    case WCR_T_SEG: return P.loss_flags & WCR_LOSS_SEG;
    default: return P.loss_flags & WCR_LOSS_DIMS;
    }
}

__device__ __forceinline__ bool finitef_(float x) { return fabsf(x) <= 3.402823466e38f; }

// uncertainty scale (roi_heads.py:1077): sqrt(2) * exp(-u), or 1 when confidence is off
__device__ __forceinline__ float uncert_sf(const WcrLossIO &io, const WcrLossParams &P, int i)
{
    return P.use_confidence > 0.f ? 1.41421356f * expf(-io.uncert[i]) : 1.f;
}

// One CTA: pose scalar, safe means (roi_heads.py:1307-1315), weights.
__global__ void __launch_bounds__(256)
loss_reduce_kernel(const WcrLossIO io, const WcrLossParams P)
{
    __shared__ float s_a[8], s_b[8];
    __shared__ float s_pose;
    const int tid = threadIdx.x;
    // pose scalar: sum over images with != 1 boxes, / (fail + 1)   (:451-464)
    float psum = 0.f, fail = 0.f;
    for (int im = tid; im < io.n_images; im += blockDim.x) {
        const int ni = io.img_off[im + 1] - io.img_off[im];
        if (ni == 1) fail += 1.f; else psum += io.img_pose[im];
    }
    block_sum2(psum, fail, s_a, s_b);
    const bool pose_present = (P.loss_flags & WCR_LOSS_POSE_ALIGN) && ((int)fail != io.n_images);
    if (tid == 0) s_pose = psum * 1.f / (fail + 1.f);
    __syncthreads();
    const float pose = s_pose;
    if (P.loss_flags & WCR_LOSS_POSE_ALIGN)
        for (int i = tid; i < io.n; i += blockDim.x) io.terms[(size_t)(WCR_T_POSE - 1) * io.n + i] = pose;
    // dim_loss returns None when any prior std is NaN (:620-623)
    float nan_std = 0.f, dummy = 0.f;
    if (P.loss_flags & WCR_LOSS_DIMS)
        for (int i = tid; i < 3 * io.n; i += blockDim.x) nan_std += (io.prior_std[i] != io.prior_std[i]) ? 1.f : 0.f;
    block_sum2(nan_std, dummy, s_a, s_b);
    __syncthreads();
    for (int t = 0; t < WCR_LOSS_TERMS; ++t) {
        float sum = 0.f, cnt = 0.f;
        bool present = term_enabled(P, t);
        if (t == WCR_T_POSE) present = pose_present;
        if (t >= WCR_T_DIMS_W && t <= WCR_T_DIMS_L) present = present && nan_std == 0.f;
        if (present) {
            for (int i = tid; i < io.n; i += blockDim.x) {
                const float v = t == WCR_T_UNCERT ? io.uncert[i]
                                                  : io.terms[(size_t)(t - 1) * io.n + i] * uncert_sf(io, P, i);
                if (finitef_(v)) { sum += v; cnt += 1.f; }
            }
        }
        block_sum2(sum, cnt, s_a, s_b);
        if (tid == 0) {
            io.present[t] = present ? 1 : 0;
            io.stats[t] = cnt;
            io.losses[t] = !present ? 0.f : (cnt > 0.f ? (sum / cnt) * term_weight(P, t) : nanf_());
        }
        __syncthreads();
    }
    if (tid == 0) {
        io.stats[WCR_STAT_POSE] = pose; io.stats[WCR_STAT_FAIL] = fail;
        io.present[WCR_LOSS_TERMS] = io.stats[WCR_STAT_TRACE_ERR] != 0.f;
        io.present[WCR_LOSS_TERMS + 1] = io.stats[WCR_STAT_SHAPE_ERR] != 0.f;
    }
}


// ---- 'pose_ground2': normal_to_rotation (roi_heads.py:686-696) -------------------------------
// The reference repeats the per-image normal once per box and normalises t0 and t1 by the
// FROBENIUS norm of the whole (n, 3) matrix (torch.norm without dim), so the rows are not unit
// length when n > 1 -- reproduced. `.any() < 0.001` is evaluated on a bool: the y-axis
// branch is taken only when every t0 is exactly zero. One CTA; images are few.
__global__ void __launch_bounds__(256)
ground_rot_kernel(const float *__restrict__ normals, const int32_t *__restrict__ img_off, int n_images,
                  float *__restrict__ gt_rot)
{
    __shared__ float s_a[8], s_b[8];
    const int tid = threadIdx.x;
    float ss = 0.f, any = 0.f;
    for (int im = tid; im < n_images; im += blockDim.x) {
        const float ni = (float)(img_off[im + 1] - img_off[im]);
        const float ny = normals[3 * im + 1], nz = normals[3 * im + 2];
        const float dot = (0.f * 0.f + nz * nz) + ny * ny;     // t0 = n x (1,0,0) = (0, nz, -ny)
        ss += ni * dot;
        if (ni > 0.f && dot != 0.f) any = 1.f;
    }
    block_sum2(ss, any, s_a, s_b);
    const bool ybranch = !(any > 0.f);
    if (ybranch) {
        ss = 0.f;
        float dummy = 0.f;
        for (int im = tid; im < n_images; im += blockDim.x) {
            const float ni = (float)(img_off[im + 1] - img_off[im]);
            const float nx = normals[3 * im], nz = normals[3 * im + 2];
            ss += ni * ((nz * nz + 0.f) + nx * nx);            // t0 = n x (0,1,0) = (-nz, 0, nx)
        }
        block_sum2(ss, dummy, s_a, s_b);
    }
    const float norm0 = sqrtf(ss);
    float s1 = 0.f, dummy = 0.f;
    for (int im = tid; im < n_images; im += blockDim.x) {
        const float ni = (float)(img_off[im + 1] - img_off[im]);
        const float nx = normals[3 * im], ny = normals[3 * im + 1], nz = normals[3 * im + 2];
        float t0[3];
        if (ybranch) { t0[0] = (0.f - nz) / norm0; t0[1] = 0.f / norm0; t0[2] = nx / norm0; }
        else         { t0[0] = 0.f / norm0; t0[1] = nz / norm0; t0[2] = (0.f - ny) / norm0; }
        // torch.cross(n, t0)
        const float c0 = fmaf(ny, t0[2], -(nz * t0[1]));
        const float c1 = fmaf(nz, t0[0], -(nx * t0[2]));
        const float c2 = fmaf(nx, t0[1], -(ny * t0[0]));
        float *G = gt_rot + 9 * im;
        G[0] = t0[0]; G[1] = t0[1]; G[2] = t0[2];
        G[3] = c0; G[4] = c1; G[5] = c2;                       // normalised below
        G[6] = nx; G[7] = ny; G[8] = nz;
        s1 += ni * ((c0 * c0 + c1 * c1) + c2 * c2);
    }
    block_sum2(s1, dummy, s_a, s_b);
    const float norm1 = sqrtf(s1);
    for (int im = tid; im < n_images; im += blockDim.x)
#pragma unroll
        for (int k = 3; k < 6; ++k) gt_rot[9 * im + k] /= norm1;
}

// ---- sliced multi-target radix select ---------------------------------------------------------
// torch.median of a depth patch (pseudo_gt_z_box_loss, roi_heads.py:598) and the two
// torch.quantile calls of depth_range_loss (:676) are order statistics of up to 307 k depth
// values per box. One CTA per box (the first version) is bound by its longest box: a
// clipped projected box can cover the whole image while most cover a few thousand pixels.
// Here a box's source region is cut into slices of SEL_SLICE elements; a persistent grid
// strides over the (box, slice) items of all boxes, each item histograms one 8-bit digit of
// the order-preserving key into shared memory and adds its non-empty bins to the box's
// global histogram; the CTA that retires a box's last slice narrows the key prefix of every
// target rank and clears the histogram. Four launches = four digits. Up to four target
// ranks are followed at once, one histogram per distinct prefix.
constexpr int SEL_SLICE = 16384;
constexpr int SEL_THREADS = 256;

struct SelState {            // 32 words
    int r0, r1, c0, c1;      // source region of the depth map: rows [r0, r1) x cols [c0, c1)
    int image;               // index into the image table
    int mask;                // mask index, or -1: every pixel of the region takes part
    int slices;              // work items of this box; 0 = nothing to select
    int T;                   // targets: 1 (median) or 4 (two interpolated quantiles)
    int arrivals, nan, ngroups, cnt;
    unsigned prefix[4], rank[4], gprefix[4];
    int group[4];
    float w_hi, w_lo;        // lerp weights of the two quantiles
    int rlo, rhi;            // mask row range gathered by the count pass: mh - min_row, max_row + 1
};
static_assert(sizeof(SelState) == 128, "SelState layout");

__device__ __forceinline__ uint32_t float_key(float f);
__device__ __forceinline__ float key_float(uint32_t k);

// a[i0:i1] as Python slices it: negative indices wrap once, then clamp to [0, n]
__device__ __forceinline__ int py_slice_index(long long i, int n)
{
    if (i < 0) i += n;
    return (int)(i < 0 ? 0 : (i > n ? n : i));
}

// torch lerp (the interpolation of torch.quantile's default 'linear' mode)
__device__ __forceinline__ float torch_lerp(float a, float b, float w)
{
    const float diff = b - a;
    return fabsf(w) < 0.5f ? a + w * diff : b - diff * (1.f - w);
}

// exclusive prefix of states[b].slices over all boxes -> item_prefix[n + 1]; one CTA of 1024
__device__ void sel_plan(const SelState *states, int n, int *item_prefix)
{
    __shared__ int s_warp[32];
    __shared__ int s_carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int b = base + tid;
        const int v = b < n ? states[b].slices : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += x;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const int before = s_carry + (wid ? s_warp[wid - 1] : 0) + incl - v;
        if (b < n) item_prefix[b] = before;
        __syncthreads();
        if (tid == 1023) s_carry = before + v;
        __syncthreads();
    }
    if (tid == 0) item_prefix[n] = s_carry;
}

__device__ __forceinline__ void sel_arm(SelState &st, int elements)
{
    st.slices = (elements + SEL_SLICE - 1) / SEL_SLICE;
    st.arrivals = 0; st.nan = 0; st.ngroups = 1;
#pragma unroll
    for (int t = 0; t < 4; ++t) { st.prefix[t] = 0u; st.gprefix[t] = 0u; st.group[t] = 0; }
}

// pseudo_gt_z_box_loss (:585-606): region + rank of the lower median of every in-image box;
// the out-of-image fallback and the degenerate cases are answered here directly.
__global__ void __launch_bounds__(1024)
sel_init_median_kernel(const WcrImage *__restrict__ images, const int32_t *__restrict__ img_off, int n_images,
                       int n, const float *__restrict__ boxes, float *__restrict__ out_gt,
                       int *__restrict__ out_inside, SelState *__restrict__ states, int *__restrict__ item_prefix)
{
    for (int b = threadIdx.x; b < n; b += blockDim.x) {
        SelState st;
        memset(&st, 0, sizeof(st));
        st.image = image_of(img_off, n_images, b);
        st.mask = -1; st.T = 1;
        const WcrImage im = images[st.image];
        // Boxes.clip((h, w)), area() > 0 on the clipped float box, then .long() slicing
        const float x1 = clamp_keep_nan(boxes[4 * b + 0], 0.f, (float)im.W), y1 = clamp_keep_nan(boxes[4 * b + 1], 0.f, (float)im.H);
        const float x2 = clamp_keep_nan(boxes[4 * b + 2], 0.f, (float)im.W), y2 = clamp_keep_nan(boxes[4 * b + 3], 0.f, (float)im.H);
        if (x1 != x1 || y1 != y1 || x2 != x2 || y2 != y2) {
            // a NaN box (cube behind the camera): the reference indexes with an undefined
            // .long() here; defined as "no pseudo depth" (NaN, masked by the safe mean)
            out_gt[b] = nanf_(); out_inside[b] = 0;
        } else if (!((x2 - x1) * (y2 - y1) > 0.f)) {
            // fallback of :600-606: centre lookup of the CLIPPED box with the 10-px clamp
            out_gt[b] = depth_at(im, (x1 + x2) / 2.f, (y1 + y2) / 2.f); out_inside[b] = 0;
        } else {
            st.c0 = (int)x1; st.c1 = (int)x2; st.r0 = (int)y1; st.r1 = (int)y2;
            st.cnt = (st.c1 - st.c0) * (st.r1 - st.r0);
            out_inside[b] = 1;
            if (st.cnt <= 0) out_gt[b] = nanf_();   // torch.median of an empty slice raises in the reference
            else { sel_arm(st, st.cnt); st.rank[0] = (unsigned)((st.cnt - 1) / 2); }
        }
        states[b] = st;
    }
    __syncthreads();
    sel_plan(states, n, item_prefix);
}

// depth_range_loss: masked pixel count and row range of every box's mask; grid (chunks, n)
constexpr int SEL_COUNT_CHUNK = 16384;
__global__ void __launch_bounds__(256)
mask_count_kernel(const uint8_t *__restrict__ masks, int mh, int mw, const int32_t *__restrict__ mask_idx,
                  SelState *__restrict__ states)
{
    const int b = blockIdx.y, tid = threadIdx.x;
    const int total = mh * mw;
    const int lo = blockIdx.x * SEL_COUNT_CHUNK, hi = min(total, lo + SEL_COUNT_CHUNK);
    const uint8_t *m = masks + (size_t)mask_idx[b] * total;
    int cnt = 0, rmin = mh, rmax = -1;
    auto visit = [&](int idx) { ++cnt; const int r = idx / mw; rmin = min(rmin, r); rmax = max(rmax, r); };
    if ((((uintptr_t)(m + lo)) & 15) == 0) {
        const int nvec = (hi - lo) / 16;
        const uint4 *mv = reinterpret_cast<const uint4 *>(m + lo);
        for (int v = tid; v < nvec; v += blockDim.x) {
            const uint4 w = __ldg(mv + v);
            if ((w.x | w.y | w.z | w.w) == 0u) continue;
            const unsigned ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if ((ww[q] >> (8 * e)) & 255u) visit(lo + 16 * v + 4 * q + e);
        }
        for (int i = lo + 16 * nvec + tid; i < hi; i += blockDim.x) if (m[i]) visit(i);
    } else {
        for (int i = lo + tid; i < hi; i += blockDim.x) if (m[i]) visit(i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        rmin = min(rmin, __shfl_xor_sync(0xffffffffu, rmin, o));
        rmax = max(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
    }
    if ((tid & 31) == 0 && cnt) {
        atomicAdd(&states[b].cnt, cnt);
        atomicMax(&states[b].rlo, mh - rmin);
        atomicMax(&states[b].rhi, rmax + 1);
    }
}

__global__ void __launch_bounds__(1024)
sel_init_quantile_kernel(const WcrImage *__restrict__ images, const int32_t *__restrict__ img_off, int n_images,
                         int n, int mh, int mw, const int32_t *__restrict__ mask_idx,
                         const float *__restrict__ gt_boxes, float *__restrict__ out_range,
                         float *__restrict__ stats, SelState *__restrict__ states, int *__restrict__ item_prefix)
{
    for (int b = threadIdx.x; b < n; b += blockDim.x) {
        SelState st = states[b];            // zeroed, then cnt / rlo / rhi from mask_count_kernel
        st.image = image_of(img_off, n_images, b);
        st.T = 4; st.slices = 0;
        const WcrImage im = images[st.image];
        int elements = 0;
        if (im.H != mh || im.W != mw) {
            out_range[b] = nanf_(); stats[WCR_STAT_SHAPE_ERR] = 1.f;
        } else {
            if (st.cnt > 0) {               // depth_map[segmentation_mask]
                st.mask = mask_idx[b];
                st.r0 = mh - st.rlo; st.r1 = st.rhi; st.c0 = 0; st.c1 = mw;
                elements = (st.r1 - st.r0) * mw;
            } else {                        // segmentation failed: the GT box patch (:674-675)
                const float *g = gt_boxes + 4 * b;
                st.mask = -1;
                st.c0 = py_slice_index((long long)g[0], mw); st.c1 = py_slice_index((long long)g[2], mw);
                st.r0 = py_slice_index((long long)g[1], mh); st.r1 = py_slice_index((long long)g[3], mh);
                st.cnt = max(st.r1 - st.r0, 0) * max(st.c1 - st.c0, 0);
                elements = st.cnt;
            }
            if (st.cnt <= 0) out_range[b] = nanf_();   // torch.quantile of an empty tensor raises; NaN here
            else {
                sel_arm(st, elements);
                const float r9 = 0.9f * (float)(st.cnt - 1), r1 = 0.1f * (float)(st.cnt - 1);
                const float b9 = floorf(r9), b1 = floorf(r1);
                st.w_hi = r9 - b9; st.w_lo = r1 - b1;
                st.rank[0] = (unsigned)b9; st.rank[1] = (unsigned)ceilf(r9);
                st.rank[2] = (unsigned)b1; st.rank[3] = (unsigned)ceilf(r1);
            }
        }
        states[b] = st;
    }
    __syncthreads();
    sel_plan(states, n, item_prefix);
}

// One digit of every box's select. mode 0: out[b] = the selected value (median);
// mode 1: out[b] = lerp(hi pair) - lerp(lo pair) (q90 - q10). NG = how many distinct key
// prefixes a box may follow in this launch (1: median, and digit 0 of anything; 4 otherwise).
template <int NG>
__global__ void __launch_bounds__(SEL_THREADS)
sel_pass_kernel(SelState *__restrict__ states, const int *__restrict__ item_prefix, int n,
                unsigned *__restrict__ hist /* [n][4][256] */, const WcrImage *__restrict__ images,
                const uint8_t *__restrict__ masks, int mh, int mw, int pass, int mode, float *__restrict__ out)
{
    __shared__ unsigned sh[NG][256];
    __shared__ int s_last;
    const int tid = threadIdx.x;
    const int total = item_prefix[n];
    const int shift = 24 - 8 * pass;
    const unsigned himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
    for (int item = blockIdx.x; item < total; item += gridDim.x) {
        int b = 0;
        {
            int hi = n;   // item_prefix[b] <= item < item_prefix[hi]
            while (hi - b > 1) {
                const int mid = (b + hi) >> 1;
                if (item_prefix[mid] <= item) b = mid; else hi = mid;
            }
        }
        const int s = item - item_prefix[b];
        SelState *sp = states + b;
        const int r0 = sp->r0, c0 = sp->c0, ncols = sp->c1 - sp->c0, elements = (sp->r1 - sp->r0) * ncols;
        const int ng = NG == 1 ? 1 : sp->ngroups, slices = sp->slices, mk = sp->mask;
        unsigned gp[NG];
#pragma unroll
        for (int g = 0; g < NG; ++g) gp[g] = sp->gprefix[g];
        const WcrImage im = images[sp->image];
        const uint8_t *m = mk >= 0 ? masks + (size_t)mk * mh * mw : nullptr;
        for (int i = tid; i < ng * 256; i += SEL_THREADS) (&sh[0][0])[i] = 0u;
        __syncthreads();
        // per-thread run-length encoding per group: neighbouring depths share their leading
        // bytes, plain shared atomics would serialise on one or two bins
        unsigned last[NG], run[NG];
#pragma unroll
        for (int g = 0; g < NG; ++g) { last[g] = ~0u; run[g] = 0u; }
        bool saw_nan = false;
        // branch-free: a predicated shared atomic only when the thread's run of equal bins ends
        auto feed = [&](float f) {
            saw_nan |= f != f;
            const uint32_t u = __float_as_uint(f);
            const uint32_t k = u ^ ((uint32_t)((int32_t)u >> 31) | 0x80000000u);   // float_key
            const unsigned hi = k & himask, bin = (k >> shift) & 255u;
#pragma unroll
            for (int g = 0; g < NG; ++g) {
                const bool match = (NG == 1 || g < ng) && hi == gp[g];
                const bool same = match && bin == last[g];
                if (match && !same && run[g]) atomicAdd(&sh[g][last[g]], run[g]);
                run[g] = same ? run[g] + 1u : (match ? 1u : run[g]);
                last[g] = match ? bin : last[g];
            }
        };
        const int lo = s * SEL_SLICE, hi = min(elements, lo + SEL_SLICE);
        const float *dbase = im.depth + (size_t)r0 * im.depth_pitch + c0;
        const bool vec_mask = m && c0 == 0 && ncols == mw && (mw & 3) == 0 && (((uintptr_t)m) & 3) == 0;
        if (vec_mask) {
            // masked rows of full width: 4 pixels per 32-bit mask word, most words are empty
            const uint32_t *mw32 = reinterpret_cast<const uint32_t *>(m + (size_t)r0 * mw);
            const int wpr = mw >> 2;                       // words per row
            const int w_lo = lo >> 2, w_hi = (hi + 3) >> 2;  // lo is a multiple of SEL_SLICE
            for (int w0 = w_lo + tid; w0 < w_hi; w0 += 2 * SEL_THREADS) {
                uint32_t bits[2];
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int w = w0 + SEL_THREADS * q;
                    bits[q] = w < w_hi ? __ldg(mw32 + w) : 0u;
                }
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    if (!bits[q]) continue;
                    const int w = w0 + SEL_THREADS * q;
                    const int r = w / wpr, c = (w - r * wpr) << 2;
                    const float *row = dbase + (size_t)r * im.depth_pitch + c;
                    float f[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) f[e] = (bits[q] >> (8 * e)) & 255u ? __ldg(row + e) : 0.f;
#pragma unroll
                    for (int e = 0; e < 4; ++e) if ((bits[q] >> (8 * e)) & 255u) feed(f[e]);
                }
            }
        } else {
            // 4 CONSECUTIVE elements per thread per round (the run-length encoder then sees
            // neighbouring pixels; a warp still covers 128 contiguous floats), loads issued
            // before use; (row, col) advance incrementally by the CTA stride
            const int stride = 4 * SEL_THREADS;
            const int dq = stride / ncols, dc = stride - dq * ncols;
            int i0 = lo + 4 * tid;
            int r = i0 / ncols, c = i0 - r * ncols;
            for (; i0 < hi; i0 += stride) {
                const float *p = dbase + (size_t)r * im.depth_pitch + c;
                if (!m && c + 3 < ncols && i0 + 3 < hi) {      // 4 pixels of one row: one address
                    const float f0 = __ldg(p), f1 = __ldg(p + 1), f2 = __ldg(p + 2), f3 = __ldg(p + 3);
                    feed(f0); feed(f1); feed(f2); feed(f3);
                } else {
                    float f[4];
                    bool take[4];
                    int rr = r, cc = c;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        take[q] = i0 + q < hi;
                        if (take[q] && m) take[q] = m[(size_t)(r0 + rr) * mw + c0 + cc] != 0;
                        f[q] = take[q] ? __ldg(dbase + (size_t)rr * im.depth_pitch + cc) : 0.f;
                        if (++cc == ncols) { cc = 0; ++rr; }
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) if (take[q]) feed(f[q]);
                }
                r += dq; c += dc;
                if (c >= ncols) { c -= ncols; ++r; }
            }
        }
#pragma unroll
        for (int g = 0; g < NG; ++g) if (run[g]) atomicAdd(&sh[g][last[g]], run[g]);
        if (pass == 0 && saw_nan) sp->nan = 1;
        __syncthreads();
        unsigned *gh = hist + (size_t)b * 1024;
        for (int i = tid; i < ng * 256; i += SEL_THREADS) {
            const unsigned v = (&sh[0][0])[i];
            if (v) atomicAdd(gh + i, v);
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) s_last = atomicAdd(&sp->arrivals, 1) == slices - 1;
        __syncthreads();
        if (s_last) {   // this CTA retired the box's last slice: narrow every target by one digit
            __threadfence();
            const int T = sp->T;
            // stage the box's histograms in shared memory (one coalesced L2 read), then one
            // warp per target: 8 bins per lane, warp prefix, the lane that holds the rank
            for (int i = tid; i < ng * 256; i += SEL_THREADS) (&sh[0][0])[i] = __ldcg(gh + i);
            __syncthreads();
            const int wid = tid >> 5, lane = tid & 31;
            if (wid < T) {
                const unsigned *h = sh[NG == 1 ? 0 : sp->group[wid]] + 8 * lane;
                unsigned c[8], sum = 0u;
#pragma unroll
                for (int q = 0; q < 8; ++q) { c[q] = h[q]; sum += c[q]; }
                unsigned incl = sum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
                    if (lane >= o) incl += t;
                }
                const unsigned rank = sp->rank[wid];
                const unsigned before = incl - sum;
                // first lane whose inclusive count exceeds the rank (the last lane if none does:
                // cannot happen while rank < number of elements under the prefix)
                const unsigned hit = __ballot_sync(0xffffffffu, incl > rank);
                const int owner = hit ? __ffs(hit) - 1 : 31;
                if (lane == owner) {
                    unsigned acc = before;
                    int q = 0;
                    for (; q < 7; ++q) {
                        if (acc + c[q] > rank) break;
                        acc += c[q];
                    }
                    sp->rank[wid] = rank - acc;
                    sp->prefix[wid] |= (unsigned)(8 * lane + q) << shift;
                }
            }
            __syncthreads();
            for (int i = tid; i < ng * 256; i += SEL_THREADS) gh[i] = 0u;
            if (tid == 0) {
                int g_n = 0;
                for (int t = 0; t < T; ++t) {
                    int g = 0;
                    for (; g < g_n; ++g) if (sp->gprefix[g] == sp->prefix[t]) break;
                    if (g == g_n) sp->gprefix[g_n++] = sp->prefix[t];
                    sp->group[t] = g;
                }
                sp->ngroups = g_n;
                sp->arrivals = 0;
                if (pass == 3) {
                    float v;
                    if (mode == 0) v = key_float(sp->prefix[0]);
                    else v = torch_lerp(key_float(sp->prefix[0]), key_float(sp->prefix[1]), sp->w_hi) -
                             torch_lerp(key_float(sp->prefix[2]), key_float(sp->prefix[3]), sp->w_lo);
                    out[b] = sp->nan ? nanf_() : v;
                }
            }
        }
        __syncthreads();
    }
}

// ---- 'segmentation': segment_loss (roi_heads.py:420-443) ------------------------------------------
// y_hat = convex_hull(mask shape, clamped projected corners): Jarvis march, then the SOFT raster
// prod_e min(max(val_e, 0), 1) (utils.py:456-488), which is differentiable in the hull vertices;
// loss_i = mean over pixels of sigmoid_focal_loss(inputs = GT mask as 0/1 "logits", targets =
// raster) -- the argument order of the reference, kept. The reference builds an (n, H, W) raster
// with ~8 full-image torch ops per edge per box; here one CTA per box evaluates every pixel once
// (early exit at the first edge with val < 0), forward and backward share the code. Gradient
// conventions of autograd: binary max/min split ties (val == 0 or val == 1 exactly -- common,
// the clamped corners and the pixel centres are integers) in half.
struct FocalConst { float p[2], c[2]; };   // sigmoid(x), log(1 + exp(-x)) for x = 0, 1

__device__ __forceinline__ FocalConst focal_const()
{
    FocalConst k;
    k.p[0] = 0.5f; k.p[1] = 1.f / (1.f + expf(-1.f));
    k.c[0] = logf(2.f); k.c[1] = logf(1.f + expf(-1.f));
    return k;
}
// torchvision sigmoid_focal_loss(inputs = x in {0, 1}, targets = t), alpha 0.25, gamma 2; optionally d/dt
__device__ __forceinline__ float focal_xt(const FocalConst &k, int x, float t, float *ddt)
{
    const float p = k.p[x], xf = (float)x;
    const float ce = (1.f - t) * xf + k.c[x];                  // BCE-with-logits, max_val = 0 for x >= 0
    const float p_t = p * t + (1.f - p) * (1.f - t);
    const float m = 1.f - p_t;
    const float loss = ce * (m * m);
    const float alpha_t = 0.25f * t + (1.f - 0.25f) * (1.f - t);
    if (ddt) *ddt = -0.5f * loss + alpha_t * (-xf * (m * m) + ce * (2.f * m) * (1.f - 2.f * p));
    return alpha_t * loss;
}

template <bool BWD>
__global__ void __launch_bounds__(256)
seg_loss_kernel(const WcrLossIO io, const WcrLossParams P, const float *__restrict__ corners_in /* [n][16] or NULL */,
                float *__restrict__ out_raw /* [n] */, float *__restrict__ out_dcorner /* [n][16] */)
{
    __shared__ float s_hx[8], s_hy[8];
    __shared__ int s_ci[8], s_m;
    __shared__ double s_red[8];
    __shared__ float s_g[8][16];
    const int i = blockIdx.x, tid = threadIdx.x;
    const int H = io.mask_h, W = io.mask_w;
    if (tid == 0) {
        float px[8], py[8], hx[8], hy[8];
        bool bad = false;
        if (corners_in) {           // stand-alone segment_loss: the caller's (already clamped) corners
#pragma unroll
            for (int c = 0; c < 8; ++c) { px[c] = corners_in[16 * (size_t)i + 2 * c]; py[c] = corners_in[16 * (size_t)i + 2 * c + 1]; }
        } else {
            const Decoded d = decode_box(io, P, i);
            float lo0, hi0, lo1, hi1, cxy[16];
            clamp_bounds(io.clamp[2 * i], lo0, hi0);
            clamp_bounds(io.clamp[2 * i + 1], lo1, hi1);
            project_cube(d.cube, d.K, lo0, hi0, lo1, hi1, cxy);
            // second clamp, roi_heads.py:951-952: x into [0, int(c0 - 1)], y into [0, int(c1 - 1)]
            const float mx = (float)(int)(io.clamp[2 * i] - 1.f), my = (float)(int)(io.clamp[2 * i + 1] - 1.f);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                px[c] = clamp_keep_nan(cxy[2 * c], 0.f, mx);
                py[c] = clamp_keep_nan(cxy[2 * c + 1], 0.f, my);
            }
        }
#pragma unroll
        for (int c = 0; c < 8; ++c) bad |= px[c] != px[c] || py[c] != py[c];
        int m = -1;
        if (!bad) m = jarvis_march(Strided{px, 1}, Strided{py, 1}, 8, Strided{hx, 1}, Strided{hy, 1});
        s_m = m;
        for (int v = 0; v < max(m, 0); ++v) {
            s_hx[v] = hx[v]; s_hy[v] = hy[v];
            int ci = 0;
            for (int c = 7; c >= 0; --c) if (px[c] == hx[v] && py[c] == hy[v]) ci = c;   // px/py carry the bump
            s_ci[v] = ci;
        }
    }
    __syncthreads();
    const int m = s_m;
    const FocalConst fk = focal_const();
    const uint8_t *mask = io.masks + (size_t)io.mask_idx[i] * H * W;
    const int N = H * W;
    float v1x[8], v1y[8], ex[8], ey[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const int e2 = e + 1 < m ? e + 1 : 0;
        v1x[e] = e < m ? s_hx[e] : 0.f; v1y[e] = e < m ? s_hy[e] : 0.f;
        ex[e] = e < m ? s_hx[e2] - s_hx[e] : 0.f; ey[e] = e < m ? s_hy[e2] - s_hy[e] : 0.f;
    }
    double acc = 0.0;
    float g[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) g[k] = 0.f;
    for (int idx = tid; idx < N; idx += blockDim.x) {
        const int y = idx / W, x = idx - y * W;
        const int xm = mask[idx] != 0;
        const float X = (float)x, Y = (float)y;
        float t = m > 0 ? 1.f : 0.f, il[8];
        bool dead = m <= 0;
        int zeros = 0, e0 = 0;
        float vmin = 3.0e38f;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            if (e >= m || dead) { il[e] = 1.f; continue; }
            const float val = (X - v1x[e]) * ey[e] - (Y - v1y[e]) * ex[e];
            if (!(val >= 0.f)) { dead = true; t = 0.f; il[e] = 0.f; continue; }    // val < 0 (or NaN): outside
            if (val == 0.f) { ++zeros; e0 = e; }
            vmin = fminf(vmin, val);
            il[e] = fminf(val, 1.f);
            t = t * il[e];
        }
        if (!BWD) {
            acc += (double)focal_xt(fk, xm, t, nullptr);
        } else if (!dead && zeros <= 1 && vmin <= 1.f) {
            float ddt;
            focal_xt(fk, xm, t, &ddt);
            const float up = ddt / (float)N;
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                if (e >= m) continue;
                const float val = (X - v1x[e]) * ey[e] - (Y - v1y[e]) * ex[e];
                float w;                                   // d clamp(val, 0, 1) / d val with torch's tie split
                if (zeros == 1) w = e == e0 ? 0.5f : 0.f;  // every other factor multiplies a zero
                else w = val < 1.f ? 1.f : (val == 1.f ? 0.5f : 0.f);
                if (w == 0.f) continue;
                float others = 1.f;
#pragma unroll
                for (int o = 0; o < 8; ++o) if (o < m && o != e) others *= il[o];
                const float ge = up * w * others;
                const int e2 = e + 1 < m ? e + 1 : 0;
                const float v2x = v1x[e] + ex[e], v2y = v1y[e] + ey[e];
                // val = (X - v1x)(v2y - v1y) - (Y - v1y)(v2x - v1x)
                g[2 * e] += ge * (Y - v2y);      g[2 * e + 1] += ge * (v2x - X);
                g[2 * e2] += ge * -(Y - v1y[e]); g[2 * e2 + 1] += ge * (X - v1x[e]);
            }
        }
    }
    const int lane = tid & 31, wid = tid >> 5;
    if (!BWD) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) s_red[wid] = acc;
        __syncthreads();
        if (tid == 0) {
            double s = 0.0;
            for (int w = 0; w < 8; ++w) s += s_red[w];
            out_raw[i] = (float)(s / (double)N);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            float v = g[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) s_g[wid][k] = v;
        }
        __syncthreads();
        if (tid < 16) {
            float out = 0.f;
            const int c = tid >> 1, ax = tid & 1;           // gradient of corner c, axis ax
            for (int v = 0; v < m; ++v)
                if (s_ci[v] == c) {
                    float s = 0.f;
                    for (int w = 0; w < 8; ++w) s += s_g[w][2 * v + ax];
                    out += s;
                }
            out_dcorner[16 * (size_t)i + tid] = out;
        }
    }
}

// --------------------------------------------------------------- backward --
// upstream coefficient of the per-box scaled loss l_{t,i} = raw_{t,i} * sf_i:
//   d(total)/d(l_{t,i}) = g_t * w_t * [l finite] / n_valid_t      (safe mean, :1307-1315)
__device__ __forceinline__ float term_coef(const WcrLossIO &io, const WcrLossParams &P, int t, float l)
{
    if (!io.present[t] || !finitef_(l)) return 0.f;
    const float nv = io.stats[t];
    return nv > 0.f ? io.grad_losses[t] * term_weight(P, t) / nv : 0.f;
}

// One CTA per image: d(total)/d(R_ego) from the pose-alignment term.
//   P = sum_im L_im / (fail+1), L_im = mean_{i>j}(1 - |cos_ij|), cos_ij = (R_i : R_j - 1)/2
__global__ void __launch_bounds__(128)
loss_pose_bwd_kernel(const WcrLossIO io, const WcrLossParams P)
{
    __shared__ float s_a[8], s_b[8];
    const int im = blockIdx.x;
    const int b0 = io.img_off[im], ni = io.img_off[im + 1] - b0;
    // dP = sum_i coef_i * sf_i  (raw pose term of every box is the same scalar P)
    const float pose = io.stats[WCR_STAT_POSE];
    float dP = 0.f, dummy = 0.f;
    for (int i = threadIdx.x; i < io.n; i += blockDim.x) {
        const float sf = uncert_sf(io, P, i);
        dP += term_coef(io, P, WCR_T_POSE, pose * sf) * sf;
    }
    block_sum2(dP, dummy, s_a, s_b);
    const float npairs = 0.5f * (float)ni * (float)(ni - 1);
    const float scale = ni > 1 ? dP / (npairs * (io.stats[WCR_STAT_FAIL] + 1.f)) : 0.f;
    for (int i = threadIdx.x; i < ni; i += blockDim.x) {
        float Ri[9], g[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) { Ri[t] = io.R_ego[9 * (size_t)(b0 + i) + t]; g[t] = 0.f; }
        for (int j = 0; j < ni; ++j) {
            if (j == i) continue;
            const float *Rj = io.R_ego + 9 * (size_t)(b0 + j);
            const float c = pair_cos(Ri, Rj);
            const float dc = -sgn(c) * scale * 0.5f;   // d(1-|cos|)/dcos * d cos/d(R_i:R_j)
#pragma unroll
            for (int t = 0; t < 9; ++t) g[t] += dc * Rj[t];
        }
#pragma unroll
        for (int t = 0; t < 9; ++t) io.d_R_pose[9 * (size_t)(b0 + i) + t] = g[t];
    }
}

// Gradient gp of a loss w.r.t. ONE projected coordinate (x if isx else y) of corner c, pushed back
// through p = (f * A + c0 * Z) / Z, corner = R (s * h) + centre, centre = z * ((cube_xy - c0) / f, 1).
__device__ __forceinline__ void corner_backprop(const Decoded &d, int c, bool isx, float gp, float Xc, float Yc,
                                                float Zc, float hx, float hy, float hz, float *dR, float *dl,
                                                float &dcx, float &dcy, float &dz)
{
    // p = (f*A + c0*Z)/Z : dA = gp*f/Z ; dZ = gp*(c0/Z - num/Z^2)
    const float f = isx ? d.K.fx : d.K.fy, c0 = isx ? d.K.cx : d.K.cy;
    const float Av = isx ? Xc : Yc;
    const float num = fmaf(c0, Zc, f * Av);
    const float dA = gp * f / Zc;
    const float dZ = gp * c0 / Zc - gp * num / (Zc * Zc);
    float sx, sy, sz;
    corner_signs(c, sx, sy, sz);
    const float l[3] = {sx * hx, sy * hy, sz * hz};
    const int ra = isx ? 0 : 1;                     // row of R feeding A
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        dR[3 * ra + k] += dA * l[k];
        dR[6 + k] += dZ * l[k];
        dl[k] += (d.R[3 * ra + k] * dA + d.R[6 + k] * dZ) * (k == 0 ? sx : (k == 1 ? sy : sz));
    }
    // centre: cube[0] or cube[1] gets dA, cube[2] gets dZ
    if (isx) {   // x3d = z*(cube_x-cx)/fx
        dcx += dA * d.z / d.K.fx;
        dz += dA * (d.cube_x - d.K.cx) / d.K.fx;
    } else {
        dcy += dA * d.z / d.K.fy;
        dz += dA * (d.cube_y - d.K.cy) / d.K.fy;
    }
    dz += dZ;
}

__global__ void __launch_bounds__(128)
loss_box_bwd_kernel(const WcrLossIO io, const WcrLossParams P)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= io.n) return;
    const Decoded d = decode_box(io, P, i);
    const int img = image_of(io.img_off, io.n_images, i);
    const float sf = uncert_sf(io, P, i);
    float du = 0.f;                       // d/d uncert
    float dz = 0.f, dcx = 0.f, dcy = 0.f; // d/d cube_z, cube_x, cube_y
    float ddims[3] = {0.f, 0.f, 0.f};     // d/d (w, h, l)
    float dR[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) dR[t] = 0.f;
    auto coef = [&](int t, float raw) {   // -> d/d raw, and accumulates d/d uncert
        const float l = raw * sf;
        const float c = term_coef(io, P, t, l);
        if (c != 0.f && P.use_confidence > 0.f) du -= c * l;
        return c * sf;
    };

    // ---- 'iou': GIoU -> projected box -> arg-extremum corners -> cube ------------
    if ((P.loss_flags & WCR_LOSS_IOU) && io.present[WCR_T_IOU]) {
        float lo0, hi0, lo1, hi1;
        clamp_bounds(io.clamp[2 * i], lo0, hi0);
        clamp_bounds(io.clamp[2 * i + 1], lo1, hi1);
        const float hx = d.dims[2] / 2.f, hy = d.dims[1] / 2.f, hz = d.dims[0] / 2.f;
        float X[8], Y[8], Z[8], px[8], py[8], cxs[8], cys[8];
        int ix1 = 0, ix2 = 0, iy1 = 0, iy2 = 0;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            float sx, sy, sz;
            corner_signs(c, sx, sy, sz);
            const float lx = sx * hx, ly = sy * hy, lz = sz * hz;
            X[c] = fmaf(d.R[2], lz, fmaf(d.R[1], ly, d.R[0] * lx)) + d.cube[0];
            Y[c] = fmaf(d.R[5], lz, fmaf(d.R[4], ly, d.R[3] * lx)) + d.cube[1];
            Z[c] = fmaf(d.R[8], lz, fmaf(d.R[7], ly, d.R[6] * lx)) + d.cube[2];
            px[c] = fmaf(d.K.cx, Z[c], d.K.fx * X[c]) / Z[c];
            py[c] = fmaf(d.K.cy, Z[c], d.K.fy * Y[c]) / Z[c];
            cxs[c] = clamp_keep_nan(px[c], lo0, hi0);
            cys[c] = clamp_keep_nan(py[c], lo1, hi1);
            if (c > 0) {   // strict comparisons keep the FIRST arg-extremum (torch.min/max over dim)
                if (cxs[c] < cxs[ix1]) ix1 = c;
                if (cxs[c] > cxs[ix2]) ix2 = c;
                if (cys[c] < cys[iy1]) iy1 = c;
                if (cys[c] > cys[iy2]) iy2 = c;
            }
        }
        const float proj[4] = {cxs[ix1], cys[iy1], cxs[ix2], cys[iy2]};
        float dproj[4];
        const float raw = giou_loss(io.gt_boxes + 4 * i, proj, dproj);
        const float g = coef(WCR_T_IOU, raw);
        if (g != 0.f) {
            const int sel[4] = {ix1, iy1, ix2, iy2};
            float dl[3] = {0.f, 0.f, 0.f};   // d/d (lx, ly, lz) magnitudes hx hy hz, signed below
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int c = sel[e];
                const bool isx = (e == 0 || e == 2);
                const float pv = isx ? px[c] : py[c];
                const float lo = isx ? lo0 : lo1, hi = isx ? hi0 : hi1;
                if (!(pv >= lo && pv <= hi)) continue;          // clamp backward: closed interval
                corner_backprop(d, c, isx, g * dproj[e], X[c], Y[c], Z[c], hx, hy, hz, dR, dl, dcx, dcy, dz);
            }
            ddims[2] += 0.5f * dl[0];   // hx = l/2
            ddims[1] += 0.5f * dl[1];   // hy = h/2
            ddims[0] += 0.5f * dl[2];   // hz = w/2
        }
    }
    // ---- 'segmentation': d raw / d (second-clamped corners) comes from seg_loss_kernel<true> ----
    if ((P.loss_flags & WCR_LOSS_SEG) && io.present[WCR_T_SEG]) {
        const float g = coef(WCR_T_SEG, io.terms[(size_t)(WCR_T_SEG - 1) * io.n + i]);
        if (g != 0.f) {
            float lo0, hi0, lo1, hi1;
            clamp_bounds(io.clamp[2 * i], lo0, hi0);
            clamp_bounds(io.clamp[2 * i + 1], lo1, hi1);
            const float mx = (float)(int)(io.clamp[2 * i] - 1.f), my = (float)(int)(io.clamp[2 * i + 1] - 1.f);
            const float hx = d.dims[2] / 2.f, hy = d.dims[1] / 2.f, hz = d.dims[0] / 2.f;
            float dl[3] = {0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                float sx, sy, sz;
                corner_signs(c, sx, sy, sz);
                const float lx = sx * hx, ly = sy * hy, lz = sz * hz;
                const float Xc = fmaf(d.R[2], lz, fmaf(d.R[1], ly, d.R[0] * lx)) + d.cube[0];
                const float Yc = fmaf(d.R[5], lz, fmaf(d.R[4], ly, d.R[3] * lx)) + d.cube[1];
                const float Zc = fmaf(d.R[8], lz, fmaf(d.R[7], ly, d.R[6] * lx)) + d.cube[2];
                const float pxy[2] = {fmaf(d.K.cx, Zc, d.K.fx * Xc) / Zc, fmaf(d.K.cy, Zc, d.K.fy * Yc) / Zc};
#pragma unroll
                for (int ax = 0; ax < 2; ++ax) {
                    const float gp = g * io.seg_dcorner[16 * (size_t)i + 2 * c + ax];
                    const float lo = ax ? lo1 : lo0, hi = ax ? hi1 : hi0, m2 = ax ? my : mx;
                    const float pv = pxy[ax];
                    if (gp == 0.f || !(pv >= lo && pv <= hi)) continue;   // first clamp (spaces.py:241-242), closed
                    if (!(pv >= 0.f && pv <= m2)) continue;               // second clamp (roi_heads.py:951-952)
                    corner_backprop(d, c, ax == 0, gp, Xc, Yc, Zc, hx, hy, hz, dR, dl, dcx, dcy, dz);
                }
            }
            ddims[2] += 0.5f * dl[0]; ddims[1] += 0.5f * dl[1]; ddims[0] += 0.5f * dl[2];
        }
    }
    // ---- 'pose_alignment' ---------------------------------------------------------
    if ((P.loss_flags & WCR_LOSS_POSE_ALIGN) && io.present[WCR_T_POSE]) {
        coef(WCR_T_POSE, io.stats[WCR_STAT_POSE]);   // only the uncertainty part; R part comes from the pair kernel
#pragma unroll
        for (int t = 0; t < 9; ++t) dR[t] += io.d_R_pose[9 * (size_t)i + t];
    }
    // ---- 'pose_ground' --------------------------------------------------------------
    if ((P.loss_flags & WCR_LOSS_POSE_GROUND) && !(P.loss_flags & WCR_LOSS_POSE_GROUND2) &&
        io.present[WCR_T_NORMAL]) {
        float an[3], bn[3], nb;
        const float c = cos_sim3(io.normals + 3 * img, d.R + 3, an, bn, &nb);
        const float conf = io.ground_conf[img];
        const float g = coef(WCR_T_NORMAL, (1.f - fabsf(c)) * conf);
        if (g != 0.f) {
            const float dc = -sgn(c) * conf * g;
            // bn = p / ||p||: dp = (dbn - bn (bn . dbn)) / ||p||, dbn = dc * an
            const float dot = dc * ((an[0] * bn[0] + an[1] * bn[1]) + an[2] * bn[2]);
#pragma unroll
            for (int k = 0; k < 3; ++k) dR[3 + k] += (dc * an[k] - bn[k] * dot) / nb;
        }
    }
    // ---- 'z_pseudo_gt_patch' (median of the depth patch; sign saved by position) --------
    if ((P.loss_flags & WCR_LOSS_Z_PATCH) && io.present[WCR_T_PSEUDO_Z]) {
        const float g = coef(WCR_T_PSEUDO_Z, io.terms[(size_t)(WCR_T_PSEUDO_Z - 1) * io.n + i]);
        dz += g * io.z_sign[i];
    } else
    // ---- 'z_pseudo_gt_center' ---------------------------------------------------------
    if ((P.loss_flags & WCR_LOSS_Z_CENTER) && io.present[WCR_T_PSEUDO_Z]) {
        const float diff = d.z - depth_at(io.images[img], d.cube_x, d.cube_y);
        const float g = coef(WCR_T_PSEUDO_Z, fabsf(diff));
        dz += g * sgn(diff);
    }
    // ---- 'pose_ground2' (replaces the cosine term in the same slot) -----------------------
    if ((P.loss_flags & WCR_LOSS_POSE_GROUND2) && io.present[WCR_T_NORMAL]) {
        const float *G = io.gt_rot + 9 * img;
        const float conf = io.ground_conf[img];
        const float g = coef(WCR_T_NORMAL, ground2_term(d.R, G, conf, nullptr));
#pragma unroll
        for (int t = 0; t < 9; ++t) dR[t] += g * conf * -0.5f * G[t];
    }
    // ---- 'z' (arg-min over constant offsets: no gradient to the cube, only to uncert) -------
    if ((P.loss_flags & WCR_LOSS_Z_SEARCH) && io.present[WCR_T_Z])
        coef(WCR_T_Z, io.terms[(size_t)(WCR_T_Z - 1) * io.n + i]);
    // ---- 'depth' ------------------------------------------------------------------------------
    if ((P.loss_flags & WCR_LOSS_DEPTH) && io.present[WCR_T_DEPTH]) {
        int cmax, cmin;
        const float pred = corner_depth_range(d, cmax, cmin);
        const float diff = io.depth_gt[i] - pred;
        const float g = coef(WCR_T_DEPTH, fabsf(diff)) * -sgn(diff);   // d|gt - pred| / d pred
        if (g != 0.f) {
            const float h[3] = {d.dims[2] / 2.f, d.dims[1] / 2.f, d.dims[0] / 2.f};
            float sa[3], sb[3];
            corner_signs(cmax, sa[0], sa[1], sa[2]);
            corner_signs(cmin, sb[0], sb[1], sb[2]);
            // Z_c = R[2,:] . (s_c * h) + z : the centre's z cancels in max - min
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                dR[6 + k] += g * (sa[k] - sb[k]) * h[k];
                const float dh = g * d.R[6 + k] * (sa[k] - sb[k]);
                ddims[2 - k] += 0.5f * dh;                             // hx = l/2, hy = h/2, hz = w/2
            }
        }
    }
    // ---- 'dims' -------------------------------------------------------------------------
    if ((P.loss_flags & WCR_LOSS_DIMS) && io.present[WCR_T_DIMS_W]) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float mu = io.prior_mean[3 * i + k], sd = io.prior_std[3 * i + k];
            const float s = fabsf(d.dims[k] - mu) / sd - 1.0f;
            const float g = coef(WCR_T_DIMS_W + k, fmaxf(s, 0.f));
            const float w = s > 0.f ? 1.f : (s == 0.f ? 0.5f : 0.f);   // torch.max(a, zeros) tie rule
            ddims[k] += g * w * sgn(d.dims[k] - mu) / sd;
        }
    }
    // ---- 'uncert' term itself: use_confidence * safe_mean(u) -------------------------------
    if (io.present[WCR_T_UNCERT]) du += term_coef(io, P, WCR_T_UNCERT, io.uncert[i]);

    // ---- back through the decode ----------------------------------------------------------
    io.d_uncert[i] = du;
    io.d_deltas[2 * i] = dcx * d.sw;
    io.d_deltas[2 * i + 1] = dcy * d.sh;
    const float v2r = io.v2r[i];
    float dzr = dz * v2r;
    if (P.z_type == 1) { const float sg = d.zf / 100.f; dzr *= 100.f * sg * (1.f - sg); }
    else if (P.z_type == 2) dzr *= d.zf;
    io.d_z_raw[i] = dzr;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float raw = io.dims_raw[3 * i + k];
        float g;
        if (P.dims_func == 1) {
            const float pm = io.prior_mean[3 * i + k], ps = io.prior_std[3 * i + k];
            const float dmin = fmaxf(pm - 3.f * ps, 0.f), dmax = pm + 3.f * ps;
            const float sg = sigmoidf_(raw);
            g = ddims[k] * (dmax - dmin) * sg * (1.f - sg);
        } else {
            g = raw <= 5.f ? ddims[k] * d.dims[k] : 0.f;   // clip(max=5) passes gradient on x <= 5
        }
        io.d_dims_raw[3 * i + k] = g;
    }
    // R_ego = M R_view (where angle > 0) -> d R_view = M^T dR
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            io.d_pose[9 * i + 3 * r + c] = d.validM
                ? (d.M[r] * dR[c] + d.M[3 + r] * dR[3 + c]) + d.M[6 + r] * dR[6 + c]
                : dR[3 * r + c];
}

}  // namespace wcr

static int check_io(const WcrLossIO *io, const WcrLossParams *p, bool bwd)
{
    WCR_REQUIRE(io && p, "wcr_weak_loss: null pointer");
    WCR_REQUIRE(io->n >= 0 && io->n_images >= 1, "wcr_weak_loss: bad sizes");
    WCR_REQUIRE(io->deltas && io->z_raw && io->dims_raw && io->pose && io->uncert && io->src_boxes &&
                io->K4 && io->v2r && io->prior_mean && io->prior_std && io->gt_boxes && io->clamp &&
                io->img_off && io->images && io->normals && io->ground_conf,
                "wcr_weak_loss: null input pointer");
    WCR_REQUIRE(io->losses && io->present && io->terms && io->R_ego && io->proj && io->stats &&
                io->img_pose, "wcr_weak_loss: null output / workspace pointer");
    if (bwd)
        WCR_REQUIRE(io->grad_losses && io->d_deltas && io->d_z_raw && io->d_dims_raw && io->d_pose &&
                    io->d_uncert && io->d_R_pose, "wcr_weak_loss_bwd: null gradient pointer");
    WCR_REQUIRE(p->z_type >= 0 && p->z_type <= 2 && p->dims_func >= 0 && p->dims_func <= 1,
                "wcr_weak_loss: unsupported z_type / dims_func");
    if (p->loss_flags & WCR_LOSS_Z_PATCH)
        WCR_REQUIRE(io->z_dec && io->z_gt && io->z_inside && io->z_sign,
                    "wcr_weak_loss: 'z_pseudo_gt_patch' needs the z_dec / z_gt / z_inside / z_sign buffers");
    if (p->loss_flags & WCR_LOSS_DEPTH)
        WCR_REQUIRE(io->masks && io->mask_idx && io->depth_gt && io->mask_h > 0 && io->mask_w > 0,
                    "wcr_weak_loss: 'depth' needs masks, mask_idx, depth_gt and the mask shape");
    if (p->loss_flags & WCR_LOSS_SEG) {
        WCR_REQUIRE(io->masks && io->mask_idx && io->mask_h > 0 && io->mask_w > 0,
                    "wcr_weak_loss: 'segmentation' needs masks, mask_idx and the mask shape");
        if (bwd) WCR_REQUIRE(io->seg_dcorner, "wcr_weak_loss_bwd: 'segmentation' needs seg_dcorner");
    }
    if (p->loss_flags & (WCR_LOSS_DEPTH | WCR_LOSS_Z_PATCH))
        WCR_REQUIRE(io->select_ws, "wcr_weak_loss: 'depth' / 'z_pseudo_gt_patch' need select_ws "
                                   "(wcr_select_ws_bytes(n) bytes)");
    if (p->loss_flags & WCR_LOSS_POSE_GROUND2)
        WCR_REQUIRE(io->gt_rot, "wcr_weak_loss: 'pose_ground2' needs the gt_rot workspace");
    return WCR_OK;
}

// torch.linspace(0.0, 4.9, 50) in fp32 (roi_heads.py:547): step = (end - start) / (steps - 1);
// the lower half counts up from start, the upper half counts down from end (ATen's
// symmetric formulation), each as ONE fused multiply-add -- both ATen's vectorised CPU
// kernel and its CUDA kernel contract it. tests/test_abi_cpu.py compares this table with
// torch.linspace bit for bit.
extern "C" void wcr_z_loss_steps(float out[WCR_Z_LOSS_STEPS])
{
    const int steps = WCR_Z_LOSS_STEPS;
    const float start = 0.0f, end = (float)((steps - 1) / 10.0);
    const float step = (end - start) / (float)(steps - 1);
    for (int i = 0; i < steps; ++i)
        out[i] = i < steps / 2 ? fmaf(step, (float)i, start) : fmaf(-step, (float)(steps - i - 1), end);
}

namespace wcr {
struct SelWs { SelState *states; unsigned *hist; int *item_prefix; };
static SelWs carve_select(void *ws, int n)
{
    char *p = (char *)ws;
    SelWs w;
    w.states = (SelState *)p; p += (size_t)n * sizeof(SelState);
    w.hist = (unsigned *)p;   p += (size_t)n * 4096;
    w.item_prefix = (int *)p;
    return w;
}
static int run_select_passes(const SelWs &w, int n, const WcrImage *images, const uint8_t *masks, int mh,
                             int mw, int mode, float *out, cudaStream_t st)
{
    const int grid = sm_count() * 8;
    for (int pass = 0; pass < 4; ++pass) {
        // the median follows one rank; two quantiles follow up to four once their prefixes part
        if (mode == 0 || pass == 0)
            sel_pass_kernel<1><<<grid, SEL_THREADS, 0, st>>>(w.states, w.item_prefix, n, w.hist, images, masks,
                                                            mh, mw, pass, mode, out);
        else
            sel_pass_kernel<4><<<grid, SEL_THREADS, 0, st>>>(w.states, w.item_prefix, n, w.hist, images, masks,
                                                            mh, mw, pass, mode, out);
        WCR_CHECK_LAUNCH("sel_pass_kernel");
    }
    return WCR_OK;
}
// lower median of the depth patch under every box -> out_gt / out_inside
static int launch_box_median(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                             const float *boxes, float *out_gt, int *out_inside, void *ws, cudaStream_t st)
{
    const SelWs w = carve_select(ws, n);
    cudaError_t e = cudaMemsetAsync(w.hist, 0, (size_t)n * 4096, st);
    if (e != cudaSuccess) { set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    sel_init_median_kernel<<<1, 1024, 0, st>>>(images, img_off, n_images, n, boxes, out_gt, out_inside,
                                              w.states, w.item_prefix);
    WCR_CHECK_LAUNCH("sel_init_median_kernel");
    return run_select_passes(w, n, images, nullptr, 0, 0, 0, out_gt, st);
}
// q90 - q10 of the depth under every box's mask -> out_range
static int launch_mask_quantile(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                                const uint8_t *masks, int mh, int mw, const int32_t *mask_idx,
                                const float *gt_boxes, float *out_range, float *stats, void *ws, cudaStream_t st)
{
    const SelWs w = carve_select(ws, n);
    cudaError_t e = cudaMemsetAsync(ws, 0, (size_t)n * (sizeof(SelState) + 4096), st);
    if (e != cudaSuccess) { set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    const dim3 grid((mh * mw + SEL_COUNT_CHUNK - 1) / SEL_COUNT_CHUNK, n);
    mask_count_kernel<<<grid, 256, 0, st>>>(masks, mh, mw, mask_idx, w.states);
    WCR_CHECK_LAUNCH("mask_count_kernel");
    sel_init_quantile_kernel<<<1, 1024, 0, st>>>(images, img_off, n_images, n, mh, mw, mask_idx, gt_boxes,
                                                out_range, stats, w.states, w.item_prefix);
    WCR_CHECK_LAUNCH("sel_init_quantile_kernel");
    return run_select_passes(w, n, images, masks, mh, mw, 1, out_range, st);
}
}  // namespace wcr

extern "C" int64_t wcr_select_ws_bytes(int n)
{
    return n < 0 ? 0 : (int64_t)n * (int64_t)(sizeof(wcr::SelState) + 4096) + (int64_t)(n + 1) * 4 + 256;
}

extern "C" int wcr_weak_loss_fwd(const WcrLossIO *h_io, const WcrLossParams *h_params, void *stream)
{
    WCR_CHECK_DEVICE();
    int rc = check_io(h_io, h_params, false);
    if (rc != WCR_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(h_io->stats, 0, WCR_LOSS_STATS * sizeof(float), st);
    if (e != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    if (h_params->loss_flags & WCR_LOSS_POSE_GROUND2) {
        wcr::ground_rot_kernel<<<1, 256, 0, st>>>(h_io->normals, h_io->img_off, h_io->n_images, h_io->gt_rot);
        WCR_CHECK_LAUNCH("ground_rot_kernel");
    }
    if (h_io->n > 0 && (h_params->loss_flags & WCR_LOSS_DEPTH)) {
        // masks of another shape than their depth maps (padded batch tensors): the caller hands in
        // the depth maps resized to the mask shape (wcr_resize_bilinear_ac, roi_heads.py:672)
        rc = wcr::launch_mask_quantile(h_io->depth_images ? h_io->depth_images : h_io->images,
                                       h_io->img_off, h_io->n_images, h_io->n, h_io->masks,
                                       h_io->mask_h, h_io->mask_w, h_io->mask_idx, h_io->gt_boxes,
                                       h_io->depth_gt, h_io->stats, h_io->select_ws, st);
        if (rc != WCR_OK) return rc;
    }
    if (h_io->n > 0) {
        wcr::ZSteps zs;
        wcr_z_loss_steps(zs.v);
        wcr::loss_box_fwd_kernel<<<(h_io->n + 127) / 128, 128, 0, st>>>(*h_io, *h_params, zs);
        WCR_CHECK_LAUNCH("loss_box_fwd_kernel");
    }
    if (h_io->n > 0 && (h_params->loss_flags & WCR_LOSS_Z_PATCH)) {
        // roi_heads.py:1010-1011: pseudo depth = median of the depth patch under the PROJECTED box
        rc = wcr::launch_box_median(h_io->images, h_io->img_off, h_io->n_images, h_io->n, h_io->proj,
                                    h_io->z_gt, h_io->z_inside, h_io->select_ws, st);
        if (rc != WCR_OK) return rc;
        wcr::box_z_loss_kernel<<<(h_io->n_images + 127) / 128, 128, 0, st>>>(
            h_io->img_off, h_io->n_images, h_io->z_gt, h_io->z_inside, h_io->z_dec,
            h_io->terms + (size_t)(WCR_T_PSEUDO_Z - 1) * h_io->n, h_io->z_sign);
        WCR_CHECK_LAUNCH("box_z_loss_kernel");
    }
    if (h_io->n > 0 && (h_params->loss_flags & WCR_LOSS_SEG)) {
        wcr::seg_loss_kernel<false><<<h_io->n, 256, 0, st>>>(
            *h_io, *h_params, nullptr, h_io->terms + (size_t)(WCR_T_SEG - 1) * h_io->n, nullptr);
        WCR_CHECK_LAUNCH("seg_loss_kernel");
    }
    wcr::loss_pose_fwd_kernel<<<h_io->n_images, 256, 0, st>>>(*h_io);
    WCR_CHECK_LAUNCH("loss_pose_fwd_kernel");
    wcr::loss_reduce_kernel<<<1, 256, 0, st>>>(*h_io, *h_params);
    WCR_CHECK_LAUNCH("loss_reduce_kernel");
    return WCR_OK;
}

extern "C" int wcr_masked_depth_range(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                                      const uint8_t *masks, int mask_h, int mask_w, const int32_t *mask_idx,
                                      const float *gt_boxes, float *out_range, float *stats, void *select_ws,
                                      void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && img_off && masks && mask_idx && gt_boxes && out_range && stats && select_ws,
                "wcr_masked_depth_range: null pointer");
    WCR_REQUIRE(n >= 0 && n_images >= 1 && mask_h > 0 && mask_w > 0, "wcr_masked_depth_range: bad sizes");
    cudaError_t e = cudaMemsetAsync(stats, 0, WCR_LOSS_STATS * sizeof(float), (cudaStream_t)stream);
    if (e != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    if (n == 0) return WCR_OK;
    return wcr::launch_mask_quantile(images, img_off, n_images, n, masks, mask_h, mask_w, mask_idx, gt_boxes,
                                     out_range, stats, select_ws, (cudaStream_t)stream);
}

extern "C" int wcr_weak_loss_bwd(const WcrLossIO *h_io, const WcrLossParams *h_params, void *stream)
{
    WCR_CHECK_DEVICE();
    int rc = check_io(h_io, h_params, true);
    if (rc != WCR_OK) return rc;
    if (h_io->n == 0) return WCR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (h_params->loss_flags & WCR_LOSS_POSE_ALIGN) {
        wcr::loss_pose_bwd_kernel<<<h_io->n_images, 128, 0, st>>>(*h_io, *h_params);
        WCR_CHECK_LAUNCH("loss_pose_bwd_kernel");
    }
    if (h_params->loss_flags & WCR_LOSS_SEG) {
        wcr::seg_loss_kernel<true><<<h_io->n, 256, 0, st>>>(*h_io, *h_params, nullptr, nullptr, h_io->seg_dcorner);
        WCR_CHECK_LAUNCH("seg_loss_kernel");
    }
    wcr::loss_box_bwd_kernel<<<(h_io->n + 127) / 128, 128, 0, st>>>(*h_io, *h_params);
    WCR_CHECK_LAUNCH("loss_box_bwd_kernel");
    return WCR_OK;
}

// ============================================================================
// Stand-alone loss helpers behind the reference's method names (roi_heads.py:445
// pose_loss, :575 pseudo_gt_z_box_loss, :612 dim_loss, :635 pseudo_gt_z_point_loss):
// one forward and one backward launch each, for callers that do not use the fused
// block above. Same device functions, same rounding order.
// ============================================================================
namespace wcr {

// out[0] = sum_images L_im / (fail + 1), out[1] = fail, out[2] = n_images  (:451-464)
__global__ void __launch_bounds__(256)
pose_reduce_kernel(const int32_t *__restrict__ img_off, int n_images, const float *__restrict__ img_pose,
                   float *__restrict__ out)
{
    __shared__ float s_a[8], s_b[8];
    float psum = 0.f, fail = 0.f;
    for (int im = threadIdx.x; im < n_images; im += blockDim.x) {
        const int ni = img_off[im + 1] - img_off[im];
        if (ni == 1) fail += 1.f; else psum += img_pose[im];
    }
    block_sum2(psum, fail, s_a, s_b);
    if (threadIdx.x == 0) { out[0] = psum * 1.f / (fail + 1.f); out[1] = fail; out[2] = (float)n_images; }
}

// d loss / d R_i = grad / (npairs_im * (fail + 1)) * sum_{j != i} -sign(cos_ij) / 2 * R_j
__global__ void __launch_bounds__(128)
pose_bwd_kernel(const float *__restrict__ R, const int32_t *__restrict__ img_off,
                const float *__restrict__ fwd_out, const float *__restrict__ grad, float *__restrict__ d_R)
{
    const int im = blockIdx.x;
    const int b0 = img_off[im], ni = img_off[im + 1] - b0;
    const float npairs = 0.5f * (float)ni * (float)(ni - 1);
    const float scale = ni > 1 ? grad[0] / (npairs * (fwd_out[1] + 1.f)) : 0.f;
    for (int i = threadIdx.x; i < ni; i += blockDim.x) {
        float Ri[9], g[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) { Ri[t] = R[9 * (size_t)(b0 + i) + t]; g[t] = 0.f; }
        for (int j = 0; j < ni; ++j) {
            if (j == i) continue;
            const float *Rj = R + 9 * (size_t)(b0 + j);
            const float dc = -sgn(pair_cos(Ri, Rj)) * scale * 0.5f;
#pragma unroll
            for (int t = 0; t < 9; ++t) g[t] += dc * Rj[t];
        }
#pragma unroll
        for (int t = 0; t < 9; ++t) d_R[9 * (size_t)(b0 + i) + t] = g[t];
    }
}

// |pred_z - depth[clamp(y), clamp(x)]| per box (:635-656); sign kept for the backward.
__global__ void __launch_bounds__(128)
pseudo_z_point_kernel(const WcrImage *__restrict__ images, const int32_t *__restrict__ img_off,
                      int n_images, int n, const float *__restrict__ pred_xy,
                      const float *__restrict__ pred_z, float *__restrict__ out_loss,
                      float *__restrict__ out_sign)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const WcrImage im = images[image_of(img_off, n_images, i)];
    const float gt = depth_at(im, pred_xy[2 * i], pred_xy[2 * i + 1]);
    const float d = pred_z[i] - gt;
    out_loss[i] = fabsf(d);       // smooth_l1(beta = 0) == |x| (:352-353)
    out_sign[i] = sgn(d);         // nan -> 0, like torch's sign in the L1 backward
}

__global__ void __launch_bounds__(256)
l1_bwd_kernel(const float *__restrict__ sign, const float *__restrict__ grad, int n, float *__restrict__ d)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d[i] = grad[i] * sign[i];
}

// dim_loss (:612-633): max(|d - mu| / sigma - 1, 0) per axis; nan_flag counts NaN stds.
__global__ void __launch_bounds__(256)
dim_loss_kernel(const float *__restrict__ mean, const float *__restrict__ std_, const float *__restrict__ dims,
                int n, float *__restrict__ out /* [3][n] */, int *__restrict__ nan_flag)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 3 * n) return;
    const int i = idx / 3, k = idx - 3 * i;
    const float s = std_[idx];
    if (s != s) atomicAdd(nan_flag, 1);
    const float a = fabsf(dims[idx] - mean[idx]) / s - 1.0f;
    out[(size_t)k * n + i] = (a != a) ? a : fmaxf(a, 0.f);   // torch.max(a, 0) keeps NaN
}

// backward of the above: binary max splits a tie evenly, |.| has gradient sign(.)
__global__ void __launch_bounds__(256)
dim_loss_bwd_kernel(const float *__restrict__ mean, const float *__restrict__ std_,
                    const float *__restrict__ dims, const float *__restrict__ grad /* [3][n] */, int n,
                    float *__restrict__ d_dims)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 3 * n) return;
    const int i = idx / 3, k = idx - 3 * i;
    const float diff = dims[idx] - mean[idx];
    const float a = fabsf(diff) / std_[idx] - 1.0f;
    const float w = a > 0.f ? 1.f : (a == 0.f ? 0.5f : 0.f);
    d_dims[idx] = grad[(size_t)k * n + i] * w * sgn(diff) / std_[idx];
}

// ---- order-preserving integer image of a float (the key of the radix select above).
// pseudo_gt_z_box_loss (:575-610) takes the LOWER median (torch.median: element (cnt-1)/2
// of the sorted patch) and NaN when the patch holds a NaN.
__device__ __forceinline__ uint32_t float_key(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// Per image: the reference appends the in-image boxes' medians first and the
// out-of-image fallbacks after them (:595-606), then pairs that list with pred_z BY
// POSITION -- reproduced here. One thread per image (boxes per image are few).
__global__ void __launch_bounds__(128)
box_z_loss_kernel(const int32_t *__restrict__ img_off, int n_images, const float *__restrict__ gt,
                  const int *__restrict__ inside, const float *__restrict__ pred_z,
                  float *__restrict__ out_loss, float *__restrict__ out_sign)
{
    const int im = blockIdx.x * blockDim.x + threadIdx.x;
    if (im >= n_images) return;
    const int b0 = img_off[im], b1 = img_off[im + 1];
    int pos = b0;
    for (int phase = 1; phase >= 0; --phase)
        for (int b = b0; b < b1; ++b)
            if (inside[b] == phase) {
                const float d = pred_z[pos] - gt[b];
                out_loss[pos] = fabsf(d);
                out_sign[pos] = sgn(d);
                ++pos;
            }
}

}  // namespace wcr

extern "C" int wcr_pose_loss_fwd(const float *R, const int32_t *img_off, int n_images, float *img_pose,
                                 float *out3, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(R && img_off && img_pose && out3, "wcr_pose_loss_fwd: null pointer");
    WCR_REQUIRE(n_images >= 1, "wcr_pose_loss_fwd: needs at least one image");
    WcrLossIO io;
    memset(&io, 0, sizeof(io));
    io.n_images = n_images; io.R_ego = const_cast<float *>(R); io.img_off = img_off; io.img_pose = img_pose;
    wcr::loss_pose_fwd_kernel<<<n_images, 256, 0, (cudaStream_t)stream>>>(io);
    WCR_CHECK_LAUNCH("loss_pose_fwd_kernel");
    wcr::pose_reduce_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(img_off, n_images, img_pose, out3);
    WCR_CHECK_LAUNCH("pose_reduce_kernel");
    return WCR_OK;
}

extern "C" int wcr_pose_loss_bwd(const float *R, const int32_t *img_off, int n_images,
                                 const float *fwd_out3, const float *grad1, float *d_R, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(R && img_off && fwd_out3 && grad1 && d_R, "wcr_pose_loss_bwd: null pointer");
    WCR_REQUIRE(n_images >= 1, "wcr_pose_loss_bwd: needs at least one image");
    wcr::pose_bwd_kernel<<<n_images, 128, 0, (cudaStream_t)stream>>>(R, img_off, fwd_out3, grad1, d_R);
    WCR_CHECK_LAUNCH("pose_bwd_kernel");
    return WCR_OK;
}

extern "C" int wcr_pseudo_z_point_loss(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                                       const float *pred_xy, const float *pred_z, float *out_loss,
                                       float *out_sign, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && img_off && pred_xy && pred_z && out_loss && out_sign,
                "wcr_pseudo_z_point_loss: null pointer");
    WCR_REQUIRE(n_images >= 1 && n >= 0, "wcr_pseudo_z_point_loss: bad sizes");
    if (n == 0) return WCR_OK;
    wcr::pseudo_z_point_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        images, img_off, n_images, n, pred_xy, pred_z, out_loss, out_sign);
    WCR_CHECK_LAUNCH("pseudo_z_point_kernel");
    return WCR_OK;
}

extern "C" int wcr_pseudo_z_box_loss(const WcrImage *images, const int32_t *img_off, int n_images, int n,
                                     const float *boxes, const float *pred_z, float *out_loss,
                                     float *out_sign, float *ws_gt, int32_t *ws_inside, void *select_ws,
                                     void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && img_off && boxes && pred_z && out_loss && out_sign && ws_gt && ws_inside && select_ws,
                "wcr_pseudo_z_box_loss: null pointer");
    WCR_REQUIRE(n_images >= 1 && n >= 0, "wcr_pseudo_z_box_loss: bad sizes");
    if (n == 0) return WCR_OK;
    int rc = wcr::launch_box_median(images, img_off, n_images, n, boxes, ws_gt, ws_inside, select_ws,
                                    (cudaStream_t)stream);
    if (rc != WCR_OK) return rc;
    wcr::box_z_loss_kernel<<<(n_images + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        img_off, n_images, ws_gt, ws_inside, pred_z, out_loss, out_sign);
    WCR_CHECK_LAUNCH("box_z_loss_kernel");
    return WCR_OK;
}

extern "C" int wcr_l1_loss_bwd(const float *sign, const float *grad, int n, float *d_pred, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(sign && grad && d_pred, "wcr_l1_loss_bwd: null pointer");
    WCR_REQUIRE(n >= 0, "wcr_l1_loss_bwd: bad size");
    if (n == 0) return WCR_OK;
    wcr::l1_bwd_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(sign, grad, n, d_pred);
    WCR_CHECK_LAUNCH("l1_bwd_kernel");
    return WCR_OK;
}

extern "C" int wcr_dim_loss_fwd(const float *prior_mean, const float *prior_std, const float *dims, int n,
                                float *out3n, int32_t *nan_flag, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(prior_mean && prior_std && dims && out3n && nan_flag, "wcr_dim_loss_fwd: null pointer");
    WCR_REQUIRE(n >= 0, "wcr_dim_loss_fwd: bad size");
    cudaError_t e = cudaMemsetAsync(nan_flag, 0, sizeof(int32_t), (cudaStream_t)stream);
    if (e != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    if (n == 0) return WCR_OK;
    wcr::dim_loss_kernel<<<(3 * n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(prior_mean, prior_std, dims, n,
                                                                              out3n, nan_flag);
    WCR_CHECK_LAUNCH("dim_loss_kernel");
    return WCR_OK;
}

extern "C" int wcr_dim_loss_bwd(const float *prior_mean, const float *prior_std, const float *dims,
                                const float *grad3n, int n, float *d_dims, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(prior_mean && prior_std && dims && grad3n && d_dims, "wcr_dim_loss_bwd: null pointer");
    WCR_REQUIRE(n >= 0, "wcr_dim_loss_bwd: bad size");
    if (n == 0) return WCR_OK;
    wcr::dim_loss_bwd_kernel<<<(3 * n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(prior_mean, prior_std, dims,
                                                                                  grad3n, n, d_dims);
    WCR_CHECK_LAUNCH("dim_loss_bwd_kernel");
    return WCR_OK;
}

extern "C" int wcr_segment_loss_fwd(const uint8_t *masks, int mask_h, int mask_w, const int32_t *mask_idx,
                                    const float *corners, int n, float *out_loss, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(masks && mask_idx && corners && out_loss, "wcr_segment_loss_fwd: null pointer");
    WCR_REQUIRE(n >= 0 && mask_h > 0 && mask_w > 0, "wcr_segment_loss_fwd: bad sizes");
    if (n == 0) return WCR_OK;
    WcrLossIO io;
    WcrLossParams P;
    memset(&io, 0, sizeof(io));
    memset(&P, 0, sizeof(P));
    io.n = n; io.masks = masks; io.mask_h = mask_h; io.mask_w = mask_w; io.mask_idx = mask_idx;
    wcr::seg_loss_kernel<false><<<n, 256, 0, (cudaStream_t)stream>>>(io, P, corners, out_loss, nullptr);
    WCR_CHECK_LAUNCH("seg_loss_kernel");
    return WCR_OK;
}

extern "C" int wcr_segment_loss_bwd(const uint8_t *masks, int mask_h, int mask_w, const int32_t *mask_idx,
                                    const float *corners, int n, float *d_corners, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(masks && mask_idx && corners && d_corners, "wcr_segment_loss_bwd: null pointer");
    WCR_REQUIRE(n >= 0 && mask_h > 0 && mask_w > 0, "wcr_segment_loss_bwd: bad sizes");
    if (n == 0) return WCR_OK;
    WcrLossIO io;
    WcrLossParams P;
    memset(&io, 0, sizeof(io));
    memset(&P, 0, sizeof(P));
    io.n = n; io.masks = masks; io.mask_h = mask_h; io.mask_w = mask_w; io.mask_idx = mask_idx;
    wcr::seg_loss_kernel<true><<<n, 256, 0, (cudaStream_t)stream>>>(io, P, corners, nullptr, d_corners);
    WCR_CHECK_LAUNCH("seg_loss_kernel");
    return WCR_OK;
}
