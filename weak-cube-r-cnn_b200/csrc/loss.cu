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

__global__ void box_median_kernel(const WcrImage *__restrict__ images, const int32_t *__restrict__ img_off,
                                  int n_images, const float *__restrict__ boxes, float *__restrict__ out_gt,
                                  int *__restrict__ out_inside);
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

// ---- 'depth': q90 - q10 of the depth under the object's mask (roi_heads.py:669-679) ----------
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

// One CTA per box: the four order statistics torch.quantile(x, 0.9) and torch.quantile(x, 0.1)
// interpolate between, found by ONE 4-pass 8-bit radix select that follows up to four key
// prefixes at once (a histogram per distinct prefix), over either the masked pixels of the
// box's image or -- when the mask is empty -- the GT box patch. No sort, no compaction buffer:
// the reference sorts every masked depth vector twice per box.
__global__ void __launch_bounds__(256)
mask_quantile_kernel(const WcrImage *__restrict__ images, const int32_t *__restrict__ img_off, int n_images,
                     const uint8_t *__restrict__ masks, int mh, int mw, const int32_t *__restrict__ mask_idx,
                     const float *__restrict__ gt_boxes, float *__restrict__ out_range, float *__restrict__ stats)
{
    __shared__ unsigned hist[4][256];
    __shared__ unsigned s_prefix[4], s_rank[4], s_gprefix[4];
    __shared__ int s_group[4], s_ngroups, s_cnt, s_rmin, s_rmax, s_nan;
    const int b = blockIdx.x, tid = threadIdx.x;
    const WcrImage im = images[image_of(img_off, n_images, b)];
    if (im.H != mh || im.W != mw) {
        if (tid == 0) { out_range[b] = nanf_(); stats[WCR_STAT_SHAPE_ERR] = 1.f; }
        return;
    }
    const uint8_t *m = masks + (size_t)mask_idx[b] * mh * mw;
    if (tid == 0) { s_cnt = 0; s_rmin = mh; s_rmax = -1; s_nan = 0; }
    __syncthreads();
    // -- pass A: how many pixels, which rows ----------------------------------------------
    {
        int cnt = 0, rmin = mh, rmax = -1;
        const int total = mh * mw;
        const int head = min(total, (int)((16 - ((uintptr_t)m & 15)) & 15));
        const int nvec = (total - head) / 16;
        auto visit = [&](int idx) { ++cnt; const int r = idx / mw; rmin = min(rmin, r); rmax = max(rmax, r); };
        for (int i = tid; i < head; i += blockDim.x) if (m[i]) visit(i);
        for (int i = head + 16 * nvec + tid; i < total; i += blockDim.x) if (m[i]) visit(i);
        const uint4 *mv = reinterpret_cast<const uint4 *>(m + head);
        for (int v = tid; v < nvec; v += blockDim.x) {
            const uint4 w = __ldg(mv + v);
            if ((w.x | w.y | w.z | w.w) == 0u) continue;
            const unsigned ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
                for (int e = 0; e < 4; ++e)
                    if ((ww[q] >> (8 * e)) & 255u) visit(head + 16 * v + 4 * q + e);
        }
        if (cnt) { atomicAdd(&s_cnt, cnt); atomicMin(&s_rmin, rmin); atomicMax(&s_rmax, rmax); }
    }
    __syncthreads();
    int cnt = s_cnt;
    const bool use_mask = cnt > 0;
    // source region: rows [r0, r1) x cols [c0, c1) of the depth map, masked or not
    int r0 = s_rmin, r1 = s_rmax + 1, c0 = 0, c1 = mw;
    if (!use_mask) {   // segmentation failed: depth_map[y1.long():y2.long(), x1.long():x2.long()] (:674-675)
        const float *g = gt_boxes + 4 * b;
        c0 = py_slice_index((long long)g[0], mw); c1 = py_slice_index((long long)g[2], mw);
        r0 = py_slice_index((long long)g[1], mh); r1 = py_slice_index((long long)g[3], mh);
        cnt = max(r1 - r0, 0) * max(c1 - c0, 0);
    }
    if (cnt <= 0) {    // torch.quantile of an empty tensor raises in the reference; NaN here
        if (tid == 0) out_range[b] = nanf_();
        return;
    }
    float w9, w1;
    {
        const float r9 = 0.9f * (float)(cnt - 1), r1f = 0.1f * (float)(cnt - 1);
        const float b9 = floorf(r9), b1 = floorf(r1f);
        w9 = r9 - b9; w1 = r1f - b1;
        if (tid == 0) {
            s_rank[0] = (unsigned)b9; s_rank[1] = (unsigned)ceilf(r9);
            s_rank[2] = (unsigned)b1; s_rank[3] = (unsigned)ceilf(r1f);
#pragma unroll
            for (int t = 0; t < 4; ++t) { s_prefix[t] = 0u; s_group[t] = 0; }
            s_gprefix[0] = 0u; s_ngroups = 1;
        }
    }
    const int ncols = c1 - c0, region = (r1 - r0) * ncols;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        for (int i = tid; i < 4 * 256; i += blockDim.x) (&hist[0][0])[i] = 0u;
        __syncthreads();
        const int ng = s_ngroups;
        const unsigned gp0 = s_gprefix[0], gp1 = s_gprefix[1], gp2 = s_gprefix[2], gp3 = s_gprefix[3];
        const unsigned himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
        // per-thread run-length encoding per group: depths under one mask share their leading
        // bytes, plain shared atomics would serialise on one or two bins
        unsigned last[4] = {~0u, ~0u, ~0u, ~0u}, run[4] = {0u, 0u, 0u, 0u};
        auto feed = [&](float f) {
            if (pass == 0 && f != f) s_nan = 1;
            const uint32_t k = float_key(f);
            const unsigned hi = k & himask, bin = (k >> shift) & 255u;
            const unsigned gp[4] = {gp0, gp1, gp2, gp3};
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                if (g >= ng || hi != gp[g]) continue;
                if (bin == last[g]) { ++run[g]; continue; }
                if (run[g]) atomicAdd(&hist[g][last[g]], run[g]);
                last[g] = bin; run[g] = 1u;
            }
        };
        for (int i = tid; i < region; i += blockDim.x) {
            const int r = i / ncols, c = i - r * ncols;
            const int rr = r0 + r, cc = c0 + c;
            if (use_mask && !m[(size_t)rr * mw + cc]) continue;
            feed(__ldg(im.depth + (size_t)rr * im.depth_pitch + cc));
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) if (run[g]) atomicAdd(&hist[g][last[g]], run[g]);
        __syncthreads();
        if (tid == 0) {
            for (int t = 0; t < 4; ++t) {
                const unsigned *h = hist[s_group[t]];
                unsigned rank = s_rank[t], acc = 0u;
                int bin = 0;
                for (; bin < 255; ++bin) {
                    if (acc + h[bin] > rank) break;
                    acc += h[bin];
                }
                s_rank[t] = rank - acc;
                s_prefix[t] |= (unsigned)bin << shift;
            }
            int g_n = 0;
            for (int t = 0; t < 4; ++t) {
                int g = 0;
                for (; g < g_n; ++g) if (s_gprefix[g] == s_prefix[t]) break;
                if (g == g_n) s_gprefix[g_n++] = s_prefix[t];
                s_group[t] = g;
            }
            s_ngroups = g_n;
        }
        __syncthreads();
    }
    if (tid == 0) {
        const float q9 = torch_lerp(key_float(s_prefix[0]), key_float(s_prefix[1]), w9);
        const float q1 = torch_lerp(key_float(s_prefix[2]), key_float(s_prefix[3]), w1);
        out_range[b] = s_nan ? nanf_() : q9 - q1;
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
                const float gp = g * dproj[e];
                // p = (f*A + c0*Z)/Z : dA = gp*f/Z ; dZ = gp*(c0/Z - num/Z^2)
                const float f = isx ? d.K.fx : d.K.fy, c0 = isx ? d.K.cx : d.K.cy;
                const float Av = isx ? X[c] : Y[c];
                const float num = fmaf(c0, Z[c], f * Av);
                const float dA = gp * f / Z[c];
                const float dZ = gp * c0 / Z[c] - gp * num / (Z[c] * Z[c]);
                float sx, sy, sz;
                corner_signs(c, sx, sy, sz);
                const float l[3] = {sx * hx, sy * hy, sz * hz};
                const int ra = isx ? 0 : 1;                     // row of R feeding A
                // corner = R l + centre
                if (isx) dcx += 0.f;                             // (centre grads below)
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    dR[3 * ra + k] += dA * l[k];
                    dR[6 + k] += dZ * l[k];
                    dl[k] += (d.R[3 * ra + k] * dA + d.R[6 + k] * dZ) * (k == 0 ? sx : (k == 1 ? sy : sz));
                }
                // centre: cube[0] or cube[1] gets dA, cube[2] gets dZ
                const float zc = d.z;
                if (isx) {   // x3d = z*(cube_x-cx)/fx
                    dcx += dA * zc / d.K.fx;
                    dz += dA * (d.cube_x - d.K.cx) / d.K.fx;
                } else {
                    dcy += dA * zc / d.K.fy;
                    dz += dA * (d.cube_y - d.K.cy) / d.K.fy;
                }
                dz += dZ;
            }
            ddims[2] += 0.5f * dl[0];   // hx = l/2
            ddims[1] += 0.5f * dl[1];   // hy = h/2
            ddims[0] += 0.5f * dl[2];   // hz = w/2
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
        wcr::mask_quantile_kernel<<<h_io->n, 256, 0, st>>>(h_io->images, h_io->img_off, h_io->n_images,
                                                           h_io->masks, h_io->mask_h, h_io->mask_w,
                                                           h_io->mask_idx, h_io->gt_boxes, h_io->depth_gt,
                                                           h_io->stats);
        WCR_CHECK_LAUNCH("mask_quantile_kernel");
    }
    if (h_io->n > 0) {
        wcr::ZSteps zs;
        wcr_z_loss_steps(zs.v);
        wcr::loss_box_fwd_kernel<<<(h_io->n + 127) / 128, 128, 0, st>>>(*h_io, *h_params, zs);
        WCR_CHECK_LAUNCH("loss_box_fwd_kernel");
    }
    if (h_io->n > 0 && (h_params->loss_flags & WCR_LOSS_Z_PATCH)) {
        // roi_heads.py:1010-1011: pseudo depth = median of the depth patch under the PROJECTED box
        wcr::box_median_kernel<<<h_io->n, 256, 0, st>>>(h_io->images, h_io->img_off, h_io->n_images,
                                                        h_io->proj, h_io->z_gt, h_io->z_inside);
        WCR_CHECK_LAUNCH("box_median_kernel");
        wcr::box_z_loss_kernel<<<(h_io->n_images + 127) / 128, 128, 0, st>>>(
            h_io->img_off, h_io->n_images, h_io->z_gt, h_io->z_inside, h_io->z_dec,
            h_io->terms + (size_t)(WCR_T_PSEUDO_Z - 1) * h_io->n, h_io->z_sign);
        WCR_CHECK_LAUNCH("box_z_loss_kernel");
    }
    wcr::loss_pose_fwd_kernel<<<h_io->n_images, 256, 0, st>>>(*h_io);
    WCR_CHECK_LAUNCH("loss_pose_fwd_kernel");
    wcr::loss_reduce_kernel<<<1, 256, 0, st>>>(*h_io, *h_params);
    WCR_CHECK_LAUNCH("loss_reduce_kernel");
    return WCR_OK;
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

// ---- pseudo_gt_z_box_loss (:575-610): lower median of the depth patch under each
// predicted 2D box. One CTA per box: 4-pass radix select (8 bits per pass, shared
// histogram) over the order-preserving integer image of the floats; the patch is
// re-read from L2 each pass. torch.median returns the LOWER median (element
// (cnt-1)/2 of the sorted patch) and NaN when the patch holds a NaN.
__device__ __forceinline__ uint32_t float_key(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k)
{
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__global__ void __launch_bounds__(256)
box_median_kernel(const WcrImage *__restrict__ images, const int32_t *__restrict__ img_off, int n_images,
                  const float *__restrict__ boxes /* [n][4] */, float *__restrict__ out_gt,
                  int *__restrict__ out_inside)
{
    __shared__ unsigned hist[256];
    __shared__ unsigned s_prefix, s_rank;
    __shared__ int s_nan;
    const int b = blockIdx.x;
    const WcrImage im = images[image_of(img_off, n_images, b)];
    // Boxes.clip((h, w)), area() > 0 on the clipped float box, then .long() slicing
    const float x1 = clamp_keep_nan(boxes[4 * b + 0], 0.f, (float)im.W), y1 = clamp_keep_nan(boxes[4 * b + 1], 0.f, (float)im.H);
    const float x2 = clamp_keep_nan(boxes[4 * b + 2], 0.f, (float)im.W), y2 = clamp_keep_nan(boxes[4 * b + 3], 0.f, (float)im.H);
    const bool inside = (x2 - x1) * (y2 - y1) > 0.f;
    if (x1 != x1 || y1 != y1 || x2 != x2 || y2 != y2) {
        // a NaN box (cube behind the camera): the reference indexes with an undefined
        // .long() here; defined as "no pseudo depth" (NaN, masked by the safe mean)
        if (threadIdx.x == 0) { out_gt[b] = nanf_(); out_inside[b] = 0; }
        return;
    }
    if (!inside) {
        // fallback of :600-606: centre lookup of the CLIPPED box with the 10-px clamp
        if (threadIdx.x == 0) {
            out_gt[b] = depth_at(im, (x1 + x2) / 2.f, (y1 + y2) / 2.f);
            out_inside[b] = 0;
        }
        return;
    }
    const int u0 = (int)x1, u1 = (int)x2, v0 = (int)y1, v1 = (int)y2;
    const int ncols = u1 - u0, cnt = ncols * (v1 - v0);
    if (cnt <= 0) {  // torch.median of an empty slice raises in the reference; NaN here
        if (threadIdx.x == 0) { out_gt[b] = nanf_(); out_inside[b] = 1; }
        return;
    }
    if (threadIdx.x == 0) { s_prefix = 0u; s_rank = (unsigned)((cnt - 1) / 2); s_nan = 0; }
    const float inv_ncols = 1.0f / (float)ncols;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        hist[threadIdx.x] = 0u;
        __syncthreads();
        const unsigned prefix = s_prefix;
        const unsigned himask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
        // Depth values of a patch share their leading digits, so a plain shared-memory
        // histogram serialises on one or two bins (2.5 ms for 1156 boxes, measured). Each
        // thread run-length encodes its own stream instead: one atomic per change of bin.
        unsigned last = 0xffffffffu, run = 0u;
        // 4 independent loads in flight per thread (a clipped projected box can cover the
        // whole image: one CTA then streams 307 k pixels per pass and is latency bound)
        for (int i0 = threadIdx.x; i0 < cnt; i0 += 1024) {
            float f[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int idx = i0 + 256 * q;
                int r = __float2int_rz(((float)idx + 0.5f) * inv_ncols), c = idx - r * ncols;
                if (c < 0) { --r; c += ncols; } else if (c >= ncols) { ++r; c -= ncols; }
                f[q] = idx < cnt ? __ldg(im.depth + (size_t)(v0 + r) * im.depth_pitch + u0 + c) : 0.f;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (i0 + 256 * q >= cnt) break;
                if (pass == 0 && f[q] != f[q]) s_nan = 1;
                const uint32_t k = float_key(f[q]);
                if ((k & himask) != prefix) continue;
                const unsigned bin = (k >> shift) & 255u;
                if (bin == last) { ++run; continue; }
                if (run) atomicAdd(&hist[last], run);
                last = bin; run = 1u;
            }
        }
        if (run) atomicAdd(&hist[last], run);
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned rank = s_rank, acc = 0u;
            int bin = 0;
            for (; bin < 255; ++bin) {
                if (acc + hist[bin] > rank) break;
                acc += hist[bin];
            }
            s_rank = rank - acc;
            s_prefix = prefix | ((unsigned)bin << shift);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        out_gt[b] = s_nan ? nanf_() : key_float(s_prefix);
        out_inside[b] = 1;
    }
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
                                     float *out_sign, float *ws_gt, int32_t *ws_inside, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && img_off && boxes && pred_z && out_loss && out_sign && ws_gt && ws_inside,
                "wcr_pseudo_z_box_loss: null pointer");
    WCR_REQUIRE(n_images >= 1 && n >= 0, "wcr_pseudo_z_box_loss: bad sizes");
    if (n == 0) return WCR_OK;
    wcr::box_median_kernel<<<n, 256, 0, (cudaStream_t)stream>>>(images, img_off, n_images, boxes, ws_gt, ws_inside);
    WCR_CHECK_LAUNCH("box_median_kernel");
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
