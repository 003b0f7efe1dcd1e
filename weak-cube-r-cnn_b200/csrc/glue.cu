// glue.cu -- the callers and data formats either side of the hot path.
//
//   rotation_6d_to_matrix  cube_head.py:174-176 (pytorch3d: Gram-Schmidt of two 3-vectors), fwd + bwd
//   class gather           roi_heads.py:823-840 (x[fg_inds, box_classes, ...]), fwd + bwd
//   per-box constants      roi_heads.py:748-781 (Ks_scaled_per_box, virtual_to_real, im_sizes)
//   NEAREST resize         dataset_mapper.py:126-145 (PIL Image.resize(..., Image.NEAREST) of the
//                          depth map and the ground mask to the augmented image size)
//
// All latency-bound helper kernels (a few KB to a few MB); no roofline claim.
#include <math.h>
#include "common.cuh"

namespace wcr {

// ---- rotation_6d_to_matrix --------------------------------------------------------------
// b1 = a1 / max(|a1|, eps); u = a2 - (b1 . a2) b1; b2 = u / max(|u|, eps); b3 = b1 x b2
__global__ void __launch_bounds__(128)
rot6d_fwd_kernel(const float *__restrict__ d6, int n, float *__restrict__ R)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *a = d6 + 6 * (size_t)i;
    const float a1[3] = {a[0], a[1], a[2]}, a2[3] = {a[3], a[4], a[5]};
    const float n1 = fmaxf(sqrtf((a1[0] * a1[0] + a1[1] * a1[1]) + a1[2] * a1[2]), 1e-12f);
    const float b1[3] = {a1[0] / n1, a1[1] / n1, a1[2] / n1};
    const float d = (b1[0] * a2[0] + b1[1] * a2[1]) + b1[2] * a2[2];
    const float u[3] = {a2[0] - d * b1[0], a2[1] - d * b1[1], a2[2] - d * b1[2]};
    const float n2 = fmaxf(sqrtf((u[0] * u[0] + u[1] * u[1]) + u[2] * u[2]), 1e-12f);
    const float b2[3] = {u[0] / n2, u[1] / n2, u[2] / n2};
    float *o = R + 9 * (size_t)i;
    o[0] = b1[0]; o[1] = b1[1]; o[2] = b1[2];
    o[3] = b2[0]; o[4] = b2[1]; o[5] = b2[2];
    o[6] = fmaf(b1[1], b2[2], -(b1[2] * b2[1]));
    o[7] = fmaf(b1[2], b2[0], -(b1[0] * b2[2]));
    o[8] = fmaf(b1[0], b2[1], -(b1[1] * b2[0]));
}

__device__ __forceinline__ void cross3(const float *a, const float *b, float *o)
{
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ float dot3(const float *a, const float *b)
{
    return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2];
}

__global__ void __launch_bounds__(128)
rot6d_bwd_kernel(const float *__restrict__ d6, const float *__restrict__ gR, int n, float *__restrict__ gd6)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *a = d6 + 6 * (size_t)i;
    const float *G = gR + 9 * (size_t)i;
    const float a1[3] = {a[0], a[1], a[2]}, a2[3] = {a[3], a[4], a[5]};
    const float r1 = sqrtf(dot3(a1, a1)), n1 = fmaxf(r1, 1e-12f);
    const float b1[3] = {a1[0] / n1, a1[1] / n1, a1[2] / n1};
    const float d = dot3(b1, a2);
    const float u[3] = {a2[0] - d * b1[0], a2[1] - d * b1[1], a2[2] - d * b1[2]};
    const float r2 = sqrtf(dot3(u, u)), n2 = fmaxf(r2, 1e-12f);
    const float b2[3] = {u[0] / n2, u[1] / n2, u[2] / n2};
    const float G1[3] = {G[0], G[1], G[2]}, G2[3] = {G[3], G[4], G[5]}, G3[3] = {G[6], G[7], G[8]};
    float gb1[3], gb2[3], t[3];
    cross3(b2, G3, t);                 // b3 = b1 x b2 : d/db1 = b2 x G3, d/db2 = G3 x b1
#pragma unroll
    for (int k = 0; k < 3; ++k) gb1[k] = G1[k] + t[k];
    cross3(G3, b1, t);
#pragma unroll
    for (int k = 0; k < 3; ++k) gb2[k] = G2[k] + t[k];
    // b2 = u / max(|u|, eps): the clamp passes no gradient to the norm below eps
    float gu[3];
    const float s2 = r2 > 1e-12f ? dot3(b2, gb2) : 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) gu[k] = (gb2[k] - b2[k] * s2) / n2;
    // u = a2 - d b1, d = b1 . a2
    const float gd = -dot3(gu, b1);
    float ga2[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        ga2[k] = gu[k] + gd * b1[k];
        gb1[k] += -d * gu[k] + gd * a2[k];
    }
    const float s1 = r1 > 1e-12f ? dot3(b1, gb1) : 0.f;
    float *o = gd6 + 6 * (size_t)i;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        o[k] = (gb1[k] - b1[k] * s1) / n1;
        o[3 + k] = ga2[k];
    }
}

// ---- class gather: out[i][:] = src[i][cls[i]][:] ---------------------------------------------
__global__ void __launch_bounds__(256)
class_gather_kernel(const float *__restrict__ src, int n, int C, int k, const int64_t *__restrict__ cls,
                    float *__restrict__ out)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * k) return;
    const int i = idx / k, j = idx - i * k;
    const int64_t c = cls[i];
    out[idx] = (c >= 0 && c < C) ? src[((size_t)i * C + c) * k + j] : nanf_();
}
__global__ void __launch_bounds__(256)
class_scatter_kernel(const float *__restrict__ g, int n, int C, int k, const int64_t *__restrict__ cls,
                     float *__restrict__ d_src /* zeroed */)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * k) return;
    const int i = idx / k, j = idx - i * k;
    const int64_t c = cls[i];
    if (c >= 0 && c < C) d_src[((size_t)i * C + c) * k + j] = g[idx];
}

// ---- per-box constants (roi_heads.py:748-781, 810-814; math_util.py:732-743) --------------------
__global__ void __launch_bounds__(128)
box_constants_kernel(const float *__restrict__ Ks /* [B][9] */, const float *__restrict__ ratio /* [B] */,
                     const float *__restrict__ dims /* [B][2] im_current_dims */,
                     const int32_t *__restrict__ img_off, int n_images, int n, float virtual_focal,
                     int virtual_depth, float *__restrict__ K4, float *__restrict__ v2r, float *__restrict__ clamp)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int lo = 0, hi = n_images;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (img_off[mid] <= i) lo = mid; else hi = mid;
    }
    const float *K = Ks + 9 * lo;
    const float r = ratio[lo];
    K4[4 * i + 0] = K[0] / r; K4[4 * i + 1] = K[4] / r;     // Ks[i] / im_scales_ratio[i]
    K4[4 * i + 2] = K[2] / r; K4[4 * i + 3] = K[5] / r;
    const float h_net = dims[2 * lo];                       // im_current_dims[i][0]
    // compute_virtual_scale_from_focal_spaces(f = Ks[1,1], H = h_net * ratio, f0 = virtual focal,
    // H0 = h_net) = (H0 * f) / (f0 * H)                       (roi_heads.py:775-779)
    v2r[i] = virtual_depth ? (h_net * K[4]) / (virtual_focal * (h_net * r)) : 1.0f;
    clamp[2 * i] = dims[2 * lo]; clamp[2 * i + 1] = dims[2 * lo + 1];
}

// ---- NEAREST resize through PIL's coordinate tables --------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
resize_gather_kernel(const T *__restrict__ src, int in_h, int in_w, int64_t src_pitch,
                     const int32_t *__restrict__ xtab, const int32_t *__restrict__ ytab,
                     T *__restrict__ dst, int out_h, int out_w, int64_t dst_pitch)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= out_w || y >= out_h) return;
    const int xi = xtab[x], yi = ytab[y];
    T v = T(0);                                              // PIL fills uncovered pixels with 0
    if (xi >= 0 && xi < in_w && yi >= 0 && yi < in_h) v = src[(size_t)yi * src_pitch + xi];
    dst[(size_t)y * dst_pitch + x] = v;
}

}  // namespace wcr

extern "C" int wcr_rotation_6d_fwd(const float *d6, int n, float *out_R, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(d6 && out_R, "wcr_rotation_6d_fwd: null pointer");
    WCR_REQUIRE(n >= 0, "wcr_rotation_6d_fwd: bad size");
    if (n == 0) return WCR_OK;
    wcr::rot6d_fwd_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(d6, n, out_R);
    WCR_CHECK_LAUNCH("rot6d_fwd_kernel");
    return WCR_OK;
}

extern "C" int wcr_rotation_6d_bwd(const float *d6, const float *grad_R, int n, float *d_d6, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(d6 && grad_R && d_d6, "wcr_rotation_6d_bwd: null pointer");
    WCR_REQUIRE(n >= 0, "wcr_rotation_6d_bwd: bad size");
    if (n == 0) return WCR_OK;
    wcr::rot6d_bwd_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(d6, grad_R, n, d_d6);
    WCR_CHECK_LAUNCH("rot6d_bwd_kernel");
    return WCR_OK;
}

extern "C" int wcr_class_gather_fwd(const float *src, int n, int C, int k, const int64_t *classes, float *out,
                                    void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(src && classes && out, "wcr_class_gather_fwd: null pointer");
    WCR_REQUIRE(n >= 0 && C >= 1 && k >= 1, "wcr_class_gather_fwd: bad sizes");
    if (n == 0) return WCR_OK;
    wcr::class_gather_kernel<<<(n * k + 255) / 256, 256, 0, (cudaStream_t)stream>>>(src, n, C, k, classes, out);
    WCR_CHECK_LAUNCH("class_gather_kernel");
    return WCR_OK;
}

extern "C" int wcr_class_gather_bwd(const float *grad_out, int n, int C, int k, const int64_t *classes,
                                    float *d_src, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(grad_out && classes && d_src, "wcr_class_gather_bwd: null pointer");
    WCR_REQUIRE(n >= 0 && C >= 1 && k >= 1, "wcr_class_gather_bwd: bad sizes");
    cudaError_t e = cudaMemsetAsync(d_src, 0, (size_t)n * C * k * sizeof(float), (cudaStream_t)stream);
    if (e != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    if (n == 0) return WCR_OK;
    wcr::class_scatter_kernel<<<(n * k + 255) / 256, 256, 0, (cudaStream_t)stream>>>(grad_out, n, C, k, classes,
                                                                                  d_src);
    WCR_CHECK_LAUNCH("class_scatter_kernel");
    return WCR_OK;
}

extern "C" int wcr_box_constants(const float *Ks, const float *im_scales_ratio, const float *im_current_dims,
                                 const int32_t *img_off, int n_images, int n, float virtual_focal,
                                 int virtual_depth, float *out_K4, float *out_v2r, float *out_clamp, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(Ks && im_scales_ratio && im_current_dims && img_off && out_K4 && out_v2r && out_clamp,
                "wcr_box_constants: null pointer");
    WCR_REQUIRE(n_images >= 1 && n >= 0, "wcr_box_constants: bad sizes");
    if (n == 0) return WCR_OK;
    wcr::box_constants_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        Ks, im_scales_ratio, im_current_dims, img_off, n_images, n, virtual_focal, virtual_depth, out_K4,
        out_v2r, out_clamp);
    WCR_CHECK_LAUNCH("box_constants_kernel");
    return WCR_OK;
}

// PIL ImagingScaleAffine (Geometry.c), the NEAREST path of Image.resize: the source index of
// output pixel x is COORD(xo), xo = a0 * 0.5 accumulated by += a0 IN DOUBLE, a0 = in / out,
// COORD(v) = v < 0 ? -1 : (int) v. The running sum is reproduced literally -- (x + 0.5) * a0
// rounds differently when the product is within an ulp of an integer.
extern "C" void wcr_pil_nearest_table(int in_size, int out_size, int32_t *table)
{
    const double a0 = (double)in_size / (double)out_size;
    double xo = 0.0 + a0 * 0.5;
    for (int x = 0; x < out_size; ++x) {
        table[x] = xo < 0.0 ? -1 : (int)xo;
        xo += a0;
    }
}

extern "C" int wcr_resize_nearest(const void *src, int elem_size, int in_h, int in_w, int64_t src_pitch,
                                  const int32_t *xtab, const int32_t *ytab, void *dst, int out_h, int out_w,
                                  int64_t dst_pitch, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(src && xtab && ytab && dst, "wcr_resize_nearest: null pointer");
    WCR_REQUIRE(elem_size == 4 || elem_size == 1, "wcr_resize_nearest: element size must be 4 (f32) or 1 (u8 / bool)");
    WCR_REQUIRE(in_h >= 1 && in_w >= 1 && out_h >= 1 && out_w >= 1 && out_h <= 65535, "wcr_resize_nearest: bad sizes");
    const dim3 grid((out_w + 255) / 256, out_h);
    if (elem_size == 4)
        wcr::resize_gather_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(
            (const float *)src, in_h, in_w, src_pitch, xtab, ytab, (float *)dst, out_h, out_w, dst_pitch);
    else
        wcr::resize_gather_kernel<uint8_t><<<grid, 256, 0, (cudaStream_t)stream>>>(
            (const uint8_t *)src, in_h, in_w, src_pitch, xtab, ytab, (uint8_t *)dst, out_h, out_w, dst_pitch);
    WCR_CHECK_LAUNCH("resize_gather_kernel");
    return WCR_OK;
}

// F.interpolate(mode='bilinear', align_corners=True) of one f32 plane (depth_range_loss resizes the
// depth map to the mask shape with it, roi_heads.py:672). ATen's upsample_bilinear2d:
//   scale = (in - 1) / (out - 1) (0 when out == 1), src = scale * dst, i0 = (int)src,
//   i1 = i0 + (i0 < in - 1), l1 = src - i0, l0 = 1 - l1,
//   out = l0y * (l0x * v00 + l1x * v01) + l1y * (l0x * v10 + l1x * v11)
namespace wcr {
__global__ void __launch_bounds__(256)
resize_bilinear_ac_kernel(const float *__restrict__ src, int in_h, int in_w, int64_t src_pitch,
                          float *__restrict__ dst, int out_h, int out_w, int64_t dst_pitch)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= out_w) return;
    const float sy = out_h > 1 ? (float)(in_h - 1) / (float)(out_h - 1) : 0.f;
    const float sx = out_w > 1 ? (float)(in_w - 1) / (float)(out_w - 1) : 0.f;
    const float fy = sy * (float)y, fx = sx * (float)x;
    const int y0 = (int)fy, x0 = (int)fx;
    const int y1 = y0 + (y0 < in_h - 1 ? 1 : 0), x1 = x0 + (x0 < in_w - 1 ? 1 : 0);
    const float l1y = fy - (float)y0, l0y = 1.f - l1y, l1x = fx - (float)x0, l0x = 1.f - l1x;
    const float v00 = __ldg(src + (size_t)y0 * src_pitch + x0), v01 = __ldg(src + (size_t)y0 * src_pitch + x1);
    const float v10 = __ldg(src + (size_t)y1 * src_pitch + x0), v11 = __ldg(src + (size_t)y1 * src_pitch + x1);
    dst[(size_t)y * dst_pitch + x] = l0y * (l0x * v00 + l1x * v01) + l1y * (l0x * v10 + l1x * v11);
}
}  // namespace wcr

extern "C" int wcr_resize_bilinear_ac(const float *src, int in_h, int in_w, int64_t src_pitch, float *dst,
                                      int out_h, int out_w, int64_t dst_pitch, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(src && dst, "wcr_resize_bilinear_ac: null pointer");
    WCR_REQUIRE(in_h >= 1 && in_w >= 1 && out_h >= 1 && out_w >= 1 && out_h <= 65535,
                "wcr_resize_bilinear_ac: bad sizes");
    const dim3 grid((out_w + 255) / 256, out_h);
    wcr::resize_bilinear_ac_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, in_h, in_w, src_pitch, dst,
                                                                            out_h, out_w, dst_pitch);
    WCR_CHECK_LAUNCH("resize_bilinear_ac_kernel");
    return WCR_OK;
}
