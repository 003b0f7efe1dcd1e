// blocks.cu -- block-level entry points behind the reference's own helper names
// (SURVEY A2.1-A2.5, A3.2-A3.8): what ProposalNetwork/utils/{spaces,conversions,
// utils}.py compute one torch op at a time, one launch each here. The fused path
// (score.cu, loss.cu) uses the same device functions (cube_math.cuh, hull.cuh,
// sample.cuh); these kernels exist so that the drop-in Python names of
// INTEGRATION.md run on the device without any torch arithmetic.
#include "cube_math.cuh"
#include "hull.cuh"

namespace wcr {

// get_cuboid_verts_faces (math_util.py:198-217) for n cubes -> [n][8][3].
__global__ void __launch_bounds__(256)
cube_corners_kernel(const float *__restrict__ cubes, int64_t n, float *__restrict__ out)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    float c[15];
#pragma unroll
    for (int i = 0; i < 15; ++i) c[i] = __ldg(cubes + j * 15 + i);
    const float hx = c[5] / 2.f, hy = c[4] / 2.f, hz = c[3] / 2.f;
    const float *R = c + 6;
    float *dst = out + j * 24;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float sx, sy, sz;
        corner_signs(i, sx, sy, sz);
        const float lx = sx * hx, ly = sy * hy, lz = sz * hz;
        dst[3 * i + 0] = fmaf(R[2], lz, fmaf(R[1], ly, R[0] * lx)) + c[0];
        dst[3 * i + 1] = fmaf(R[5], lz, fmaf(R[4], ly, R[3] * lx)) + c[1];
        dst[3 * i + 2] = fmaf(R[8], lz, fmaf(R[7], ly, R[6] * lx)) + c[2];
    }
}

// Cubes.get_bube_corners (spaces.py:224-245) and cubes_to_box (conversions.py:25-48).
__global__ void __launch_bounds__(256)
bube_corners_kernel(const float *__restrict__ cubes, int64_t n, Intr K, float clamp0, float clamp1,
                    float *__restrict__ out_corners, float *__restrict__ out_boxes)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    float c[15], cxy[16];
#pragma unroll
    for (int i = 0; i < 15; ++i) c[i] = __ldg(cubes + j * 15 + i);
    float lo0, hi0, lo1, hi1;
    clamp_bounds(clamp0, lo0, hi0);
    clamp_bounds(clamp1, lo1, hi1);
    const ProjBox pb = project_cube(c, K, lo0, hi0, lo1, hi1, cxy);
    if (out_corners) {
#pragma unroll
        for (int i = 0; i < 16; ++i) out_corners[j * 16 + i] = cxy[i];
    }
    if (out_boxes) {
        out_boxes[j * 4 + 0] = pb.x1; out_boxes[j * 4 + 1] = pb.y1;
        out_boxes[j * 4 + 2] = pb.x2; out_boxes[j * 4 + 3] = pb.y2;
    }
}

// iou_2d (utils.py:181-187): one GT box against P proposal boxes.
__global__ void __launch_bounds__(256)
iou_2d_kernel(const float *__restrict__ gt, const float *__restrict__ boxes, int64_t P,
              float *__restrict__ out)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= P) return;
    const float4 b = __ldg(reinterpret_cast<const float4 *>(boxes) + j);
    const ProjBox pb = {b.x, b.y, b.z, b.w};
    out[j] = iou_2d(__ldg(gt), __ldg(gt + 1), __ldg(gt + 2), __ldg(gt + 3), pb);
}

// convex_hull (utils.py:490-493): every CTA marches the (<= 32) points once in
// shared memory, then fills its 32x8 pixel tile with the soft product mask.
__global__ void __launch_bounds__(256)
convex_hull_kernel(const float *__restrict__ coords, int n, int H, int W, float *__restrict__ out,
                   int64_t pitch)
{
    __shared__ float s_p[64], s_h[64];
    __shared__ int s_m;
    if (threadIdx.x == 0) {
        for (int i = 0; i < n; ++i) { s_p[i] = coords[2 * i]; s_p[32 + i] = coords[2 * i + 1]; }
        const Strided px = {s_p, 1}, py = {s_p + 32, 1}, hx = {s_h, 1}, hy = {s_h + 32, 1};
        s_m = jarvis_march(px, py, n, hx, hy);
    }
    __syncthreads();
    const int X = blockIdx.x * 32 + (threadIdx.x & 31), Y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (X >= W || Y >= H) return;
    const Strided hx = {s_h, 1}, hy = {s_h + 32, 1};
    out[(int64_t)Y * pitch + X] = s_m < 1 ? 0.f : fill_value(hx, hy, s_m, (float)X, (float)Y);
}

// mask_iou / mod_mask_iou (utils.py:225-245): intersection = sum(seg * bube) in
// fp64 (the reference sums fp32 with torch's pairwise order; agreement is to fp32
// rounding), union = count(seg != 0 or bube != 0). acc[0] = intersection (double),
// acc[1] = union (int64 bits), acc[2] = retired-block ticket.
template <typename SegT>
__global__ void __launch_bounds__(256)
mask_iou_kernel(const SegT *__restrict__ seg, const float *__restrict__ bube, int64_t numel,
                int mode, double *__restrict__ acc, float *__restrict__ out)
{
    double inter = 0.0;
    long long uni = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float s = (float)seg[i], b = bube[i];
        inter += (double)(s * b);
        uni += (s != 0.f || b != 0.f) ? 1 : 0;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        inter += __shfl_xor_sync(0xffffffffu, inter, o);
        uni += __shfl_xor_sync(0xffffffffu, uni, o);
    }
    __shared__ int s_last;
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(acc, inter);
        atomicAdd(reinterpret_cast<unsigned long long *>(acc + 1), (unsigned long long)uni);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
        s_last = atomicAdd(reinterpret_cast<unsigned int *>(acc + 2), 1u) == gridDim.x - 1;
    __syncthreads();
    if (s_last && threadIdx.x == 0) {
        __threadfence();
        const double it = *reinterpret_cast<volatile double *>(acc);
        const long long un = *reinterpret_cast<volatile long long *>(acc + 1);
        const float fi = (float)it, fu = (float)un;
        float r = 0.f;
        if (it != 0.0) r = mode == 1 ? ((((fi * fi) * fi) * fi) * fi) / fu : fi / fu;
        out[0] = r;
    }
}

// sample_normal_in_range (utils.py:41-59), Philox edition: element (m, j) is a
// normal(mean_m, std_m) redrawn until it lies in [lo, hi_m] (bounded number of
// tries, then clamped; the reference redraws whole tensors up to 10 000 times).
__global__ void __launch_bounds__(256)
sample_normal_kernel(const float *__restrict__ means, const float *__restrict__ stds, int M,
                     int count, int has_thresh, float lo, const float *__restrict__ hi,
                     uint64_t seed, int max_tries, float *__restrict__ out)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)M * count) return;
    const int m = (int)(idx / count), j = (int)(idx - (int64_t)m * count);
    const float mean = __ldg(means + m), sd = __ldg(stds + m);
    const float l = has_thresh ? lo : -INFINITY, h = has_thresh ? __ldg(hi + m) : INFINITY;
    float s = mean;
    bool ok = false;
    for (int blk = 0; !ok && blk * 4 < max_tries; ++blk) {
        uint32_t r[4];
        Philox::gen(seed, (uint32_t)j, (uint32_t)m, 0x534e4952u /* "SNIR" */, (uint32_t)blk, r);
#pragma unroll
        for (int t = 0; t < 2 && !ok; ++t) {
            const float rad = sqrtf(-2.f * logf(u01(r[2 * t])));
            float sn, cs;
            sincosf(6.283185307179586f * u01(r[2 * t + 1]), &sn, &cs);
            s = mean + sd * (rad * cs);
            ok = s >= l && s <= h;
            if (!ok) { s = mean + sd * (rad * sn); ok = s >= l && s <= h; }
        }
    }
    if (!ok && has_thresh) s = fminf(fmaxf(s, l), h);
    out[idx] = s;
}

// orthobasis_from_normal_t (utils.py:119-131) -> [T][3][3], columns [x, n, y].
__global__ void __launch_bounds__(256)
orthobasis_kernel(const float *__restrict__ normal, const float *__restrict__ yaw, int T,
                  float *__restrict__ out)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= T) return;
    const YawBasis yb = make_yaw_basis(__ldg(normal), __ldg(normal + 1), __ldg(normal + 2));
    float R[9];
    yaw_to_R(yb, __ldg(yaw + j), R);
#pragma unroll
    for (int i = 0; i < 9; ++i) out[(int64_t)j * 9 + i] = R[i];
}

// vectorized_linspace (utils.py:169-176): arange(steps) * ((end-start)/(steps-1)) + start.
__global__ void __launch_bounds__(256)
linspace_kernel(const float *__restrict__ start, const float *__restrict__ end, int M, int steps,
                float *__restrict__ out)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)M * steps) return;
    const int m = (int)(idx / steps), j = (int)(idx - (int64_t)m * steps);
    const float s = __ldg(start + m);
    const float spacing = (__ldg(end + m) - s) / (float)(steps - 1);
    out[idx] = (float)j * spacing + s;
}

inline unsigned blocks_for(int64_t n) { return (unsigned)((n + 255) / 256); }

}  // namespace wcr

extern "C" int wcr_cube_corners(const float *cubes, int64_t n, float *out_corners, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(cubes && out_corners, "wcr_cube_corners: null pointer");
    WCR_REQUIRE(n >= 0 && n < ((int64_t)1 << 31), "wcr_cube_corners: bad size");
    if (n == 0) return WCR_OK;
    wcr::cube_corners_kernel<<<wcr::blocks_for(n), 256, 0, (cudaStream_t)stream>>>(cubes, n, out_corners);
    WCR_CHECK_LAUNCH("cube_corners_kernel");
    return WCR_OK;
}

extern "C" int wcr_bube_corners(const float *cubes, int64_t n, float fx, float fy, float cx, float cy,
                                float clamp0, float clamp1, float *out_corners, float *out_boxes,
                                void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(cubes && (out_corners || out_boxes), "wcr_bube_corners: null pointer");
    WCR_REQUIRE(n >= 0 && n < ((int64_t)1 << 31), "wcr_bube_corners: bad size");
    if (n == 0) return WCR_OK;
    const wcr::Intr K = {fx, fy, cx, cy};
    wcr::bube_corners_kernel<<<wcr::blocks_for(n), 256, 0, (cudaStream_t)stream>>>(
        cubes, n, K, clamp0, clamp1, out_corners, out_boxes);
    WCR_CHECK_LAUNCH("bube_corners_kernel");
    return WCR_OK;
}

extern "C" int wcr_iou_2d(const float *gt_box, const float *boxes, int64_t P, float *out_iou,
                          void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(gt_box && boxes && out_iou, "wcr_iou_2d: null pointer");
    WCR_REQUIRE(P >= 0 && P < ((int64_t)1 << 31), "wcr_iou_2d: bad size");
    WCR_REQUIRE(((uintptr_t)boxes & 15) == 0, "wcr_iou_2d: boxes not 16-byte aligned");
    if (P == 0) return WCR_OK;
    wcr::iou_2d_kernel<<<wcr::blocks_for(P), 256, 0, (cudaStream_t)stream>>>(gt_box, boxes, P, out_iou);
    WCR_CHECK_LAUNCH("iou_2d_kernel");
    return WCR_OK;
}

extern "C" int wcr_convex_hull(const float *coords, int n_points, int H, int W, float *out_mask,
                               int64_t pitch, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(coords && out_mask, "wcr_convex_hull: null pointer");
    WCR_REQUIRE(n_points >= 1 && n_points <= 32, "wcr_convex_hull: 1..32 points supported");
    WCR_REQUIRE(H >= 0 && W >= 0 && pitch >= W, "wcr_convex_hull: bad sizes");
    if (H == 0 || W == 0) return WCR_OK;
    WCR_REQUIRE((H + 7) / 8 <= 65535, "wcr_convex_hull: H too large");
    dim3 grid((W + 31) / 32, (H + 7) / 8);
    wcr::convex_hull_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(coords, n_points, H, W, out_mask, pitch);
    WCR_CHECK_LAUNCH("convex_hull_kernel");
    return WCR_OK;
}

extern "C" int wcr_mask_iou(const void *seg, int seg_is_float, const float *bube, int64_t numel,
                            int mode, float *out, void *ws /* 32 bytes */, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(seg && bube && out && ws, "wcr_mask_iou: null pointer");
    WCR_REQUIRE(numel >= 0 && (mode == 0 || mode == 1), "wcr_mask_iou: bad arguments");
    WCR_REQUIRE(((uintptr_t)ws & 7) == 0, "wcr_mask_iou: workspace not 8-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(ws, 0, 32, st);
    if (e != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    int64_t want = (numel + 255) / 256;
    const unsigned grid = (unsigned)(want < 1 ? 1 : (want > 592 ? 592 : want));   // 4 CTAs x 148 SMs
    if (seg_is_float)
        wcr::mask_iou_kernel<float><<<grid, 256, 0, st>>>((const float *)seg, bube, numel, mode, (double *)ws, out);
    else
        wcr::mask_iou_kernel<uint8_t><<<grid, 256, 0, st>>>((const uint8_t *)seg, bube, numel, mode, (double *)ws, out);
    WCR_CHECK_LAUNCH("mask_iou_kernel");
    return WCR_OK;
}

extern "C" int wcr_sample_normal_in_range(const float *means, const float *stds, int M, int count,
                                          int has_thresholds, float threshold_low,
                                          const float *threshold_high, uint64_t seed, int max_tries,
                                          float *out, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(means && stds && out, "wcr_sample_normal_in_range: null pointer");
    WCR_REQUIRE(M >= 0 && count >= 0 && max_tries >= 1, "wcr_sample_normal_in_range: bad sizes");
    WCR_REQUIRE(!has_thresholds || threshold_high, "wcr_sample_normal_in_range: threshold_high missing");
    if ((int64_t)M * count == 0) return WCR_OK;
    wcr::sample_normal_kernel<<<wcr::blocks_for((int64_t)M * count), 256, 0, (cudaStream_t)stream>>>(
        means, stds, M, count, has_thresholds, threshold_low, threshold_high, seed, max_tries, out);
    WCR_CHECK_LAUNCH("sample_normal_kernel");
    return WCR_OK;
}

extern "C" int wcr_orthobasis_from_normal(const float *normal, const float *yaw, int T, float *out_R,
                                          void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(normal && yaw && out_R, "wcr_orthobasis_from_normal: null pointer");
    WCR_REQUIRE(T >= 0, "wcr_orthobasis_from_normal: bad size");
    if (T == 0) return WCR_OK;
    wcr::orthobasis_kernel<<<wcr::blocks_for(T), 256, 0, (cudaStream_t)stream>>>(normal, yaw, T, out_R);
    WCR_CHECK_LAUNCH("orthobasis_kernel");
    return WCR_OK;
}

extern "C" int wcr_vectorized_linspace(const float *start, const float *end, int M, int steps,
                                       float *out, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(start && end && out, "wcr_vectorized_linspace: null pointer");
    WCR_REQUIRE(M >= 0 && steps >= 0, "wcr_vectorized_linspace: bad sizes");
    if ((int64_t)M * steps == 0) return WCR_OK;
    wcr::linspace_kernel<<<wcr::blocks_for((int64_t)M * steps), 256, 0, (cudaStream_t)stream>>>(
        start, end, M, steps, out);
    WCR_CHECK_LAUNCH("linspace_kernel");
    return WCR_OK;
}
