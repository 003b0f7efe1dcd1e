// blocks2.cu -- second set of block-level entry points:
//   backward of Cubes.get_all_corners / get_bube_corners / cubes_to_box (the reference uses
//     them INSIDE autograd, spaces.py:224-245 "we must clamp and then stack ...", call site
//     roi_heads.py:947-960): the drop-in Cubes methods carry an autograd.Function on top of these;
//   custom_mapping (utils.py:207-223), randn_orthobasis_torch (:61-68), gt_in_norm_range
//     (:148-152), is_gt_included (:257-298);
//   stand-alone z_loss (roi_heads.py:530-573) and normal_to_rotation (:686-696).
// Same conventions as blocks.cu: one launch each, -fmad=false, explicit fmaf() only.
#include "cube_math.cuh"

namespace wcr {

// d/d cube of sum_i <g_i, corner_i>, corner_i = R (s_i * half) + c (math_util.py:198-217).
__device__ __forceinline__ void corner_grad_accumulate(const float *c, int i, float gX, float gY, float gZ,
                                                       float d[15])
{
    float sx, sy, sz;
    corner_signs(i, sx, sy, sz);
    const float l[3] = {sx * (c[5] / 2.f), sy * (c[4] / 2.f), sz * (c[3] / 2.f)};
    const float *R = c + 6;
    const float g[3] = {gX, gY, gZ};
    d[0] += gX; d[1] += gY; d[2] += gZ;
    float dl[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            d[6 + 3 * r + k] += g[r] * l[k];
            dl[k] += g[r] * R[3 * r + k];
        }
    d[5] += dl[0] * sx * 0.5f;   // local x extent is l (cube[5])
    d[4] += dl[1] * sy * 0.5f;   // local y extent is h
    d[3] += dl[2] * sz * 0.5f;   // local z extent is w
}

__global__ void __launch_bounds__(256)
cube_corners_bwd_kernel(const float *__restrict__ cubes, const float *__restrict__ grad, int64_t n,
                        float *__restrict__ d_cubes)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    float c[15], d[15];
#pragma unroll
    for (int i = 0; i < 15; ++i) { c[i] = __ldg(cubes + j * 15 + i); d[i] = 0.f; }
    const float *g = grad + j * 24;
#pragma unroll
    for (int i = 0; i < 8; ++i) corner_grad_accumulate(c, i, g[3 * i], g[3 * i + 1], g[3 * i + 2], d);
#pragma unroll
    for (int i = 0; i < 15; ++i) d_cubes[j * 15 + i] = d[i];
}

// Backward of get_bube_corners (K @ corners, divide by z, clamp; spaces.py:231-243) and of the
// min / max of cubes_to_box (conversions.py:36-42), torch conventions: clamp passes the gradient
// on the closed interval [lo, hi]; min / max over the 8 corners send it to the FIRST extremum.
__global__ void __launch_bounds__(256)
bube_corners_bwd_kernel(const float *__restrict__ cubes, int64_t n, Intr K, float clamp0, float clamp1,
                        const float *__restrict__ g_corners, const float *__restrict__ g_boxes,
                        float *__restrict__ d_cubes)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    float c[15], d[15], cxy[16];
#pragma unroll
    for (int i = 0; i < 15; ++i) { c[i] = __ldg(cubes + j * 15 + i); d[i] = 0.f; }
    float lo0, hi0, lo1, hi1;
    clamp_bounds(clamp0, lo0, hi0);
    clamp_bounds(clamp1, lo1, hi1);
    project_cube(c, K, lo0, hi0, lo1, hi1, cxy);
    float gx[8], gy[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        gx[i] = g_corners ? g_corners[j * 16 + 2 * i] : 0.f;
        gy[i] = g_corners ? g_corners[j * 16 + 2 * i + 1] : 0.f;
    }
    if (g_boxes) {
        int ix1 = 0, ix2 = 0, iy1 = 0, iy2 = 0;
#pragma unroll
        for (int i = 1; i < 8; ++i) {
            if (cxy[2 * i] < cxy[2 * ix1]) ix1 = i;
            if (cxy[2 * i] > cxy[2 * ix2]) ix2 = i;
            if (cxy[2 * i + 1] < cxy[2 * iy1 + 1]) iy1 = i;
            if (cxy[2 * i + 1] > cxy[2 * iy2 + 1]) iy2 = i;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            gx[i] += (i == ix1 ? g_boxes[j * 4 + 0] : 0.f) + (i == ix2 ? g_boxes[j * 4 + 2] : 0.f);
            gy[i] += (i == iy1 ? g_boxes[j * 4 + 1] : 0.f) + (i == iy2 ? g_boxes[j * 4 + 3] : 0.f);
        }
    }
    const float hx = c[5] / 2.f, hy = c[4] / 2.f, hz = c[3] / 2.f;
    const float *R = c + 6;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        float sx, sy, sz;
        corner_signs(i, sx, sy, sz);
        const float lx = sx * hx, ly = sy * hy, lz = sz * hz;
        const float X = fmaf(R[2], lz, fmaf(R[1], ly, R[0] * lx)) + c[0];
        const float Y = fmaf(R[5], lz, fmaf(R[4], ly, R[3] * lx)) + c[1];
        const float Z = fmaf(R[8], lz, fmaf(R[7], ly, R[6] * lx)) + c[2];
        const float px = fmaf(K.cx, Z, K.fx * X) / Z, py = fmaf(K.cy, Z, K.fy * Y) / Z;
        const float gpx = (px >= lo0 && px <= hi0) ? gx[i] : 0.f;   // NaN: no gradient
        const float gpy = (py >= lo1 && py <= hi1) ? gy[i] : 0.f;
        // p = (c Z + f X) / Z: autograd of the reference divides num = K @ corner by Z
        const float gX = gpx * K.fx / Z, gY = gpy * K.fy / Z;
        const float numx = fmaf(K.cx, Z, K.fx * X), numy = fmaf(K.cy, Z, K.fy * Y);
        const float gZ = gpx * (K.cx / Z - numx / (Z * Z)) + gpy * (K.cy / Z - numy / (Z * Z));
        if (gpx != 0.f || gpy != 0.f) corner_grad_accumulate(c, i, gX, gY, gZ, d);
    }
#pragma unroll
    for (int i = 0; i < 15; ++i) d_cubes[j * 15 + i] = d[i];
}

// custom_mapping (utils.py:207-223): x <= 0 -> 0, else 1 / (1 + (x / (1 - x)) ** (-beta)).
// The reference evaluates Python floats (double): so does this kernel (n is a handful of values).
__global__ void __launch_bounds__(256)
custom_mapping_kernel(const float *__restrict__ x, int64_t n, double beta, float *__restrict__ out)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const double v = (double)x[j];
    out[j] = v <= 0.0 ? 0.f : (float)(1.0 / (1.0 + pow(v / (1.0 - v), -beta)));
}

// randn_orthobasis_torch (utils.py:61-68) on given normal draws z [n][3][3] (row-major rows
// z0 z1 z2): rows normalised, z0 = cross(z1, z2) / |.|, z1 = cross(z2, z0) / |.|.
// torch rounding: norm = sqrt(sum of squares in index order), cross = fma(a1, b2, -(a2 b1)).
__device__ __forceinline__ float norm3(float a, float b, float c) { return sqrtf((a * a + b * b) + c * c); }

__global__ void __launch_bounds__(256)
randn_orthobasis_kernel(const float *__restrict__ z_in, int64_t n, uint64_t seed, float *__restrict__ out)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    float z[9];
    if (z_in) {
#pragma unroll
        for (int i = 0; i < 9; ++i) z[i] = z_in[j * 9 + i];
    } else {   // native: Box-Muller on Philox4x32-10, counter (basis lo, basis hi, "ORTH", block)
#pragma unroll
        for (int blk = 0; blk < 3; ++blk) {
            uint32_t r[4];
            Philox::gen(seed, (uint32_t)j, (uint32_t)((uint64_t)j >> 32), 0x4f525448u, (uint32_t)blk, r);
            const float r0 = sqrtf(-2.f * logf(u01(r[0]))), r1 = sqrtf(-2.f * logf(u01(r[2])));
            float s0, c0, s1, c1;
            sincosf(6.283185307179586f * u01(r[1]), &s0, &c0);
            sincosf(6.283185307179586f * u01(r[3]), &s1, &c1);
            z[3 * blk] = r0 * c0; z[3 * blk + 1] = r0 * s0; z[3 * blk + 2] = r1 * c1;
        }
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float m = norm3(z[3 * r], z[3 * r + 1], z[3 * r + 2]);
        z[3 * r] /= m; z[3 * r + 1] /= m; z[3 * r + 2] /= m;
    }
    float a0 = cross_term(z[4], z[8], z[5], z[7]), a1 = cross_term(z[5], z[6], z[3], z[8]),
          a2 = cross_term(z[3], z[7], z[4], z[6]);
    float m = norm3(a0, a1, a2);
    z[0] = a0 / m; z[1] = a1 / m; z[2] = a2 / m;
    a0 = cross_term(z[7], z[2], z[8], z[1]); a1 = cross_term(z[8], z[0], z[6], z[2]);
    a2 = cross_term(z[6], z[1], z[7], z[0]);
    m = norm3(a0, a1, a2);
    z[3] = a0 / m; z[4] = a1 / m; z[5] = a2 / m;
#pragma unroll
    for (int i = 0; i < 9; ++i) out[j * 9 + i] = z[i];
}

// gt_in_norm_range (utils.py:148-152): (gt - lo) / |hi - lo|, element-wise over n (range, gt) pairs.
__global__ void __launch_bounds__(256)
gt_in_norm_range_kernel(const float *__restrict__ lo, const float *__restrict__ hi,
                        const float *__restrict__ gt, int64_t n, float *__restrict__ out)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    out[j] = (gt[j] - lo[j]) / fabsf(hi[j] - lo[j]);
}

// is_gt_included (utils.py:257-298) for n GT cubes at once. Per cube: a bit mask of the reasons
// the GT cannot be reached (0 == included) and the three centre distances the reference prints.
//   ranges [n][6] = x_lo x_hi y_lo y_hi z_lo z_hi; priors [n][6] = w(mean,std) h(..) l(..)
//   bits: 1 x, 2 y, 4 z, 8 w-, 16 w+, 32 h-, 64 h+, 128 l-, 256 l+
__global__ void __launch_bounds__(256)
is_gt_included_kernel(const float *__restrict__ gt_cubes, const float *__restrict__ ranges,
                      const float *__restrict__ priors, int64_t n, int32_t *__restrict__ out_bits,
                      float *__restrict__ out_by)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const float *c = gt_cubes + j * 15, *r = ranges + j * 6, *p = priors + j * 6;
    const float stds_away = 1.5f;
    int bits = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float v = c[a], lo = r[2 * a], hi = r[2 * a + 1];
        float by = 0.f;
        if (!(lo < v && v < hi)) {
            bits |= 1 << a;
            by = v < lo ? fabsf(lo - v) : fabsf(v - hi);
        }
        out_by[j * 3 + a] = by;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float dim = c[3 + a], mean = p[2 * a], sd = p[2 * a + 1];
        if (dim < mean - stds_away * sd) bits |= 8 << (2 * a);
        if (dim > mean + stds_away * sd) bits |= 16 << (2 * a);
    }
    out_bits[j] = bits;
}

// ROIHeads3DScore.z_loss (roi_heads.py:530-573) on given cubes: 50 offsets of 0.1 m along z,
// arg-min of |gt_area - projected area| (first minimum; zero areas count as 1e7), score =
// |z - z_best| / 2; 0.1 * 50 / 2 when the projected centre is "outside" -- with the reference's
// chained comparison kept ((lower <= centre) is a BOOL compared with the upper bound, :543-544).
struct ZStepTable { float v[WCR_Z_LOSS_STEPS]; };

__global__ void __launch_bounds__(128)
z_loss_kernel(const float *__restrict__ gt_boxes, const float *__restrict__ cubes,
              const float *__restrict__ K4, const float *__restrict__ clamp,
              const float *__restrict__ proj, int n, const ZStepTable zs, float *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float cube[15];
#pragma unroll
    for (int t = 0; t < 15; ++t) cube[t] = cubes[(size_t)i * 15 + t];
    const Intr K = {K4[4 * i], K4[4 * i + 1], K4[4 * i + 2], K4[4 * i + 3]};
    float lo0, hi0, lo1, hi1;
    clamp_bounds(clamp[2 * i], lo0, hi0);
    clamp_bounds(clamp[2 * i + 1], lo1, hi1);
    const float *g = gt_boxes + 4 * i, *p = proj + 4 * i;
    const float gt_area = (g[2] - g[0]) * (g[3] - g[1]);
    const float pcx = (p[0] + p[2]) / 2.f, pcy = (p[1] + p[3]) / 2.f;
    const float pred_area = (p[2] - p[0]) * (p[3] - p[1]);
    const float bx = (g[0] - 50.f <= pcx) ? 1.f : 0.f, by = (g[1] - 50.f <= pcy) ? 1.f : 0.f;
    const bool within = (bx <= g[2] + 50.f) && (by <= g[3] + 50.f);
    float score = 5.0f;
    if (within) {
        const bool smaller = gt_area < pred_area;
        const float z = cube[2];
        float best = 0.f, best_z = z;
        for (int k = 0; k < WCR_Z_LOSS_STEPS; ++k) {
            const float zk = smaller ? z + zs.v[k] : z - zs.v[k];
            cube[2] = zk;
            const ProjBox mb = project_cube(cube, K, lo0, hi0, lo1, hi1, nullptr);
            float area = (mb.x2 - mb.x1) * (mb.y2 - mb.y1);
            area = area + (area == 0.f ? 10000000.f : 0.f);
            const float diff = fabsf(gt_area - area);
            if (k == 0 || (best == best && (diff != diff || diff < best))) { best = diff; best_z = zk; }
        }
        score = fabsf(z - best_z);
    }
    out[i] = score / 2.f;
}

// ROIHeads3DScore.normal_to_rotation (roi_heads.py:686-696), rows [t0, t1, normal]. Reference
// quirks kept: `bmm(...).flatten().any() < 0.001` tests a BOOL (the y-axis fallback is taken only
// when EVERY t0 is exactly zero), and torch.norm(t0) is the Frobenius norm of the WHOLE (n, 3)
// matrix, not a per-row norm. Two passes: norms (one CTA), then rows.
__global__ void __launch_bounds__(256)
normal_to_rotation_kernel(const float *__restrict__ normal, int n, float *__restrict__ out)
{
    __shared__ float s_red[256];
    __shared__ int s_any;
    const int tid = threadIdx.x;
    if (tid == 0) s_any = 0;
    __syncthreads();
    // t0 = cross(normal, [1,0,0]) = (0, nz, -ny); does any row have a non-zero squared length?
    int any = 0;
    for (int i = tid; i < n; i += 256) {
        const float ny = normal[3 * i + 1], nz = normal[3 * i + 2];
        const float t0y = cross_term(nz, 1.f, normal[3 * i], 0.f), t0z = cross_term(normal[3 * i], 0.f, ny, 1.f);
        const float sq = (0.f * 0.f + t0y * t0y) + t0z * t0z;
        any |= (sq != 0.f);
    }
    if (any) s_any = 1;
    __syncthreads();
    const bool use_y = !s_any;   // (any() < 0.001) <=> any() is False
    auto t0_of = [&](int i, float t[3]) {
        const float nx = normal[3 * i], ny = normal[3 * i + 1], nz = normal[3 * i + 2];
        if (!use_y) { t[0] = cross_term(ny, 0.f, nz, 0.f); t[1] = cross_term(nz, 1.f, nx, 0.f); t[2] = cross_term(nx, 0.f, ny, 1.f); }
        else { t[0] = cross_term(ny, 0.f, nz, 1.f); t[1] = cross_term(nz, 0.f, nx, 0.f); t[2] = cross_term(nx, 1.f, ny, 0.f); }
    };
    auto block_sum = [&](float v) {
        s_red[tid] = v;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) { if (tid < o) s_red[tid] += s_red[tid + o]; __syncthreads(); }
        const float r = s_red[0];
        __syncthreads();
        return r;
    };
    float acc = 0.f;
    for (int i = tid; i < n; i += 256) { float t[3]; t0_of(i, t); acc += (t[0] * t[0] + t[1] * t[1]) + t[2] * t[2]; }
    const float n0 = sqrtf(block_sum(acc));
    acc = 0.f;
    for (int i = tid; i < n; i += 256) {
        float t[3]; t0_of(i, t);
        t[0] /= n0; t[1] /= n0; t[2] /= n0;
        const float nx = normal[3 * i], ny = normal[3 * i + 1], nz = normal[3 * i + 2];
        const float u0 = cross_term(ny, t[2], nz, t[1]), u1 = cross_term(nz, t[0], nx, t[2]), u2 = cross_term(nx, t[1], ny, t[0]);
        acc += (u0 * u0 + u1 * u1) + u2 * u2;
    }
    const float n1 = sqrtf(block_sum(acc));
    for (int i = tid; i < n; i += 256) {
        float t[3]; t0_of(i, t);
        t[0] /= n0; t[1] /= n0; t[2] /= n0;
        const float nx = normal[3 * i], ny = normal[3 * i + 1], nz = normal[3 * i + 2];
        const float u0 = cross_term(ny, t[2], nz, t[1]), u1 = cross_term(nz, t[0], nx, t[2]), u2 = cross_term(nx, t[1], ny, t[0]);
        float *o = out + 9 * i;
        o[0] = t[0]; o[1] = t[1]; o[2] = t[2];
        o[3] = u0 / n1; o[4] = u1 / n1; o[5] = u2 / n1;
        o[6] = nx; o[7] = ny; o[8] = nz;
    }
}

}  // namespace wcr

#define WCR_GRID(n) (unsigned)(((n) + 255) / 256)

extern "C" int wcr_cube_corners_bwd(const float *cubes, const float *grad_corners, int64_t n, float *d_cubes,
                                    void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(cubes && grad_corners && d_cubes, "wcr_cube_corners_bwd: null pointer");
    WCR_REQUIRE(n >= 0 && n <= 0x7fffffffLL * 64, "wcr_cube_corners_bwd: bad size");
    if (n == 0) return WCR_OK;
    wcr::cube_corners_bwd_kernel<<<WCR_GRID(n), 256, 0, (cudaStream_t)stream>>>(cubes, grad_corners, n, d_cubes);
    WCR_CHECK_LAUNCH("cube_corners_bwd_kernel");
    return WCR_OK;
}

extern "C" int wcr_bube_corners_bwd(const float *cubes, int64_t n, float fx, float fy, float cx, float cy,
                                    float clamp0, float clamp1, const float *grad_corners,
                                    const float *grad_boxes, float *d_cubes, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(cubes && d_cubes && (grad_corners || grad_boxes), "wcr_bube_corners_bwd: null pointer");
    WCR_REQUIRE(n >= 0 && n <= 0x7fffffffLL * 64, "wcr_bube_corners_bwd: bad size");
    if (n == 0) return WCR_OK;
    const wcr::Intr K = {fx, fy, cx, cy};
    wcr::bube_corners_bwd_kernel<<<WCR_GRID(n), 256, 0, (cudaStream_t)stream>>>(cubes, n, K, clamp0, clamp1,
                                                                                grad_corners, grad_boxes, d_cubes);
    WCR_CHECK_LAUNCH("bube_corners_bwd_kernel");
    return WCR_OK;
}

extern "C" int wcr_custom_mapping(const float *x, int64_t n, double beta, float *out, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(x && out && n >= 0, "wcr_custom_mapping: bad arguments");
    if (n == 0) return WCR_OK;
    wcr::custom_mapping_kernel<<<WCR_GRID(n), 256, 0, (cudaStream_t)stream>>>(x, n, beta, out);
    WCR_CHECK_LAUNCH("custom_mapping_kernel");
    return WCR_OK;
}

extern "C" int wcr_randn_orthobasis(const float *z, int64_t n, uint64_t seed, float *out, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(out && n >= 0, "wcr_randn_orthobasis: bad arguments");
    if (n == 0) return WCR_OK;
    wcr::randn_orthobasis_kernel<<<WCR_GRID(n), 256, 0, (cudaStream_t)stream>>>(z, n, seed, out);
    WCR_CHECK_LAUNCH("randn_orthobasis_kernel");
    return WCR_OK;
}

extern "C" int wcr_gt_in_norm_range(const float *lo, const float *hi, const float *gt, int64_t n, float *out,
                                    void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(lo && hi && gt && out && n >= 0, "wcr_gt_in_norm_range: bad arguments");
    if (n == 0) return WCR_OK;
    wcr::gt_in_norm_range_kernel<<<WCR_GRID(n), 256, 0, (cudaStream_t)stream>>>(lo, hi, gt, n, out);
    WCR_CHECK_LAUNCH("gt_in_norm_range_kernel");
    return WCR_OK;
}

extern "C" int wcr_is_gt_included(const float *gt_cubes, const float *ranges, const float *priors, int64_t n,
                                  int32_t *out_bits, float *out_by, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(gt_cubes && ranges && priors && out_bits && out_by && n >= 0, "wcr_is_gt_included: bad arguments");
    if (n == 0) return WCR_OK;
    wcr::is_gt_included_kernel<<<WCR_GRID(n), 256, 0, (cudaStream_t)stream>>>(gt_cubes, ranges, priors, n,
                                                                              out_bits, out_by);
    WCR_CHECK_LAUNCH("is_gt_included_kernel");
    return WCR_OK;
}

extern "C" int wcr_z_loss(const float *gt_boxes, const float *cubes, const float *K4, const float *clamp,
                          const float *proj_boxes, int n, float *out, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(gt_boxes && cubes && K4 && clamp && proj_boxes && out && n >= 0, "wcr_z_loss: bad arguments");
    if (n == 0) return WCR_OK;
    wcr::ZStepTable zs;
    wcr_z_loss_steps(zs.v);
    wcr::z_loss_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(gt_boxes, cubes, K4, clamp, proj_boxes,
                                                                          n, zs, out);
    WCR_CHECK_LAUNCH("z_loss_kernel");
    return WCR_OK;
}

extern "C" int wcr_normal_to_rotation(const float *normal, int n, float *out, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(normal && out && n >= 0, "wcr_normal_to_rotation: bad arguments");
    if (n == 0) return WCR_OK;
    wcr::normal_to_rotation_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(normal, n, out);
    WCR_CHECK_LAUNCH("normal_to_rotation_kernel");
    return WCR_OK;
}
