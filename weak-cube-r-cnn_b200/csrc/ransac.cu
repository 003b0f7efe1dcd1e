// ransac.cu -- RANSAC ground plane + up vector (north_star step 1, second half).
// Reference: ProposalNetwork/utils/plane.py:79-134 (Plane_torch.fit_parallel),
// roi_heads.py:508-525 (axis fix-ups).
//
// The reference materialises a (1000 x N) distance matrix four times over; here
// each thread owns two hypotheses (planes in registers); a CTA tests them against one
// 1024-point shared-memory tile of the image and adds the inlier counts atomically.
// Rounding order is the reference's (see oracle/csrc/wcr_oracle.c):
//   cross  = fma(a1, b2, -(a2*b1))           (torch.cross on CPU)
//   norm   = sqrt(fma(z,z,fma(y,y,x*x)))     (torch.norm p=2)
//   k      = -((c0*q0 + c1*q1) + c2*q2)      q = sample point 1 (plane.py:112)
//   dist   = (((a*x + b*y) + c*z) + d) / sqrt((a*a + b*b) + c*c)
// The library is built with -fmad=false so none of the separately rounded
// products above is contracted.
#include "common.cuh"

namespace wcr {

constexpr int RANSAC_TPB = 128;    // threads per CTA
constexpr int RANSAC_HPT = 4;      // hypotheses per thread (registers)
constexpr int RANSAC_HPB = RANSAC_TPB * RANSAC_HPT;
constexpr int RANSAC_TILE = 1024;  // points per CTA (one 24 KB shared-memory tile, coordinates doubled)

__device__ __forceinline__ float cross_c(float a1, float b2, float a2, float b1)
{
    const float t = a2 * b1;
    return fmaf(a1, b2, -t);
}

// three distinct indices in [0, n) from one Philox block (partial Fisher-Yates)
__device__ __forceinline__ void philox_triple(uint64_t seed, uint32_t image, uint32_t iter,
                                              int n, int &i0, int &i1, int &i2)
{
    uint32_t r[4];
    Philox::gen(seed, iter, image, 0x52414e53u /* "RANS" */, 0u, r);
    i0 = (int)__umulhi(r[0], (uint32_t)n);
    i1 = (int)__umulhi(r[1], (uint32_t)(n - 1));
    i2 = (int)__umulhi(r[2], (uint32_t)(n - 2));
    if (i1 >= i0) ++i1;
    const int lo = min(i0, i1), hi = max(i0, i1);
    if (i2 >= lo) ++i2;
    if (i2 >= hi) ++i2;
}

// grid = (ceil(iters / RANSAC_HPB), n_images, ceil(cap / RANSAC_TILE)): a CTA counts
// the inliers of 512 hypotheses (4 per thread, planes in registers) among one tile of
// 1024 points of one image and adds them to out_counts (zeroed by the caller). Cutting
// the points as well as the hypotheses keeps every SM busy although images hold
// between a few hundred and cap ground points; CTAs past the image's point count exit.
__global__ void __launch_bounds__(RANSAC_TPB)
ransac_count_kernel(const float4 *__restrict__ pts, const int *__restrict__ count, int cap,
                    const int32_t *__restrict__ sample_idx, int iters, float thresh,
                    uint64_t seed, float4 *__restrict__ eqs, int *__restrict__ out_counts,
                    float k_one, float k_negzero)
{
    // coordinates stored twice, (x,x,y,y) and (z,z): the 64-bit operands of the packed test below
    __shared__ float4 tile_xy[RANSAC_TILE];
    __shared__ float2 tile_z[RANSAC_TILE];
    const int img = blockIdx.y;
    const int n = count[img];
    const int base = blockIdx.z * RANSAC_TILE;
    if (base >= n && !(blockIdx.z == 0)) return;  // tile 0 always runs: it publishes the planes
    const float4 *P = pts + (size_t)img * cap;
    const int m = max(0, min(RANSAC_TILE, n - base));
    for (int j = threadIdx.x; j < m; j += RANSAC_TPB) {
        const float4 p = P[base + j];
        tile_xy[j] = make_float4(p.x, p.x, p.y, p.y);
        tile_z[j] = make_float2(p.z, p.z);
    }

    float a[RANSAC_HPT], b[RANSAC_HPT], c[RANSAC_HPT], d[RANSAC_HPT], den[RANSAC_HPT];
    bool live[RANSAC_HPT];
#pragma unroll
    for (int t = 0; t < RANSAC_HPT; ++t) {
        const int h = blockIdx.x * RANSAC_HPB + t * RANSAC_TPB + threadIdx.x;
        live[t] = h < iters;
        a[t] = b[t] = c[t] = d[t] = nanf_(); den[t] = 1.f;
        if (live[t] && n >= 3) {
            int i0, i1, i2;
            if (sample_idx) {
                const int32_t *s = sample_idx + ((size_t)img * iters + h) * 3;
                i0 = s[0]; i1 = s[1]; i2 = s[2];
            } else {
                philox_triple(seed, (uint32_t)img, (uint32_t)h, n, i0, i1, i2);
            }
            const float4 p0 = P[i0], p1 = P[i1], p2 = P[i2];
            const float ax = p1.x - p0.x, ay = p1.y - p0.y, az = p1.z - p0.z;
            const float bx = p2.x - p0.x, by = p2.y - p0.y, bz = p2.z - p0.z;
            float cx = cross_c(ay, bz, az, by);
            float cy = cross_c(az, bx, ax, bz);
            float cz = cross_c(ax, by, ay, bx);
            const float nrm = sqrtf(fmaf(cz, cz, fmaf(cy, cy, cx * cx)));
            cx = cx / nrm; cy = cy / nrm; cz = cz / nrm;  // collinear triple -> NaN, never an inlier
            a[t] = cx; b[t] = cy; c[t] = cz;
            d[t] = -((cx * p1.x + cy * p1.y) + cz * p1.z);
            den[t] = sqrtf((cx * cx + cy * cy) + cz * cz);
        }
        if (live[t] && blockIdx.z == 0) eqs[(size_t)img * iters + h] = make_float4(a[t], b[t], c[t], d[t]);
    }
    __syncthreads();

    // float counters (exact: a tile holds 1024 points): FSET.BF + FADD run at full rate,
    // FSETP + predicated IADD at a quarter of it on sm_100 (profiles/fp32_micro_r1.jsonl)
    float cnt[RANSAC_HPT];
#pragma unroll
    for (int t = 0; t < RANSAC_HPT; ++t) cnt[t] = 0.f;
    // The reference tests fl(|dist| / den) <= thresh with den = |normal| ~ 1 (plane.py:118-121).
    // x -> fl(x / den) is monotone (den > 0, round to nearest), so there is a largest float T with
    // fl(T / den) <= thresh, and |dist| <= T decides EXACTLY like the division does. T is found once
    // per hypothesis (start at fl(thresh * den), walk a few ulps); the per-point test then costs
    // 3 FMUL + 3 FADD + FSET + FADD instead of an IEEE division on top. A hypothesis whose walk does
    // not settle (den infinite or not positive) keeps the division.
    float thr[RANSAC_HPT];
    bool fast = true;
#pragma unroll
    for (int t = 0; t < RANSAC_HPT; ++t) {
        float g = thresh * den[t];
        bool ok = den[t] > 0.f && den[t] < 3.0e38f && g > 0.f && g < 3.0e38f;
        if (ok) {
            int it = 0;
            while (it < 8 && g / den[t] > thresh) { g = __int_as_float(__float_as_int(g) - 1); ++it; }
            while (it < 16) {
                const float up = __int_as_float(__float_as_int(g) + 1);
                if (!(up / den[t] <= thresh)) break;
                g = up; ++it;
            }
            ok = it < 16 && g / den[t] <= thresh;
        }
        thr[t] = g;
        // NaN den <=> NaN normal <=> NaN dist: never an inlier on either path
        fast = fast && (ok || den[t] != den[t]);
    }
    if (fast) {
        // Two hypotheses per FFMA2. A product is fma(a, x, -0) and a sum fma(s, 1, t): both round
        // once, exactly like FMUL / FADD (x*1 is exact; p + -0 == p for every p incl. signed
        // zeros), so the chain keeps the reference's roundings at 3.25 issue slots per test
        // instead of 8. The constants arrive as kernel arguments so that ptxas cannot fold the
        // fma back into mul/add (it expands mul.rn.f32x2 / add.rn.f32x2 to scalar code).
        const u64 ONE = pack2(k_one, k_one), NZ = pack2(k_negzero, k_negzero);
        u64 A[RANSAC_HPT / 2], B[RANSAC_HPT / 2], C[RANSAC_HPT / 2], D[RANSAC_HPT / 2], CNT[RANSAC_HPT / 2];
#pragma unroll
        for (int q = 0; q < RANSAC_HPT / 2; ++q) {
            A[q] = pack2(a[2 * q], a[2 * q + 1]); B[q] = pack2(b[2 * q], b[2 * q + 1]);
            C[q] = pack2(c[2 * q], c[2 * q + 1]); D[q] = pack2(d[2 * q], d[2 * q + 1]);
            CNT[q] = pack2(0.f, 0.f);
        }
#pragma unroll 8
        for (int j = 0; j < m; ++j) {
            const float4 xy = tile_xy[j];
            const float2 zz = tile_z[j];
            const u64 X = pack2(xy.x, xy.y), Y = pack2(xy.z, xy.w), Z = pack2(zz.x, zz.y);
#pragma unroll
            for (int q = 0; q < RANSAC_HPT / 2; ++q) {
                const u64 m1 = ffma2(A[q], X, NZ), m2 = ffma2(B[q], Y, NZ), m3 = ffma2(C[q], Z, NZ);
                const u64 ds = ffma2(ffma2(ffma2(m1, ONE, m2), ONE, m3), ONE, D[q]);
                float d0, d1;
                unpack2(ds, d0, d1);
                const float i0 = (fabsf(d0) <= thr[2 * q]) ? 1.f : 0.f;
                const float i1 = (fabsf(d1) <= thr[2 * q + 1]) ? 1.f : 0.f;
                CNT[q] = ffma2(pack2(i0, i1), ONE, CNT[q]);
            }
        }
#pragma unroll
        for (int q = 0; q < RANSAC_HPT / 2; ++q) unpack2(CNT[q], cnt[2 * q], cnt[2 * q + 1]);
    } else {
#pragma unroll 4
        for (int j = 0; j < m; ++j) {
            const float px = tile_xy[j].x, py = tile_xy[j].z, pz = tile_z[j].x;
#pragma unroll
            for (int t = 0; t < RANSAC_HPT; ++t) {
                const float dist = (((a[t] * px + b[t] * py) + c[t] * pz) + d[t]) / den[t];
                cnt[t] += (fabsf(dist) <= thresh) ? 1.f : 0.f;
            }
        }
    }
#pragma unroll
    for (int t = 0; t < RANSAC_HPT; ++t) {
        const int h = blockIdx.x * RANSAC_HPB + t * RANSAC_TPB + threadIdx.x;
        if (live[t] && cnt[t] > 0.f) atomicAdd(out_counts + (size_t)img * iters + h, (int)cnt[t]);
    }
}

// One CTA per image: first arg-max of the counts (torch.argmax returns the first
// maximal index, plane.py:128), negate (plane.py:134), fix-ups
// (roi_heads.py:514-525), optional inlier mask of the winner.
__global__ void __launch_bounds__(256)
ransac_select_kernel(const float4 *__restrict__ pts, const int *__restrict__ count, int cap,
                     const float4 *__restrict__ eqs, const int *__restrict__ counts, int iters,
                     float thresh, float *__restrict__ out_eq, float *__restrict__ out_normal,
                     int *__restrict__ out_best, uint8_t *__restrict__ out_mask)
{
    __shared__ int s_cnt[256];
    __shared__ int s_idx[256];
    const int img = blockIdx.x;
    const int *C = counts + (size_t)img * iters;
    int best_c = -1, best_i = 0x7fffffff;
    for (int i = threadIdx.x; i < iters; i += 256) {
        const int c = C[i];
        if (c > best_c) { best_c = c; best_i = i; }  // strided ascending: keeps first max
    }
    s_cnt[threadIdx.x] = best_c; s_idx[threadIdx.x] = best_i;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            const int c2 = s_cnt[threadIdx.x + o], i2 = s_idx[threadIdx.x + o];
            const int c1 = s_cnt[threadIdx.x], i1 = s_idx[threadIdx.x];
            if (c2 > c1 || (c2 == c1 && i2 < i1)) { s_cnt[threadIdx.x] = c2; s_idx[threadIdx.x] = i2; }
        }
        __syncthreads();
    }
    const int best = s_idx[0];
    const float4 e = eqs[(size_t)img * iters + best];
    if (threadIdx.x == 0) {
        out_best[img] = best;
        out_eq[img * 4 + 0] = -e.x; out_eq[img * 4 + 1] = -e.y;
        out_eq[img * 4 + 2] = -e.z; out_eq[img * 4 + 3] = -e.w;
        float nx = -e.x, ny = -e.y, nz = -e.z;
        if (fabsf(nz) > fabsf(ny)) {  // back wall: n[[0,2,1]] * [1,1,-1]
            const float t = ny; ny = nz; nz = -t;
        }
        if (fabsf(nx) > fabsf(ny)) {  // side wall: n[[2,0,1]] * [-1,1,1]
            const float ox = nx, oy = ny, oz = nz;
            nx = -oz; ny = ox; nz = oy;
        }
        if (ny < 0.f) { nx = -nx; ny = -ny; nz = -nz; }
        out_normal[img * 3 + 0] = nx; out_normal[img * 3 + 1] = ny; out_normal[img * 3 + 2] = nz;
    }
    if (out_mask) {
        const int n = count[img];
        const float4 *P = pts + (size_t)img * cap;
        const float den = sqrtf((e.x * e.x + e.y * e.y) + e.z * e.z);
        for (int j = threadIdx.x; j < n; j += 256) {
            const float4 p = P[j];
            const float dist = (((e.x * p.x + e.y * p.y) + e.z * p.z) + e.w) / den;
            out_mask[(size_t)img * cap + j] = (fabsf(dist) <= thresh) ? 1 : 0;
        }
    }
}

}  // namespace wcr

extern "C" int wcr_ransac_ground(const float *pts, const int *count, int cap, int n_images,
                                 const int32_t *sample_idx, int iters, float thresh,
                                 uint64_t seed, float *out_eq, float *out_normal,
                                 int *out_best, int *out_counts, uint8_t *out_inlier_mask,
                                 float *ws_eqs, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(pts && count && out_eq && out_normal && out_best && out_counts && ws_eqs,
                "wcr_ransac_ground: null pointer");
    WCR_REQUIRE(n_images >= 0 && iters >= 1 && cap >= 1, "wcr_ransac_ground: bad sizes");
    WCR_REQUIRE(n_images <= 65535, "wcr_ransac_ground: too many images for one call");
    WCR_REQUIRE((((uintptr_t)pts | (uintptr_t)ws_eqs) & 15) == 0,
                "wcr_ransac_ground: pts/ws_eqs not 16-byte aligned");
    if (n_images == 0) return WCR_OK;
    WCR_REQUIRE((cap + wcr::RANSAC_TILE - 1) / wcr::RANSAC_TILE <= 65535, "wcr_ransac_ground: cap too large");
    cudaError_t me = cudaMemsetAsync(out_counts, 0, (size_t)n_images * iters * sizeof(int), (cudaStream_t)stream);
    if (me != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(me)); return WCR_ELAUNCH; }
    dim3 grid((iters + wcr::RANSAC_HPB - 1) / wcr::RANSAC_HPB, n_images,
              (cap + wcr::RANSAC_TILE - 1) / wcr::RANSAC_TILE);
    wcr::ransac_count_kernel<<<grid, wcr::RANSAC_TPB, 0, (cudaStream_t)stream>>>(
        (const float4 *)pts, count, cap, sample_idx, iters, thresh, seed, (float4 *)ws_eqs,
        out_counts, 1.0f, -0.0f);
    WCR_CHECK_LAUNCH("ransac_count_kernel");
    wcr::ransac_select_kernel<<<n_images, 256, 0, (cudaStream_t)stream>>>(
        (const float4 *)pts, count, cap, (const float4 *)ws_eqs, out_counts, iters, thresh,
        out_eq, out_normal, out_best, out_inlier_mask);
    WCR_CHECK_LAUNCH("ransac_select_kernel");
    return WCR_OK;
}

// ---- allocation-free variant (no counterpart in the reference; added
// for callers that cannot cudaMalloc: CUDA graph capture, serving loops) -------
//
// The hypothesis workspace (ws_eqs of wcr_ransac_ground) lives in a file-scope device array of
// WCR_RANSAC_STATIC_MAX_HYP float4 = 1 MiB, held by every GPU that loads the library until the
// process exits. In HOST code `g_ransac_static_eqs` is only a stand-in for the symbol, never a
// device address: the address is resolved once per device with cudaGetSymbolAddress (that device
// current) and that pointer is what the kernels receive.
namespace wcr {

static __device__ __align__(256) float4 g_ransac_static_eqs[WCR_RANSAC_STATIC_MAX_HYP];

constexpr int MAX_DEVICES = 64;
static float4 *g_ransac_static_ptr[MAX_DEVICES];   // per device, 0 = not resolved yet

// -> device address of the scratch array on the CURRENT device, or nullptr (error set)
static float4 *ransac_static_scratch()
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_error("cudaGetDevice: %s", cudaGetErrorString(e)); return nullptr; }
    if (dev < 0 || dev >= MAX_DEVICES) { set_error("device ordinal %d out of range", dev); return nullptr; }
    float4 *p = __atomic_load_n(&g_ransac_static_ptr[dev], __ATOMIC_ACQUIRE);
    if (p) return p;
    void *sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_ransac_static_eqs);   // no allocation: the array is part of the module image
    if (e != cudaSuccess || sym == nullptr) {
        set_error("cudaGetSymbolAddress(g_ransac_static_eqs) on device %d: %s", dev,
                  cudaGetErrorString(e));
        return nullptr;
    }
    // two threads racing here resolve the same address; the store is idempotent
    __atomic_store_n(&g_ransac_static_ptr[dev], (float4 *)sym, __ATOMIC_RELEASE);
    return (float4 *)sym;
}

}  // namespace wcr

extern "C" int wcr_ransac_ground_static(const float *pts, const int *count, int cap, int n_images,
                                        const int32_t *sample_idx, int iters, float thresh,
                                        uint64_t seed, float *out_eq, float *out_normal,
                                        int *out_best, int *out_counts, uint8_t *out_inlier_mask,
                                        void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(pts && count && out_eq && out_normal && out_best && out_counts,
                "wcr_ransac_ground_static: null pointer");
    WCR_REQUIRE(n_images >= 0 && iters >= 1 && cap >= 1, "wcr_ransac_ground_static: bad sizes");
    WCR_REQUIRE(n_images <= 65535, "wcr_ransac_ground_static: too many images for one call");
    WCR_REQUIRE((int64_t)n_images * iters <= (int64_t)WCR_RANSAC_STATIC_MAX_HYP,
                "wcr_ransac_ground_static: n_images * iters = %lld exceeds the static scratch of %d "
                "hypotheses (WCR_RANSAC_STATIC_MAX_HYP); split the shard or use wcr_ransac_ground",
                (long long)n_images * iters, WCR_RANSAC_STATIC_MAX_HYP);
    WCR_REQUIRE(((uintptr_t)pts & 15) == 0, "wcr_ransac_ground_static: pts not 16-byte aligned");
    if (n_images == 0) return WCR_OK;
    WCR_REQUIRE((cap + wcr::RANSAC_TILE - 1) / wcr::RANSAC_TILE <= 65535,
                "wcr_ransac_ground_static: cap too large");
    float4 *eqs = wcr::ransac_static_scratch();
    if (!eqs) return WCR_ELAUNCH;
    // Nothing in the scratch has to start at zero: tile 0 of the count kernel writes every
    // hypothesis (NaN for images with < 3 points) before the select kernel reads it, so the
    // previous call's contents are dead. The counts live in the caller's out_counts and are
    // cleared here, as in wcr_ransac_ground.
    cudaError_t me = cudaMemsetAsync(out_counts, 0, (size_t)n_images * iters * sizeof(int), (cudaStream_t)stream);
    if (me != cudaSuccess) { wcr::set_error("cudaMemsetAsync: %s", cudaGetErrorString(me)); return WCR_ELAUNCH; }
    dim3 grid((iters + wcr::RANSAC_HPB - 1) / wcr::RANSAC_HPB, n_images,
              (cap + wcr::RANSAC_TILE - 1) / wcr::RANSAC_TILE);
    wcr::ransac_count_kernel<<<grid, wcr::RANSAC_TPB, 0, (cudaStream_t)stream>>>(
        (const float4 *)pts, count, cap, sample_idx, iters, thresh, seed, eqs, out_counts, 1.0f, -0.0f);
    WCR_CHECK_LAUNCH("ransac_count_kernel");
    wcr::ransac_select_kernel<<<n_images, 256, 0, (cudaStream_t)stream>>>(
        (const float4 *)pts, count, cap, eqs, out_counts, iters, thresh,
        out_eq, out_normal, out_best, out_inlier_mask);
    WCR_CHECK_LAUNCH("ransac_select_kernel");
    return WCR_OK;
}
