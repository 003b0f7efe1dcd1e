// unproject.cu -- depth map -> point cloud (north_star step 1, first half).
//
//  * unproject_ground_kernel : strided, image-centre frame, ordered compaction of
//    ground-mask pixels.  Reference: roi_heads.py:474-502.
//  * unproject_cloud_kernel  : dense, K-principal-point frame,
//    HBM-bound elementwise: 4 B read + 16 B written per pixel, 4 pixels per thread.
#include "common.cuh"

namespace wcr {

// One CTA per image. 1024 threads walk the strided grid in row-major order,
// 1024 candidates per step; a ballot/popc block scan keeps the scan order the
// reference's boolean-mask indexing produces (roi_heads.py:489-491).
__global__ void __launch_bounds__(1024)
unproject_ground_kernel(const WcrImage *__restrict__ images, int use_nth,
                        float4 *__restrict__ out_pts, int cap, int *__restrict__ out_count)
{
    __shared__ int warp_tot[32];
    __shared__ int warp_excl[32];
    __shared__ int step_total;
    const WcrImage im = images[blockIdx.x];
    const int Hs = (im.H + use_nth - 1) / use_nth;
    const int Ws = (im.W + use_nth - 1) / use_nth;
    const int total = Hs * Ws;
    // K[0,0] // use_nth : float floor-division (roi_heads.py:479)
    const float fx = floorf(im.fx / (float)use_nth);
    const float fy = floorf(im.fy / (float)use_nth);
    const float cx = (float)Ws * 0.5f, cy = (float)Hs * 0.5f;  // roi_heads.py:482
    float4 *dst = out_pts + (size_t)blockIdx.x * cap;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int base = 0;  // replicated in every thread

    for (int start = 0; start < total; start += 1024) {
        const int i = start + threadIdx.x;
        bool keep = false;
        float x = 0.f, y = 0.f, z = 0.f;
        if (i < total) {
            const int vs = i / Ws, us = i - vs * Ws;
            const int v = vs * use_nth, u = us * use_nth;
            keep = im.ground ? (im.ground[(size_t)v * im.ground_pitch + u] > 0) : true;
            if (keep) {
                z = __ldg(im.depth + (size_t)v * im.depth_pitch + u);
                x = (((float)us - cx) * z) / fx;  // roi_heads.py:484
                y = (((float)vs - cy) * z) / fy;  // roi_heads.py:485
            }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, keep);
        const int before = __popc(bal & ((1u << lane) - 1u));
        if (lane == 0) warp_tot[wid] = __popc(bal);
        __syncthreads();
        if (wid == 0) {
            const int t = warp_tot[lane];
            int incl = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int n = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += n;
            }
            warp_excl[lane] = incl - t;
            if (lane == 31) step_total = incl;
        }
        __syncthreads();
        if (keep && base + warp_excl[wid] + before < cap)
            dst[base + warp_excl[wid] + before] = make_float4(x, y, z, 0.f);
        base += step_total;
    }
    if (threadIdx.x == 0) out_count[blockIdx.x] = base < cap ? base : cap;
}

// Dense cloud, HBM bound: 4 B read + 16 B written per pixel. A warp takes 128 consecutive
// pixels of a row, lane l the pixels l, l+32, l+64, l+96: every load instruction of the warp
// reads one full 128-byte line and every STG.128 writes 512 contiguous bytes (four full
// lines), so no sector is ever written in two pieces; each thread keeps 80 B in flight.
// grid = (ceil(Hs * ceil(Ws/128) / 8), n_images), 8 warps per CTA.
__device__ __forceinline__ float4 lift_pixel(float z, int u, int v, const WcrImage &im)
{
    if (z > 0.f && z <= 3.402823466e38f) {  // finite and positive
        float4 p;
        p.x = (z * ((float)u - im.cx)) / im.fx;  // roi_heads.py:927
        p.y = (z * ((float)v - im.cy)) / im.fy;  // roi_heads.py:928
        p.z = z;
        p.w = 0.f;
        return p;
    }
    const float q = nanf_();
    return make_float4(q, q, q, 0.f);
}

__global__ void __launch_bounds__(256)
unproject_cloud_kernel(const WcrImage *__restrict__ images, int stride,
                       float4 *__restrict__ out_xyz, int64_t plane_elems)
{
    const WcrImage im = images[blockIdx.y];
    const int Hs = (im.H + stride - 1) / stride;
    const int Ws = (im.W + stride - 1) / stride;
    const int spans = (Ws + 127) >> 7;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= Hs * spans) return;
    const int vs = w / spans, us0 = ((w - vs * spans) << 7) + lane;
    const int v = vs * stride;
    const float *row = im.depth + (size_t)v * im.depth_pitch;
    float4 *dst = out_xyz + (size_t)blockIdx.y * plane_elems + (size_t)vs * Ws;
    float z[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int us = us0 + 32 * k;
        z[k] = us < Ws ? __ldcs(row + (size_t)us * stride) : 0.f;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int us = us0 + 32 * k;
        if (us < Ws) __stcs(dst + us, lift_pixel(z[k], us * stride, v, im));
    }
}

}  // namespace wcr

extern "C" int wcr_unproject_ground(const WcrImage *images, int n_images, int use_nth,
                                    float *out_pts, int cap, int *out_count, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && out_pts && out_count, "wcr_unproject_ground: null pointer");
    WCR_REQUIRE(n_images >= 0 && use_nth >= 1 && cap >= 1, "wcr_unproject_ground: bad sizes");
    WCR_REQUIRE(((uintptr_t)out_pts & 15) == 0, "wcr_unproject_ground: out_pts not 16-byte aligned");
    if (n_images == 0) return WCR_OK;
    wcr::unproject_ground_kernel<<<n_images, 1024, 0, (cudaStream_t)stream>>>(
        images, use_nth, (float4 *)out_pts, cap, out_count);
    WCR_CHECK_LAUNCH("unproject_ground_kernel");
    return WCR_OK;
}

extern "C" int wcr_unproject_cloud(const WcrImage *images, int n_images, int stride,
                                   int max_H, int max_W, float *out_xyz,
                                   int64_t plane_elems, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && out_xyz, "wcr_unproject_cloud: null pointer");
    WCR_REQUIRE(n_images >= 0 && stride >= 1 && max_H >= 1 && max_W >= 1,
                "wcr_unproject_cloud: bad sizes");
    const int Hs = (max_H + stride - 1) / stride, Ws = (max_W + stride - 1) / stride;
    WCR_REQUIRE(plane_elems >= (int64_t)Hs * Ws, "wcr_unproject_cloud: plane_elems < Hs*Ws");
    WCR_REQUIRE(((uintptr_t)out_xyz & 15) == 0, "wcr_unproject_cloud: out_xyz not 16-byte aligned");
    WCR_REQUIRE(Hs <= 65535 && n_images <= 65535, "wcr_unproject_cloud: grid too large");
    if (n_images == 0) return WCR_OK;
    const int64_t warps = (int64_t)Hs * ((Ws + 127) / 128);
    WCR_REQUIRE((warps + 7) / 8 < ((int64_t)1 << 31), "wcr_unproject_cloud: image too large");
    dim3 grid((unsigned)((warps + 7) / 8), n_images);
    wcr::unproject_cloud_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
        images, stride, (float4 *)out_xyz, plane_elems);
    WCR_CHECK_LAUNCH("unproject_cloud_kernel");
    return WCR_OK;
}
