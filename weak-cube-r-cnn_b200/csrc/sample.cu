// sample.cu -- pseudo depth per box + per-box cube sampling (north_star step 2).
#include "sample.cuh"

namespace wcr {

// Depth at the GT box centre with the reference's 10-px clamp
// (roi_heads.py:645-652 applied to Boxes.get_centers()).
__global__ void __launch_bounds__(128)
box_center_depth_kernel(const WcrImage *__restrict__ images, const WcrBox *__restrict__ boxes,
                        int n_boxes, float *__restrict__ out_z0)
{
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_boxes) return;
    const WcrBox bx = boxes[b];
    const WcrImage im = images[bx.image];
    float x = (bx.x1 + bx.x2) / 2.f, y = (bx.y1 + bx.y2) / 2.f;
    x = clamp_keep_nan(x, 10.f, (float)(im.W - 11));
    y = clamp_keep_nan(y, 10.f, (float)(im.H - 11));
    const int xi = (int)x, yi = (int)y;  // .long(): truncation toward zero
    out_z0[b] = im.depth[(size_t)yi * im.depth_pitch + xi];
}

// grid = (n_boxes, ceil(N/256)); one proposal per thread.
__global__ void __launch_bounds__(256)
sample_cubes_kernel(const WcrImage *__restrict__ images, const WcrBox *__restrict__ boxes,
                    const float *__restrict__ priors, const float *__restrict__ normals,
                    const float *__restrict__ z0, WcrSampleParams sp, int N,
                    const float *__restrict__ replay, uint64_t seed, float *__restrict__ out_cubes)
{
    const int b = blockIdx.x;
    const int j = blockIdx.y * blockDim.x + threadIdx.x;
    if (j >= N) return;
    const WcrBox bx = boxes[b];
    const WcrImage im = images[bx.image];
    const BoxCtx ctx = make_box_ctx(im, bx, priors, normals, z0[b], sp);
    float cube[15];
    sample_cube(ctx, replay, N, seed, (uint32_t)b, (uint32_t)j, cube);
    float *dst = out_cubes + ((size_t)b * N + j) * 15;
#pragma unroll
    for (int t = 0; t < 15; ++t) dst[t] = cube[t];
}

}  // namespace wcr

namespace wcr {
// host-side launcher shared with score.cu (library-side sampling of wcr_score_topk)
cudaError_t launch_sample_cubes(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                                const float *priors, const float *normals, const float *z0,
                                const WcrSampleParams &sp, int N, const float *replay,
                                uint64_t seed, float *out_cubes, cudaStream_t st)
{
    dim3 grid(n_boxes, (N + 255) / 256);
    sample_cubes_kernel<<<grid, 256, 0, st>>>(images, boxes, priors, normals, z0, sp, N, replay,
                                              seed, out_cubes);
    return cudaGetLastError();
}
}  // namespace wcr

extern "C" int wcr_box_center_depth(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                                    float *out_z0, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && boxes && out_z0, "wcr_box_center_depth: null pointer");
    WCR_REQUIRE(n_boxes >= 0, "wcr_box_center_depth: bad sizes");
    if (n_boxes == 0) return WCR_OK;
    wcr::box_center_depth_kernel<<<(n_boxes + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        images, boxes, n_boxes, out_z0);
    WCR_CHECK_LAUNCH("box_center_depth_kernel");
    return WCR_OK;
}

extern "C" int wcr_sample_cubes(const WcrImage *images, const WcrBox *boxes, int n_boxes,
                                const float *priors, const float *normals, const float *z0,
                                const WcrSampleParams *h_params, int N, const float *replay,
                                uint64_t seed, float *out_cubes, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(images && boxes && priors && normals && z0 && h_params && out_cubes,
                "wcr_sample_cubes: null pointer");
    WCR_REQUIRE(n_boxes >= 0 && N >= 1, "wcr_sample_cubes: bad sizes");
    WCR_REQUIRE((N + 255) / 256 <= 65535, "wcr_sample_cubes: N too large");
    if (n_boxes == 0) return WCR_OK;
    wcr::launch_sample_cubes(images, boxes, n_boxes, priors, normals, z0, *h_params, N, replay,
                             seed, out_cubes, (cudaStream_t)stream);
    WCR_CHECK_LAUNCH("sample_cubes_kernel");
    return WCR_OK;
}

namespace wcr {
__global__ void philox_kat_kernel(uint64_t seed, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                  uint32_t *out)
{
    uint32_t r[4];
    Philox::gen(seed, c0, c1, c2, c3, r);
    out[0] = r[0]; out[1] = r[1]; out[2] = r[2]; out[3] = r[3];
}
}  // namespace wcr

extern "C" int wcr_philox_block(uint64_t key, uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                uint32_t *out4, void *stream)
{
    WCR_CHECK_DEVICE();
    WCR_REQUIRE(out4, "wcr_philox_block: null pointer");
    wcr::philox_kat_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(key, c0, c1, c2, c3, out4);
    WCR_CHECK_LAUNCH("philox_kat_kernel");
    return WCR_OK;
}
