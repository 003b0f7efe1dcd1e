// sample.cuh -- per-box cube sampler shared by wcr_sample_cubes and the fused
// in-kernel sampling path of wcr_score_topk (north_star step 2).
//
// Two sources of the 7 sampled attributes (u v z w h l yaw) of proposal j:
//   replay : values drawn by the reference's own sample_normal_in_range
//            (utils.py:41-59) under a seeded torch generator, passed in;
//   native : Philox4x32-10, counter = (proposal, box, attribute, block), key =
//            seed; Box-Muller normals; truncated by rejection with a bounded
//            number of tries (the reference loops up to 10 000 full redraws,
//            utils.py:50-57), then clamped into the range.
#pragma once
#include "cube_math.cuh"

namespace wcr {

// Per-box constants of the sampler.
struct BoxCtx {
    Intr K;
    YawBasis yb;
    float mean[6], std[6], lo[6], hi[6];  // u v z w h l
    float yaw_lo, yaw_hi;
    int max_tries;
};

__device__ __forceinline__ BoxCtx make_box_ctx(const WcrImage &im, const WcrBox &bx,
                                               const float *__restrict__ priors,
                                               const float *__restrict__ normals, float z0,
                                               const WcrSampleParams &sp)
{
    BoxCtx c;
    c.K.fx = im.fx; c.K.fy = im.fy; c.K.cx = im.cx; c.K.cy = im.cy;
    const float *n = normals + 3 * bx.image;
    c.yb = make_yaw_basis(n[0], n[1], n[2]);
    const float bw = bx.x2 - bx.x1, bh = bx.y2 - bx.y1;
    c.mean[0] = (bx.x1 + bx.x2) / 2.f; c.std[0] = sp.sigma_xy * bw; c.lo[0] = bx.x1; c.hi[0] = bx.x2;
    c.mean[1] = (bx.y1 + bx.y2) / 2.f; c.std[1] = sp.sigma_xy * bh; c.lo[1] = bx.y1; c.hi[1] = bx.y2;
    c.mean[2] = z0; c.std[2] = sp.sigma_z * z0; c.lo[2] = z0 * sp.z_lo; c.hi[2] = z0 * sp.z_hi;
    const float *pm = priors + (size_t)bx.cls * 6;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        c.mean[3 + d] = pm[d]; c.std[3 + d] = pm[3 + d];
        c.lo[3 + d] = sp.dim_lo; c.hi[3 + d] = pm[d] + sp.dim_hi_sigmas * pm[3 + d];
    }
    c.yaw_lo = sp.yaw_lo; c.yaw_hi = sp.yaw_hi; c.max_tries = sp.max_tries;
    return c;
}

// One truncated normal. 4 candidates per Philox block.
__device__ __forceinline__ float philox_trunc_normal(uint64_t seed, uint32_t j, uint32_t box,
                                                     uint32_t attr, float mean, float std, float lo,
                                                     float hi, int max_tries)
{
    float s = mean;
    for (int blk = 0; blk * 4 < max_tries; ++blk) {
        uint32_t r[4];
        Philox::gen(seed, j, box, attr, (uint32_t)blk, r);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float rad = sqrtf(-2.f * logf(u01(r[2 * h])));
            float sn, cs;
            sincosf(6.283185307179586f * u01(r[2 * h + 1]), &sn, &cs);
            s = mean + std * (rad * cs);
            if (s >= lo && s <= hi) return s;
            s = mean + std * (rad * sn);
            if (s >= lo && s <= hi) return s;
        }
    }
    return fminf(fmaxf(s, lo), hi);
}

// The 7 attributes of proposal j of box `box` (global box index in the shard).
__device__ __forceinline__ void draw_attrs(const BoxCtx &c, const float *__restrict__ replay,
                                           int N, uint64_t seed, uint32_t box, uint32_t j,
                                           float a[7])
{
    if (replay) {
        const float *r = replay + (size_t)box * 7 * N + j;
#pragma unroll
        for (int t = 0; t < 7; ++t) a[t] = r[(size_t)t * N];
        return;
    }
#pragma unroll
    for (int t = 0; t < 6; ++t)
        a[t] = philox_trunc_normal(seed, j, box, (uint32_t)t, c.mean[t], c.std[t], c.lo[t], c.hi[t],
                                   c.max_tries);
    uint32_t r[4];
    Philox::gen(seed, j, box, 6u, 0u, r);
    a[6] = c.yaw_lo + (c.yaw_hi - c.yaw_lo) * ((float)(r[0] >> 8) * (1.0f / 16777216.0f));
}

__device__ __forceinline__ void sample_cube(const BoxCtx &c, const float *__restrict__ replay, int N,
                                            uint64_t seed, uint32_t box, uint32_t j, float cube[15])
{
    float a[7];
    draw_attrs(c, replay, N, seed, box, j, a);
    assemble_cube(c.K, c.yb, a[0], a[1], a[2], a[3], a[4], a[5], a[6], cube);
}

}  // namespace wcr
