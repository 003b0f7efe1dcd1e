// common.cuh -- shared helpers for libwcr_b200 (sm_100a only).
//
// Build flags that matter for parity (see build.py): -fmad=false, so the ONLY
// fused multiply-adds in the binary are the explicit fmaf()/__fmaf_rn()/
// fma.rn.f32x2 sites; division and sqrt are IEEE (nvcc defaults, no fast-math).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "wcr.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libwcr_b200 is written for sm_100a only"
#endif

namespace wcr {

void set_error(const char *fmt, ...);
int  check_device();           // WCR_OK or WCR_EARCH
void count_launch(int n = 1);  // bookkeeping for wcr_launch_count()
int  sm_count();

#define WCR_REQUIRE(cond, ...)                       \
    do {                                             \
        if (!(cond)) {                               \
            wcr::set_error(__VA_ARGS__);             \
            return WCR_EINVAL;                       \
        }                                            \
    } while (0)

#define WCR_CHECK_DEVICE()                           \
    do {                                             \
        int _rc = wcr::check_device();               \
        if (_rc != WCR_OK) return _rc;               \
    } while (0)

#define WCR_CHECK_LAUNCH(name)                                              \
    do {                                                                    \
        cudaError_t _e = cudaGetLastError();                                \
        if (_e != cudaSuccess) {                                            \
            wcr::set_error("%s: %s", name, cudaGetErrorString(_e));         \
            return WCR_ELAUNCH;                                             \
        }                                                                   \
        wcr::count_launch();                                                \
    } while (0)

// ---- Philox4x32-10 (Salmon et al. 2011), counter-based RNG -----------------
struct Philox {
    static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    static constexpr uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    __host__ __device__ static inline void round(uint32_t c[4], uint32_t k0, uint32_t k1)
    {
#ifdef __CUDA_ARCH__
        uint32_t hi0 = __umulhi(M0, c[0]), hi1 = __umulhi(M1, c[2]);
#else
        uint32_t hi0 = (uint32_t)(((uint64_t)M0 * c[0]) >> 32);
        uint32_t hi1 = (uint32_t)(((uint64_t)M1 * c[2]) >> 32);
#endif
        uint32_t lo0 = M0 * c[0], lo1 = M1 * c[2];
        uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    }
    // out[4] = philox(counter = (c0,c1,c2,c3), key = seed)
    __host__ __device__ static inline void gen(uint64_t seed, uint32_t c0, uint32_t c1,
                                               uint32_t c2, uint32_t c3, uint32_t out[4])
    {
        uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
        uint32_t c[4] = {c0, c1, c2, c3};
#pragma unroll
        for (int i = 0; i < 10; ++i) {
            round(c, k0, k1);
            k0 += W0; k1 += W1;
        }
        out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
    }
};

// uniform in (0,1]: never 0 so log() is finite
__host__ __device__ inline float u01(uint32_t x) { return ((x >> 8) + 1) * (1.0f / 16777216.0f); }

__device__ __forceinline__ float nanf_() { return __int_as_float(0x7fc00000); }

// ---- packed FP32 (Blackwell FFMA2): two independent IEEE fma.rn per instruction --------------
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi)
{
    u64 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(u64 v, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c)
{
    u64 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
__device__ __forceinline__ u64 fadd2(u64 a, u64 b)
{
    u64 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}

}  // namespace wcr
