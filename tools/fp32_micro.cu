// fp32_micro.cu -- B200 FP32-pipe microbenchmarks that decide the inner loop of
// the fused scoring kernel (DESIGN.md "Scoring kernel: choosing the test form").
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -lineinfo \
//        -o tools/fp32_micro tools/fp32_micro.cu && tools/fp32_micro
//
// Part 1: raw pipe rates (FFMA, FFMA2, FFMA+FSETP mixes) -> measured FP32 peak.
// Part 2: candidate point-in-cube inner loops, thread = T proposals, points
//         broadcast from shared memory; reports point-tests/s.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { \
    printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1);} } while (0)

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
__device__ __forceinline__ float fmax3abs(float a, float b, float c)
{
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(fabsf(a)), "f"(fabsf(b)), "f"(fabsf(c)));
    return r;
}

// ------------------------------------------------------------------ part 1 --
template <int MODE>
__global__ void __launch_bounds__(256) pipe_kernel(float *out, int iters, float s)
{
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
    const float m = 1.0f + s * 1e-7f, c = s * 1e-7f;
    int cnt = 0;
    if (MODE == 0) {  // 16 independent FFMA chains
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
    } else if (MODE == 1) {  // 8 independent FFMA2 chains
        u64 p[8];
        const u64 mm = pack2(m, m), cc = pack2(c, c);
#pragma unroll
        for (int i = 0; i < 8; ++i) p[i] = pack2(a[2 * i], a[2 * i + 1]);
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = ffma2(p[i], mm, cc);
#pragma unroll
        for (int i = 0; i < 8; ++i) unpack2(p[i], a[2 * i], a[2 * i + 1]);
    } else if (MODE == 2) {  // 12 FFMA + 4 FSETP/predicated add per step
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 12; ++i) a[i] = fmaf(a[i], m, c);
#pragma unroll
            for (int i = 0; i < 4; ++i) cnt += (fabsf(a[i]) <= a[12 + i]) ? 1 : 0;
        }
    } else if (MODE == 3) {  // 6 FFMA2 (=12 FMA) + 4 FSETP/add
        u64 p[6];
        const u64 mm = pack2(m, m), cc = pack2(c, c);
#pragma unroll
        for (int i = 0; i < 6; ++i) p[i] = pack2(a[2 * i], a[2 * i + 1]);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 6; ++i) p[i] = ffma2(p[i], mm, cc);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float lo, hi; unpack2(p[i], lo, hi);
                cnt += (fabsf(lo) <= a[12 + i]) ? 1 : 0;
            }
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) unpack2(p[i], a[2 * i], a[2 * i + 1]);
    } else if (MODE == 4) {  // FSETP + add only (16 per step)
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 16; ++i) cnt += (fabsf(a[i]) <= m * (float)(it + i)) ? 1 : 0;
    } else if (MODE == 5) {  // FMNMX3 only
        for (int it = 0; it < iters; ++it)
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmax3abs(a[i], a[(i + 1) & 15], m);
    }
    float sum = (float)cnt;
#pragma unroll
    for (int i = 0; i < 16; ++i) sum += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
}

// ------------------------------------------------------------------ part 2 --
// proposal state layouts are synthetic here (values irrelevant, op mix exact)
constexpr int NP = 2048;  // points per smem tile

// V1: pre-dotted, 9 FFMA + 3 FSETP + add.  points float4 (x,y,z,-)
template <int T>
__global__ void __launch_bounds__(256) pic_v1(const float4 *pts, int reps, int *out)
{
    __shared__ float4 sp[NP];
    for (int i = threadIdx.x; i < NP; i += 256) sp[i] = pts[i];
    __syncthreads();
    float r[T][9], o[T][3], h[T][3];
    int cnt[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
        cnt[t] = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) r[t][k] = 0.01f * (threadIdx.x + t + k);
#pragma unroll
        for (int k = 0; k < 3; ++k) { o[t][k] = 0.5f + k; h[t][k] = 0.3f + 0.1f * t; }
    }
    for (int rep = 0; rep < reps; ++rep)
#pragma unroll 2
        for (int j = 0; j < NP; ++j) {
            const float4 p = sp[j];
#pragma unroll
            for (int t = 0; t < T; ++t) {
                const float qx = fmaf(r[t][0], p.x, fmaf(r[t][1], p.y, fmaf(r[t][2], p.z, -o[t][0])));
                const float qy = fmaf(r[t][3], p.x, fmaf(r[t][4], p.y, fmaf(r[t][5], p.z, -o[t][1])));
                const float qz = fmaf(r[t][6], p.x, fmaf(r[t][7], p.y, fmaf(r[t][8], p.z, -o[t][2])));
                cnt[t] += (fabsf(qx) <= h[t][0] && fabsf(qy) <= h[t][1] && fabsf(qz) <= h[t][2]) ? 1 : 0;
            }
        }
    int s = 0;
#pragma unroll
    for (int t = 0; t < T; ++t) s += cnt[t];
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

// V2: row-factored, 6 FFMA + 3 FSETP + add.  points float2 (a_u, z)
template <int T>
__global__ void __launch_bounds__(256) pic_v2(const float4 *pts, int reps, int *out)
{
    __shared__ float2 sp[NP];
    for (int i = threadIdx.x; i < NP; i += 256) sp[i] = make_float2(pts[i].x, pts[i].z);
    __syncthreads();
    float r[T][3], g[T][3], o[T][3], h[T][3];
    int cnt[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
        cnt[t] = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            r[t][k] = 0.01f * (threadIdx.x + t + k); g[t][k] = 0.02f * (t + k);
            o[t][k] = 0.5f + k; h[t][k] = 0.3f + 0.1f * t;
        }
    }
    for (int rep = 0; rep < reps; ++rep)
#pragma unroll 2
        for (int j = 0; j < NP; j += 2) {
            const float4 pp = *reinterpret_cast<const float4 *>(&sp[j]);
            const float av[2] = {pp.x, pp.z}, zv[2] = {pp.y, pp.w};
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int t = 0; t < T; ++t) {
                    const float qx = fmaf(zv[e], fmaf(r[t][0], av[e], g[t][0]), -o[t][0]);
                    const float qy = fmaf(zv[e], fmaf(r[t][1], av[e], g[t][1]), -o[t][1]);
                    const float qz = fmaf(zv[e], fmaf(r[t][2], av[e], g[t][2]), -o[t][2]);
                    cnt[t] += (fabsf(qx) <= h[t][0] && fabsf(qy) <= h[t][1] && fabsf(qz) <= h[t][2]) ? 1 : 0;
                }
        }
    int s = 0;
#pragma unroll
    for (int t = 0; t < T; ++t) s += cnt[t];
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

// V3: row-factored + FFMA2 over proposal pairs. points (a,a,z,z) float4.
// CMP=0: 3 FSETP chain; CMP=1: unit-cube form, FMNMX3(|.|) <= 1
template <int T2, int CMP>
__global__ void __launch_bounds__(256) pic_v3(const float4 *pts, int reps, int *out)
{
    __shared__ float4 sp[NP];
    for (int i = threadIdx.x; i < NP; i += 256) sp[i] = make_float4(pts[i].x, pts[i].x, pts[i].z, pts[i].z);
    __syncthreads();
    u64 r[T2][3], g[T2][3], no[T2][3];
    float h[T2][2][3];
    int cnt[T2][2];
#pragma unroll
    for (int t = 0; t < T2; ++t) {
        cnt[t][0] = cnt[t][1] = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            r[t][k] = pack2(0.01f * (threadIdx.x + t + k), 0.011f * (threadIdx.x + t + k));
            g[t][k] = pack2(0.02f * (t + k), 0.021f * (t + k));
            no[t][k] = pack2(-0.5f - k, -0.6f - k);
            h[t][0][k] = 0.3f + 0.1f * t; h[t][1][k] = 0.35f + 0.1f * t;
        }
    }
    for (int rep = 0; rep < reps; ++rep)
#pragma unroll 2
        for (int j = 0; j < NP; ++j) {
            const float4 p = sp[j];
            const u64 aa = pack2(p.x, p.y), zz = pack2(p.z, p.w);
#pragma unroll
            for (int t = 0; t < T2; ++t) {
                float q[3][2];
#pragma unroll
                for (int k = 0; k < 3; ++k) unpack2(ffma2(zz, ffma2(r[t][k], aa, g[t][k]), no[t][k]), q[k][0], q[k][1]);
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    if (CMP == 0)
                        cnt[t][e] += (fabsf(q[0][e]) <= h[t][e][0] && fabsf(q[1][e]) <= h[t][e][1] &&
                                      fabsf(q[2][e]) <= h[t][e][2]) ? 1 : 0;
                    else
                        cnt[t][e] += (fmax3abs(q[0][e], q[1][e], q[2][e]) <= 1.0f) ? 1 : 0;
                }
            }
        }
    int s = 0;
#pragma unroll
    for (int t = 0; t < T2; ++t) s += cnt[t][0] + cnt[t][1];
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

// V4: pre-dotted (9 FMA) + FFMA2 over proposal pairs. points (x,x,y,y),(z,z,-,-)
template <int T2, int CMP>
__global__ void __launch_bounds__(256) pic_v4(const float4 *pts, int reps, int *out)
{
    __shared__ float4 sp[NP];
    __shared__ float2 sz[NP];
    for (int i = threadIdx.x; i < NP; i += 256) {
        sp[i] = make_float4(pts[i].x, pts[i].x, pts[i].y, pts[i].y);
        sz[i] = make_float2(pts[i].z, pts[i].z);
    }
    __syncthreads();
    u64 r[T2][9], no[T2][3];
    float h[T2][2][3];
    int cnt[T2][2];
#pragma unroll
    for (int t = 0; t < T2; ++t) {
        cnt[t][0] = cnt[t][1] = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) r[t][k] = pack2(0.01f * (threadIdx.x + t + k), 0.011f * (threadIdx.x + t + k));
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            no[t][k] = pack2(-0.5f - k, -0.6f - k);
            h[t][0][k] = 0.3f + 0.1f * t; h[t][1][k] = 0.35f + 0.1f * t;
        }
    }
    for (int rep = 0; rep < reps; ++rep)
#pragma unroll 2
        for (int j = 0; j < NP; ++j) {
            const float4 p = sp[j];
            const float2 pz = sz[j];
            const u64 xx = pack2(p.x, p.y), yy = pack2(p.z, p.w), zz = pack2(pz.x, pz.y);
#pragma unroll
            for (int t = 0; t < T2; ++t) {
                float q[3][2];
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    unpack2(ffma2(r[t][3 * k], xx, ffma2(r[t][3 * k + 1], yy, ffma2(r[t][3 * k + 2], zz, no[t][k]))),
                            q[k][0], q[k][1]);
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    if (CMP == 0)
                        cnt[t][e] += (fabsf(q[0][e]) <= h[t][e][0] && fabsf(q[1][e]) <= h[t][e][1] &&
                                      fabsf(q[2][e]) <= h[t][e][2]) ? 1 : 0;
                    else
                        cnt[t][e] += (fmax3abs(q[0][e], q[1][e], q[2][e]) <= 1.0f) ? 1 : 0;
                }
            }
        }
    int s = 0;
#pragma unroll
    for (int t = 0; t < T2; ++t) s += cnt[t][0] + cnt[t][1];
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

template <typename F>
static float time_ms(F launch, int n = 5)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    launch(); launch();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int i = 0; i < n; ++i) {
        CK(cudaEventRecord(a));
        launch();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main()
{
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const int sms = prop.multiProcessorCount;
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", prop.name, sms, clk);
    const int grid = sms * 8;
    float *fo; int *io; float4 *pts;
    CK(cudaMalloc(&fo, (size_t)grid * 256 * 4)); CK(cudaMalloc(&io, (size_t)grid * 256 * 4));
    std::vector<float4> hp(NP);
    for (int i = 0; i < NP; ++i) hp[i] = make_float4(0.001f * i - 1.f, 0.002f * i - 2.f, 1.f + 0.001f * i, 0.f);
    CK(cudaMalloc(&pts, NP * sizeof(float4)));
    CK(cudaMemcpy(pts, hp.data(), NP * sizeof(float4), cudaMemcpyHostToDevice));

    const int iters = 4096;
    const char *names[6] = {"ffma_x16", "ffma2_x8", "12ffma+4fsetp", "6ffma2+4fsetp", "fsetp_x16", "fmnmx3_x16"};
    const double fma_per_it[6] = {16, 16, 12, 12, 0, 0};
    const double inst_per_it[6] = {16, 8, 16 + 4, 6 + 4 + 4, 32, 16};
    for (int m = 0; m < 6; ++m) {
        float ms = 0;
        switch (m) {
        case 0: ms = time_ms([&] { pipe_kernel<0><<<grid, 256>>>(fo, iters, 1.f); }); break;
        case 1: ms = time_ms([&] { pipe_kernel<1><<<grid, 256>>>(fo, iters, 1.f); }); break;
        case 2: ms = time_ms([&] { pipe_kernel<2><<<grid, 256>>>(fo, iters, 1.f); }); break;
        case 3: ms = time_ms([&] { pipe_kernel<3><<<grid, 256>>>(fo, iters, 1.f); }); break;
        case 4: ms = time_ms([&] { pipe_kernel<4><<<grid, 256>>>(fo, iters, 1.f); }); break;
        case 5: ms = time_ms([&] { pipe_kernel<5><<<grid, 256>>>(fo, iters, 1.f); }); break;
        }
        const double thr = (double)grid * 256 * iters;
        printf("{\"pipe\": \"%s\", \"ms\": %.4f, \"tflops\": %.2f, \"src_ops_per_clk_per_sm\": %.1f}\n", names[m], ms,
               2.0 * fma_per_it[m] * thr / (ms * 1e-3) / 1e12,
               inst_per_it[m] * thr / (ms * 1e-3) / ((double)clk * 1e3) / sms);
    }

    const int reps = 8;
#define RUN_PIC(label, kern, T)                                                             \
    {                                                                                       \
        float ms = time_ms([&] { kern<<<grid, 256>>>(pts, reps, io); });                    \
        double tests = (double)grid * 256 * (T) * NP * reps;                                \
        printf("{\"pic\": \"%s\", \"T\": %d, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", \
               label, (T), ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); \
    }
    RUN_PIC("v1_9ffma_fsetp", pic_v1<2>, 2);
    RUN_PIC("v1_9ffma_fsetp", pic_v1<4>, 4);
    RUN_PIC("v1_9ffma_fsetp", pic_v1<8>, 8);
    RUN_PIC("v2_6ffma_fsetp", pic_v2<2>, 2);
    RUN_PIC("v2_6ffma_fsetp", pic_v2<4>, 4);
    RUN_PIC("v2_6ffma_fsetp", pic_v2<8>, 8);
    { float ms = time_ms([&] { pic_v3<1, 0><<<grid, 256>>>(pts, reps, io); }); double tests = (double)grid * 256 * 2 * NP * reps;
      printf("{\"pic\": \"v3_6ffma2_fsetp\", \"T\": 2, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); }
    { float ms = time_ms([&] { pic_v3<2, 0><<<grid, 256>>>(pts, reps, io); }); double tests = (double)grid * 256 * 4 * NP * reps;
      printf("{\"pic\": \"v3_6ffma2_fsetp\", \"T\": 4, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); }
    { float ms = time_ms([&] { pic_v3<4, 0><<<grid, 256>>>(pts, reps, io); }); double tests = (double)grid * 256 * 8 * NP * reps;
      printf("{\"pic\": \"v3_6ffma2_fsetp\", \"T\": 8, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); }
    { float ms = time_ms([&] { pic_v3<2, 1><<<grid, 256>>>(pts, reps, io); }); double tests = (double)grid * 256 * 4 * NP * reps;
      printf("{\"pic\": \"v3_6ffma2_fmnmx3\", \"T\": 4, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); }
    { float ms = time_ms([&] { pic_v3<4, 1><<<grid, 256>>>(pts, reps, io); }); double tests = (double)grid * 256 * 8 * NP * reps;
      printf("{\"pic\": \"v3_6ffma2_fmnmx3\", \"T\": 8, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); }
    { float ms = time_ms([&] { pic_v4<2, 0><<<grid, 256>>>(pts, reps, io); }); double tests = (double)grid * 256 * 4 * NP * reps;
      printf("{\"pic\": \"v4_9ffma2_fsetp\", \"T\": 4, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); }
    { float ms = time_ms([&] { pic_v4<2, 1><<<grid, 256>>>(pts, reps, io); }); double tests = (double)grid * 256 * 4 * NP * reps;
      printf("{\"pic\": \"v4_9ffma2_fmnmx3\", \"T\": 4, \"ms\": %.4f, \"gtests_per_s\": %.1f, \"alg_tflops_18\": %.2f}\n", ms, tests / (ms * 1e-3) / 1e9, 18.0 * tests / (ms * 1e-3) / 1e12); }
    return 0;
}
