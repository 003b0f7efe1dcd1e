// abi.cu -- error state, device check, launch bookkeeping for libwcr_b200.
#include <stdarg.h>
#include <string.h>
#include <atomic>
#include "common.cuh"

namespace wcr {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};  // process-wide: autograd runs backward on its own thread

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

static int g_checked_dev = -1;
static int g_checked_rc = WCR_OK;
static int g_sm_count = 0;

int check_device()
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        set_error("cudaGetDevice: %s", cudaGetErrorString(e));
        return WCR_ELAUNCH;
    }
    if (dev == g_checked_dev) {
        if (g_checked_rc != WCR_OK) set_error("device %d is not sm_100 (B200)", dev);
        return g_checked_rc;
    }
    int major = 0, minor = 0, sms = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    g_checked_dev = dev;
    g_sm_count = sms;
    if (major != 10 || minor != 0) {
        g_checked_rc = WCR_EARCH;
        set_error("device %d is sm_%d%d; libwcr_b200 only runs on sm_100 (B200)", dev, major, minor);
    } else {
        g_checked_rc = WCR_OK;
    }
    return g_checked_rc;
}

int sm_count()
{
    if (g_sm_count == 0) check_device();
    return g_sm_count > 0 ? g_sm_count : 148;
}

}  // namespace wcr

extern "C" {

int wcr_abi_version(void) { return WCR_ABI_VERSION; }
const char *wcr_last_error(void) { return wcr::g_err; }
int wcr_device_check(void) { return wcr::check_device(); }
int64_t wcr_launch_count(void) { return wcr::g_launches.load(std::memory_order_relaxed); }
void wcr_reset_launch_count(void) { wcr::g_launches.store(0, std::memory_order_relaxed); }

}  // extern "C"

// FP32 pipe probe: 16 independent FFMA chains per thread. bench.py times it with
// CUDA events to get the FP32 peak of THIS device under THIS run's clocks (the
// roofline denominator of the scoring kernel; MEASURED_PEAKS.json has no FP32 figure).
namespace wcr {
__global__ void __launch_bounds__(256) ffma_probe_kernel(float *out, int iters, float s)
{
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
    const float m = 1.0f + s * 1e-7f, c = s * 1e-7f;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) sum += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
}
}  // namespace wcr

extern "C" int64_t wcr_probe_ffma(float *out, int64_t out_elems, int iters, void *stream)
{
    if (wcr::check_device() != WCR_OK) return WCR_EARCH;
    const int grid = wcr::sm_count() * 8;
    if (!out || out_elems < (int64_t)grid * 256 || iters < 1) {
        wcr::set_error("wcr_probe_ffma: out must hold %d floats", grid * 256);
        return WCR_EINVAL;
    }
    wcr::ffma_probe_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(out, iters, 1.f);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { wcr::set_error("ffma_probe_kernel: %s", cudaGetErrorString(e)); return WCR_ELAUNCH; }
    return (int64_t)grid * 256 * (int64_t)iters * 16 * 2;  // FLOPs enqueued
}

extern "C" int64_t wcr_sizeof(int which)
{
    switch (which) {
    case 0: return (int64_t)sizeof(WcrImage);
    case 1: return (int64_t)sizeof(WcrBox);
    case 2: return (int64_t)sizeof(WcrSampleParams);
    case 3: return (int64_t)sizeof(WcrScoreParams);
    case 4: return (int64_t)sizeof(WcrLossParams);
    case 5: return (int64_t)sizeof(WcrLossIO);
    default: return 0;
    }
}
