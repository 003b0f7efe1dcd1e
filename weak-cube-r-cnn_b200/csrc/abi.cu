// abi.cu -- error state, device check, launch bookkeeping for libwcr_b200.
#include <stdarg.h>
#include <string.h>
#include "common.cuh"

namespace wcr {

static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void count_launch(int n) { g_launches += n; }

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
int64_t wcr_launch_count(void) { return wcr::g_launches; }
void wcr_reset_launch_count(void) { wcr::g_launches = 0; }

}  // extern "C"
