// abi.cu - error reporting and device queries of the C ABI (include/qb200.h).
#include <stdarg.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void qb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int qb_sm_count() {
    static thread_local int cached_dev = -1, cached_sms = 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (dev != cached_dev) {
        int sms = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
        cached_dev = dev;
        cached_sms = sms;
    }
    return cached_sms;
}

extern "C" const char* qb_last_error(void) { return g_err; }

extern "C" int qb_version(void) { return 100; }

extern "C" int qb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* l2_bytes) {
    int n = 0;
    QB_CUDA(cudaGetDeviceCount(&n));
    QB_CHECK_ARG(device >= 0 && device < n, "qb_device_info: device %d out of range (%d devices)", device, n);
    int sms = 0, maj = 0, min = 0, l2 = 0;
    QB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    QB_CUDA(cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, device));
    QB_CUDA(cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, device));
    QB_CUDA(cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, device));
    if (sm_count) *sm_count = sms;
    if (cc_major) *cc_major = maj;
    if (cc_minor) *cc_minor = min;
    if (l2_bytes) *l2_bytes = (size_t)l2;
    if (maj != 10) {
        qb_set_error("qb_device_info: device %d is sm_%d%d; this library is built for sm_100a only", device, maj, min);
        return QB_ERR_UNSUPPORTED;
    }
    return QB_OK;
}
