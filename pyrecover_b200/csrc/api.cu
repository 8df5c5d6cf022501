// Error plumbing and device queries of libpyrecover_b200 (C ABI: include/pyrecover_b200.h).
#include <stdarg.h>

#include <atomic>

#include "common.cuh"

static thread_local char g_err[1024] = "";
static std::atomic<uint64_t> g_launches{0};

void pr_count_launch(int n) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

void pr_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int pr_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

extern "C" const char *pr_last_error(void) { return g_err; }

extern "C" int pr_abi_version(void) { return PR_ABI_VERSION; }

extern "C" int pr_device_info(int *sm_count, int *cc_major, int *cc_minor) {
    int dev = 0;
    PR_CUDA(cudaGetDevice(&dev));
    int sm = 0, major = 0, minor = 0;
    PR_CUDA(cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev));
    PR_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    PR_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
    if (sm_count) *sm_count = sm;
    if (cc_major) *cc_major = major;
    if (cc_minor) *cc_minor = minor;
    return PR_OK;
}

extern "C" uint64_t pr_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
