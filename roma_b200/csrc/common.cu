// Error reporting, device checks and ABI bookkeeping for libroma_b200.so.
#include "common.cuh"
#include <atomic>

namespace roma {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

static std::atomic<unsigned long long> g_launches{0};

int check_launch(const char* what, int launches) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return (int)e;
    }
    g_launches.fetch_add((unsigned long long)launches, std::memory_order_relaxed);
    return 0;
}

int sm_count() {
    // immutable per-device attribute; cached per (thread, device) is overkill: cache per device id
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

}  // namespace roma

extern "C" int roma_abi_version(void) { return ROMA_B200_ABI_VERSION; }

extern "C" unsigned long long roma_kernel_launches(void) { return roma::g_launches.load(std::memory_order_relaxed); }

extern "C" const char* roma_last_error_string(void) { return roma::g_err; }

extern "C" int roma_check_device(int device) {
    int major = 0, minor = 0;
    cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
    if (e != cudaSuccess) return roma::fail((int)e, "roma_check_device: %s", cudaGetErrorString(e));
    e = cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device);
    if (e != cudaSuccess) return roma::fail((int)e, "roma_check_device: %s", cudaGetErrorString(e));
    if (major != 10)
        return roma::fail(ROMA_ERR_ARCH, "device %d is sm_%d%d; libroma_b200 is built for sm_100a only", device, major, minor);
    return ROMA_OK;
}
