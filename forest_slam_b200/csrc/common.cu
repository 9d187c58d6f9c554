// common.cu — error strings, version, device queries for the C ABI (include/forestslam_b200.h).
#include <stdarg.h>
#include <atomic>
#include "common.cuh"

namespace fs {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
long long launches() { return g_launches.load(std::memory_order_relaxed); }

int sm_count() {
    static int cached = 0;
    if (cached) return cached;
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
    cached = n;
    return n;
}

}  // namespace fs

extern "C" {

int fs_version(void) { return 100; /* 0.1.0 */ }

const char* fs_last_error_string(void) { return fs::g_err; }

const char* fs_status_string(int s) {
    switch (s) {
        case FS_OK: return "FS_OK";
        case FS_ERR_BAD_ARG: return "FS_ERR_BAD_ARG (bad argument)";
        case FS_ERR_WORKSPACE: return "FS_ERR_WORKSPACE (workspace too small)";
        case FS_ERR_CUDA: return "FS_ERR_CUDA (CUDA error)";
        case FS_ERR_UNSUPPORTED: return "FS_ERR_UNSUPPORTED (unsupported size or device)";
        default: return "unknown status";
    }
}

int fs_device_sm_count(void) { return fs::sm_count(); }

long long fs_launch_count(void) { return fs::launches(); }

}  // extern "C"
