// common.cuh — shared helpers for the forestslam_b200 C-ABI library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/forestslam_b200.h"

namespace fs {

// thread-local last-error text returned by fs_last_error_string()
void set_error(const char* fmt, ...);

inline int check_cuda(cudaError_t e, const char* what) {
    if (e != cudaSuccess) {
        set_error("%s: %s", what, cudaGetErrorString(e));
        return FS_ERR_CUDA;
    }
    return FS_OK;
}

#define FS_CUDA(call)                                         \
    do {                                                      \
        int _s = fs::check_cuda((call), #call);               \
        if (_s != FS_OK) return _s;                           \
    } while (0)

#define FS_LAUNCH_CHECK(name)                                 \
    do {                                                      \
        fs::count_launch();                                   \
        int _s = fs::check_cuda(cudaGetLastError(), name);    \
        if (_s != FS_OK) return _s;                           \
    } while (0)

#define FS_REQUIRE(cond, msg)                                 \
    do {                                                      \
        if (!(cond)) {                                        \
            fs::set_error("bad argument: %s", msg);           \
            return FS_ERR_BAD_ARG;                            \
        }                                                     \
    } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Opt a kernel into `bytes` of dynamic shared memory.  The attribute is sticky per device, so it is only raised when a
// larger request arrives: after warm-up no call is made, which keeps the launch sequence capturable in a CUDA graph.
// (Per-device bookkeeping: one slot per CUDA device, 64 is far above any single node.)
template <class K>
inline int ensure_dyn_smem(K kernel, int* granted /* [64], zero-initialised static */, size_t bytes) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    if ((int)bytes <= granted[dev]) return FS_OK;
    int s = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes),
                       "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
    if (s == FS_OK) granted[dev] = (int)bytes;
    return s;
}

int sm_count();
void count_launch();  // every kernel launch of the library bumps fs_launch_count()

// carve a workspace pointer
struct Carver {
    char* base;
    size_t off;
    explicit Carver(void* p) : base((char*)p), off(0) {}
    template <class T>
    T* take(size_t n) {
        off = align_up(off, 256);
        T* r = (T*)(base + off);
        off += n * sizeof(T);
        return r;
    }
};

}  // namespace fs
