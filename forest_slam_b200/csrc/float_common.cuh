// float_common.cuh — pieces shared by the two A1 (float descriptor) matchers:
// the exact fp32 CUDA-core kernel (float_fp32.cu) and the tcgen05 tensor-core kernel (float_tc.cu).
#pragma once
#include "common.cuh"

namespace fs {

// (value, index) ordering of superglue.py:268 made deterministic: larger value wins, lower index wins ties
__device__ __forceinline__ bool better(float a, int ia, float b, int ib) {
    return a > b || (a == b && ia < ib);
}

struct Top2 {
    float best, second;
    int idx;
    __device__ __forceinline__ void init() {
        best = -INFINITY;
        second = -INFINITY;
        idx = 0x7fffffff;
    }
    // candidate with a column index that may be smaller or larger than the current best's
    __device__ __forceinline__ void push(float v, int i) {
        if (better(v, i, best, idx)) {
            second = best;
            best = v;
            idx = i;
        } else {
            second = fmaxf(second, v);
        }
    }
    __device__ __forceinline__ void merge(float ob, float os, int oi) {
        if (better(ob, oi, best, idx)) {
            second = fmaxf(best, os);
            best = ob;
            idx = oi;
        } else {
            second = fmaxf(second, ob);
        }
    }
};

// per-direction scratch: best value / runner-up value / best index for every row of every pair
struct FloatRowTop2 {
    float* best;
    float* second;
    int32_t* idx;
};

// mutual check + threshold + ratio + -1 fill (superglue.py:270-278 on the raw contraction)
int launch_float_finalize(int n_max, const int32_t* counts, const int32_t* pair_a, const int32_t* pair_b,
                          int n_pairs, FloatRowTop2 d0, const int32_t* idx1, float scale, float threshold,
                          float ratio, int32_t* matches0, float* score0, float* second0, int32_t* matches1,
                          int32_t* n_matches, cudaStream_t stream, int cross_check = 1);

}  // namespace fs
