// float_tc.cu — A1 on the 5th-generation tensor cores (tcgen05 + TMEM), with a fused row top-2 epilogue.
// (operand preparation below; the tcgen05 contraction kernel lands in the next milestone)
#include <cuda_bf16.h>
#include "float_common.cuh"

namespace {

// [F, 256, n_max] fp32 channel-major  ->  [F, n_max, 512] bf16, row = [hi(256) | lo(256)],
// hi = bf16(x), lo = bf16(x - hi).  hi*hi + hi*lo + lo*hi reproduces the fp32 product to ~2^-16 relative.
// Transposes through shared memory so that both the read (along n) and the write (along k) are coalesced.
__global__ void __launch_bounds__(256)
k_float_prepare(const float* __restrict__ desc, int n_max, const int32_t* __restrict__ counts,
                __nv_bfloat16* __restrict__ ops) {
    __shared__ float tile[32][33];
    const int f = blockIdx.z;
    const int n0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    const int cnt = min(counts[f], n_max);
    const float* src = desc + (size_t)f * FS_FLOAT_DIM * n_max;
    __nv_bfloat16* dst = ops + (size_t)f * n_max * 512;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int r = ty; r < 32; r += 8) {
        const int n = n0 + tx;
        tile[r][tx] = (n < cnt) ? src[(size_t)(k0 + r) * n_max + n] : 0.f;
    }
    __syncthreads();
    for (int r = ty; r < 32; r += 8) {
        const int n = n0 + r;
        if (n >= n_max) continue;
        const float x = tile[tx][r];
        const __nv_bfloat16 hi = __float2bfloat16_rn(x);
        const __nv_bfloat16 lo = __float2bfloat16_rn(x - __bfloat162float(hi));
        dst[(size_t)n * 512 + k0 + tx] = hi;
        dst[(size_t)n * 512 + 256 + k0 + tx] = lo;
    }
}

}  // namespace

extern "C" int fs_float_prepare(const float* desc_f32, int n_frames, int n_max, const int32_t* counts,
                                void* operands_bf16, fs_stream_t stream) {
    FS_REQUIRE(n_frames >= 0 && n_max >= 0, "negative size");
    if (n_frames == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(desc_f32 && counts && operands_bf16, "null pointer");
    FS_REQUIRE(n_frames <= 65535, "n_frames too large for one launch");
    dim3 grid((n_max + 31) / 32, FS_FLOAT_DIM / 32, n_frames);
    k_float_prepare<<<grid, 256, 0, (cudaStream_t)stream>>>(desc_f32, n_max, counts, (__nv_bfloat16*)operands_bf16);
    FS_LAUNCH_CHECK("k_float_prepare");
    return FS_OK;
}

extern "C" int fs_float_match(const void*, int, int, const int32_t*, const int32_t*, const int32_t*, int, float,
                              float, float, int32_t*, float*, float*, int32_t*, int32_t*, void*, size_t,
                              fs_stream_t) {
    fs::set_error("fs_float_match: tcgen05 kernel not built yet");
    return FS_ERR_UNSUPPORTED;
}
