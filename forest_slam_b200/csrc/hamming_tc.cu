// hamming_tc.cu — A2 on the tensor cores: 256-bit Hamming distance as a dense contraction.
//
// Same contract as hamming.cu (cv2.BFMatcher(NORM_HAMMING, crossCheck).match, stereo_slam.py:85, :234, :242; distance
// ORB_SLAM3/src/ORBmatcher.cc:2058-2074; k=2 ratio ORB_SLAM3/src/Frame.cc:1144-1151), different pipe.  hamming.cu runs
// at ~90 % of the POPC pipe (16 lanes/clk/SM), which caps it at ~7.7 us per 2048 x 2048 pair.  Here every bit becomes
// a +-1 fp16 operand, so
//     dot(a, b) = (#equal bits) - (#different bits) = 256 - 2 * hamming(a, b)
// is an EXACT integer in the fp32 accumulator of tcgen05.mma (256 products of +-1), and arg-min distance = arg-max dot.
// The contraction + row top-3 epilogue is k_float_top3_tc (float_top3.cu) unchanged, one pass per direction: its packed
// key (score bits << 12 | 4095 - column) orders by (distance, lowest index) exactly as BFMatcher does, the reverse
// direction gives every train row's best query for crossCheck, and the keys decode to exact integers, so there is no
// certification / rescoring stage.  This is not a reshaping of byte work into a GEMM for its own sake: the N0 x N1 x 256
// bit comparisons ARE a contraction, and the POPC pipe is the slower place to do them on this machine.
//
//   fs_hamming_prepare  : [F, n_max, 32] u8 -> operand bank [F, n_max + 1, 512] fp16 (layout of float_ops.cuh: hi = +-1,
//                         lo = 0, statistics row = |row|^2 = 256), once per FRAME;
//   fs_hamming_match_tc : k_float_top3_tc (both directions) + k_hamming_tc_finalize (decode, crossCheck, ratio, -1 fill).
#include <cuda.h>
#include <stdlib.h>
#include <cuda_fp16.h>
#include "float_ops.cuh"

namespace {

// thread = (frame, row, byte): 8 bits -> 8 halves (one 16-byte store into hi) + 8 zero halves into lo
__global__ void __launch_bounds__(256)
k_hamming_expand(const uint8_t* __restrict__ desc, int n_frames, int n_max, const int32_t* __restrict__ counts,
                 __half* __restrict__ operands) {
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t total = (size_t)n_frames * fs::op_rows(n_max) * 32;
    if (g >= total) return;
    const int byte = (int)(g & 31);
    const size_t frow = g >> 5;
    const int f = (int)(frow / fs::op_rows(n_max)), row = (int)(frow % fs::op_rows(n_max));
    __half* dst = operands + frow * fs::OP_ROW_HALFS;
    uint4 hi = make_uint4(0u, 0u, 0u, 0u);
    if (row == n_max) {
        // statistics row (float_ops.cuh: FrameStats): max |hi row|^2 = 256 exactly, lo part empty
        if (byte == 0) hi = make_uint4(__float_as_uint(256.f), 0u, 0u, 0u);
    } else if (row < min(counts[f], n_max)) {
        const uint32_t b = desc[((size_t)f * n_max + row) * 32 + byte];
        const uint32_t P = 0x3C00u, M = 0xBC00u;  // fp16 +1 / -1
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t lo16 = ((b >> (2 * i)) & 1u) ? P : M, hi16 = ((b >> (2 * i + 1)) & 1u) ? P : M;
            w[i] = lo16 | (hi16 << 16);
        }
        hi = make_uint4(w[0], w[1], w[2], w[3]);
    }
    *reinterpret_cast<uint4*>(dst + byte * 8) = hi;
    *reinterpret_cast<uint4*>(dst + 256 + byte * 8) = make_uint4(0u, 0u, 0u, 0u);
}

// key -> exact Hamming distance.  The key holds v' = fma(dot, sc, KEY_BIAS) rounded to 2^-24; consecutive integer dots
// are ~2000 ulps apart (sc ~ 2.4e-4), so rounding the decoded score recovers dot exactly.
__device__ __forceinline__ int key_distance(unsigned int key, const fs::PairScale& ps) {
    const int dot = __float2int_rn(fs::key_score(key, ps));
    return (256 - dot) >> 1;
}

// thread = (pair, query): same decisions and outputs as k_hamming_finalize (hamming.cu)
__global__ void __launch_bounds__(256)
k_hamming_tc_finalize(const void* __restrict__ operands, int n_max, const int32_t* __restrict__ counts,
                      const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b, int n_pairs,
                      const uint4* __restrict__ keys, int cross_check, float ratio, int32_t* __restrict__ matches0,
                      int32_t* __restrict__ dist0, int32_t* __restrict__ second0, int32_t* __restrict__ matches1,
                      int32_t* __restrict__ n_matches) {
    const int p = blockIdx.y;
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int fa = pair_a[p], fb = pair_b[p];
    const int n0 = min(counts[fa], n_max), n1 = min(counts[fb], n_max);
    bool ok = false;
    if (q < n_max) {
        const size_t o = (size_t)p * n_max + q;
        int m = -1, d = INT32_MAX, s = INT32_MAX;
        if (q < n0 && n1 > 0) {
            const fs::PairScale ps = fs::pair_scale(operands, n_max, fa, fb);
            const uint4 k = keys[o];  // direction 0: rows of a against b
            const int t = fs::key_col(k.x);
            d = key_distance(k.x, ps);
            // runner-up VALUE of the row = larger of (second chunk maximum, best chunk runner-up); keys order by score
            const unsigned int k2 = max(k.y >> 12, k.w >> 12);
            const bool has2 = k2 != 0u;
            if (has2) s = key_distance(k2 << 12, ps);
            ok = true;
            if (cross_check) {
                const uint4 kc = keys[((size_t)n_pairs + p) * n_max + t];  // direction 1: row t of b against a
                ok = fs::key_col(kc.x) == q;
            }
            if (ratio > 0.f) ok = ok && has2 && ((double)d < (double)s * (double)ratio);
            if (ok) {
                m = t;
                if (matches1) atomicMin((unsigned int*)matches1 + (size_t)p * n_max + t, (unsigned int)q);
            }
        }
        matches0[o] = m;
        if (dist0) dist0[o] = d;
        if (second0) second0[o] = s;
    }
    if (n_matches) {
        const unsigned bal = __ballot_sync(0xffffffffu, ok);
        if ((threadIdx.x & 31) == 0 && bal) atomicAdd(n_matches + p, __popc(bal));
    }
}

// Single-pass form: thread = (pair, train row t).  The slab results of column t hold, per slab of 16 query rows,
// (129 + dot / 2) << 4 | 15 - m  of the slab's best row (row = 32 * (slab >> 1) + 2 m + (slab & 1)); the largest
// (value, lowest row) over the slabs is the train row's best query exactly as the reverse pass would report it.
// Written in the key format of direction 1 (only the column field is read by the finalize kernel).
__global__ void __launch_bounds__(256)
k_hamming_colmerge(int n_max, const int32_t* __restrict__ counts, const int32_t* __restrict__ pair_a,
                   const int32_t* __restrict__ pair_b, const uint32_t* __restrict__ colraw, int cpitch, int nslab,
                   uint4* __restrict__ keys1) {
    const int p = blockIdx.y, t = blockIdx.x * 256 + threadIdx.x;
    const int n0 = min(counts[pair_a[p]], n_max), n1 = min(counts[pair_b[p]], n_max);
    if (t >= n1) return;
    const int nsl = ((n0 + 31) >> 5) * 2;
    const uint32_t* src = colraw + (size_t)p * nslab * cpitch + t;
    uint32_t best = 0;
    for (int s0 = 0; s0 < nsl; s0 += 8) {
        uint32_t x[8];
#pragma unroll
        for (int u = 0; u < 8; u++) x[u] = (s0 + u < nsl) ? __ldg(src + (size_t)(s0 + u) * cpitch) : 0u;
#pragma unroll
        for (int u = 0; u < 8; u++) {
            const uint32_t h = x[u] & 0xFFFFu;
            const int sl = s0 + u;
            const int row = (sl >> 1) * 32 + 2 * (15 - (int)(h & 15u)) + (sl & 1);
            best = max(best, h ? (((h >> 4) << 12) | (uint32_t)(4095 - row)) : 0u);
        }
    }
    keys1[(size_t)p * n_max + t] = make_uint4(best, 0u, 0u, 0u);
}

// FS_TUNE_HAMMING_SINGLE_PASS=0 keeps fs_hamming_match_tc on the two-direction kernel (A/B measurements, tests); read on
// every call (host side only)
bool hamming_single_pass_enabled() {
    const char* e = getenv("FS_TUNE_HAMMING_SINGLE_PASS");
    return !(e && e[0] == '0');
}

}  // namespace

extern "C" size_t fs_hamming_operand_bytes(int n_frames, int n_max) { return fs_float_operand_bytes(n_frames, n_max); }

extern "C" int fs_hamming_prepare(const uint8_t* desc, int n_frames, int n_max, const int32_t* counts, void* operands,
                                  fs_stream_t stream_) {
    FS_REQUIRE(n_frames >= 0 && n_max >= 0, "negative size");
    if (n_frames == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(desc && counts && operands, "null pointer");
    FS_REQUIRE(((uintptr_t)operands & 15) == 0, "operands must be 16-byte aligned");
    const size_t total = (size_t)n_frames * fs::op_rows(n_max) * 32;
    FS_REQUIRE((total + 255) / 256 < (1ull << 31), "bank too large for one launch");
    k_hamming_expand<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream_>>>(
        desc, n_frames, n_max, counts, static_cast<__half*>(operands));
    FS_LAUNCH_CHECK("k_hamming_expand");
    return FS_OK;
}

extern "C" size_t fs_hamming_match_tc_workspace_bytes(int n_pairs, int n_max) {
    if (n_pairs <= 0 || n_max <= 0) return 256;
    // keys of both directions + (single-pass form) the slab results of the column side
    return fs::align_up(2 * (size_t)n_pairs * n_max * sizeof(uint4), 256) +
           fs::align_up(fs::float_top3_colraw_elems(n_pairs, n_max) * 4, 256) + 256;
}

extern "C" int fs_hamming_match_tc(const void* operands, int n_frames, int n_max, const int32_t* counts,
                                   const int32_t* pair_a, const int32_t* pair_b, int n_pairs, int cross_check,
                                   float ratio, int32_t* matches0, int32_t* dist0, int32_t* second0,
                                   int32_t* matches1, int32_t* n_matches, void* workspace, size_t workspace_bytes,
                                   fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_frames >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(operands && counts && pair_a && pair_b && matches0, "null pointer");
    FS_REQUIRE(((uintptr_t)operands & 15) == 0, "operands must be 16-byte aligned");
    if (n_max > fs::KEY_MAX_COLS) {
        fs::set_error("fs_hamming_match_tc: n_max %d > %d (use fs_hamming_match)", n_max, fs::KEY_MAX_COLS);
        return FS_ERR_UNSUPPORTED;
    }
    FS_REQUIRE((size_t)n_pairs * n_max < (1ull << 31), "n_pairs * n_max too large for one call");
    const size_t need = fs_hamming_match_tc_workspace_bytes(n_pairs, n_max);
    if (!workspace || workspace_bytes < need) {
        fs::set_error("fs_hamming_match_tc: workspace %zu < %zu", workspace_bytes, need);
        return FS_ERR_WORKSPACE;
    }
    CUtensorMap tmap, tmap_half;
    int st = fs::encode_operand_map(operands, n_frames, n_max, &tmap);
    if (st == FS_OK) st = fs::encode_operand_map(operands, n_frames, n_max, &tmap_half, 64);
    if (st != FS_OK) return st;
    fs::Carver ws(workspace);
    const size_t n = (size_t)n_pairs * n_max;
    uint4* keys = ws.take<uint4>(2 * n);
    if (matches1) FS_CUDA(cudaMemsetAsync(matches1, 0xFF, n * sizeof(int32_t), stream));
    if (n_matches) FS_CUDA(cudaMemsetAsync(n_matches, 0, (size_t)n_pairs * sizeof(int32_t), stream));
    // crossCheck: single pass (direction 0 on the tensor cores, the train rows' best queries from the slab results);
    // without it only direction 0 is needed at all
    uint32_t* colraw = ws.take<uint32_t>(fs::float_top3_colraw_elems(n_pairs, n_max));
    const bool sp = cross_check && hamming_single_pass_enabled();
    st = fs::launch_float_top3(tmap, tmap_half, operands, n_max, counts, pair_a, pair_b, n_pairs, keys, stream, 0, nullptr,
                               sp ? colraw : nullptr, true, cross_check ? 2 : 1);
    if (st != FS_OK) return st;
    if (sp) {
        const int cpitch = (n_max + 127) / 128 * 128, nslab = (n_max + 31) / 32 * 2;
        for (int p0 = 0; p0 < n_pairs; p0 += 32768) {
            const int np = min(32768, n_pairs - p0);
            k_hamming_colmerge<<<dim3((n_max + 255) / 256, np), 256, 0, stream>>>(
                n_max, counts, pair_a + p0, pair_b + p0, colraw + (size_t)p0 * nslab * cpitch, cpitch, nslab,
                keys + n + (size_t)p0 * n_max);
            FS_LAUNCH_CHECK("k_hamming_colmerge");
        }
    }
    for (int p0 = 0; p0 < n_pairs; p0 += 32768) {
        const int np = min(32768, n_pairs - p0);
        const size_t o = (size_t)p0 * n_max;
        dim3 grid((n_max + 255) / 256, np);
        // keys of direction 1 sit n_pairs * n_max entries after direction 0: pass the slab-shifted base and the TOTAL
        // pair count so that the kernel's ((n_pairs + p) * n_max) offset still lands in direction 1
        k_hamming_tc_finalize<<<grid, 256, 0, stream>>>(operands, n_max, counts, pair_a + p0, pair_b + p0, n_pairs,
                                                        keys + o, cross_check, ratio, matches0 + o,
                                                        dist0 ? dist0 + o : nullptr, second0 ? second0 + o : nullptr,
                                                        matches1 ? matches1 + o : nullptr,
                                                        n_matches ? n_matches + p0 : nullptr);
        FS_LAUNCH_CHECK("k_hamming_tc_finalize");
    }
    return FS_OK;
}
