// float_tc.cu — A1 on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), with a fused row top-2 / column top-1
// epilogue.  Replaces the tail of SuperGlue.forward that mono_slam.py:100-103 / stereo_slam.py:210-213 consume:
//   scores = einsum('bdn,bdm->bnm', d0, d1)   superglue.py:259      (dense N0 x N1 x 256 contraction)
//   max0, max1 = scores.max(2), scores.max(1) ; mutual check ; threshold ; -1 fill        superglue.py:268-278
//
// Precision: every fp32 descriptor component x is split into two fp16 numbers, hi = fp16(x), lo = fp16(x - hi)
// (|x| <= 1, so lo is an fp16 subnormal carrying 2^-25 absolute precision), and the score is accumulated in fp32 as
// hi.hi + lo.hi + hi.lo — three kind::f16 MMAs per k-step into the SAME TMEM accumulator.  The dropped lo.lo term is
// below 2^-30; the result is fp32-grade (~1e-7), so arg-max / ratio decisions match the exact-fp32 kernel on every row
// the oracle does not flag as a numerical tie.
//
// One CTA = one (frame pair, 128-query tile).  Warp roles (192 threads):
//   warp 0 : TMA producer — the A tile (128 x 512 fp16 = 128 KB, hi|lo) is loaded ONCE and stays resident; the train
//            descriptors stream through a 2-stage ring of [128 x 64] hi + lo boxes (128-byte swizzle);
//   warp 1 : MMA issuer — one thread issues 48 tcgen05.mma (128x128x16) per 128-column tile into one of two TMEM
//            accumulators (2 x 128 columns), tcgen05.commit releases ring stages and publishes accumulators;
//   warps 2-5 : epilogue — tcgen05.ld 32x32b (thread = query row): row top-2 (value, index) lives in registers across
//            the whole sweep; the column maximum of the tile is a redux.sync over a packed (ordered score | lane) key,
//            merged across the four warps in shared memory and across CTAs with one 64-bit atomicMax per column
//            (order independent => deterministic; lowest index wins ties).
#include <cuda.h>
#include <cuda_fp16.h>
#include "float_ops.cuh"
#include "sm100_ptx.cuh"

namespace {

using namespace sm100;

// [F, 256, n_max] fp32 channel-major  ->  [F, n_max + 1, 512] fp16, row = [hi(256) | lo(256)], last row = statistics.
// One block = 32 keypoints x all 256 channels, in four passes of 64 channels through shared memory so that the read
// (along n) and the write (along k, half2 per lane = 128 B per warp) are both coalesced.  The per-row sums of hi^2
// and lo^2 are accumulated in a fixed order; their frame maxima go to the statistics row (float_ops.cuh).
__global__ void __launch_bounds__(256)
k_float_prepare(const float* __restrict__ desc, int n_max, const int32_t* __restrict__ counts,
                __half* __restrict__ ops) {
    __shared__ float tile[64][33];
    __shared__ float red[2][8];
    const int f = blockIdx.y;
    const int n0 = blockIdx.x * 32;
    const int cnt = min(counts[f], n_max);
    const float* src = desc + (size_t)f * FS_FLOAT_DIM * n_max;
    __half* dst = ops + (size_t)f * fs::op_rows(n_max) * fs::OP_ROW_HALFS;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float h2[4] = {0.f, 0.f, 0.f, 0.f}, l2[4] = {0.f, 0.f, 0.f, 0.f};  // rows warp*4 + i, this lane's channels
    for (int kt = 0; kt < FS_FLOAT_DIM / 64; kt++) {
        __syncthreads();
        for (int r = warp; r < 64; r += 8) {
            const int n = n0 + lane;
            tile[r][lane] = (n < cnt) ? src[(size_t)(kt * 64 + r) * n_max + n] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int rl = warp * 4 + i, n = n0 + rl;
            if (n >= n_max) continue;
            const float x0 = tile[2 * lane][rl], x1 = tile[2 * lane + 1][rl];
            const __half hi0 = __float2half_rn(x0), hi1 = __float2half_rn(x1);
            const __half lo0 = __float2half_rn(x0 - __half2float(hi0)), lo1 = __float2half_rn(x1 - __half2float(hi1));
            __half2* out = reinterpret_cast<__half2*>(dst + (size_t)n * fs::OP_ROW_HALFS + kt * 64) + lane;
            out[0] = __halves2half2(hi0, hi1);
            out[128] = __halves2half2(lo0, lo1);  // + 256 halfs
            const float a0 = __half2float(hi0), a1 = __half2float(hi1), b0 = __half2float(lo0), b1 = __half2float(lo1);
            h2[i] = fmaf(a1, a1, fmaf(a0, a0, h2[i]));
            l2[i] = fmaf(b1, b1, fmaf(b0, b0, l2[i]));
        }
    }
    float mh = 0.f, ml = 0.f;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float a = h2[i], b = l2[i];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        mh = fmaxf(mh, a);
        ml = fmaxf(ml, b);
    }
    if (lane == 0) {
        red[0][warp] = mh;
        red[1][warp] = ml;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) {
            mh = fmaxf(mh, red[0][w]);
            ml = fmaxf(ml, red[1][w]);
        }
        unsigned int* st = reinterpret_cast<unsigned int*>(dst + (size_t)n_max * fs::OP_ROW_HALFS);
        atomicMax(st, __float_as_uint(mh));
        atomicMax(st + 1, __float_as_uint(ml));
    }
}

// zero the statistics rows before k_float_prepare accumulates its maxima
__global__ void k_float_prepare_clear(int n_max, __half* __restrict__ ops) {
    uint4* row = reinterpret_cast<uint4*>(ops + ((size_t)blockIdx.x * fs::op_rows(n_max) + n_max) * fs::OP_ROW_HALFS);
    if (threadIdx.x < fs::OP_ROW_BYTES / 16) row[threadIdx.x] = make_uint4(0, 0, 0, 0);
}

// ------------------------------------------------------------------------------------------------ main kernel
constexpr int TC_BM = 128, TC_BN = 128, TC_BK = 64, TC_KCH = FS_FLOAT_DIM / TC_BK;  // 4 k-chunks of 64
constexpr int TC_STAGES = 2;
constexpr int TC_THREADS = 192;
constexpr int TC_TILE_BYTES = TC_BM * TC_BK * 2;                 // 16 KB: one [128 x 64] fp16 box
constexpr int TC_A_BYTES = 2 * TC_KCH * TC_TILE_BYTES;           // 128 KB: hi + lo, 4 k-chunks each
constexpr int TC_STAGE_BYTES = 2 * TC_TILE_BYTES;                // 32 KB: B hi + lo of one k-chunk
constexpr int TC_TMEM_COLS = 256;                                // two 128-column fp32 accumulators
constexpr uint32_t TC_IDESC = make_idesc_f16(TC_BM, TC_BN, /*F16*/ 0);

struct TCBarriers {
    uint64_t a_full;
    uint64_t full[TC_STAGES], empty[TC_STAGES];
    uint64_t tmem_full[2], tmem_empty[2];
    uint32_t tmem_base;
    uint32_t pad;
};
constexpr int TC_SMEM_BYTES = 1024 /*alignment slack*/ + TC_A_BYTES + TC_STAGES * TC_STAGE_BYTES +
                              2 * 4 * TC_BN * 4 /*column partials*/ + 256 /*barriers*/;

__device__ __forceinline__ uint32_t f2ord(float v) {
    const uint32_t b = __float_as_uint(v);
    return b ^ ((uint32_t)((int32_t)b >> 31) | 0x80000000u);
}

// STORE = true: the same contraction with a plain epilogue that writes scale * score to scores_out [P, n_max, n_max]
// (fs_float_scores: the score matrix log_optimal_transport consumes, superglue.py:259-260).
template <bool STORE>
__global__ void __launch_bounds__(TC_THREADS, 1)
k_float_match_tc(const __grid_constant__ CUtensorMap tmap, int n_max, const int32_t* __restrict__ counts,
                 const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b,
                 float* __restrict__ row_best, float* __restrict__ row_second, int32_t* __restrict__ row_idx,
                 unsigned long long* __restrict__ col_best, float* __restrict__ scores_out, float scale) {
    const int p = blockIdx.y;
    const int fa = pair_a[p], fb = pair_b[p];
    const int n0 = min(counts[fa], n_max), n1 = min(counts[fb], n_max);
    const int m0 = blockIdx.x * TC_BM;
    if (m0 >= n0) return;
    const int ntiles = (n1 + TC_BN - 1) / TC_BN;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                                    // [hi|lo][kc] boxes of 16 KB
    uint8_t* sB = smem + TC_A_BYTES;                       // [stage][hi|lo] boxes of 16 KB
    uint32_t* colpart = reinterpret_cast<uint32_t*>(sB + TC_STAGES * TC_STAGE_BYTES);  // [2][4][128]
    TCBarriers& bar = *reinterpret_cast<TCBarriers*>(colpart + 2 * 4 * TC_BN);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        prefetch_tensormap(&tmap);
        mbar_init(&bar.a_full, 1);
        for (int s = 0; s < TC_STAGES; s++) {
            mbar_init(&bar.full[s], 1);
            mbar_init(&bar.empty[s], 1);
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(&bar.tmem_full[b], 1);
            mbar_init(&bar.tmem_empty[b], 4);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        tmem_alloc(&bar.tmem_base, TC_TMEM_COLS);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bar.tmem_base;

    // TMA and MMA warps loop converged and elect one lane per asynchronous instruction, so that addresses and descriptors
    // stay in uniform registers (float_top3.cu has the measurement: under `if (lane == 0)` every UTCHMMA / UTMALDG sits in
    // an ELECT / R2UR.BROADCAST / BRA.U.ANY loop of ~13 dependent instructions)
    if (warp == 0) {
        // ===================================================== TMA producer
        if (ntiles > 0) {
            if (elect_one()) {
                mbar_arrive_expect_tx(&bar.a_full, TC_A_BYTES);
                for (int h = 0; h < 2; h++)
                    for (int kc = 0; kc < TC_KCH; kc++)
                        tma_load_3d(sA + (h * TC_KCH + kc) * TC_TILE_BYTES, &tmap, &bar.a_full, h * 256 + kc * TC_BK, m0, fa);
            }
            __syncwarp();
            int it = 0;
            for (int t = 0; t < ntiles; t++)
                for (int kc = 0; kc < TC_KCH; kc++, it++) {
                    const int s = it % TC_STAGES;
                    mbar_wait(&bar.empty[s], ((it / TC_STAGES) & 1) ^ 1);
                    if (elect_one()) {
                        mbar_arrive_expect_tx(&bar.full[s], TC_STAGE_BYTES);
                        uint8_t* dst = sB + s * TC_STAGE_BYTES;
                        tma_load_3d(dst, &tmap, &bar.full[s], kc * TC_BK, t * TC_BN, fb);
                        tma_load_3d(dst + TC_TILE_BYTES, &tmap, &bar.full[s], 256 + kc * TC_BK, t * TC_BN, fb);
                    }
                    __syncwarp();
                }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer
        if (ntiles > 0) {
            mbar_wait(&bar.a_full, 0);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(sA), b_addr = smem_u32(sB);
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);  // (tells the compiler it is warp-uniform)
            int it = 0;
            for (int t = 0; t < ntiles; t++) {
                const int buf = t & 1;
                mbar_wait(&bar.tmem_empty[buf], ((t >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d = tmem_u + buf * TC_BN;
                for (int kc = 0; kc < TC_KCH; kc++, it++) {
                    const int s = it % TC_STAGES;
                    mbar_wait(&bar.full[s], (it / TC_STAGES) & 1);
                    tc_fence_after();
                    const uint64_t a_hi = make_desc_k_sw128(a_addr + kc * TC_TILE_BYTES);
                    const uint64_t a_lo = make_desc_k_sw128(a_addr + (TC_KCH + kc) * TC_TILE_BYTES);
                    const uint64_t b_hi = make_desc_k_sw128(b_addr + s * TC_STAGE_BYTES);
                    const uint64_t b_lo = make_desc_k_sw128(b_addr + s * TC_STAGE_BYTES + TC_TILE_BYTES);
                    if (elect_one()) {
#pragma unroll
                        for (int ks = 0; ks < TC_BK / 16; ks++) {
                            const uint64_t o = (uint64_t)(ks * 32 >> 4);  // +32 bytes per K = 16 step inside the swizzle row
                            umma_f16(d, a_hi + o, b_hi + o, TC_IDESC, (kc | ks) != 0);
                            umma_f16(d, a_lo + o, b_hi + o, TC_IDESC, 1);
                            umma_f16(d, a_hi + o, b_lo + o, TC_IDESC, 1);
                        }
                        umma_commit(&bar.empty[s]);  // ring stage free once these MMAs have read it
                    }
                    __syncwarp();
                }
                if (elect_one()) umma_commit(&bar.tmem_full[buf]);  // accumulator complete
                __syncwarp();
            }
        }
    } else {
        // ===================================================== epilogue: warps 2..5, thread = query row
        const int quarter = warp & 3;  // the TMEM lane quarter this warp may read
        const int row_in_tile = quarter * 32 + lane;
        const int row = m0 + row_in_tile;
        const bool row_ok = row < n0;
        const int etid = (warp - 2) * 32 + lane;  // 0..127
        float best = -INFINITY, second = -INFINITY;
        int bidx = 0x7fffffff;
        for (int t = 0; t < ntiles; t++) {
            const int buf = t & 1;
            mbar_wait(&bar.tmem_full[buf], (t >> 1) & 1);
            tc_fence_after();
            uint32_t* part = colpart + (buf * 4 + quarter) * TC_BN;
#pragma unroll 1
            for (int c = 0; c < TC_BN / 32; c++) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + buf * TC_BN + c * 32, r);
                tmem_ld_wait();
                if (c == TC_BN / 32 - 1) {  // accumulator fully read: hand it back to the MMA warp
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bar.tmem_empty[buf]);
                }
                const int col0 = t * TC_BN + c * 32;
                if constexpr (STORE) {
                    // thread = row: 32 consecutive columns = one full 128-byte line per thread.  Columns >= n1 are
                    // products with zero operand rows (stored as 0, never read by the consumer).
                    if (row_ok) {
                        float4* dst = reinterpret_cast<float4*>(scores_out + ((size_t)p * n_max + row) * n_max + col0);
#pragma unroll
                        for (int j = 0; j < 8; j++)
                            __stcs(dst + j, make_float4(__uint_as_float(r[4 * j]) * scale, __uint_as_float(r[4 * j + 1]) * scale,
                                                        __uint_as_float(r[4 * j + 2]) * scale,
                                                        __uint_as_float(r[4 * j + 3]) * scale));
                    }
                    continue;
                }
                const int nvalid = n1 - col0;  // columns of this chunk that exist
                uint32_t mykey = 0;
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    float v = __uint_as_float(r[j]);
                    if (j >= nvalid) v = -INFINITY;
                    // row top-2, lowest column wins ties (columns arrive in ascending order)
                    const bool gt1 = v > best, gt2 = v > second;
                    second = gt1 ? best : (gt2 ? v : second);
                    bidx = gt1 ? (col0 + j) : bidx;
                    best = gt1 ? v : best;
                    // column maximum over the warp's 32 rows: ordered score in the high 27 bits, 31 - lane below
                    uint32_t key = row_ok ? ((f2ord(v) & ~31u) | (uint32_t)(31 - lane)) : 0u;
                    const uint32_t m = __reduce_max_sync(0xffffffffu, key);
                    if (lane == j) mykey = m;
                }
                part[c * 32 + lane] = mykey;
            }
            if constexpr (STORE) continue;
            named_bar_sync(1, 128);
            {
                // merge the four warps' partials of column `etid`, one 64-bit atomicMax per (column, CTA)
                const int col = t * TC_BN + etid;
                uint32_t bk = 0;
                int brow = 0;
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const uint32_t k = colpart[(buf * 4 + q) * TC_BN + etid];
                    if ((k & ~31u) > (bk & ~31u)) {
                        bk = k;
                        brow = q * 32 + 31 - (int)(k & 31u);
                    }
                }
                if (col < n1 && bk != 0) {
                    const unsigned long long key64 =
                        ((unsigned long long)(bk & ~31u) << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)(m0 + brow));
                    atomicMax(col_best + (size_t)p * n_max + col, key64);
                }
            }
        }
        if (!STORE && row_ok) {
            const size_t o = (size_t)p * n_max + row;
            row_best[o] = best;
            row_second[o] = second;
            row_idx[o] = n1 > 0 ? bidx : -1;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TC_TMEM_COLS);
    }
}

// unpack the packed column keys into the int32 index array the shared finalize kernel reads
__global__ void __launch_bounds__(256)
k_unpack_cols(const unsigned long long* __restrict__ col_best, size_t n, int32_t* __restrict__ idx1) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long k = col_best[i];
    idx1[i] = k ? (int32_t)(0xFFFFFFFFu - (uint32_t)(k & 0xFFFFFFFFull)) : -1;
}

}  // namespace

namespace fs {
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int encode_operand_map(const void* operands, int n_frames, int n_max, CUtensorMap* out, int box_rows) {
    static PFN_tmapEncodeTiled fn = nullptr;
    if (!fn) {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        FS_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q));
        if (!sym || q != cudaDriverEntryPointSuccess) {
            fs::set_error("cuTensorMapEncodeTiled is not available in this driver");
            return FS_ERR_UNSUPPORTED;
        }
        fn = (PFN_tmapEncodeTiled)sym;
    }
    const cuuint64_t gdim[3] = {512, (cuuint64_t)fs::op_rows(n_max), (cuuint64_t)n_frames};
    const cuuint64_t gstride[2] = {512 * 2, (cuuint64_t)fs::op_rows(n_max) * 512 * 2};
    const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(operands), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fs::set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        return FS_ERR_CUDA;
    }
    return FS_OK;
}

}  // namespace fs

extern "C" int fs_float_prepare(const float* desc_f32, int n_frames, int n_max, const int32_t* counts,
                                void* operands, fs_stream_t stream) {
    FS_REQUIRE(n_frames >= 0 && n_max >= 0, "negative size");
    if (n_frames == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(desc_f32 && counts && operands, "null pointer");
    FS_REQUIRE(n_frames <= 65535, "n_frames too large for one launch");
    k_float_prepare_clear<<<n_frames, 64, 0, (cudaStream_t)stream>>>(n_max, (__half*)operands);
    FS_LAUNCH_CHECK("k_float_prepare_clear");
    dim3 grid((n_max + 31) / 32, n_frames);
    k_float_prepare<<<grid, 256, 0, (cudaStream_t)stream>>>(desc_f32, n_max, counts, (__half*)operands);
    FS_LAUNCH_CHECK("k_float_prepare");
    return FS_OK;
}

extern "C" size_t fs_float_match_split3_workspace_bytes(int n_pairs, int n_max) {
    if (n_pairs <= 0 || n_max <= 0) return 256;
    const size_t n = (size_t)n_pairs * n_max;
    // best, second, idx0, idx1 (4 B each) + packed column keys (8 B)
    return 4 * fs::align_up(n * 4, 256) + fs::align_up(n * 8, 256) + 256;
}

extern "C" int fs_float_match_split3(const void* operands, int n_frames, int n_max, const int32_t* counts,
                              const int32_t* pair_a, const int32_t* pair_b, int n_pairs, float scale, float threshold,
                              float ratio, int32_t* matches0, float* score0, float* second0, int32_t* matches1,
                              int32_t* n_matches, void* workspace, size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_frames >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(operands && counts && pair_a && pair_b && matches0, "null pointer");
    FS_REQUIRE(((uintptr_t)operands & 15) == 0, "operands must be 16-byte aligned");
    const size_t need = fs_float_match_split3_workspace_bytes(n_pairs, n_max);
    if (!workspace || workspace_bytes < need) {
        fs::set_error("fs_float_match_split3: workspace %zu < %zu", workspace_bytes, need);
        return FS_ERR_WORKSPACE;
    }
    CUtensorMap tmap;
    int st = fs::encode_operand_map(operands, n_frames, n_max, &tmap);
    if (st != FS_OK) return st;
    fs::Carver ws(workspace);
    const size_t n = (size_t)n_pairs * n_max;
    fs::FloatRowTop2 d0{ws.take<float>(n), ws.take<float>(n), ws.take<int32_t>(n)};
    int32_t* idx1 = ws.take<int32_t>(n);
    unsigned long long* col_best = ws.take<unsigned long long>(n);
    FS_CUDA(cudaMemsetAsync(col_best, 0, n * sizeof(unsigned long long), stream));
    {
        static int granted[64];
        const int st_ = fs::ensure_dyn_smem(k_float_match_tc<false>, granted, (size_t)(TC_SMEM_BYTES));
        if (st_ != FS_OK) return st_;
    }
    for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
        const int np = min(65535, n_pairs - p0);
        const size_t o = (size_t)p0 * n_max;
        dim3 grid((n_max + TC_BM - 1) / TC_BM, np);
        k_float_match_tc<false><<<grid, TC_THREADS, TC_SMEM_BYTES, stream>>>(
            tmap, n_max, counts, pair_a + p0, pair_b + p0, d0.best + o, d0.second + o, d0.idx + o, col_best + o,
            nullptr, 1.f);
        FS_LAUNCH_CHECK("k_float_match_tc");
    }
    k_unpack_cols<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(col_best, n, idx1);
    FS_LAUNCH_CHECK("k_unpack_cols");
    return fs::launch_float_finalize(n_max, counts, pair_a, pair_b, n_pairs, d0, idx1, scale, threshold, ratio, matches0,
                                     score0, second0, matches1, n_matches, stream);
}

// Score matrices for the SuperGlue head: scores[p, i, j] = scale * <d_a[i], d_b[j]> (superglue.py:259-260), fp32-grade
// (split-fp16, three MMAs per k-step), written once to HBM for fs_sinkhorn_match.
extern "C" int fs_float_scores(const void* operands, int n_frames, int n_max, const int32_t* counts,
                               const int32_t* pair_a, const int32_t* pair_b, int n_pairs, float scale, float* scores,
                               fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_frames >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(operands && counts && pair_a && pair_b && scores, "null pointer");
    FS_REQUIRE(n_max % TC_BN == 0, "n_max must be a multiple of 128");
    FS_REQUIRE(((uintptr_t)operands & 15) == 0 && ((uintptr_t)scores & 15) == 0, "buffers must be 16-byte aligned");
    CUtensorMap tmap;
    int st = fs::encode_operand_map(operands, n_frames, n_max, &tmap);
    if (st != FS_OK) return st;
    {
        static int granted[64];
        const int st_ = fs::ensure_dyn_smem(k_float_match_tc<true>, granted, (size_t)(TC_SMEM_BYTES));
        if (st_ != FS_OK) return st_;
    }
    for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
        const int np = min(65535, n_pairs - p0);
        dim3 grid((n_max + TC_BM - 1) / TC_BM, np);
        k_float_match_tc<true><<<grid, TC_THREADS, TC_SMEM_BYTES, stream>>>(
            tmap, n_max, counts, pair_a + p0, pair_b + p0, nullptr, nullptr, nullptr, nullptr,
            scores + (size_t)p0 * n_max * n_max, scale);
        FS_LAUNCH_CHECK("k_float_match_tc<store>");
    }
    return FS_OK;
}
