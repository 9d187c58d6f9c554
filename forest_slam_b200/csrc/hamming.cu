// hamming.cu — A2: brute-force 256-bit Hamming matching with mutual-NN (crossCheck) and ratio test.
//
// Replaces cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(d0, d1)
//   (reference: ros_ws/src/stereo_slam.py:85, :234, :242; distance = ORB_SLAM3/src/ORBmatcher.cc:2058-2074;
//    k=2 + ratio variant = ORB_SLAM3/src/Frame.cc:1144-1151).
//
// Design (B200): ONE pass over the N0 x N1 distance matrix serves both directions.
//   * a CTA owns a tile of 256 queries (2 per thread, 128-bit loads, descriptors live in registers)
//     and streams the train descriptors of the pair through shared memory in 8 KB chunks
//     (every lane reads the same train row: LDS.128 broadcast);
//   * per distance: 8 x (LOP3.xor + POPC) + adds; the row top-2 is a branch-free packed update
//       p = dist << 20 | train ;  second = min(second, max(p, best)) ;  best = min(best, p)
//     (lowest train index wins ties because the index sits in the low bits);
//   * the column best (needed by crossCheck) is a warp-wide redux.sync min over
//       dist << 20 | query, staged per warp in shared memory and merged with one global
//     atomicMin per (train, CTA).  min over packed integers is order independent => deterministic.
// The kernel is bound by the POPC pipe (16 lanes/clk/SM), not by HBM: see DESIGN.md.
#include "common.cuh"

namespace {

constexpr int HM_THREADS = 128;
constexpr int HM_QPT = 2;
constexpr int HM_QTILE = HM_THREADS * HM_QPT;  // 256 queries per CTA
constexpr int HM_TCHUNK = 256;                 // train rows per shared-memory chunk
constexpr int HM_WARPS = HM_THREADS / 32;
constexpr int IDX_BITS = 20;
constexpr uint32_t IDX_MASK = (1u << IDX_BITS) - 1;
constexpr uint32_t NONE = 0xFFFFFFFFu;

__device__ __forceinline__ int ham256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(HM_THREADS)
k_hamming_top2(const uint4* __restrict__ desc, int n_max, const int32_t* __restrict__ counts,
               const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b,
               uint32_t* __restrict__ row_best, uint32_t* __restrict__ row_second,
               uint32_t* __restrict__ col_best) {
    __shared__ uint4 s_tr[HM_TCHUNK * 2];
    __shared__ uint32_t s_col[HM_WARPS][HM_TCHUNK];

    const int p = blockIdx.y;
    const int fa = pair_a[p], fb = pair_b[p];
    const int n0 = min(counts[fa], n_max), n1 = min(counts[fb], n_max);
    const int qbase = blockIdx.x * HM_QTILE;
    if (qbase >= n0) return;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint4* qa = desc + (size_t)fa * n_max * 2;
    const uint4* tb = desc + (size_t)fb * n_max * 2;

    uint4 q0[HM_QPT], q1[HM_QPT];
    uint32_t qid[HM_QPT], best[HM_QPT], second[HM_QPT];
#pragma unroll
    for (int j = 0; j < HM_QPT; ++j) {
        int q = qbase + tid + j * HM_THREADS;
        bool ok = q < n0;
        int qq = ok ? q : qbase;
        q0[j] = __ldg(qa + 2 * qq);
        q1[j] = __ldg(qa + 2 * qq + 1);
        qid[j] = ok ? (uint32_t)q : NONE;  // NONE poisons the column key of padded queries
        best[j] = NONE;
        second[j] = NONE;
    }

    for (int tbase = 0; tbase < n1; tbase += HM_TCHUNK) {
        const int nt = min(HM_TCHUNK, n1 - tbase);
        __syncthreads();  // previous chunk fully consumed
        for (int i = tid; i < nt * 2; i += HM_THREADS) s_tr[i] = __ldg(tb + 2 * tbase + i);
        __syncthreads();

        uint32_t mycol = NONE;
        for (int t0 = 0; t0 < nt; t0 += 32) {
            const int tend = min(32, nt - t0);
#pragma unroll 4
            for (int tt = 0; tt < tend; ++tt) {
                const int t = t0 + tt;
                const uint4 b0 = s_tr[2 * t], b1 = s_tr[2 * t + 1];
                uint32_t ckey = NONE;
#pragma unroll
                for (int j = 0; j < HM_QPT; ++j) {
                    const uint32_t d = (uint32_t)ham256(q0[j], q1[j], b0, b1) << IDX_BITS;
                    const uint32_t pk = d | (uint32_t)(tbase + t);
                    second[j] = min(second[j], max(pk, best[j]));
                    best[j] = min(best[j], pk);
                    ckey = min(ckey, d | qid[j]);  // qid == NONE saturates the key for padded queries
                }
                const uint32_t w = __reduce_min_sync(0xffffffffu, ckey);
                if (lane == tt) mycol = w;
            }
            if (lane < tend) s_col[warp][t0 + lane] = mycol;
        }
        __syncthreads();
        for (int t = tid; t < nt; t += HM_THREADS) {
            uint32_t m = s_col[0][t];
#pragma unroll
            for (int w = 1; w < HM_WARPS; ++w) m = min(m, s_col[w][t]);
            atomicMin(col_best + (size_t)p * n_max + tbase + t, m);
        }
    }

#pragma unroll
    for (int j = 0; j < HM_QPT; ++j) {
        int q = qbase + tid + j * HM_THREADS;
        if (q < n0) {
            row_best[(size_t)p * n_max + q] = best[j];
            row_second[(size_t)p * n_max + q] = second[j];
        }
    }
}

__global__ void __launch_bounds__(256)
k_hamming_finalize(int n_max, const int32_t* __restrict__ counts, const int32_t* __restrict__ pair_a,
                   const int32_t* __restrict__ pair_b, const uint32_t* __restrict__ row_best,
                   const uint32_t* __restrict__ row_second, const uint32_t* __restrict__ col_best,
                   int cross_check, float ratio, int32_t* __restrict__ matches0, int32_t* __restrict__ dist0,
                   int32_t* __restrict__ second0, int32_t* __restrict__ matches1,
                   int32_t* __restrict__ n_matches) {
    const int p = blockIdx.y;
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int n0 = min(counts[pair_a[p]], n_max), n1 = min(counts[pair_b[p]], n_max);
    bool ok = false;
    if (q < n_max) {
        const size_t o = (size_t)p * n_max + q;
        int m = -1, d = INT32_MAX, s = INT32_MAX;
        if (q < n0 && n1 > 0) {
            const uint32_t b = row_best[o], sc = row_second[o];
            const int t = (int)(b & IDX_MASK);
            d = (int)(b >> IDX_BITS);
            const bool has2 = sc != NONE;
            if (has2) s = (int)(sc >> IDX_BITS);
            ok = true;
            if (cross_check) ok = (col_best[(size_t)p * n_max + t] & IDX_MASK) == (uint32_t)q;
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

}  // namespace

extern "C" size_t fs_hamming_match_workspace_bytes(int n_pairs, int n_max) {
    if (n_pairs <= 0 || n_max <= 0) return 256;
    return 3 * fs::align_up((size_t)n_pairs * n_max * sizeof(uint32_t), 256) + 256;
}

extern "C" int fs_hamming_match(const uint8_t* desc, int n_frames, int n_max, const int32_t* counts,
                                const int32_t* pair_a, const int32_t* pair_b, int n_pairs, int cross_check,
                                float ratio, int32_t* matches0, int32_t* dist0, int32_t* second0,
                                int32_t* matches1, int32_t* n_matches, void* workspace,
                                size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_frames >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(desc && counts && pair_a && pair_b && matches0, "null pointer");
    FS_REQUIRE(n_max <= (int)IDX_MASK, "n_max exceeds 2^20-1");
    FS_REQUIRE(n_pairs <= 65535 * 64, "n_pairs too large for one launch");
    FS_REQUIRE(((uintptr_t)desc & 15) == 0, "desc must be 16-byte aligned");
    if (workspace_bytes < fs_hamming_match_workspace_bytes(n_pairs, n_max) || !workspace) {
        fs::set_error("fs_hamming_match: workspace %zu < %zu", workspace_bytes,
                      fs_hamming_match_workspace_bytes(n_pairs, n_max));
        return FS_ERR_WORKSPACE;
    }
    fs::Carver ws(workspace);
    const size_t n = (size_t)n_pairs * n_max;
    uint32_t* row_best = ws.take<uint32_t>(n);
    uint32_t* row_second = ws.take<uint32_t>(n);
    uint32_t* col_best = ws.take<uint32_t>(n);
    FS_CUDA(cudaMemsetAsync(col_best, 0xFF, n * sizeof(uint32_t), stream));
    if (matches1) FS_CUDA(cudaMemsetAsync(matches1, 0xFF, n * sizeof(int32_t), stream));
    if (n_matches) FS_CUDA(cudaMemsetAsync(n_matches, 0, (size_t)n_pairs * sizeof(int32_t), stream));

    // grid.y is limited to 65535: split very long pair lists (loop closure) into slabs
    for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
        const int np = min(65535, n_pairs - p0);
        const size_t o = (size_t)p0 * n_max;
        dim3 g1((n_max + HM_QTILE - 1) / HM_QTILE, np);
        k_hamming_top2<<<g1, HM_THREADS, 0, stream>>>((const uint4*)desc, n_max, counts, pair_a + p0,
                                                      pair_b + p0, row_best + o, row_second + o, col_best + o);
        FS_LAUNCH_CHECK("k_hamming_top2");
        dim3 g2((n_max + 255) / 256, np);
        k_hamming_finalize<<<g2, 256, 0, stream>>>(n_max, counts, pair_a + p0, pair_b + p0, row_best + o,
                                                   row_second + o, col_best + o, cross_check, ratio,
                                                   matches0 + o, dist0 ? dist0 + o : nullptr,
                                                   second0 ? second0 + o : nullptr,
                                                   matches1 ? matches1 + o : nullptr,
                                                   n_matches ? n_matches + p0 : nullptr);
        FS_LAUNCH_CHECK("k_hamming_finalize");
    }
    return FS_OK;
}
