// float_fp32.cu — A1 on CUDA cores in exact fp32: dense 256-d contraction with a fused row top-2
// epilogue, run once per direction, then the mutual / threshold / ratio finalisation.
//
// Reference: SuperGluePretrainedNetwork/models/superglue.py:259-260 (einsum 'bdn,bdm->bnm') and
// :268-278 (row max, column max, mutual check, threshold, -1 fill); descriptor layout [256, N] from
// superpoint.py:90-91; call sites mono_slam.py:100-103, stereo_slam.py:210-213.
//
// This variant exists for descriptor counts that are not multiples of the tensor-core tile and as
// the in-library cross-check of float_tc.cu.  Accumulation is a k-ascending fp32 FMA chain.
#include "float_common.cuh"

namespace {

constexpr int TM = 64, TN = 64, TK = 16, D = FS_FLOAT_DIM;

// rows come from frame X (pair_x), columns from frame Y (pair_y)
__global__ void __launch_bounds__(256)
k_float_rowtop2_fp32(const float* __restrict__ desc, int n_max, const int32_t* __restrict__ counts,
                     const int32_t* __restrict__ pair_x, const int32_t* __restrict__ pair_y,
                     float* __restrict__ out_best, float* __restrict__ out_second,
                     int32_t* __restrict__ out_idx) {
    __shared__ __align__(16) float As[TK][TM];
    __shared__ __align__(16) float Bs[TK][TN];
    const int p = blockIdx.y;
    const int fx = pair_x[p], fy = pair_y[p];
    const int nx = min(counts[fx], n_max), ny = min(counts[fy], n_max);
    const int row0 = blockIdx.x * TM;
    if (row0 >= nx) return;
    const float* X = desc + (size_t)fx * D * n_max;
    const float* Y = desc + (size_t)fy * D * n_max;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;

    fs::Top2 top[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) top[r].init();

    for (int col0 = 0; col0 < ny; col0 += TN) {
        float acc[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;

        for (int k0 = 0; k0 < D; k0 += TK) {
            __syncthreads();
            // 16 x 64 floats each: 1024 elements, 4 per thread, coalesced along n
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = tid + i * 256, kk = e >> 6, nn = e & 63;
                const int rx = row0 + nn, cy = col0 + nn;
                As[kk][nn] = rx < nx ? X[(size_t)(k0 + kk) * n_max + rx] : 0.f;
                Bs[kk][nn] = cy < ny ? Y[(size_t)(k0 + kk) * n_max + cy] : 0.f;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < TK; ++kk) {
                const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
                const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(av[r], bv[c], acc[r][c]);
            }
        }
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int col = col0 + tx * 4 + c;
                if (col < ny) top[r].push(acc[r][c], col);
            }
    }
    // merge the 16 column-interleaved partial top-2 of every row (16 consecutive lanes share ty)
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, top[r].best, o);
            const float os = __shfl_xor_sync(0xffffffffu, top[r].second, o);
            const int oi = __shfl_xor_sync(0xffffffffu, top[r].idx, o);
            top[r].merge(ob, os, oi);
        }
        const int row = row0 + ty * 4 + r;
        if (tx == 0 && row < nx) {
            const size_t o = (size_t)p * n_max + row;
            out_best[o] = top[r].best;
            out_second[o] = top[r].second;
            out_idx[o] = ny > 0 ? top[r].idx : -1;
        }
    }
}

__global__ void __launch_bounds__(256)
k_float_finalize(int n_max, const int32_t* __restrict__ counts, const int32_t* __restrict__ pair_a,
                 const int32_t* __restrict__ pair_b, const float* __restrict__ best0,
                 const float* __restrict__ sec0, const int32_t* __restrict__ idx0,
                 const int32_t* __restrict__ idx1, float scale, float threshold, float ratio, int cross_check,
                 int32_t* __restrict__ matches0, float* __restrict__ score0, float* __restrict__ second0,
                 int32_t* __restrict__ matches1, int32_t* __restrict__ n_matches) {
    const int p = blockIdx.y;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const int n0 = min(counts[pair_a[p]], n_max), n1 = min(counts[pair_b[p]], n_max);
    bool ok = false;
    if (n < n_max) {
        const size_t o = (size_t)p * n_max + n;
        int m = -1;
        float sb = -INFINITY, ss = -INFINITY;
        if (n < n0 && n1 > 0) {
            const float b = best0[o], s = sec0[o];
            const int i0 = idx0[o];
            sb = __fmul_rn(b, scale);
            ss = __fmul_rn(s, scale);
            ok = !cross_check || idx1[(size_t)p * n_max + i0] == n;      // mutual0, superglue.py:270 (knn use: no check)
            ok = ok && (sb > threshold);                  // valid0, superglue.py:275
            if (ratio > 0.f) {
                // Lowe ratio on the L2 distance of unit vectors, d = sqrt(max(0, 2 - 2 cos)); strict <
                const float db = __fsqrt_rn(fmaxf(0.f, __fsub_rn(2.f, __fmul_rn(2.f, b))));
                const float ds = __fsqrt_rn(fmaxf(0.f, __fsub_rn(2.f, __fmul_rn(2.f, s))));
                ok = ok && (s > -INFINITY) && (db < __fmul_rn(ratio, ds));
            }
            if (ok) {
                m = i0;
                if (matches1 && cross_check) matches1[(size_t)p * n_max + i0] = n;  // mutual => unique writer
            }
        }
        matches0[o] = m;
        if (score0) score0[o] = sb;
        if (second0) second0[o] = ss;
    }
    if (n_matches) {
        const unsigned bal = __ballot_sync(0xffffffffu, ok);
        if ((threadIdx.x & 31) == 0 && bal) atomicAdd(n_matches + p, __popc(bal));
    }
}

}  // namespace

namespace fs {

int launch_float_finalize(int n_max, const int32_t* counts, const int32_t* pair_a, const int32_t* pair_b,
                          int n_pairs, FloatRowTop2 d0, const int32_t* idx1, float scale, float threshold,
                          float ratio, int32_t* matches0, float* score0, float* second0, int32_t* matches1,
                          int32_t* n_matches, cudaStream_t stream, int cross_check) {
    const size_t n = (size_t)n_pairs * n_max;
    if (matches1) FS_CUDA(cudaMemsetAsync(matches1, 0xFF, n * sizeof(int32_t), stream));
    if (n_matches) FS_CUDA(cudaMemsetAsync(n_matches, 0, (size_t)n_pairs * sizeof(int32_t), stream));
    for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
        const int np = min(65535, n_pairs - p0);
        const size_t o = (size_t)p0 * n_max;
        dim3 g((n_max + 255) / 256, np);
        k_float_finalize<<<g, 256, 0, stream>>>(n_max, counts, pair_a + p0, pair_b + p0, d0.best + o,
                                                d0.second + o, d0.idx + o, idx1 + o, scale, threshold, ratio,
                                                cross_check, matches0 + o, score0 ? score0 + o : nullptr,
                                                second0 ? second0 + o : nullptr,
                                                matches1 ? matches1 + o : nullptr,
                                                n_matches ? n_matches + p0 : nullptr);
        FS_LAUNCH_CHECK("k_float_finalize");
    }
    return FS_OK;
}

}  // namespace fs

extern "C" size_t fs_float_match_workspace_bytes(int n_pairs, int n_max) {
    if (n_pairs <= 0 || n_max <= 0) return 256;
    // two directions x (best, second, idx)
    return 6 * fs::align_up((size_t)n_pairs * n_max * 4, 256) + 256;
}

extern "C" int fs_float_match_fp32(const float* desc, int n_frames, int n_max, const int32_t* counts,
                                   const int32_t* pair_a, const int32_t* pair_b, int n_pairs, float scale,
                                   float threshold, float ratio, int32_t* matches0, float* score0,
                                   float* second0, int32_t* matches1, int32_t* n_matches, void* workspace,
                                   size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_frames >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(desc && counts && pair_a && pair_b && matches0, "null pointer");
    if (!workspace || workspace_bytes < fs_float_match_workspace_bytes(n_pairs, n_max)) {
        fs::set_error("fs_float_match_fp32: workspace %zu < %zu", workspace_bytes,
                      fs_float_match_workspace_bytes(n_pairs, n_max));
        return FS_ERR_WORKSPACE;
    }
    fs::Carver ws(workspace);
    const size_t n = (size_t)n_pairs * n_max;
    fs::FloatRowTop2 d0{ws.take<float>(n), ws.take<float>(n), ws.take<int32_t>(n)};
    fs::FloatRowTop2 d1{ws.take<float>(n), ws.take<float>(n), ws.take<int32_t>(n)};
    for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
        const int np = min(65535, n_pairs - p0);
        const size_t o = (size_t)p0 * n_max;
        dim3 g((n_max + TM - 1) / TM, np);
        k_float_rowtop2_fp32<<<g, 256, 0, stream>>>(desc, n_max, counts, pair_a + p0, pair_b + p0, d0.best + o,
                                                    d0.second + o, d0.idx + o);
        FS_LAUNCH_CHECK("k_float_rowtop2_fp32(dir0)");
        k_float_rowtop2_fp32<<<g, 256, 0, stream>>>(desc, n_max, counts, pair_b + p0, pair_a + p0, d1.best + o,
                                                    d1.second + o, d1.idx + o);
        FS_LAUNCH_CHECK("k_float_rowtop2_fp32(dir1)");
    }
    return fs::launch_float_finalize(n_max, counts, pair_a, pair_b, n_pairs, d0, d1.idx, scale, threshold,
                                     ratio, matches0, score0, second0, matches1, n_matches, stream);
}
