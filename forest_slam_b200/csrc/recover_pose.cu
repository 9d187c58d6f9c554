// recover_pose.cu — A5: pose from the essential matrix.
// Replaces  _, R, t, _ = cv2.recoverPose(E, mkpts0, mkpts1, focal=K0[0,0], pp=(K0[0,2], K0[1,2]))  mono_slam.py:112
// The reference passes no mask, so ALL correspondences are triangulated for each of the four (R, t)
// combinations: 4 K independent 4x4 Jacobi SVDs per frame pair — on the CPU this is the single largest cost
// of the mono path.  One CTA per pair; thread 0 decomposes E, then every thread owns points.
#include "common.cuh"
#include "posemath.cuh"

namespace {

constexpr int RP_THREADS = 256;

__global__ void __launch_bounds__(RP_THREADS)
k_recover_pose(const double* __restrict__ E_in, const float* __restrict__ mkpts0, const float* __restrict__ mkpts1,
               const int32_t* __restrict__ n_corr, const int32_t* __restrict__ status_in, int n_max, double focal,
               double ppx, double ppy, double dist, double* __restrict__ R_out, double* __restrict__ t_out,
               uint8_t* __restrict__ mask_out, int32_t* __restrict__ n_good) {
    __shared__ double sR[2][9];
    __shared__ double st[3];
    __shared__ int s_cnt[4];
    __shared__ int s_best;
    const int p = blockIdx.x, tid = threadIdx.x;
    const int n = min(n_corr[p], n_max);
    uint8_t* mask = mask_out + (size_t)p * n_max;
    if (status_in && status_in[p] != FS_PAIR_OK) {
        for (int i = tid; i < n_max; i += RP_THREADS) mask[i] = 0;
        if (tid < 9) R_out[9 * (size_t)p + tid] = (tid % 4 == 0) ? 1.0 : 0.0;
        if (tid < 3) t_out[3 * (size_t)p + tid] = 0.0;
        if (tid == 0) n_good[p] = 0;
        return;
    }
    if (tid == 0) {
        double E[9];
        for (int i = 0; i < 9; i++) E[i] = E_in[9 * (size_t)p + i];
        cvm::decompose_essential(E, sR[0], sR[1], st);
    }
    if (tid < 4) s_cnt[tid] = 0;
    __syncthreads();
    double R1[9], R2[9], tp[3], tm[3];
#pragma unroll
    for (int i = 0; i < 9; i++) {
        R1[i] = sR[0][i];
        R2[i] = sR[1][i];
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        tp[i] = st[i];
        tm[i] = -st[i];
    }
    const float2* a = reinterpret_cast<const float2*>(mkpts0) + (size_t)p * n_max;
    const float2* b = reinterpret_cast<const float2*>(mkpts1) + (size_t)p * n_max;
    int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    for (int i = tid; i < n; i += RP_THREADS) {
        const float2 pa = a[i], pb = b[i];
        const double x1 = ((double)pa.x - ppx) / focal, y1 = ((double)pa.y - ppy) / focal;
        const double x2 = ((double)pb.x - ppx) / focal, y2 = ((double)pb.y - ppy) / focal;
        // P1 = [R1|t], P2 = [R2|t], P3 = [R1|-t], P4 = [R2|-t]
        const bool m0 = cvm::cheirality(R1, tp, x1, y1, x2, y2, dist);
        const bool m1 = cvm::cheirality(R2, tp, x1, y1, x2, y2, dist);
        const bool m2 = cvm::cheirality(R1, tm, x1, y1, x2, y2, dist);
        const bool m3 = cvm::cheirality(R2, tm, x1, y1, x2, y2, dist);
        c0 += m0; c1 += m1; c2 += m2; c3 += m3;
        mask[i] = (uint8_t)(m0 | (m1 << 1) | (m2 << 2) | (m3 << 3));  // scratch: 4 mask bits, resolved below
    }
    c0 = __reduce_add_sync(0xffffffffu, c0);
    c1 = __reduce_add_sync(0xffffffffu, c1);
    c2 = __reduce_add_sync(0xffffffffu, c2);
    c3 = __reduce_add_sync(0xffffffffu, c3);
    if ((tid & 31) == 0) {
        atomicAdd(&s_cnt[0], c0);
        atomicAdd(&s_cnt[1], c1);
        atomicAdd(&s_cnt[2], c2);
        atomicAdd(&s_cnt[3], c3);
    }
    __syncthreads();
    if (tid == 0) {
        const int g1 = s_cnt[0], g2 = s_cnt[1], g3 = s_cnt[2], g4 = s_cnt[3];
        int best;
        if (g1 >= g2 && g1 >= g3 && g1 >= g4) best = 0;
        else if (g2 >= g1 && g2 >= g3 && g2 >= g4) best = 1;
        else if (g3 >= g1 && g3 >= g2 && g3 >= g4) best = 2;
        else best = 3;
        s_best = best;
        n_good[p] = s_cnt[best];
    }
    __syncthreads();
    const int best = s_best;
    if (tid < 9) R_out[9 * (size_t)p + tid] = (best & 1) ? R2[tid] : R1[tid];
    if (tid < 3) t_out[3 * (size_t)p + tid] = (best & 2) ? tm[tid] : tp[tid];
    for (int i = tid; i < n_max; i += RP_THREADS) mask[i] = (i < n && ((mask[i] >> best) & 1)) ? 255 : 0;
}

}  // namespace

extern "C" int fs_recover_pose(const double* E, const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                               const int32_t* status_in, int n_pairs, int n_max, double focal, double pp_x,
                               double pp_y, double distance_thresh, double* R, double* t, uint8_t* mask,
                               int32_t* n_good, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(E && mkpts0 && mkpts1 && n_corr && R && t && mask && n_good, "null pointer");
    FS_REQUIRE(focal != 0, "focal must be non-zero");
    k_recover_pose<<<n_pairs, RP_THREADS, 0, (cudaStream_t)stream>>>(E, mkpts0, mkpts1, n_corr, status_in, n_max,
                                                                    focal, pp_x, pp_y, distance_thresh, R, t, mask,
                                                                    n_good);
    FS_LAUNCH_CHECK("k_recover_pose");
    return FS_OK;
}
