// ransac_essential.cu — A4: essential-matrix RANSAC with OpenCV's exact loop semantics, one CTA per frame
// pair, everything (sampler, minimal solver, scoring, adaptive stop) on the device in ONE launch.
//
// Replaces  E, mask = cv2.findEssentialMat(mkpts0, mkpts1, focal=K0[0,0], pp=(K0[0,2], K0[1,2]),
//                                          method=cv2.RANSAC, prob=0.999, threshold=1.0)    mono_slam.py:111
// (OpenCV: calib3d/five-point.cpp findEssentialMat + EMEstimatorCallback, calib3d/ptsetreg.cpp
//  RANSACPointSetRegistrator::run / getSubset / findInliers / RANSACUpdateNumIters.)
//
// Structure of one CTA (256 threads, 8 warps), per round of 32 RANSAC iterations:
//   sample : thread 0 replays cv::RNG(2^64-1) and draws the next 32 five-subsets (sequential by nature);
//   solve  : 32 threads, one five-point solve each (fivepoint.cuh) -> <= 10 candidate E per iteration in smem;
//   score  : ONE WARP PER HYPOTHESIS over all correspondences, which live in shared memory as normalised
//            doubles (SoA); Sampson error in OpenCV's operation order (no FMA), cast to float32, compared
//            with (float)thr^2, counted with ballot/popc;
//   select : thread 0 walks the round's counts in OpenCV's sequential order (strictly-greater update,
//            adaptive iteration cap) so that the result is the one the serial loop would have produced.
// Work beyond the adaptive cap inside the last round is discarded, never observed.
#include "common.cuh"
#include "fivepoint.cuh"

namespace {

constexpr int ER_THREADS = 256;
constexpr int ER_WARPS = ER_THREADS / 32;
constexpr int ER_CHUNK = 32;
constexpr int ER_MAXMODELS = 10;

struct SampsonE {
    double e[9];
};

using cvm::sampson_f32;

// warp-cooperative inlier count of one hypothesis
__device__ __forceinline__ int warp_count_inliers(const double* E, const double* px1, const double* py1,
                                                  const double* px2, const double* py2, int n, float thr2,
                                                  int lane) {
    int cnt = 0;
    for (int i = lane; i < n; i += 32) cnt += sampson_f32(E, px1[i], py1[i], px2[i], py2[i]) <= thr2 ? 1 : 0;
    return __reduce_add_sync(0xffffffffu, cnt);
}

struct ERShared {
    int nmodels[ER_CHUNK];
    int counts[ER_CHUNK][ER_MAXMODELS];
    int subsets[ER_CHUNK][5];
    int hyp_offset[ER_CHUNK + 1];
    int niters, best, done_iters, stop;
    double bestE[9];
};

__device__ __forceinline__ void stage_points(const float* mk0, const float* mk1, int n, double focal, double ppx,
                                             double ppy, double* px1, double* py1, double* px2, double* py2) {
    // findEssentialMat: points converted to double, (x - cx) / fx, (y - cy) / fy with fx = fy = focal
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float2 a = reinterpret_cast<const float2*>(mk0)[i];
        const float2 b = reinterpret_cast<const float2*>(mk1)[i];
        px1[i] = ((double)a.x - ppx) / focal;
        py1[i] = ((double)a.y - ppy) / focal;
        px2[i] = ((double)b.x - ppx) / focal;
        py2[i] = ((double)b.y - ppy) / focal;
    }
}

__global__ void __launch_bounds__(ER_THREADS)
k_essential_ransac(const float* __restrict__ mkpts0, const float* __restrict__ mkpts1,
                   const int32_t* __restrict__ n_corr, int n_max, double focal, double ppx, double ppy,
                   double prob, double threshold, int max_iters, double* __restrict__ E_out,
                   uint8_t* __restrict__ mask_out, int32_t* __restrict__ n_inliers,
                   int32_t* __restrict__ iters_run, int32_t* __restrict__ status) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* px1 = reinterpret_cast<double*>(smem_raw);
    double* py1 = px1 + n_max;
    double* px2 = py1 + n_max;
    double* py2 = px2 + n_max;
    double* models = py2 + n_max;  // [ER_CHUNK][ER_MAXMODELS][9]
    ERShared& sh = *reinterpret_cast<ERShared*>(models + ER_CHUNK * ER_MAXMODELS * 9);

    const int p = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(n_corr[p], n_max);
    uint8_t* mask = mask_out + (size_t)p * n_max;
    double* Eo = E_out + 9 * (size_t)p;

    if (n < 5) {  // cv2 returns (None, None): RANSACPointSetRegistrator::run -> count < modelPoints
        for (int i = tid; i < n_max; i += ER_THREADS) mask[i] = 0;
        if (tid < 9) Eo[tid] = 0.0;
        if (tid == 0) {
            n_inliers[p] = 0;
            if (iters_run) iters_run[p] = 0;
            status[p] = FS_PAIR_TOO_FEW;
        }
        return;
    }
    stage_points(mkpts0 + 2 * (size_t)p * n_max, mkpts1 + 2 * (size_t)p * n_max, n, focal, ppx, ppy, px1, py1,
                 px2, py2);
    double thr = threshold / ((focal + focal) / 2);
    const float thr2 = (float)(thr * thr);

    cvm::RNG rng(0xFFFFFFFFFFFFFFFFull);  // RNG rng((uint64)-1) inside run(); only thread 0 advances it
    if (tid == 0) {
        sh.niters = max(max_iters, 1);
        sh.best = 0;
        sh.done_iters = 0;
        sh.stop = 0;
    }
    __syncthreads();

    for (int base = 0;; base += ER_CHUNK) {
        const int niters = sh.niters;
        if (base >= niters || sh.stop) break;
        const int round = (n == 5) ? 1 : min(ER_CHUNK, niters - base);
        // ---- sample (getSubset): draw until 5 distinct indices, draw order kept
        if (tid == 0) {
            for (int s = 0; s < round; s++) {
                int* idx = sh.subsets[s];
                if (n == 5) {  // count == modelPoints: the kernel runs once on all points
                    for (int i = 0; i < 5; i++) idx[i] = i;
                    continue;
                }
                for (int i = 0; i < 5; i++) {
                    int v;
                    bool dup;
                    do {
                        v = rng.uniform(0, n);
                        dup = false;
                        for (int k = 0; k < i; k++) dup |= idx[k] == v;
                    } while (dup);
                    idx[i] = v;
                }
            }
        }
        __syncthreads();
        // ---- solve: one thread per iteration of the round.  The solver is a long serial fp64 chain (latency bound),
        // so the 32 solves are spread over all 8 warps (4 lanes each): every SM sub-partition interleaves two
        // independent instruction streams instead of idling behind one warp.
        {
            constexpr int LANES = ER_CHUNK / ER_WARPS;  // 4
            if (lane < LANES) {
                const int s = warp * LANES + lane;
                int cnt = 0;
                if (s < round) {
                    double q1[5][2], q2[5][2];
                    for (int i = 0; i < 5; i++) {
                        const int k = sh.subsets[s][i];
                        q1[i][0] = px1[k]; q1[i][1] = py1[k];
                        q2[i][0] = px2[k]; q2[i][1] = py2[k];
                    }
                    cnt = cvm::five_point(q1, q2, models + (size_t)s * ER_MAXMODELS * 9);
                }
                sh.nmodels[s] = cnt;
            }
        }
        __syncthreads();
        if (warp == 0) {
            // exclusive prefix of model counts -> flat hypothesis list
            const int cnt = sh.nmodels[lane];
            int incl = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            sh.hyp_offset[lane] = incl - cnt;
            if (lane == 31) sh.hyp_offset[32] = incl;
        }
        __syncthreads();
        // ---- score: one warp per hypothesis
        const int H = sh.hyp_offset[ER_CHUNK];
        for (int h = warp; h < H; h += ER_WARPS) {
            // locate (s, k): largest s with hyp_offset[s] <= h  (32 entries, one per lane)
            const unsigned le = __ballot_sync(0xffffffffu, sh.hyp_offset[lane] <= h && sh.nmodels[lane] > 0 &&
                                                               h < sh.hyp_offset[lane] + sh.nmodels[lane]);
            const int s = __ffs(le) - 1;
            const int k = h - sh.hyp_offset[s];
            double E[9];
            const double* m = models + ((size_t)s * ER_MAXMODELS + k) * 9;
#pragma unroll
            for (int i = 0; i < 9; i++) E[i] = m[i];
            const int c = warp_count_inliers(E, px1, py1, px2, py2, n, thr2, lane);
            if (lane == 0) sh.counts[s][k] = c;
        }
        __syncthreads();
        // ---- select: OpenCV's serial order over (iteration, candidate)
        if (tid == 0) {
            int nit = sh.niters, best = sh.best, done = sh.done_iters;
            for (int s = 0; s < round; s++) {
                const int it = base + s;
                if (it >= nit) break;
                for (int k = 0; k < sh.nmodels[s]; k++) {
                    const int c = sh.counts[s][k];
                    if (c > max(best, 4)) {
                        best = c;
                        const double* m = models + ((size_t)s * ER_MAXMODELS + k) * 9;
                        for (int i = 0; i < 9; i++) sh.bestE[i] = m[i];
                        nit = cvm::ransac_update_num_iters(prob, (double)(n - c) / n, 5, nit);
                    }
                }
                done = it + 1;
            }
            sh.niters = nit;
            sh.best = best;
            sh.done_iters = done;
            if (n == 5) sh.stop = 1;
        }
        __syncthreads();
    }

    // ---- outputs: the winner's mask is recomputed (identical arithmetic) instead of being stored per hypothesis
    const int best = sh.best;
    if (best == 0) {
        for (int i = tid; i < n_max; i += ER_THREADS) mask[i] = 0;
        if (tid < 9) Eo[tid] = 0.0;
        if (tid == 0) {
            n_inliers[p] = 0;
            if (iters_run) iters_run[p] = sh.done_iters;
            status[p] = FS_PAIR_NO_MODEL;
        }
        return;
    }
    double E[9];
#pragma unroll
    for (int i = 0; i < 9; i++) E[i] = sh.bestE[i];
    for (int i = tid; i < n_max; i += ER_THREADS) {
        uint8_t v = 0;
        if (i < n) v = (n == 5) ? 1 : (sampson_f32(E, px1[i], py1[i], px2[i], py2[i]) <= thr2 ? 1 : 0);
        mask[i] = v;
    }
    if (tid < 9) Eo[tid] = E[tid];
    if (tid == 0) {
        n_inliers[p] = (n == 5) ? 5 : best;
        if (iters_run) iters_run[p] = sh.done_iters;
        status[p] = FS_PAIR_OK;
    }
}

__global__ void k_ransac_subsets(const int32_t* __restrict__ n_corr, int n_pairs, int max_iters,
                                 int32_t* __restrict__ subsets) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    const int n = n_corr[p];
    int32_t* out = subsets + (size_t)p * max_iters * 5;
    if (n < 5) {
        for (int i = 0; i < max_iters * 5; i++) out[i] = -1;
        return;
    }
    cvm::RNG rng(0xFFFFFFFFFFFFFFFFull);
    for (int it = 0; it < max_iters; it++) {
        int idx[5];
        for (int i = 0; i < 5; i++) {
            if (n == 5) {
                idx[i] = i;
                continue;
            }
            int v;
            bool dup;
            do {
                v = rng.uniform(0, n);
                dup = false;
                for (int k = 0; k < i; k++) dup |= idx[k] == v;
            } while (dup);
            idx[i] = v;
        }
        for (int i = 0; i < 5; i++) out[5 * it + i] = idx[i];
    }
}

// config-4 sweep: every CTA stages the pair's correspondences once and scores a slab of hypotheses
constexpr int SC_THREADS = 256;
constexpr int SC_HYP_PER_CTA = 64;

__global__ void __launch_bounds__(SC_THREADS)
k_score_essential(const float* __restrict__ mkpts0, const float* __restrict__ mkpts1,
                  const int32_t* __restrict__ n_corr, int n_max, double focal, double ppx, double ppy,
                  double threshold, const double* __restrict__ models, int n_hyp, int32_t* __restrict__ counts) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* px1 = reinterpret_cast<double*>(smem_raw);
    double* py1 = px1 + n_max;
    double* px2 = py1 + n_max;
    double* py2 = px2 + n_max;
    const int p = blockIdx.y;
    const int n = min(n_corr[p], n_max);
    stage_points(mkpts0 + 2 * (size_t)p * n_max, mkpts1 + 2 * (size_t)p * n_max, n, focal, ppx, ppy, px1, py1,
                 px2, py2);
    __syncthreads();
    double thr = threshold / ((focal + focal) / 2);
    const float thr2 = (float)(thr * thr);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h0 = blockIdx.x * SC_HYP_PER_CTA;
    for (int h = h0 + warp; h < min(h0 + SC_HYP_PER_CTA, n_hyp); h += SC_THREADS / 32) {
        double E[9];
        const double* m = models + ((size_t)p * n_hyp + h) * 9;
#pragma unroll
        for (int i = 0; i < 9; i++) E[i] = __ldg(m + i);
        const int c = warp_count_inliers(E, px1, py1, px2, py2, n, thr2, lane);
        if (lane == 0) counts[(size_t)p * n_hyp + h] = c;
    }
}

size_t er_smem_bytes(int n_max) {
    return (size_t)n_max * 4 * sizeof(double) + (size_t)ER_CHUNK * ER_MAXMODELS * 9 * sizeof(double) +
           sizeof(ERShared) + 64;
}

}  // namespace

extern "C" size_t fs_find_essential_workspace_bytes(int, int, int) { return 256; }

extern "C" int fs_find_essential_ransac(const float* mkpts0, const float* mkpts1, const int32_t* n_corr, int n_pairs,
                                        int n_max, double focal, double pp_x, double pp_y, double prob,
                                        double threshold, int max_iters, double* E, uint8_t* mask,
                                        int32_t* n_inliers, int32_t* iters_run, int32_t* status, void*, size_t,
                                        fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(n_max >= 1, "n_max must be positive");
    FS_REQUIRE(mkpts0 && mkpts1 && n_corr && E && mask && n_inliers && status, "null pointer");
    FS_REQUIRE(prob > 0 && prob < 1, "prob must be in (0, 1)");
    FS_REQUIRE(focal != 0, "focal must be non-zero");
    const size_t smem = er_smem_bytes(n_max);
    FS_REQUIRE(smem <= 227 * 1024, "n_max too large for the shared-memory staging (limit ~6200)");
    {
        static int granted[64];
        const int st_ = fs::ensure_dyn_smem(k_essential_ransac, granted, (size_t)((int)smem));
        if (st_ != FS_OK) return st_;
    }
    k_essential_ransac<<<n_pairs, ER_THREADS, smem, (cudaStream_t)stream>>>(
        mkpts0, mkpts1, n_corr, n_max, focal, pp_x, pp_y, prob, threshold, max_iters, E, mask, n_inliers, iters_run,
        status);
    FS_LAUNCH_CHECK("k_essential_ransac");
    return FS_OK;
}

extern "C" int fs_ransac_subsets(const int32_t* n_corr, int n_pairs, int max_iters, int32_t* subsets,
                                 fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && max_iters >= 0, "negative size");
    if (n_pairs == 0 || max_iters == 0) return FS_OK;
    FS_REQUIRE(n_corr && subsets, "null pointer");
    k_ransac_subsets<<<(n_pairs + 31) / 32, 32, 0, (cudaStream_t)stream>>>(n_corr, n_pairs, max_iters, subsets);
    FS_LAUNCH_CHECK("k_ransac_subsets");
    return FS_OK;
}

extern "C" int fs_score_essential(const float* mkpts0, const float* mkpts1, const int32_t* n_corr, int n_pairs,
                                  int n_max, double focal, double pp_x, double pp_y, double threshold,
                                  const double* models_E, int n_hyp, int32_t* counts, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0 && n_hyp >= 0, "negative size");
    if (n_pairs == 0 || n_hyp == 0) return FS_OK;
    FS_REQUIRE(n_max >= 1, "n_max must be positive");
    FS_REQUIRE(mkpts0 && mkpts1 && n_corr && models_E && counts, "null pointer");
    const size_t smem = (size_t)n_max * 4 * sizeof(double);
    FS_REQUIRE(smem <= 227 * 1024, "n_max too large for the shared-memory staging");
    {
        static int granted[64];
        const int st_ = fs::ensure_dyn_smem(k_score_essential, granted, (size_t)((int)smem));
        if (st_ != FS_OK) return st_;
    }
    dim3 grid((n_hyp + SC_HYP_PER_CTA - 1) / SC_HYP_PER_CTA, n_pairs);
    k_score_essential<<<grid, SC_THREADS, smem, (cudaStream_t)stream>>>(mkpts0, mkpts1, n_corr, n_max, focal, pp_x,
                                                                      pp_y, threshold, models_E, n_hyp, counts);
    FS_LAUNCH_CHECK("k_score_essential");
    return FS_OK;
}
