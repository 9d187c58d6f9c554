// ransac_essential.cu — A4: essential-matrix RANSAC with OpenCV's exact loop semantics for a whole batch of frame
// pairs, everything (sampler, minimal solver, scoring, adaptive stop) on the device.
//
// Replaces  E, mask = cv2.findEssentialMat(mkpts0, mkpts1, focal=K0[0,0], pp=(K0[0,2], K0[1,2]),
//                                          method=cv2.RANSAC, prob=0.999, threshold=1.0)    mono_slam.py:111
// (OpenCV: calib3d/five-point.cpp findEssentialMat + EMEstimatorCallback, calib3d/ptsetreg.cpp
//  RANSACPointSetRegistrator::run / getSubset / findInliers / RANSACUpdateNumIters.)
//
// Launch sequence (ransac_grid.cuh): normalise -> rounds of {select + sample, solve, score} spread over the whole
// GPU -> finish (the winner's mask is recomputed with identical arithmetic instead of being stored per hypothesis).
//   solve  : one thread per RANSAC iteration, Nister five-point (fivepoint.cuh) -> <= 10 candidate E;
//   score  : ONE WARP PER HYPOTHESIS over the pair's correspondences, which live in shared memory as normalised
//            doubles (SoA); Sampson error in OpenCV's operation order (no FMA), cast to float32, compared with
//            (float)thr^2, counted with a warp reduction.
#include "common.cuh"
#include "fivepoint.cuh"
#include "ransac_grid.cuh"

namespace {

using cvm::sampson_f32;

// warp-cooperative inlier count of one hypothesis
__device__ __forceinline__ int warp_count_inliers(const double* E, const double* px1, const double* py1,
                                                  const double* px2, const double* py2, int n, float thr2,
                                                  int lane) {
    int cnt = 0;
    for (int i = lane; i < n; i += 32) cnt += sampson_f32(E, px1[i], py1[i], px2[i], py2[i]) <= thr2 ? 1 : 0;
    return __reduce_add_sync(0xffffffffu, cnt);
}

__device__ __forceinline__ void stage_points(const float* mk0, const float* mk1, int n, double focal, double ppx,
                                             double ppy, double* px1, double* py1, double* px2, double* py2) {
    // findEssentialMat: points converted to double, (x - cx) / fx, (y - cy) / fy with fx = fy = focal
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float2 a = reinterpret_cast<const float2*>(mk0)[i];
        const float2 b = reinterpret_cast<const float2*>(mk1)[i];
        const cvm::CvNormalise nrm(focal, ppx, ppy);
        px1[i] = nrm.x(a.x);
        py1[i] = nrm.y(a.y);
        px2[i] = nrm.x(b.x);
        py2[i] = nrm.y(b.y);
    }
}

// ------------------------------------------------------------------------------------------------ engine policy
struct EPolicy {
    static constexpr int FIRST_CHUNK = 48;         // speculative iterations per pair in round 0 (~30 are typical; profiles/r01zv)
    static constexpr int FIRST_CHUNK_MIN = 32;     // ... large batches go down to this (ransac_grid.cuh: make_plan)
    static constexpr int SOLVE_WARPS_PER_SM = 4;  // solver warps per SM (ransac_grid.cuh: run_rounds)
    static constexpr int SOLVE_WARPS_PER_SM_MAX = 6;  // ... when round 0 has more solves than 4 warps per SM have lanes
    static constexpr const char* TUNE_FIRST = "FS_TUNE_E_FIRST_CHUNK";
    static constexpr const char* TUNE_WARPS = "FS_TUNE_E_SOLVE_WARPS";
    static constexpr int MP = 5;     // model points
    static constexpr int MAXM = 10;  // candidate models per sample
    static constexpr int MD = 9;     // doubles per model
    static constexpr int SLAB = 8;   // iterations scored by one CTA (<= 80 hypotheses)
    struct Params {
        const double* pts;  // [n_pairs][4][n_max]: x1 | y1 | x2 | y2, normalised
        int n_max;
        float thr2;
    };
    static size_t stage_bytes(int n_max) { return (size_t)n_max * 4 * sizeof(double); }
    __device__ static void stage(const Params& prm, int p, int n, unsigned char* smem) {
        double* dst = reinterpret_cast<double*>(smem);
        const double* src = prm.pts + (size_t)p * 4 * prm.n_max;
        for (int c = 0; c < 4; c++)
            for (int i = threadIdx.x; i < n; i += blockDim.x) dst[c * prm.n_max + i] = src[c * prm.n_max + i];
    }
    // cv::solvePoly runs 300 Durand-Kerner sweeps on every polynomial and the bits it returns are those of sweep 300
    // (fivepoint.cuh: dk_stop_rule), so every solve runs them all: ~480k dependent-ish fp64 operations per thread, the
    // bulk of the solver.  One warp per SM sub-partition is both the latency and the fp64-issue optimum (a sweep's
    // critical path of ~420 dependent operations keeps the pipe ~3/4 busy from one warp), so no solve is deferred.
    static constexpr int DK_BUDGET = 300;
    static constexpr bool HAS_TAIL = false;
    struct Tail {
        int slot, pad;
    };
    static constexpr int SOLVE_THREADS = 64;
    static constexpr int SOLVE_SMEM_DOUBLES = 0;  // the five-point solver keeps its state in registers / local memory
    static constexpr int STATE_DOUBLES = 0;       // one launch per solve
    static constexpr int SOLVE2_THREADS = 64;
    __device__ static int solve2(const Params&, const double*, double*) { return 0; }
    __device__ static int solve(const Params& prm, int p, const int* sub, double* models, int slot, Tail* tail,
                                int* n_tail, double*, double*) {
        const double* src = prm.pts + (size_t)p * 4 * prm.n_max;
        double q1[5][2], q2[5][2];
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const int k = sub[i];
            q1[i][0] = src[k];
            q1[i][1] = src[prm.n_max + k];
            q2[i][0] = src[2 * prm.n_max + k];
            q2[i][1] = src[3 * prm.n_max + k];
        }
        cvm::FPSetup fs;
        if (!cvm::five_point_setup(q1, q2, fs)) return 0;
        cvm::Cplx roots[10];
        const int n = cvm::dk_solve(fs.coeffs, roots, DK_BUDGET);
        return cvm::five_point_models(fs, roots, n, models);
    }
    __device__ static int solve_tail(const Params&, Tail&, double*) { return 0; }
    struct Scorer {
        double E[9];
        __device__ Scorer(const Params&, const double* m) {
#pragma unroll
            for (int i = 0; i < 9; i++) E[i] = m[i];
        }
        __device__ bool inlier(const unsigned char* smem, const Params& prm, int i) const {
            const double* s = reinterpret_cast<const double*>(smem);
            return sampson_f32(E, s[i], s[prm.n_max + i], s[2 * prm.n_max + i], s[3 * prm.n_max + i]) <= prm.thr2;
        }
    };
};

__global__ void __launch_bounds__(256)
k_er_normalise(const float* __restrict__ mkpts0, const float* __restrict__ mkpts1, const int32_t* __restrict__ n_corr,
               int n_max, double focal, double ppx, double ppy, double* __restrict__ pts) {
    const int p = blockIdx.y;
    const int n = min(n_corr[p], n_max);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 a = reinterpret_cast<const float2*>(mkpts0)[(size_t)p * n_max + i];
    const float2 b = reinterpret_cast<const float2*>(mkpts1)[(size_t)p * n_max + i];
    double* dst = pts + (size_t)p * 4 * n_max;
    const cvm::CvNormalise nrm(focal, ppx, ppy);
    dst[i] = nrm.x(a.x);
    dst[n_max + i] = nrm.y(a.y);
    dst[2 * n_max + i] = nrm.x(b.x);
    dst[3 * n_max + i] = nrm.y(b.y);
}

__global__ void __launch_bounds__(256)
k_er_finish(const rg::PairState<9>* __restrict__ st, const double* __restrict__ pts, int n_max, float thr2,
            double* __restrict__ E_out, uint8_t* __restrict__ mask_out, int32_t* __restrict__ n_inliers,
            int32_t* __restrict__ iters_run, int32_t* __restrict__ status) {
    const int p = blockIdx.y;
    const rg::PairState<9>& S = st[p];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool ok = S.status == FS_PAIR_OK;
    if (blockIdx.x == 0) {
        if (threadIdx.x < 9) E_out[9 * (size_t)p + threadIdx.x] = ok ? S.best_model[threadIdx.x] : 0.0;
        if (threadIdx.x == 0) {
            n_inliers[p] = ok ? (S.all_points ? 5 : S.best) : 0;
            if (iters_run) iters_run[p] = S.status == FS_PAIR_TOO_FEW ? 0 : S.done;
            status[p] = S.status;
        }
    }
    if (i >= n_max) return;
    uint8_t v = 0;
    if (ok && i < S.n) {
        if (S.all_points) {
            v = 1;
        } else {
            double E[9];
#pragma unroll
            for (int k = 0; k < 9; k++) E[k] = S.best_model[k];
            const double* s = pts + (size_t)p * 4 * n_max;
            v = sampson_f32(E, s[i], s[n_max + i], s[2 * n_max + i], s[3 * n_max + i]) <= thr2 ? 1 : 0;
        }
    }
    mask_out[(size_t)p * n_max + i] = v;
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

}  // namespace

extern "C" size_t fs_find_essential_workspace_bytes(int n_pairs, int n_max, int max_iters) {
    const rg::Plan pl = rg::make_plan<EPolicy>(n_pairs, max_iters);
    const size_t np = (size_t)(n_pairs > 0 ? n_pairs : 1), nm = (size_t)(n_max > 0 ? n_max : 1);
    return fs::align_up(np * 4 * nm * sizeof(double), 256) + rg::Workspace<EPolicy>::bytes(n_pairs, pl) + 512;
}

extern "C" int fs_find_essential_ransac(const float* mkpts0, const float* mkpts1, const int32_t* n_corr, int n_pairs,
                                        int n_max, double focal, double pp_x, double pp_y, double prob,
                                        double threshold, int max_iters, double* E, uint8_t* mask,
                                        int32_t* n_inliers, int32_t* iters_run, int32_t* status, void* workspace,
                                        size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(n_max >= 1, "n_max must be positive");
    FS_REQUIRE(n_pairs <= 65535, "n_pairs too large for one launch (limit 65535)");
    FS_REQUIRE(mkpts0 && mkpts1 && n_corr && E && mask && n_inliers && status, "null pointer");
    FS_REQUIRE(prob > 0 && prob < 1, "prob must be in (0, 1)");
    FS_REQUIRE(focal != 0, "focal must be non-zero");
    const size_t smem = EPolicy::stage_bytes(n_max);
    FS_REQUIRE(smem <= 227 * 1024, "n_max too large for the shared-memory staging (limit ~7200)");
    const size_t need = fs_find_essential_workspace_bytes(n_pairs, n_max, max_iters);
    if (!workspace || workspace_bytes < need) {
        fs::set_error("fs_find_essential_ransac: workspace %zu < %zu", workspace_bytes, need);
        return FS_ERR_WORKSPACE;
    }
    const rg::Plan pl = rg::make_plan<EPolicy>(n_pairs, max_iters);
    FS_REQUIRE(pl.pool < (size_t)1 << 30, "n_pairs * max_iters too large");
    fs::Carver cv(workspace);
    double* pts = cv.take<double>((size_t)n_pairs * 4 * n_max);
    rg::Workspace<EPolicy> ws;
    ws.carve(cv, n_pairs, pl);
    const double thr = threshold / ((focal + focal) / 2);
    EPolicy::Params prm{pts, n_max, (float)(thr * thr)};
    dim3 pgrid((n_max + 255) / 256, n_pairs);
    k_er_normalise<<<pgrid, 256, 0, stream>>>(mkpts0, mkpts1, n_corr, n_max, focal, pp_x, pp_y, pts);
    FS_LAUNCH_CHECK("k_er_normalise");
    const int st_ = rg::run_rounds<EPolicy>(n_pairs, n_corr, n_max, /*min_points*/ 5, max_iters, prob, prm, pl, ws, smem,
                                            stream);
    if (st_ != FS_OK) return st_;
    k_er_finish<<<pgrid, 256, 0, stream>>>(ws.st, pts, n_max, prm.thr2, E, mask, n_inliers, iters_run, status);
    FS_LAUNCH_CHECK("k_er_finish");
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
