// ransac_pnp.cu — A7: PnP RANSAC (cv2.solvePnPRansac, stereo_slam.py:294-295) and the reprojection scoring sweep.
// Scoring: one warp per hypothesis over all 3D-2D correspondences staged in shared memory (SoA), projection
// with the reference's distortion coefficients in OpenCV's operation order, float32 error <= (float)thr^2.
#include "common.cuh"
#include "epnp.cuh"
#include "ransac_grid.cuh"

namespace {

constexpr int PS_THREADS = 256;
constexpr int PS_HYP_PER_CTA = 64;

__device__ __forceinline__ cvm::CamD to_cam(const FSCamera& c) {
    cvm::CamD d;
    d.fx = c.fx; d.fy = c.fy; d.cx = c.cx; d.cy = c.cy;
    for (int i = 0; i < 5; i++) d.k[i] = c.dist[i];
    return d;
}

// shared-memory staging of one pair's correspondences: X Y Z as double (projectPoints converts the float32
// object points to double), u v as float32
struct PnPStage {
    double *X, *Y, *Z;
    float *u, *v;
    __device__ void carve(unsigned char* raw, int n_max) {
        X = reinterpret_cast<double*>(raw);
        Y = X + n_max;
        Z = Y + n_max;
        u = reinterpret_cast<float*>(Z + n_max);
        v = u + n_max;
    }
    __device__ void load(const float* p3, const float* p2, int n) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            X[i] = (double)p3[3 * i];
            Y[i] = (double)p3[3 * i + 1];
            Z[i] = (double)p3[3 * i + 2];
            const float2 q = reinterpret_cast<const float2*>(p2)[i];
            u[i] = q.x;
            v[i] = q.y;
        }
    }
    __host__ __device__ static size_t bytes(int n_max) { return (size_t)n_max * (3 * sizeof(double) + 2 * sizeof(float)); }
};

__device__ __forceinline__ int warp_count_reproj(const double* R, const double* t, const cvm::CamD& cam,
                                                 const PnPStage& st, int n, float thr2, int lane) {
    int cnt = 0;
    for (int i = lane; i < n; i += 32) {
        float pu, pv;
        cvm::project_point_f32(R, t, cam, st.X[i], st.Y[i], st.Z[i], pu, pv);
        cnt += cvm::reproj_err_f32(st.u[i], st.v[i], pu, pv) <= thr2 ? 1 : 0;
    }
    return __reduce_add_sync(0xffffffffu, cnt);
}

__global__ void __launch_bounds__(PS_THREADS)
k_score_pnp(const float* __restrict__ points3d, const float* __restrict__ pts2d, const int32_t* __restrict__ n_corr,
            int n_max, FSCamera cam_, double reproj_error, const double* __restrict__ models, int n_hyp,
            int32_t* __restrict__ counts) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    PnPStage st;
    st.carve(smem_raw, n_max);
    const int p = blockIdx.y;
    const int n = min(n_corr[p], n_max);
    st.load(points3d + 3 * (size_t)p * n_max, pts2d + 2 * (size_t)p * n_max, n);
    __syncthreads();
    const cvm::CamD cam = to_cam(cam_);
    const float thr2 = (float)(reproj_error * reproj_error);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int h0 = blockIdx.x * PS_HYP_PER_CTA;
    for (int h = h0 + warp; h < min(h0 + PS_HYP_PER_CTA, n_hyp); h += PS_THREADS / 32) {
        const double* m = models + ((size_t)p * n_hyp + h) * 6;
        double rv[3], t[3], R[9];
        for (int i = 0; i < 3; i++) {
            rv[i] = __ldg(m + i);
            t[i] = __ldg(m + 3 + i);
        }
        cvm::rodrigues_vec2mat(rv, R);
        const int c = warp_count_reproj(R, t, cam, st, n, thr2, lane);
        if (lane == 0) counts[(size_t)p * n_hyp + h] = c;
    }
}


// ------------------------------------------------------------------------------------------------ A7 RANSAC
// Runs on the batch-wide engine of ransac_grid.cuh (rounds of select + sample / solve / score over the whole GPU):
//   solve  : one thread per RANSAC iteration, EPnP-on-5-points (epnp.cuh) -> one model (rvec | tvec);
//   score  : ONE WARP PER HYPOTHESIS over the pair's correspondences staged in shared memory: projectPoints with the
//            reference's distortion coefficients, float32 reprojection error <= (float)thr^2, warp-reduced count;
//   finish : mask of the winner (identical arithmetic) + ascending inlier index list.
struct PPolicy {
    static constexpr int FIRST_CHUNK = 192;        // speculative iterations per pair in round 0 (~390 are typical)
    static constexpr int FIRST_CHUNK_MIN = 192;    // (no batch-size dependence: the sweep is flat, profiles/README.md)
    static constexpr int SOLVE_WARPS_PER_SM = 4;  // solver warps per SM (ransac_grid.cuh: run_rounds)
    static constexpr int SOLVE_WARPS_PER_SM_MAX = 6;  // three CTAs of 64 fill the SM's shared memory (72 KB each)
    static constexpr const char* TUNE_FIRST = "FS_TUNE_P_FIRST_CHUNK";
    static constexpr const char* TUNE_WARPS = "FS_TUNE_P_SOLVE_WARPS";
    static constexpr int MP = 5;
    static constexpr int MAXM = 1;
    static constexpr int MD = 6;
    static constexpr int SLAB = 32;  // iterations (= hypotheses) scored by one CTA
    static constexpr bool HAS_TAIL = false;
    struct Tail {
        int slot;
    };
    struct Params {
        const float* points3d;  // [n_pairs][n_max][3]
        const float* pts2d;     // [n_pairs][n_max][2]
        int n_max;
        float thr2;
        cvm::CamD cam;
    };
    static size_t stage_bytes(int n_max) { return PnPStage::bytes(n_max); }
    __device__ static void stage(const Params& prm, int p, int n, unsigned char* smem) {
        PnPStage st;
        st.carve(smem, prm.n_max);
        st.load(prm.points3d + 3 * (size_t)p * prm.n_max, prm.pts2d + 2 * (size_t)p * prm.n_max, n);
    }
    // 12x12 U^T of EPnP's decomposition, per solver thread, in shared memory (72 KB per CTA of 64: three CTAs per SM)
    static constexpr int SOLVE_SMEM_DOUBLES = 144;
    // (measured: CTAs of 32 threads so that five solver warps fit an SM are SLOWER — 33.3k pairs/s at four warps, 31k at
    // five, against 37k for two CTAs of 64: a fifth warp doubles up on one SM sub-partition and stretches the critical path)
    static constexpr int SOLVE_THREADS = 64;
    // Two launches per solve: the first half (undistort, control points, M^T M, the 12x12 Jacobi SVD: 57 % of the solver)
    // needs the shared-memory matrices and is held to four warps per SM by them; the second half (three beta
    // approximations x Gauss-Newton, absolute orientation, Rodrigues) needs none and runs at eight warps per SM.
    static constexpr int STATE_DOUBLES = cvm::EPnP<5>::STATE_DOUBLES;
    static constexpr int SOLVE2_THREADS = 64;
    __device__ static int solve2(const Params& prm, const double* state, double* model) {
        double m[6];
        cvm::pnp_minimal_epnp5_finish(prm.cam, state, m);
#pragma unroll
        for (int i = 0; i < 6; i++) model[i] = m[i];
        return 1;
    }
    __device__ static int solve(const Params& prm, int p, const int* sub, double* model, int, Tail*, int*,
                                double* smem, double* state) {
        const float* p3 = prm.points3d + 3 * (size_t)p * prm.n_max;
        const float2* p2 = reinterpret_cast<const float2*>(prm.pts2d) + (size_t)p * prm.n_max;
        double obj[5][3];
        float img[5][2];
#pragma unroll
        for (int i = 0; i < 5; i++) {
            const int k = sub[i];
            obj[i][0] = (double)p3[3 * k];
            obj[i][1] = (double)p3[3 * k + 1];
            obj[i][2] = (double)p3[3 * k + 2];
            const float2 q = p2[k];
            img[i][0] = q.x;
            img[i][1] = q.y;
        }
        using SP = cvm::Strided<SOLVE_THREADS>;
        double w[12];  // the singular values: 96 B of local memory, so that the matrix alone sets the shared-memory footprint
        cvm::pnp_minimal_epnp5_basis<SP, double*>(prm.cam, obj, img, state, SP{smem}, w);
        return -1;  // finished by k_rg_solve2
    }
    __device__ static int solve_tail(const Params&, Tail&, double*) { return 0; }
    struct Scorer {
        double R[9], t[3];
        __device__ Scorer(const Params&, const double* m) {
            double rv[3];
#pragma unroll
            for (int i = 0; i < 3; i++) {
                rv[i] = m[i];
                t[i] = m[3 + i];
            }
            cvm::rodrigues_vec2mat(rv, R);
        }
        __device__ bool inlier(unsigned char* smem, const Params& prm, int i) const {
            PnPStage st;
            st.carve(smem, prm.n_max);
            float pu, pv;
            cvm::project_point_f32(R, t, prm.cam, st.X[i], st.Y[i], st.Z[i], pu, pv);
            return cvm::reproj_err_f32(st.u[i], st.v[i], pu, pv) <= prm.thr2;
        }
    };
};

// rare path (exactly 5 correspondences) kept out of line so that it does not set the register budget of k_pnp_finish
__device__ __noinline__ void epnp5_all_points(const cvm::CamD& cam, const float* p3, const float2* p2, double* m) {
    double obj[5][3];
    float img[5][2];
    for (int i = 0; i < 5; i++) {
        obj[i][0] = (double)p3[3 * i]; obj[i][1] = (double)p3[3 * i + 1]; obj[i][2] = (double)p3[3 * i + 2];
        img[i][0] = p2[i].x; img[i][1] = p2[i].y;
    }
    cvm::pnp_minimal_epnp5(cam, obj, img, m);
}

constexpr int PF_THREADS = 256;
constexpr int PF_WARPS = PF_THREADS / 32;

// one CTA per pair: outputs of the RANSAC stage (model, {0,1} mask, ascending inlier list, status)
__global__ void __launch_bounds__(PF_THREADS)
k_pnp_finish(const rg::PairState<6>* __restrict__ stt, PPolicy::Params prm, double* __restrict__ model_out,
             uint8_t* __restrict__ mask_out, int32_t* __restrict__ inlier_idx, int32_t* __restrict__ n_inliers,
             int32_t* __restrict__ iters_run, int32_t* __restrict__ status) {
    __shared__ int s_warp[33];
    __shared__ double s_model[6];
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const rg::PairState<6>& S = stt[p];
    const int n_max = prm.n_max, n = S.n;
    uint8_t* mask = mask_out + (size_t)p * n_max;
    double* mo = model_out + 6 * (size_t)p;
    const float* p3 = prm.points3d + 3 * (size_t)p * n_max;
    const float2* p2 = reinterpret_cast<const float2*>(prm.pts2d) + (size_t)p * n_max;
    if (S.all_points) {
        // model_points == npoints: solvePnP(EPNP) on all points, every point an inlier, no refit
        if (tid == 0) {
            double m[6];
            epnp5_all_points(prm.cam, p3, p2, m);
            for (int i = 0; i < 6; i++) mo[i] = m[i];
            n_inliers[p] = 5;
            if (iters_run) iters_run[p] = 0;
            status[p] = FS_PAIR_OK;
        }
        for (int i = tid; i < n_max; i += PF_THREADS) {
            mask[i] = i < 5 ? 1 : 0;
            if (inlier_idx && i < 5) inlier_idx[(size_t)p * n_max + i] = i;
        }
        return;
    }
    if (S.status != FS_PAIR_OK) {
        for (int i = tid; i < n_max; i += PF_THREADS) mask[i] = 0;
        if (tid < 6) mo[tid] = 0.0;
        if (tid == 0) {
            n_inliers[p] = 0;
            if (iters_run) iters_run[p] = S.status == FS_PAIR_TOO_FEW ? 0 : S.done;
            status[p] = S.status;
        }
        return;
    }
    if (tid < 6) s_model[tid] = S.best_model[tid];
    __syncthreads();
    double rv[3], t[3], R[9];
    for (int i = 0; i < 3; i++) {
        rv[i] = s_model[i];
        t[i] = s_model[3 + i];
    }
    cvm::rodrigues_vec2mat(rv, R);
    int basei = 0;
    for (int i0 = 0; i0 < n_max; i0 += PF_THREADS) {
        const int i = i0 + tid;
        bool in = false;
        if (i < n) {
            float pu, pv;
            cvm::project_point_f32(R, t, prm.cam, (double)p3[3 * i], (double)p3[3 * i + 1], (double)p3[3 * i + 2], pu, pv);
            const float2 q = p2[i];
            in = cvm::reproj_err_f32(q.x, q.y, pu, pv) <= prm.thr2;
        }
        if (i < n_max) mask[i] = in ? 1 : 0;
        const unsigned bal = __ballot_sync(0xffffffffu, in);
        __syncthreads();
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        if (warp == 0) {
            int v = lane < PF_WARPS ? s_warp[lane] : 0, incl = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int nb = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += nb;
            }
            s_warp[lane] = incl - v;
            if (lane == 31) s_warp[32] = incl;
        }
        __syncthreads();
        if (in && inlier_idx)
            inlier_idx[(size_t)p * n_max + basei + s_warp[warp] + __popc(bal & ((1u << lane) - 1))] = i;
        basei += s_warp[32];
    }
    if (tid < 6) mo[tid] = s_model[tid];
    if (tid == 0) {
        n_inliers[p] = S.best;
        if (iters_run) iters_run[p] = S.done;
        status[p] = FS_PAIR_OK;
    }
}


// ------------------------------------------------------------------------------------------------ A7 refit
// cv::solvePnPRansac ends with solvePnP(SOLVEPNP_ITERATIVE) on the consensus set: a Levenberg-Marquardt minimisation
// of the distorted reprojection error over the 6 pose parameters.  OpenCV starts it from a DLT pose; here it starts
// from the RANSAC winner and iterates to convergence on the same objective, so both land on the same minimum (the
// target is 1e-4 rad / 1e-4 relative, not bit-exactness).  Rotation updates are multiplicative, R <- exp(d) R.
// One CTA per pair; normal equations are reduced in a fixed order (deterministic).  Exactly 5 inliers: OpenCV's DLT
// raises and solvePnPRansac falls back to the minimal-sample model — same here.
constexpr int RF_THREADS = 256;
constexpr int RF_WARPS = RF_THREADS / 32;
constexpr int RF_NACC = 28;  // 21 (upper JtJ) + 6 (Jt e) + 1 (e^T e)
constexpr int RF_MAX_ITERS = 40;

struct RFState {
    double R[9], t[3], H[21], g[6], err;
    double Rc[9], tc[3], step2, lambda;
    double part[RF_WARPS][RF_NACC];
    double acc[RF_NACC];
    int done;
};

__device__ __forceinline__ void refit_point(const double* R, const double* t, const cvm::CamD& cam, double Xw, double Yw,
                                            double Zw, double u, double v, double* acc) {
    const double qx = R[0] * Xw + R[1] * Yw + R[2] * Zw, qy = R[3] * Xw + R[4] * Yw + R[5] * Zw,
                 qz = R[6] * Xw + R[7] * Yw + R[8] * Zw;
    const double X = qx + t[0], Y = qy + t[1], Z = qz + t[2];
    const double iz = Z != 0 ? 1. / Z : 1., x = X * iz, y = Y * iz;
    const double k1 = cam.k[0], k2 = cam.k[1], p1 = cam.k[2], p2 = cam.k[3], k3 = cam.k[4];
    const double r2 = x * x + y * y, cd = 1 + r2 * (k1 + r2 * (k2 + r2 * k3)), dcd = k1 + r2 * (2 * k2 + 3 * k3 * r2);
    const double xd = x * cd + p1 * (2 * x * y) + p2 * (r2 + 2 * x * x);
    const double yd = y * cd + p1 * (r2 + 2 * y * y) + p2 * (2 * x * y);
    const double eu = cam.fx * xd + cam.cx - u, ev = cam.fy * yd + cam.cy - v;
    const double xdx = cd + 2 * x * x * dcd + 2 * p1 * y + 6 * p2 * x, xdy = 2 * x * y * dcd + 2 * p1 * x + 2 * p2 * y;
    const double ydx = 2 * x * y * dcd + 2 * p1 * x + 2 * p2 * y, ydy = cd + 2 * y * y * dcd + 6 * p1 * y + 2 * p2 * x;
    // d(x, y)/dP
    const double xP[3] = {iz, 0., -x * iz}, yP[3] = {0., iz, -y * iz};
    double uP[3], vP[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        uP[k] = cam.fx * (xdx * xP[k] + xdy * yP[k]);
        vP[k] = cam.fy * (ydx * xP[k] + ydy * yP[k]);
    }
    // dP/d(delta) = -[q]_x, dP/dt = I
    double Ju[6], Jv[6];
    Ju[0] = -uP[1] * qz + uP[2] * qy; Ju[1] = uP[0] * qz - uP[2] * qx; Ju[2] = -uP[0] * qy + uP[1] * qx;
    Jv[0] = -vP[1] * qz + vP[2] * qy; Jv[1] = vP[0] * qz - vP[2] * qx; Jv[2] = -vP[0] * qy + vP[1] * qx;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        Ju[3 + k] = uP[k];
        Jv[3 + k] = vP[k];
    }
    int o = 0;
#pragma unroll
    for (int a = 0; a < 6; a++)
#pragma unroll
        for (int b = a; b < 6; b++) acc[o++] += Ju[a] * Ju[b] + Jv[a] * Jv[b];
#pragma unroll
    for (int a = 0; a < 6; a++) acc[21 + a] += Ju[a] * eu + Jv[a] * ev;
    acc[27] += eu * eu + ev * ev;
}

// Cholesky solve of the damped 6x6 system; returns false if not positive definite
__device__ bool solve6(const double* Hup, const double* g, double lambda, double* d) {
    double A[6][6];
    int o = 0;
    for (int a = 0; a < 6; a++)
        for (int b = a; b < 6; b++) {
            A[a][b] = A[b][a] = Hup[o++];
        }
    for (int a = 0; a < 6; a++) A[a][a] += lambda * A[a][a] + 1e-300;
    double L[6][6];
    for (int i = 0; i < 6; i++)
        for (int j = 0; j <= i; j++) {
            double s = A[i][j];
            for (int k = 0; k < j; k++) s -= L[i][k] * L[j][k];
            if (i == j) {
                if (!(s > 0)) return false;
                L[i][i] = sqrt(s);
            } else {
                L[i][j] = s / L[j][j];
            }
        }
    double y[6];
    for (int i = 0; i < 6; i++) {
        double s = -g[i];
        for (int k = 0; k < i; k++) s -= L[i][k] * y[k];
        y[i] = s / L[i][i];
    }
    for (int i = 5; i >= 0; i--) {
        double s = y[i];
        for (int k = i + 1; k < 6; k++) s -= L[k][i] * d[k];
        d[i] = s / L[i][i];
    }
    return true;
}

__global__ void __launch_bounds__(RF_THREADS)
k_pnp_refine(const float* __restrict__ points3d, const float* __restrict__ pts2d, const int32_t* __restrict__ n_corr,
             int n_max, FSCamera cam_, const double* __restrict__ model, const uint8_t* __restrict__ mask_in,
             const int32_t* __restrict__ n_inliers, const int32_t* __restrict__ status, double* __restrict__ rvec,
             double* __restrict__ tvec) {
    __shared__ RFState S;
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(n_corr[p], n_max);
    const double* m = model + 6 * (size_t)p;
    if (status[p] != FS_PAIR_OK || n_inliers[p] <= 5) {
        if (tid < 3) {
            rvec[3 * (size_t)p + tid] = m[tid];
            tvec[3 * (size_t)p + tid] = m[3 + tid];
        }
        return;
    }
    const cvm::CamD cam = to_cam(cam_);
    const float* p3 = points3d + 3 * (size_t)p * n_max;
    const float2* p2 = reinterpret_cast<const float2*>(pts2d) + (size_t)p * n_max;
    const uint8_t* mask = mask_in + (size_t)p * n_max;
    if (tid == 0) {
        double rv[3] = {m[0], m[1], m[2]};
        cvm::rodrigues_vec2mat(rv, S.Rc);
        for (int i = 0; i < 3; i++) S.tc[i] = m[3 + i];
        S.lambda = 1e-3;
        S.err = -1.0;  // no accepted state yet
        S.done = 0;
        S.step2 = 0;
    }
    __syncthreads();
    for (int it = 0; it <= RF_MAX_ITERS; it++) {
        // ---- evaluate the candidate (Rc, tc)
        double acc[RF_NACC];
#pragma unroll
        for (int k = 0; k < RF_NACC; k++) acc[k] = 0;
        double R[9], t[3];
#pragma unroll
        for (int k = 0; k < 9; k++) R[k] = S.Rc[k];
#pragma unroll
        for (int k = 0; k < 3; k++) t[k] = S.tc[k];
        for (int i = tid; i < n; i += RF_THREADS) {
            if (!mask[i]) continue;
            const float2 uv = p2[i];
            refit_point(R, t, cam, (double)p3[3 * i], (double)p3[3 * i + 1], (double)p3[3 * i + 2], (double)uv.x,
                        (double)uv.y, acc);
        }
#pragma unroll
        for (int k = 0; k < RF_NACC; k++) {
            double v = acc[k];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            if (lane == 0) S.part[warp][k] = v;
        }
        __syncthreads();
        if (tid < RF_NACC) {
            double v = 0;
            for (int w = 0; w < RF_WARPS; w++) v += S.part[w][tid];
            S.acc[tid] = v;
        }
        __syncthreads();
        // ---- accept / reject, next step
        if (tid == 0) {
            const double errc = S.acc[27];
            bool accepted = false;
            if (S.err < 0 || errc < S.err) {
                accepted = true;
                const double prev = S.err;
                for (int k = 0; k < 9; k++) S.R[k] = S.Rc[k];
                for (int k = 0; k < 3; k++) S.t[k] = S.tc[k];
                for (int k = 0; k < 21; k++) S.H[k] = S.acc[k];
                for (int k = 0; k < 6; k++) S.g[k] = S.acc[21 + k];
                S.err = errc;
                if (prev >= 0) {
                    S.lambda = fmax(S.lambda * 0.1, 1e-15);
                    if (S.step2 < 1e-26 || (prev - errc) <= 1e-15 * prev) S.done = 1;
                }
            } else {
                S.lambda *= 10.;
                if (S.lambda > 1e12) S.done = 1;
            }
            (void)accepted;
            if (!S.done && it < RF_MAX_ITERS) {
                double d[6];
                bool ok = solve6(S.H, S.g, S.lambda, d);
                while (!ok && S.lambda < 1e12) {
                    S.lambda = fmax(S.lambda * 10., 1e-6);
                    ok = solve6(S.H, S.g, S.lambda, d);
                }
                if (!ok) {
                    S.done = 1;
                } else {
                    double dR[9];
                    cvm::rodrigues_vec2mat(d, dR);
                    cvm::matmul33(dR, S.R, S.Rc);
                    for (int k = 0; k < 3; k++) S.tc[k] = S.t[k] + d[3 + k];
                    S.step2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2] + d[3] * d[3] + d[4] * d[4] + d[5] * d[5];
                }
            }
        }
        __syncthreads();
        if (S.done) break;
    }
    if (tid == 0) {
        double rv[3];
        cvm::rodrigues_mat2vec(S.R, rv);
        for (int k = 0; k < 3; k++) {
            rvec[3 * (size_t)p + k] = rv[k];
            tvec[3 * (size_t)p + k] = S.t[k];
        }
    }
}

}  // namespace

extern "C" int fs_score_pnp(const float* points3d, const float* pts2d, const int32_t* n_corr, int n_pairs, int n_max,
                            const FSCamera* cam, double reproj_error, const double* models_rt, int n_hyp,
                            int32_t* counts, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0 && n_hyp >= 0, "negative size");
    if (n_pairs == 0 || n_hyp == 0) return FS_OK;
    FS_REQUIRE(n_max >= 1, "n_max must be positive");
    FS_REQUIRE(points3d && pts2d && n_corr && cam && models_rt && counts, "null pointer");
    const size_t smem = PnPStage::bytes(n_max);
    FS_REQUIRE(smem <= 227 * 1024, "n_max too large for the shared-memory staging");
    {
        static int granted[64];
        const int st_ = fs::ensure_dyn_smem(k_score_pnp, granted, (size_t)((int)smem));
        if (st_ != FS_OK) return st_;
    }
    dim3 grid((n_hyp + PS_HYP_PER_CTA - 1) / PS_HYP_PER_CTA, n_pairs);
    k_score_pnp<<<grid, PS_THREADS, smem, (cudaStream_t)stream>>>(points3d, pts2d, n_corr, n_max, *cam, reproj_error,
                                                                models_rt, n_hyp, counts);
    FS_LAUNCH_CHECK("k_score_pnp");
    return FS_OK;
}


extern "C" size_t fs_solve_pnp_workspace_bytes(int n_pairs, int, int max_iters) {
    const rg::Plan pl = rg::make_plan<PPolicy>(n_pairs, max_iters);
    return fs::align_up((size_t)(n_pairs > 0 ? n_pairs : 1) * 6 * sizeof(double), 256) +
           rg::Workspace<PPolicy>::bytes(n_pairs, pl) + 512;
}

extern "C" int fs_solve_pnp_ransac(const float* points3d, const float* pts2d, const int32_t* n_corr, int n_pairs,
                                   int n_max, const FSCamera* cam, double reproj_error, double confidence,
                                   int max_iters, int min_points, double* rvec, double* tvec, uint8_t* mask,
                                   int32_t* inlier_idx, int32_t* n_inliers, int32_t* iters_run, int32_t* status,
                                   void* workspace, size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(n_max >= 1, "n_max must be positive");
    FS_REQUIRE(points3d && pts2d && n_corr && cam && rvec && tvec && mask && n_inliers && status, "null pointer");
    FS_REQUIRE(confidence > 0, "confidence must be in (0, 1], or > 1 to switch the adaptive iteration cap off");
    const size_t need = fs_solve_pnp_workspace_bytes(n_pairs, n_max, max_iters);
    if (!workspace || workspace_bytes < need) {
        fs::set_error("fs_solve_pnp_ransac: workspace %zu < %zu", workspace_bytes, need);
        return FS_ERR_WORKSPACE;
    }
    const rg::Plan pl = rg::make_plan<PPolicy>(n_pairs, max_iters);
    FS_REQUIRE(pl.pool < (size_t)1 << 30, "n_pairs * max_iters too large");
    fs::Carver cv(workspace);
    double* model = cv.take<double>((size_t)n_pairs * 6);
    rg::Workspace<PPolicy> ws;
    ws.carve(cv, n_pairs, pl);
    const size_t smem = PPolicy::stage_bytes(n_max);
    FS_REQUIRE(smem <= 227 * 1024, "n_max too large for the shared-memory staging (limit ~7200)");
    PPolicy::Params prm;
    prm.points3d = points3d;
    prm.pts2d = pts2d;
    prm.n_max = n_max;
    prm.thr2 = (float)(reproj_error * reproj_error);
    prm.cam.fx = cam->fx; prm.cam.fy = cam->fy; prm.cam.cx = cam->cx; prm.cam.cy = cam->cy;
    for (int i = 0; i < 5; i++) prm.cam.k[i] = cam->dist[i];
    const int st_ = rg::run_rounds<PPolicy>(n_pairs, n_corr, n_max, min_points, max_iters, confidence, prm, pl, ws, smem,
                                            stream);
    if (st_ != FS_OK) return st_;
    k_pnp_finish<<<n_pairs, PF_THREADS, 0, stream>>>(ws.st, prm, model, mask, inlier_idx, n_inliers, iters_run, status);
    FS_LAUNCH_CHECK("k_pnp_finish");
    k_pnp_refine<<<n_pairs, RF_THREADS, 0, stream>>>(points3d, pts2d, n_corr, n_max, *cam, model, mask, n_inliers,
                                                    status, rvec, tvec);
    FS_LAUNCH_CHECK("k_pnp_refine");
    return FS_OK;
}
