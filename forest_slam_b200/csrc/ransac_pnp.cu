// ransac_pnp.cu — A7: PnP RANSAC (cv2.solvePnPRansac, stereo_slam.py:294-295) and the reprojection scoring sweep.
// Scoring: one warp per hypothesis over all 3D-2D correspondences staged in shared memory (SoA), projection
// with the reference's distortion coefficients in OpenCV's operation order, float32 error <= (float)thr^2.
#include "common.cuh"
#include "pnpmath.cuh"

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
    static size_t bytes(int n_max) { return (size_t)n_max * (3 * sizeof(double) + 2 * sizeof(float)); }
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
    FS_CUDA(cudaFuncSetAttribute(k_score_pnp, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((n_hyp + PS_HYP_PER_CTA - 1) / PS_HYP_PER_CTA, n_pairs);
    k_score_pnp<<<grid, PS_THREADS, smem, (cudaStream_t)stream>>>(points3d, pts2d, n_corr, n_max, *cam, reproj_error,
                                                                models_rt, n_hyp, counts);
    FS_LAUNCH_CHECK("k_score_pnp");
    return FS_OK;
}

extern "C" size_t fs_solve_pnp_workspace_bytes(int, int, int) { return 256; }

extern "C" int fs_solve_pnp_ransac(const float*, const float*, const int32_t*, int, int, const FSCamera*, double,
                                   double, int, int, double*, double*, uint8_t*, int32_t*, int32_t*, int32_t*,
                                   int32_t*, void*, size_t, fs_stream_t) {
    fs::set_error("fs_solve_pnp_ransac: EPnP minimal solver not built yet");
    return FS_ERR_UNSUPPORTED;
}
