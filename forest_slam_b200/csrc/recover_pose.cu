// recover_pose.cu — A5: pose from the essential matrix.
// Replaces  _, R, t, _ = cv2.recoverPose(E, mkpts0, mkpts1, focal=K0[0,0], pp=(K0[0,2], K0[1,2]))  mono_slam.py:112
// The reference passes no mask, so ALL correspondences are triangulated for each of the four (R, t)
// combinations: 4 K independent 4x4 Jacobi SVDs per frame pair — on the CPU this is the single largest cost
// of the mono path.
#include "common.cuh"
#include "posemath.cuh"

namespace {

// Three launches, all batch-wide:
//   decompose : one thread per pair, decomposeEssentialMat (3x3 Jacobi SVD) -> R1, R2, t in the workspace;
//   cheirality: one thread per (point, candidate pose) — 4 K independent 4x4 Jacobi SVDs per pair are the whole cost
//               of this stage and they are latency bound, so they are spread over as many warps as the GPU holds;
//               the four mask bits of a point are merged with quad shuffles, the four counts with ballots and one
//               integer atomicAdd per (warp, candidate) (integer sums: order independent);
//   resolve   : arg-max candidate (first maximum wins, as OpenCV's if-chain), final {0,255} mask, R, t.
constexpr int RP_THREADS = 256;
constexpr int RP_PTS_PER_CTA = RP_THREADS / 4;

struct RPPair {
    double R[2][9];
    double t[3];
    int cnt[4];
    int skip;
    int pad;
};

__global__ void __launch_bounds__(128)
k_rp_decompose(const double* __restrict__ E_in, const int32_t* __restrict__ status_in, int n_pairs,
               RPPair* __restrict__ ws) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    RPPair& W = ws[p];
    for (int c = 0; c < 4; c++) W.cnt[c] = 0;
    W.skip = (status_in && status_in[p] != FS_PAIR_OK) ? 1 : 0;
    if (W.skip) return;
    double E[9], R1[9], R2[9], t[3];
    for (int i = 0; i < 9; i++) E[i] = E_in[9 * (size_t)p + i];
    cvm::decompose_essential(E, R1, R2, t);
    for (int i = 0; i < 9; i++) {
        W.R[0][i] = R1[i];
        W.R[1][i] = R2[i];
    }
    for (int i = 0; i < 3; i++) W.t[i] = t[i];
}

__global__ void __launch_bounds__(RP_THREADS)
k_rp_cheirality(const float* __restrict__ mkpts0, const float* __restrict__ mkpts1, const int32_t* __restrict__ n_corr,
                int n_max, double focal, double ppx, double ppy, double dist, RPPair* __restrict__ ws,
                uint8_t* __restrict__ mask_out) {
    const int p = blockIdx.y;
    RPPair& W = ws[p];
    if (W.skip) return;
    const int n = min(n_corr[p], n_max);
    const int i = blockIdx.x * RP_PTS_PER_CTA + (threadIdx.x >> 2);
    const int combo = threadIdx.x & 3, lane = threadIdx.x & 31;
    if (blockIdx.x * RP_PTS_PER_CTA >= n) return;
    bool m = false;
    if (i < n) {
        // P1 = [R1|t], P2 = [R2|t], P3 = [R1|-t], P4 = [R2|-t]
        double R[9], t[3];
#pragma unroll
        for (int k = 0; k < 9; k++) R[k] = W.R[combo & 1][k];
#pragma unroll
        for (int k = 0; k < 3; k++) t[k] = (combo & 2) ? -W.t[k] : W.t[k];
        const float2 pa = reinterpret_cast<const float2*>(mkpts0)[(size_t)p * n_max + i];
        const float2 pb = reinterpret_cast<const float2*>(mkpts1)[(size_t)p * n_max + i];
        const cvm::CvNormalise nrm(focal, ppx, ppy);
        const double x1 = nrm.x(pa.x), y1 = nrm.y(pa.y);
        const double x2 = nrm.x(pb.x), y2 = nrm.y(pb.y);
        m = cvm::cheirality(R, t, x1, y1, x2, y2, dist);
    }
    const unsigned bal = __ballot_sync(0xffffffffu, m);
    if (lane < 4) {
        const int c = __popc(bal & (0x11111111u << lane));
        if (c) atomicAdd(&W.cnt[lane], c);
    }
    if (combo == 0 && i < n) mask_out[(size_t)p * n_max + i] = (uint8_t)((bal >> (lane & ~3)) & 0xFu);  // 4 scratch bits
}

__global__ void __launch_bounds__(256)
k_rp_resolve(const int32_t* __restrict__ n_corr, int n_max, const RPPair* __restrict__ ws, double* __restrict__ R_out,
             double* __restrict__ t_out, uint8_t* __restrict__ mask_out, int32_t* __restrict__ n_good) {
    const int p = blockIdx.y;
    const RPPair& W = ws[p];
    const int n = min(n_corr[p], n_max);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int tid = threadIdx.x;
    if (W.skip) {
        if (i < n_max) mask_out[(size_t)p * n_max + i] = 0;
        if (blockIdx.x == 0) {
            if (tid < 9) R_out[9 * (size_t)p + tid] = (tid % 4 == 0) ? 1.0 : 0.0;
            if (tid < 3) t_out[3 * (size_t)p + tid] = 0.0;
            if (tid == 0) n_good[p] = 0;
        }
        return;
    }
    const int g1 = W.cnt[0], g2 = W.cnt[1], g3 = W.cnt[2], g4 = W.cnt[3];
    int best;
    if (g1 >= g2 && g1 >= g3 && g1 >= g4) best = 0;
    else if (g2 >= g1 && g2 >= g3 && g2 >= g4) best = 1;
    else if (g3 >= g1 && g3 >= g2 && g3 >= g4) best = 2;
    else best = 3;
    if (blockIdx.x == 0) {
        if (tid < 9) R_out[9 * (size_t)p + tid] = W.R[best & 1][tid];
        if (tid < 3) t_out[3 * (size_t)p + tid] = (best & 2) ? -W.t[tid] : W.t[tid];
        if (tid == 0) n_good[p] = W.cnt[best];
    }
    if (i < n_max) {
        uint8_t* mk = mask_out + (size_t)p * n_max + i;
        *mk = (i < n && ((*mk >> best) & 1)) ? 255 : 0;
    }
}

}  // namespace

extern "C" size_t fs_recover_pose_workspace_bytes(int n_pairs) {
    return fs::align_up((size_t)(n_pairs > 0 ? n_pairs : 1) * sizeof(RPPair), 256) + 256;
}

extern "C" int fs_recover_pose(const double* E, const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                               const int32_t* status_in, int n_pairs, int n_max, double focal, double pp_x,
                               double pp_y, double distance_thresh, double* R, double* t, uint8_t* mask,
                               int32_t* n_good, void* workspace, size_t workspace_bytes, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(n_max >= 1, "n_max must be positive");
    FS_REQUIRE(n_pairs <= 65535, "n_pairs too large for one launch (limit 65535)");
    FS_REQUIRE(E && mkpts0 && mkpts1 && n_corr && R && t && mask && n_good, "null pointer");
    FS_REQUIRE(focal != 0, "focal must be non-zero");
    const size_t need = fs_recover_pose_workspace_bytes(n_pairs);
    if (!workspace || workspace_bytes < need) {
        fs::set_error("fs_recover_pose: workspace %zu < %zu", workspace_bytes, need);
        return FS_ERR_WORKSPACE;
    }
    fs::Carver cv(workspace);
    RPPair* ws = cv.take<RPPair>((size_t)n_pairs);
    k_rp_decompose<<<(n_pairs + 127) / 128, 128, 0, stream>>>(E, status_in, n_pairs, ws);
    FS_LAUNCH_CHECK("k_rp_decompose");
    dim3 cgrid((n_max + RP_PTS_PER_CTA - 1) / RP_PTS_PER_CTA, n_pairs);
    k_rp_cheirality<<<cgrid, RP_THREADS, 0, stream>>>(mkpts0, mkpts1, n_corr, n_max, focal, pp_x, pp_y,
                                                      distance_thresh, ws, mask);
    FS_LAUNCH_CHECK("k_rp_cheirality");
    dim3 rgrid((n_max + 255) / 256, n_pairs);
    k_rp_resolve<<<rgrid, 256, 0, stream>>>(n_corr, n_max, ws, R, t, mask, n_good);
    FS_LAUNCH_CHECK("k_rp_resolve");
    return FS_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// fs_recover_pose_static: the same three launches with the per-pair scratch (candidate poses + counters) in a
// file-scope device array instead of a caller workspace, for callers that cannot allocate (CUDA graph capture, serving
// loops).  No counterpart in the reference.
// The array is shared by every call on a device and keeps the last call's contents: k_rp_decompose rewrites every field
// it uses (counters zeroed) for the pairs of the call, so nothing has to be cleared beforehand; calls that overlap on
// the GPU (different streams) race on it.
// ---------------------------------------------------------------------------------------------------------------------
#include <mutex>

namespace {

static_assert(sizeof(RPPair) == 192, "the header documents 192 B of scratch per pair");
static __device__ __align__(256) RPPair g_rp_static_scratch[FS_RECOVER_POSE_STATIC_MAX_PAIRS];

// In host code `g_rp_static_scratch` is only a stand-in symbol: its device address is looked up once per device, with
// that device current, and cached.
int rp_static_scratch(RPPair** out) {
    static std::mutex mu;
    static RPPair* addr[64] = {};
    int dev = 0;
    FS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        fs::set_error("fs_recover_pose_static: device ordinal %d outside the per-device table (64)", dev);
        return FS_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        FS_CUDA(cudaGetSymbolAddress(&p, g_rp_static_scratch));
        if (!p) {
            fs::set_error("fs_recover_pose_static: cudaGetSymbolAddress returned a null address");
            return FS_ERR_CUDA;
        }
        addr[dev] = (RPPair*)p;
    }
    *out = addr[dev];
    return FS_OK;
}

}  // namespace

extern "C" int fs_recover_pose_static_max_pairs(void) { return FS_RECOVER_POSE_STATIC_MAX_PAIRS; }

extern "C" int fs_recover_pose_static(const double* E, const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                                      const int32_t* status_in, int n_pairs, int n_max, double focal, double pp_x,
                                      double pp_y, double distance_thresh, double* R, double* t, uint8_t* mask,
                                      int32_t* n_good, fs_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs > FS_RECOVER_POSE_STATIC_MAX_PAIRS) {
        fs::set_error("fs_recover_pose_static: n_pairs %d exceeds the static scratch (%d pairs); use fs_recover_pose "
                      "with a caller workspace", n_pairs, FS_RECOVER_POSE_STATIC_MAX_PAIRS);
        return FS_ERR_BAD_ARG;
    }
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(n_max >= 1, "n_max must be positive");
    FS_REQUIRE(E && mkpts0 && mkpts1 && n_corr && R && t && mask && n_good, "null pointer");
    FS_REQUIRE(focal != 0, "focal must be non-zero");
    RPPair* ws = nullptr;
    const int s = rp_static_scratch(&ws);
    if (s != FS_OK) return s;
    k_rp_decompose<<<(n_pairs + 127) / 128, 128, 0, stream>>>(E, status_in, n_pairs, ws);
    FS_LAUNCH_CHECK("k_rp_decompose (static scratch)");
    dim3 cgrid((n_max + RP_PTS_PER_CTA - 1) / RP_PTS_PER_CTA, n_pairs);
    k_rp_cheirality<<<cgrid, RP_THREADS, 0, stream>>>(mkpts0, mkpts1, n_corr, n_max, focal, pp_x, pp_y,
                                                      distance_thresh, ws, mask);
    FS_LAUNCH_CHECK("k_rp_cheirality (static scratch)");
    dim3 rgrid((n_max + 255) / 256, n_pairs);
    k_rp_resolve<<<rgrid, 256, 0, stream>>>(n_corr, n_max, ws, R, t, mask, n_good);
    FS_LAUNCH_CHECK("k_rp_resolve (static scratch)");
    return FS_OK;
}
