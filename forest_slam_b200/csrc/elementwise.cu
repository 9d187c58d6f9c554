// elementwise.cu — A3 (match gather / compaction), A6 (stereo back-projection + depth gate) and
// A8 (pose algebra).  All of these are HBM-bound, coalesced, one-touch kernels.
//
// Reference sites (ros_ws/src/): mono_slam.py:106-108, stereo_slam.py:216-218 (A3);
// stereo_slam.py:117-121, :264-289 (A6); stereo_slam.py:298-311, mono_slam.py:115-118 (A8).
#include "common.cuh"

namespace {

// order-preserving block compaction: returns this thread's exclusive offset among flagged threads,
// and the block total in `total` (valid for all threads).  blockDim.x must be a multiple of 32, <= 1024.
__device__ __forceinline__ int block_compact(bool flag, int& total, int* s_warp /* [33] */) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    const unsigned bal = __ballot_sync(0xffffffffu, flag);
    const int in_warp = __popc(bal & ((1u << lane) - 1));
    __syncthreads();  // protect s_warp reuse across calls
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
        int v = lane < nw ? s_warp[lane] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += n;
        }
        s_warp[lane] = incl - v;  // exclusive
        if (lane == 31) s_warp[32] = incl;
    }
    __syncthreads();
    total = s_warp[32];
    return s_warp[warp] + in_warp;
}

// ---------------------------------------------------------------- A3
__global__ void __launch_bounds__(256)
k_gather_matches(const float2* __restrict__ kpts, int n_max, const int32_t* __restrict__ counts,
                 const int32_t* __restrict__ pair_a, const int32_t* __restrict__ pair_b,
                 const int32_t* __restrict__ matches0, float2* __restrict__ mk0, float2* __restrict__ mk1,
                 int32_t* __restrict__ n_corr, int32_t* __restrict__ src_idx) {
    __shared__ int s_warp[33];
    const int p = blockIdx.x;
    const int fa = pair_a[p], fb = pair_b[p];
    const int n0 = min(counts[fa], n_max), n1 = min(counts[fb], n_max);
    const float2* ka = kpts + (size_t)fa * n_max;
    const float2* kb = kpts + (size_t)fb * n_max;
    const size_t o = (size_t)p * n_max;
    int base = 0;
    for (int q0 = 0; q0 < n0; q0 += blockDim.x) {
        const int q = q0 + threadIdx.x;
        int m = -1;
        if (q < n0) m = matches0[o + q];
        const bool flag = m >= 0 && m < n1;
        int total;
        const int off = block_compact(flag, total, s_warp);
        if (flag) {
            mk0[o + base + off] = ka[q];
            mk1[o + base + off] = kb[m];
            if (src_idx) src_idx[o + base + off] = q;
        }
        base += total;
    }
    if (threadIdx.x == 0) n_corr[p] = base;
}

// ---------------------------------------------------------------- A6
struct BackprojConst {
    float c;           // (float)(fx * baseline): NumPy-1.x value-based casting demotes the f64 scalar
    float fx, fy, cx, cy;
    float zmin, zmax;
};

__device__ __forceinline__ float fix_disparity(float d) {
    // stereo_slam.py:120-121: exactly 0 and exactly -1 become 0.1
    return (d == 0.0f || d == -1.0f) ? 0.1f : d;
}

__device__ __forceinline__ bool backproject_one(float x, float y, float d, const BackprojConst& k, float& X,
                                                float& Y, float& Z) {
    Z = __fdiv_rn(k.c, fix_disparity(d));
    X = __fmul_rn(__fdiv_rn(__fsub_rn(x, k.cx), k.fx), Z);
    Y = __fmul_rn(__fdiv_rn(__fsub_rn(y, k.cy), k.fy), Z);
    return (Z > k.zmin) && (Z < k.zmax);
}

__global__ void __launch_bounds__(256)
k_backproject(const float2* __restrict__ pts, const float* __restrict__ disparity,
              const int32_t* __restrict__ map_index, int64_t n, int map_h, int map_w, BackprojConst k,
              float* __restrict__ out, uint8_t* __restrict__ valid) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float2 xy = pts[i];
        float d;
        if (map_h > 0) {
            // depth[Y.astype(int), X.astype(int)]: C truncation toward zero; NumPy would raise on
            // out-of-range indices, here they are clamped and the point is marked invalid.
            int ix = (int)xy.x, iy = (int)xy.y;
            const bool inside = ix >= 0 && ix < map_w && iy >= 0 && iy < map_h;
            ix = min(max(ix, 0), map_w - 1);
            iy = min(max(iy, 0), map_h - 1);
            const int64_t m = map_index ? map_index[i] : 0;
            d = disparity[(m * map_h + iy) * (int64_t)map_w + ix];
            if (!inside) d = 0.0f;
        } else {
            d = disparity[i];
        }
        float X, Y, Z;
        const bool ok = backproject_one(xy.x, xy.y, d, k, X, Y, Z);
        out[3 * i + 0] = X;
        out[3 * i + 1] = Y;
        out[3 * i + 2] = Z;
        valid[i] = ok ? 1 : 0;
    }
}

__global__ void __launch_bounds__(256)
k_backproject_filter(const float2* __restrict__ mk0, const float2* __restrict__ mk1,
                     const float* __restrict__ disparity, const int32_t* __restrict__ n_corr, int n_max,
                     BackprojConst k, float* __restrict__ p3, float2* __restrict__ p2,
                     int32_t* __restrict__ n_out) {
    __shared__ int s_warp[33];
    const int p = blockIdx.x;
    const int n = min(n_corr[p], n_max);
    const size_t o = (size_t)p * n_max;
    int base = 0;
    for (int i0 = 0; i0 < n; i0 += blockDim.x) {
        const int i = i0 + threadIdx.x;
        float X = 0, Y = 0, Z = 0;
        bool ok = false;
        float2 partner = make_float2(0.f, 0.f);
        if (i < n) {
            const float2 xy = mk0[o + i];
            ok = backproject_one(xy.x, xy.y, disparity[o + i], k, X, Y, Z);
            partner = mk1[o + i];
        }
        int total;
        const int off = block_compact(ok, total, s_warp);
        if (ok) {
            float* dst = p3 + 3 * (o + base + off);
            dst[0] = X;
            dst[1] = Y;
            dst[2] = Z;
            p2[o + base + off] = partner;
        }
        base += total;
    }
    if (threadIdx.x == 0) n_out[p] = base;
}

// ---------------------------------------------------------------- A8
__device__ void rodrigues_dev(const double* r, double* R) {
    // cv::Rodrigues (vector -> matrix): R = c I + (1-c) r r^T + s [r]_x, r normalised
    const double rx = r[0], ry = r[1], rz = r[2];
    const double theta = sqrt(rx * rx + ry * ry + rz * rz);
    if (theta < 2.220446049250313e-16) {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        return;
    }
    const double c = cos(theta), s = sin(theta), c1 = 1.0 - c, it = 1.0 / theta;
    const double x = rx * it, y = ry * it, z = rz * it;
    R[0] = c + c1 * x * x;     R[1] = c1 * x * y - s * z; R[2] = c1 * x * z + s * y;
    R[3] = c1 * x * y + s * z; R[4] = c + c1 * y * y;     R[5] = c1 * y * z - s * x;
    R[6] = c1 * x * z - s * y; R[7] = c1 * y * z + s * x; R[8] = c + c1 * z * z;
}

__global__ void k_rodrigues(const double* __restrict__ rvec, int n, double* __restrict__ R) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double r[3] = {rvec[3 * i], rvec[3 * i + 1], rvec[3 * i + 2]}, Rm[9];
    rodrigues_dev(r, Rm);
    for (int k = 0; k < 9; ++k) R[9 * i + k] = Rm[k];
}

// T_cum[k] = T_init * T_0 * ... * T_k  (mono_slam.py:118, stereo_slam.py:306); a prefix product is
// sequential by definition: 16 lanes of one warp each own one element of the running 4x4.
struct Pose16 {
    double m[16];
};
__global__ void k_compose(const double* __restrict__ R, const double* __restrict__ t,
                          const int32_t* __restrict__ status, int n, const Pose16 T0v, double* T_cum) {
    // the initial pose travels BY VALUE in the kernel arguments: no host buffer outlives the call (graph-capture safe)
    const double* T0 = T0v.m;
    __shared__ double s_T[16];
    __shared__ double s_M[16];
    const int e = threadIdx.x;  // 0..15
    const int i = e >> 2, j = e & 3;
    if (e < 16) s_T[e] = T0[e];
    __syncwarp();
    for (int k = 0; k < n; ++k) {
        const bool skip = status && status[k] != 0;
        if (e < 16) {
            double m;
            if (i < 3) m = j < 3 ? R[9 * k + 3 * i + j] : t[3 * k + i];
            else m = j < 3 ? 0.0 : 1.0;
            s_M[e] = m;
        }
        __syncwarp();
        double acc = 0.0;
        if (e < 16) {
            if (skip) acc = s_T[e];
            else {
#pragma unroll
                for (int q = 0; q < 4; ++q) acc += s_T[4 * i + q] * s_M[4 * q + j];
            }
        }
        __syncwarp();
        if (e < 16) {
            s_T[e] = acc;
            T_cum[16 * (size_t)k + e] = acc;
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(256)
k_transform_points(const float* __restrict__ pts, const int32_t* __restrict__ n_pts, int n_max,
                   const double* __restrict__ T, float* __restrict__ out) {
    const int p = blockIdx.y;
    const int n = min(n_pts[p], n_max);
    double M[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) M[k] = T[16 * (size_t)p + k];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float* s = pts + 3 * ((size_t)p * n_max + i);
        const double x = s[0], y = s[1], z = s[2];
        float* d = out + 3 * ((size_t)p * n_max + i);
        d[0] = (float)(M[0] * x + M[1] * y + M[2] * z + M[3]);
        d[1] = (float)(M[4] * x + M[5] * y + M[6] * z + M[7]);
        d[2] = (float)(M[8] * x + M[9] * y + M[10] * z + M[11]);
    }
}

// A6 over a frame bank: the disparity of correspondence i of pair p is looked up where the reference looks it up —
// per-keypoint (synthetic: disparity[frame][query index]) or in the frame's dense map at (int(y), int(x))
// (stereo_slam.py:274-281) — then back-projected and filtered jointly with the partner points (:288-289).
__global__ void __launch_bounds__(256)
k_backproject_filter_bank(const float2* __restrict__ mk0, const float2* __restrict__ mk1,
                          const int32_t* __restrict__ n_corr, int n_max, const float* __restrict__ disparity,
                          int per_frame, int map_h, int map_w, const int32_t* __restrict__ pair_a,
                          const int32_t* __restrict__ src_idx, BackprojConst k, float* __restrict__ p3,
                          float2* __restrict__ p2, int32_t* __restrict__ n_out) {
    __shared__ int s_warp[33];
    const int p = blockIdx.x;
    const int n = min(n_corr[p], n_max);
    const size_t o = (size_t)p * n_max;
    const float* dmap = disparity + (size_t)pair_a[p] * per_frame;
    int base = 0;
    for (int i0 = 0; i0 < n; i0 += blockDim.x) {
        const int i = i0 + threadIdx.x;
        float X = 0, Y = 0, Z = 0;
        bool ok = false;
        float2 partner = make_float2(0.f, 0.f);
        if (i < n) {
            const float2 xy = mk0[o + i];
            float d;
            if (map_h > 0) {
                int ix = (int)xy.x, iy = (int)xy.y;
                const bool inside = ix >= 0 && ix < map_w && iy >= 0 && iy < map_h;
                ix = min(max(ix, 0), map_w - 1);
                iy = min(max(iy, 0), map_h - 1);
                d = inside ? dmap[(size_t)iy * map_w + ix] : 0.0f;
            } else {
                d = dmap[min(max(src_idx[o + i], 0), n_max - 1)];
            }
            ok = backproject_one(xy.x, xy.y, d, k, X, Y, Z);
            partner = mk1[o + i];
        }
        int total;
        const int off = block_compact(ok, total, s_warp);
        if (ok) {
            float* dst = p3 + 3 * (o + base + off);
            dst[0] = X;
            dst[1] = Y;
            dst[2] = Z;
            p2[o + base + off] = partner;
        }
        base += total;
    }
    if (threadIdx.x == 0) n_out[p] = base;
}

// matches whose descriptor distance exceeds max_distance are dropped (ORB_SLAM3/src/ORBmatcher.cc:704 `bestDist <= TH_LOW`
// style gate; the loop-closure driver counts only such matches) and the pair's match count is recomputed.
__global__ void __launch_bounds__(256)
k_filter_matches_distance(int32_t* __restrict__ matches0, const int32_t* __restrict__ dist0, int n_max, int max_distance,
                          int32_t* __restrict__ n_matches) {
    __shared__ int s_cnt;
    const int p = blockIdx.x;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    int c = 0;
    for (int i = threadIdx.x; i < n_max; i += blockDim.x) {
        const size_t o = (size_t)p * n_max + i;
        int m = matches0[o];
        if (m >= 0 && dist0[o] > max_distance) matches0[o] = m = -1;
        c += m >= 0;
    }
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) n_matches[p] = s_cnt;
}

// the fixed-size per-pair result record of the one gather (sequence.py RECORD_F64 = 20 doubles):
// n_matches, n_inliers, status, n_good, R[9], t[3], iters, pad[3]
__global__ void k_pack_records(const int32_t* __restrict__ n_matches, const int32_t* __restrict__ n_inliers,
                               const int32_t* __restrict__ status, const int32_t* __restrict__ n_good,
                               const double* __restrict__ R, const double* __restrict__ t,
                               const int32_t* __restrict__ iters, int n, double* __restrict__ rec) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 20) return;
    const int p = i / 20, f = i % 20;
    double v = 0.0;
    if (f == 0) v = n_matches[p];
    else if (f == 1) v = n_inliers[p];
    else if (f == 2) v = status[p];
    else if (f == 3) v = n_good ? n_good[p] : 0;
    else if (f < 13) v = R[9 * (size_t)p + f - 4];
    else if (f < 16) v = t[3 * (size_t)p + f - 13];
    else if (f == 16) v = iters ? iters[p] : 0;
    rec[i] = v;
}

BackprojConst make_bp(const FSCamera* cam, double baseline, double zmin, double zmax) {
    BackprojConst k;
    k.c = (float)(cam->fx * baseline);
    k.fx = (float)cam->fx;
    k.fy = (float)cam->fy;
    k.cx = (float)cam->cx;
    k.cy = (float)cam->cy;
    k.zmin = (float)zmin;
    k.zmax = (float)zmax;
    return k;
}

}  // namespace

extern "C" int fs_gather_matches(const float* kpts, int n_frames, int n_max, const int32_t* counts,
                                 const int32_t* pair_a, const int32_t* pair_b, int n_pairs,
                                 const int32_t* matches0, float* mkpts0, float* mkpts1, int32_t* n_corr,
                                 int32_t* src_idx, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0 && n_frames >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(kpts && counts && pair_a && pair_b && matches0 && mkpts0 && mkpts1 && n_corr, "null pointer");
    k_gather_matches<<<n_pairs, 256, 0, (cudaStream_t)stream>>>((const float2*)kpts, n_max, counts, pair_a,
                                                                pair_b, matches0, (float2*)mkpts0,
                                                                (float2*)mkpts1, n_corr, src_idx);
    FS_LAUNCH_CHECK("k_gather_matches");
    return FS_OK;
}

extern "C" int fs_backproject_stereo(const float* pts, const float* disparity, const int32_t* map_index,
                                     int64_t n_points, int map_h, int map_w, const FSCamera* cam,
                                     double baseline, double z_min, double z_max, float* points3d,
                                     uint8_t* valid, fs_stream_t stream) {
    FS_REQUIRE(n_points >= 0 && map_h >= 0 && map_w >= 0, "negative size");
    if (n_points == 0) return FS_OK;
    FS_REQUIRE(pts && disparity && cam && points3d && valid, "null pointer");
    FS_REQUIRE(map_h == 0 || map_w > 0, "map_w must be positive when map_h > 0");
    const int threads = 256;
    int64_t blocks = (n_points + threads - 1) / threads;
    const int64_t cap = (int64_t)fs::sm_count() * 16;  // 16 resident CTAs of 256 threads per SM: grid-stride beyond
    if (blocks > cap) blocks = cap;
    k_backproject<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(
        (const float2*)pts, disparity, map_index, n_points, map_h, map_w, make_bp(cam, baseline, z_min, z_max),
        points3d, valid);
    FS_LAUNCH_CHECK("k_backproject");
    return FS_OK;
}

extern "C" int fs_backproject_filter(const float* mkpts0, const float* mkpts1, const float* disparity,
                                     const int32_t* n_corr, int n_pairs, int n_max, const FSCamera* cam,
                                     double baseline, double z_min, double z_max, float* points3d_out,
                                     float* pts2d_out, int32_t* n_out, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(mkpts0 && mkpts1 && disparity && n_corr && cam && points3d_out && pts2d_out && n_out,
               "null pointer");
    k_backproject_filter<<<n_pairs, 256, 0, (cudaStream_t)stream>>>(
        (const float2*)mkpts0, (const float2*)mkpts1, disparity, n_corr, n_max,
        make_bp(cam, baseline, z_min, z_max), points3d_out, (float2*)pts2d_out, n_out);
    FS_LAUNCH_CHECK("k_backproject_filter");
    return FS_OK;
}

extern "C" int fs_backproject_filter_bank(const float* mkpts0, const float* mkpts1, const int32_t* n_corr,
                                          int n_pairs, int n_max, const float* disparity, int n_frames, int map_h,
                                          int map_w, const int32_t* pair_a, const int32_t* src_idx,
                                          const FSCamera* cam, double baseline, double z_min, double z_max,
                                          float* points3d_out, float* pts2d_out, int32_t* n_out,
                                          fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0 && n_frames >= 0 && map_h >= 0 && map_w >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(mkpts0 && mkpts1 && disparity && n_corr && pair_a && cam && points3d_out && pts2d_out && n_out,
               "null pointer");
    FS_REQUIRE(map_h > 0 ? map_w > 0 : src_idx != nullptr, "dense maps need map_w, per-keypoint disparity needs src_idx");
    const int per_frame = map_h > 0 ? map_h * map_w : n_max;
    k_backproject_filter_bank<<<n_pairs, 256, 0, (cudaStream_t)stream>>>(
        (const float2*)mkpts0, (const float2*)mkpts1, n_corr, n_max, disparity, per_frame, map_h, map_w, pair_a,
        src_idx, make_bp(cam, baseline, z_min, z_max), points3d_out, (float2*)pts2d_out, n_out);
    FS_LAUNCH_CHECK("k_backproject_filter_bank");
    return FS_OK;
}

extern "C" int fs_filter_matches_distance(int32_t* matches0, const int32_t* dist0, int n_pairs, int n_max,
                                          int max_distance, int32_t* n_matches, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(matches0 && dist0 && n_matches, "null pointer");
    k_filter_matches_distance<<<n_pairs, 256, 0, (cudaStream_t)stream>>>(matches0, dist0, n_max, max_distance,
                                                                         n_matches);
    FS_LAUNCH_CHECK("k_filter_matches_distance");
    return FS_OK;
}

extern "C" int fs_pack_records(const int32_t* n_matches, const int32_t* n_inliers, const int32_t* status,
                               const int32_t* n_good, const double* R, const double* t, const int32_t* iters,
                               int n_pairs, double* records, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0, "negative size");
    if (n_pairs == 0) return FS_OK;
    FS_REQUIRE(n_matches && n_inliers && status && R && t && records, "null pointer");
    k_pack_records<<<(n_pairs * 20 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(n_matches, n_inliers, status, n_good,
                                                                                 R, t, iters, n_pairs, records);
    FS_LAUNCH_CHECK("k_pack_records");
    return FS_OK;
}

extern "C" int fs_rodrigues(const double* rvec, int n, double* R, fs_stream_t stream) {
    FS_REQUIRE(n >= 0, "negative size");
    if (n == 0) return FS_OK;
    FS_REQUIRE(rvec && R, "null pointer");
    k_rodrigues<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(rvec, n, R);
    FS_LAUNCH_CHECK("k_rodrigues");
    return FS_OK;
}

extern "C" int fs_compose_trajectory(const double* R, const double* t, const int32_t* status, int n,
                                     const double* T_init_host, double* T_cum, fs_stream_t stream) {
    FS_REQUIRE(n >= 0, "negative size");
    if (n == 0) return FS_OK;
    FS_REQUIRE(R && t && T_cum, "null pointer");
    Pose16 T0 = {{1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}};
    if (T_init_host) memcpy(T0.m, T_init_host, sizeof(T0.m));
    k_compose<<<1, 32, 0, (cudaStream_t)stream>>>(R, t, status, n, T0, T_cum);
    FS_LAUNCH_CHECK("k_compose");
    return FS_OK;
}

extern "C" int fs_transform_points(const float* points3d, const int32_t* n_pts, int n_pairs, int n_max,
                                   const double* T, float* out, fs_stream_t stream) {
    FS_REQUIRE(n_pairs >= 0 && n_max >= 0, "negative size");
    if (n_pairs == 0 || n_max == 0) return FS_OK;
    FS_REQUIRE(points3d && n_pts && T && out, "null pointer");
    dim3 g((n_max + 255) / 256, n_pairs);
    k_transform_points<<<g, 256, 0, (cudaStream_t)stream>>>(points3d, n_pts, n_max, T, out);
    FS_LAUNCH_CHECK("k_transform_points");
    return FS_OK;
}
