// cvmath.cuh — device restatements of the small OpenCV numerical building blocks the RANSAC / pose
// (a restatement of OpenCV arithmetic — OpenCV is Apache-2.0 licensed, (C) its contributors; nothing here is taken from /root/reference)
// kernels need: cv::RNG (multiply-with-carry), the one-sided Jacobi SVD of core/lapack.cpp including
// its deterministic completion of missing singular vectors, cv::solvePoly's Durand-Kerner iteration,
// RANSACUpdateNumIters, Rodrigues.  OpenCV is a third-party dependency of the reference that is not
// vendored under /root/reference (Dockerfile:62-67); the algorithms are restated from their published
// form and pinned stage by stage against the cv2 4.13 binary by oracle/opencv_replay.py + tests/.
//
// Files including this header are compiled with -fmad=false: OpenCV's x86-64 baseline build (SSE3)
// cannot contract a*b+c, and the float32 inlier tests downstream are sensitive to the last bit.
#pragma once
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <stdint.h>

#define CVM_HD __host__ __device__ inline

namespace cvm {

// Pixel -> normalised coordinate exactly as cv::findEssentialMat / cv::recoverPose compute it.  Their source line is
// `points.col(0) = (points.col(0) - cx) / fx`, but that is a cv::MatExpr: dividing by a scalar multiplies the expression's
// scale AND its shift by 1/fx, and the assignment runs convertTo(scale, shift), whose double kernel is a fused
// multiply-add: q = fma(x, 1/fx, -cx * (1/fx)).  (Found on an ill-conditioned 5-subset whose inlier count moved from
// 463 to 459 with (x - cx) / fx; checked against cv2 with exact rational arithmetic, tests/test_oracle_cpu.py.)
struct CvNormalise {
    double inv_f, shift_x, shift_y;
    CVM_HD CvNormalise(double focal, double ppx, double ppy) {
        inv_f = 1.0 / focal;
        shift_x = -ppx * inv_f;
        shift_y = -ppy * inv_f;
    }
    CVM_HD double x(float px) const { return fma((double)px, inv_f, shift_x); }
    CVM_HD double y(float py) const { return fma((double)py, inv_f, shift_y); }
};


// ------------------------------------------------------------------------------------------ cv::RNG
struct RNG {
    unsigned long long state;
    __host__ __device__ explicit RNG(unsigned long long s = 0xFFFFFFFFFFFFFFFFull) : state(s ? s : 0xFFFFFFFFull) {}
    __host__ __device__ unsigned next() {
        state = (unsigned long long)(unsigned)state * 4164903690ull + (unsigned)(state >> 32);
        return (unsigned)state;
    }
    __host__ __device__ int uniform(int a, int b) { return a == b ? a : a + (int)(next() % (unsigned)(b - a)); }
};

// RANSACUpdateNumIters (calib3d/ptsetreg.cpp)
CVM_HD int ransac_update_num_iters(double p, double ep, int model_points, int max_iters) {
    p = fmax(p, 0.);
    p = fmin(p, 1.);
    ep = fmax(ep, 0.);
    ep = fmin(ep, 1.);
    double num = fmax(1. - p, DBL_MIN);
    double denom = 1. - pow(1. - ep, (double)model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num);
    denom = log(denom);
    return (denom >= 0 || -num >= max_iters * (-denom)) ? max_iters : (int)rint(num / denom);
}

// core/lapack.cpp defines its own hypot (plain arithmetic, no libm): used by the Jacobi rotations
CVM_HD double hypot_cv(double a, double b) {
    a = fabs(a);
    b = fabs(b);
    if (a > b) {
        b /= a;
        return a * sqrt(1 + b * b);
    }
    if (b > 0) {
        a /= b;
        return b * sqrt(1 + a * a);
    }
    return 0;
}

// ------------------------------------------------------------------------------------------ JacobiSVD
// JacobiSVDImpl_<double> (core/lapack.cpp).  At holds N row vectors of length M (the matrix whose rows
// are orthogonalised in place), Vt the N x N accumulated rotations, W the singular values.  After the
// sweeps the rows are sorted by decreasing norm and rows 0..N1-1 of At are normalised; rows whose norm
// is <= DBL_MIN (and rows N..N1-1, the missing vectors of a FULL_UV request) are filled with the
// deterministic pseudo-random completion.  LDA is the row stride of At (>= M), At must have N1 rows.
// KEEP_V = false (with Vt == nullptr): everything a V-requesting call does to At and W (sorting swaps, completion,
// normalisation), but V itself is neither accumulated nor stored — At and W never depend on it.  EPnP's 12x12
// decomposition (cvSVD(&MtM, &D, &Ut, 0, ...): V is not an output) uses this: a third less work per rotation and
// 1152 bytes less local memory per solver thread.
// PA / PW are pointer-like: double*, or Strided<S> (element i of a thread's matrix at base[i * S]: per-thread matrices
// interleaved in shared memory, bank-conflict free when S = threads per CTA).
template <int S>
struct Strided {
    double* p;
    CVM_HD Strided operator+(int o) const { return Strided{p + (size_t)o * S}; }
    CVM_HD double& operator[](int i) const { return p[(size_t)i * S]; }
};

// WAVEFRONT = true visits the row pairs of a sweep by anti-diagonals d = i + j instead of OpenCV's (i, j) row-major
// order.  Every row still sees ITS rotations in OpenCV's order ((0,r), (1,r), .., (r-1,r), (r,r+1), .., (r,N-1) have
// increasing d) and rotations on disjoint row pairs commute exactly, so the result is bit-identical; but the up to N/2
// pairs of one anti-diagonal are independent, so their dot products (each a serial chain of M dependent additions — the
// latency that bounds the EPnP solver kernel) run as interleaved chains instead of one after the other.  Kept as a
// verified option; NOT used by the product path (measured slower, see epnp.cuh).
template <int M, int N, int N1, bool KEEP_V = true, class PA = double*, class PW = double*, bool WAVEFRONT = false>
CVM_HD void jacobi_svd(PA At, int lda, PW W, double* Vt /* N*N or nullptr */) {
    const bool as_if_v = Vt != nullptr || !KEEP_V;
    const double eps = DBL_EPSILON * 10, minval = DBL_MIN;
    const int max_iter = M > 30 ? M : 30;
    for (int i = 0; i < N; i++) {
        double sd = 0;
        for (int k = 0; k < M; k++) {
            double t = At[i * lda + k];
            sd += t * t;
        }
        W[i] = sd;
        if (KEEP_V && Vt) {
            for (int k = 0; k < N; k++) Vt[i * N + k] = 0;
            Vt[i * N + i] = 1;
        }
    }
    for (int iter = 0; iter < max_iter; iter++) {
        bool changed = false;
        constexpr int NW = WAVEFRONT ? 2 * N - 3 : N - 1;  // anti-diagonals d = 1 .. 2N-3, or rows i = 0 .. N-2
        for (int wv = 0; wv < NW; wv++) {
            constexpr int U = N / 2;  // pairs on one anti-diagonal at most
            double pw[U];
            int first, cnt;  // WAVEFRONT: pairs (first + u, d - first - u), u < cnt; else row i = wv, j = i+1 .. N-1
            if (WAVEFRONT) {
                const int d = wv + 1;
                first = d - (N - 1) > 0 ? d - (N - 1) : 0;
                cnt = (d - 1) / 2 - first + 1;
#pragma unroll
                for (int u = 0; u < U; u++) pw[u] = 0;
                for (int k = 0; k < M; k++) {
#pragma unroll
                    for (int u = 0; u < U; u++)
                        if (u < cnt) pw[u] += (At + (first + u) * lda)[k] * (At + (d - first - u) * lda)[k];
                }
            } else {
                first = wv + 1;
                cnt = N - 1 - wv;
            }
            for (int u = 0; u < cnt; u++) {
                const int i = WAVEFRONT ? first + u : wv;
                const int j = WAVEFRONT ? wv + 1 - i : first + u;
                const PA Ai = At + i * lda;
                const PA Aj = At + j * lda;
                double a = W[i], p = 0, b = W[j];
                if (WAVEFRONT) {
                    p = 0;
#pragma unroll
                    for (int q = 0; q < U; q++)
                        if (q == u) p = pw[q];  // compile-time indices: pw stays in registers
                } else {
                    for (int k = 0; k < M; k++) p += Ai[k] * Aj[k];
                }
                if (fabs(p) <= eps * sqrt(a * b)) continue;
                p *= 2;
                double beta = a - b, gamma = hypot_cv(p, beta), c, s;
                if (beta < 0) {
                    double delta = (gamma - beta) * 0.5;
                    s = sqrt(delta / gamma);
                    c = p / (gamma * s * 2);
                } else {
                    c = sqrt((gamma + beta) / (gamma * 2));
                    s = p / (gamma * c * 2);
                }
                a = b = 0;
                for (int k = 0; k < M; k++) {
                    const double t0 = c * Ai[k] + s * Aj[k];
                    const double t1 = -s * Ai[k] + c * Aj[k];
                    Ai[k] = t0;
                    Aj[k] = t1;
                    a += t0 * t0;
                    b += t1 * t1;
                }
                W[i] = a;
                W[j] = b;
                changed = true;
                if (KEEP_V && Vt) {
                    double* Vi = Vt + i * N;
                    double* Vj = Vt + j * N;
                    for (int k = 0; k < N; k++) {
                        double t0 = c * Vi[k] + s * Vj[k];
                        double t1 = -s * Vi[k] + c * Vj[k];
                        Vi[k] = t0;
                        Vj[k] = t1;
                    }
                }
            }
        }
        if (!changed) break;
    }
    for (int i = 0; i < N; i++) {
        double sd = 0;
        for (int k = 0; k < M; k++) {
            double t = At[i * lda + k];
            sd += t * t;
        }
        W[i] = sqrt(sd);
    }
    for (int i = 0; i < N - 1; i++) {
        int j = i;
        for (int k = i + 1; k < N; k++)
            if (W[j] < W[k]) j = k;
        if (i != j) {
            double t = W[i];
            W[i] = W[j];
            W[j] = t;
            if (as_if_v) {
                for (int k = 0; k < M; k++) {
                    t = At[i * lda + k];
                    At[i * lda + k] = At[j * lda + k];
                    At[j * lda + k] = t;
                }
                if (KEEP_V && Vt)
                    for (int k = 0; k < N; k++) {
                        t = Vt[i * N + k];
                        Vt[i * N + k] = Vt[j * N + k];
                        Vt[j * N + k] = t;
                    }
            }
        }
    }
    if (!as_if_v) return;
    RNG rng(0x12345678ull);
    for (int i = 0; i < N1; i++) {
        double sd = i < N ? W[i] : 0;
        for (int ii = 0; ii < 100 && sd <= minval; ii++) {
            const double val0 = 1. / M;
            for (int k = 0; k < M; k++) At[i * lda + k] = (rng.next() & 256) != 0 ? val0 : -val0;
            for (int iter = 0; iter < 2; iter++)
                for (int j = 0; j < i; j++) {
                    sd = 0;
                    for (int k = 0; k < M; k++) sd += At[i * lda + k] * At[j * lda + k];
                    double asum = 0;
                    for (int k = 0; k < M; k++) {
                        double t = At[i * lda + k] - sd * At[j * lda + k];
                        At[i * lda + k] = t;
                        asum += fabs(t);
                    }
                    asum = asum > eps * 100 ? 1 / asum : 0;
                    for (int k = 0; k < M; k++) At[i * lda + k] *= asum;
                }
            sd = 0;
            for (int k = 0; k < M; k++) {
                double t = At[i * lda + k];
                sd += t * t;
            }
            sd = sqrt(sd);
        }
        const double s = sd > minval ? 1 / sd : 0.;
        for (int k = 0; k < M; k++) At[i * lda + k] *= s;
    }
}

// cv::SVD::compute(A[R x C]) for R >= C (no transposition inside OpenCV): w[C], u[R x C] (columns = left
// singular vectors), vt[C x C].  OpenCV feeds JacobiSVD with the transpose (rows = columns of A).
template <int R, int C>
CVM_HD void svd_tall(const double* A, double* w, double* u /* R*C or null */, double* vt) {
    double At[C * R];
    for (int i = 0; i < R; i++)
        for (int j = 0; j < C; j++) At[j * R + i] = A[i * C + j];
    jacobi_svd<R, C, C>(At, R, w, vt);
    if (u)
        for (int i = 0; i < R; i++)
            for (int j = 0; j < C; j++) u[i * C + j] = At[j * R + i];
}

// ------------------------------------------------------------------------------------------ Rodrigues
CVM_HD void rodrigues_vec2mat(const double* r, double* R) {
    const double rx = r[0], ry = r[1], rz = r[2];
    const double theta = sqrt(rx * rx + ry * ry + rz * rz);
    if (theta < DBL_EPSILON) {
        R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
        return;
    }
    const double c = cos(theta), s = sin(theta), c1 = 1. - c, it = theta ? 1. / theta : 0.;
    const double x = rx * it, y = ry * it, z = rz * it;
    // R = c*I + c1*(r r^T) + s*[r]_x, element by element in that order (cvRodrigues2)
    R[0] = (c + c1 * (x * x));            R[1] = (c1 * (x * y)) + s * (-z);    R[2] = (c1 * (x * z)) + s * y;
    R[3] = (c1 * (x * y)) + s * z;        R[4] = (c + c1 * (y * y));           R[5] = (c1 * (y * z)) + s * (-x);
    R[6] = (c1 * (x * z)) + s * (-y);     R[7] = (c1 * (y * z)) + s * x;       R[8] = (c + c1 * (z * z));
}

}  // namespace cvm
