// fivepoint.cuh — Nister's five-point essential-matrix solver, as OpenCV's EMEstimatorCallback::runKernel
// (calib3d/five-point.cpp, OpenCV, Apache-2.0 licence) executes it; one thread per 5-subset.  A restatement of that
// routine's arithmetic (not of any file under /root/reference).  Replaces the minimal solver inside
//   cv2.findEssentialMat(..., method=cv2.RANSAC)          ros_ws/src/mono_slam.py:111
//
// Stage-by-stage fidelity (pinned by oracle/opencv_replay.py against cv2 4.13):
//   1. Q (5x9) with e = vec(E) row-major, e[3i+j] multiplying x2_i * x1_j;
//   2. null space = rows 5..8 of Vt of SVD::compute(Q, FULL_UV): OpenCV's Jacobi SVD completes the basis
//      with fixed pseudo-random sign vectors, so the basis (hence the hidden variable z and the order in
//      which Durand-Kerner returns the roots) is a deterministic function of the row space of Q;
//   3. 10 cubic constraints (det E = 0, 2 E E^T E - tr(E E^T) E = 0): OpenCV's getCoeffMat is GENERATED straight-line
//      code; its term and factor order is read out of the cv2 4.13.0 binary with an instruction trace
//      (oracle/cvtrace/) and kept as a table (fivepoint_cv_tables.inc), so every A[k] is bit-identical to cv2's;
//      reduced by cv::solve(LU) (what MatExpr turns A[:, :10].inv() * A[:, 10:] into);
//   4. det of the 3x3 polynomial matrix B(z) -> degree-10 polynomial, same table treatment;
//   5. cv::solvePoly's Durand-Kerner (Gauss-Seidel order, start (1+i)^k), all 300 sweeps (see dk_stop_rule);
//   6. real roots in slot order -> null vector of B(z) (SVD::solveZ) -> E = x E0 + y E1 + z E2 + E3, unit norm.
#pragma once
#include "cvmath.cuh"

namespace cvm {

struct Cplx {
    double re, im;
};
CVM_HD Cplx cmul(Cplx a, Cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
CVM_HD Cplx cdiv(Cplx a, Cplx b) {
    const double t = 1. / (b.re * b.re + b.im * b.im);
    return {(a.re * b.re + a.im * b.im) * t, (-a.re * b.im + a.im * b.re) * t};
}

#ifdef FS_PROFILE_FIVEPOINT  // scratch/ micro-benchmarks only: cycle stamps between the solver's stages
#ifdef __CUDA_ARCH__
#define FP_TICK(k) fp_prof[k] = clock64()
#define FP_SWEEPS(n) atomicAdd(&fp_sweeps[(n) < 511 ? (n) : 511], 1)
#else
#define FP_TICK(k)
#define FP_SWEEPS(n)
#endif
__device__ long long fp_prof[8];
__device__ int fp_sweeps[512];  // histogram of Durand-Kerner sweeps
#else
#define FP_TICK(k)
#define FP_SWEEPS(n)
#endif
// exact "both components are +-0" test on the integer pipe (the fp64 pipe is the scarce resource of this solver)
CVM_HD bool cplx_is_zero(Cplx d) {
#ifdef __CUDA_ARCH__
    return (((unsigned long long)__double_as_longlong(d.re) | (unsigned long long)__double_as_longlong(d.im)) << 1) == 0;
#else
    return d.re == 0 && d.im == 0;
#endif
}

// Durand-Kerner state of one polynomial (a run can be advanced in pieces: tests/host_harness steps it sweep by sweep).
struct DKState {
    Cplx r[10];
    double floor_md;
    int n;        // degree being solved
    int iter;     // sweeps done
    int stalled;
    int done;
};

// Stop rule = OpenCV's: 300 sweeps unless a sweep changes nothing at all (maxDiff <= 0).  The iterates reach their
// rounding floor after ~20-40 sweeps and then JITTER in the last bits, in about a quarter of the polynomials without
// ever entering an exact cycle, and the real roots take part in that jitter: the bits cv2 returns are those of sweep
// 300 and nothing short of running the sweeps reproduces them (round 1 stopped early on a stall heuristic and was
// wrong in the last bit of about one root in ten; measured against cv2.solvePoly).  Returns true when the run is over.
CVM_HD bool dk_stop_rule(double max_diff, int /*iter*/, double& /*floor_md*/, int& /*stalled*/) {
    return max_diff <= 0;
}

// (Measured: rolling the loop over the roots — one copy of the body, root i through select chains — is SLOWER, 11.5k
// against 7.4k cycles per sweep on one thread: fp64 operations have ~18 cycles of latency here, the sweep's dependency
// chain (Gauss-Seidel through the denominators, ~240 levels) needs the cross-root overlap only the unrolled form gives.)
// Sweeps of cv::solvePoly's Durand-Kerner loop for a fixed degree N: Gauss-Seidel updates in slot order, start
// (1+i)^k, every update in exactly OpenCV's operation order.  Fully unrolled for N known at compile time so that the
// roots live in registers.  Runs until the stop rule fires or `until_iter` sweeps have been done in total.
template <int N>
CVM_HD void dk_run(const double* coeffs, DKState& st, int until_iter) {
    Cplx r[N];
#pragma unroll
    for (int i = 0; i < N; i++) r[i] = st.r[i];
    double floor_md = st.floor_md;
    int stalled = st.stalled, iter = st.iter;
    bool done = false;
    for (; iter < until_iter && !done; iter++) {
        // numerators P(r_i): root i still holds its pre-update value when OpenCV evaluates it, so all N Horner chains
        // are independent of this sweep's updates and are advanced in lock-step (same operations per chain)
        Cplx num[N];
#pragma unroll
        for (int i = 0; i < N; i++) num[i] = {coeffs[N], 0};
        // rolled over the Horner steps (the N chains stay unrolled side by side): a third less code in the sweep, whose
        // instruction fetch was a visible stall (E-RANSAC stage 2.74 -> 2.66 ms per 512 pairs, same bits)
#pragma unroll 1
        for (int j = 0; j < N; j++) {
#pragma unroll
            for (int i = 0; i < N; i++) {
                num[i] = cmul(num[i], r[i]);
                num[i].re += coeffs[N - j - 1];
            }
        }
        double max_sq = 0;
#pragma unroll
        for (int i = 0; i < N; i++) {
            const Cplx p = r[i];
            Cplx denom = {coeffs[N], 0};
            bool coincident = false;
#pragma unroll
            for (int j = 0; j < N; j++) {
                if (j != i) {
                    const Cplx d = {p.re - r[j].re, p.im - r[j].im};
                    coincident |= cplx_is_zero(d);
                    denom = cmul(denom, d);
                }
            }
            if (coincident) {
                // coincident iterates (OpenCV's multiple-root branch) do not occur for generic data; the zero factors
                // are skipped like OpenCV does before its special-case correction
                denom = {coeffs[N], 0};
#pragma unroll
                for (int j = 0; j < N; j++) {
                    if (j != i) {
                        const Cplx d = {p.re - r[j].re, p.im - r[j].im};
                        if (!cplx_is_zero(d)) denom = cmul(denom, d);
                    }
                }
            }
            const Cplx q = cdiv(num[i], denom);
            r[i].re = p.re - q.re;
            r[i].im = p.im - q.im;
            max_sq = fmax(max_sq, q.re * q.re + q.im * q.im);
        }
        // max_i sqrt(m_i) == sqrt(max_i m_i): sqrt is correctly rounded, hence monotone
        const double max_diff = sqrt(max_sq);
#ifdef FP_DK_TRACE
        FP_DK_TRACE(iter, max_diff, r);
#endif
        done = dk_stop_rule(max_diff, iter, floor_md, stalled);
    }
#pragma unroll
    for (int i = 0; i < N; i++) st.r[i] = r[i];
    st.floor_md = floor_md;
    st.stalled = stalled;
    st.iter = iter;
    st.done = (done || iter >= 300) ? 1 : 0;
}

// cv::solvePoly for a real polynomial of degree <= 10, coeffs[k] multiplies z^k: start state.  The degree actually
// solved drops leading coefficients <= DBL_EPSILON as OpenCV does.
CVM_HD void dk_init(const double* coeffs, DKState& st) {
    int n = 10;
    for (; n > 1; n--)
        if (fabs(coeffs[n]) > DBL_EPSILON) break;
    st.n = n;
    Cplx p = {1, 0};
    const Cplx q = {1, 1};
    for (int i = 0; i < 10; i++) {
        st.r[i] = p;
        p = cmul(p, q);
    }
    st.floor_md = DBL_MAX;
    st.iter = 0;
    st.stalled = 0;
    st.done = 0;
}

// advance until the stop rule fires or `until_iter` total sweeps (OpenCV's cap is 300)
CVM_HD void dk_advance(const double* coeffs, DKState& st, int until_iter) {
    switch (st.n) {
        case 10: dk_run<10>(coeffs, st, until_iter); break;
        case 9: dk_run<9>(coeffs, st, until_iter); break;
        case 8: dk_run<8>(coeffs, st, until_iter); break;
        case 7: dk_run<7>(coeffs, st, until_iter); break;
        case 6: dk_run<6>(coeffs, st, until_iter); break;
        case 5: dk_run<5>(coeffs, st, until_iter); break;
        case 4: dk_run<4>(coeffs, st, until_iter); break;
        case 3: dk_run<3>(coeffs, st, until_iter); break;
        case 2: dk_run<2>(coeffs, st, until_iter); break;
        default: dk_run<1>(coeffs, st, until_iter); break;
    }
}

// roots in slot order once a run is over (OpenCV zeroes negligible imaginary parts)
CVM_HD int dk_finish(DKState& st, Cplx* roots) {
    FP_SWEEPS(st.iter);
    for (int i = 0; i < st.n; i++) {
        if (fabs(st.r[i].re) < 1e-100) st.r[i].re = 0;
        if (fabs(st.r[i].im) < 1e-100) st.r[i].im = 0;
        roots[i] = st.r[i];
    }
    return st.n;
}

CVM_HD int dk_solve(const double* coeffs, Cplx* roots, int max_iters = 300) {
    DKState st;
    dk_init(coeffs, st);
    dk_advance(coeffs, st, max_iters);
    return dk_finish(st, roots);
}

// Term tables of OpenCV's two GENERATED straight-line blocks (getCoeffMat, det B(z)): one copy for the host build of
// this header (tests/host_harness) and one in constant memory for the kernels (every thread of a warp walks the table
// in lock-step, so the reads are broadcasts).
#define FS_TABLE_END };
#define FS_TABLE_BEGIN(T, name, n) alignas(8) static const T name##_h[n] = {
#include "fivepoint_cv_tables.inc"
#undef FS_TABLE_BEGIN
#ifdef __CUDACC__
#define FS_TABLE_BEGIN(T, name, n) alignas(8) static __constant__ T name##_d[n] = {
#include "fivepoint_cv_tables.inc"
#undef FS_TABLE_BEGIN
#endif
#undef FS_TABLE_END
#ifdef __CUDA_ARCH__
#define FP_TAB(name) name##_d
#else
#define FP_TAB(name) name##_h
#endif

// What the stages after the polynomial need from the stages before it.
struct FPSetup {
    double EE[36];     // null-space basis: EE[a*9 + 3i + j], basis matrix a, entry (i, j)
    double B[3][13];   // B(z) in OpenCV's layout
    double coeffs[11]; // det B(z), coefficient k multiplies z^k
};

// One output of a term table: fold(+/-) of left-associated products, in the order the cv2 binary evaluates them.
// Branch-free: absent factors are the table's 1.0 leaf (x * 1.0 == x exactly), a subtracted term is added negated
// (a - p == a + (-p) exactly), the first term starts the sum.  NF = bytes per term (4: getCoeffMat, 3: det B(z)).
template <int NF>
CVM_HD double fp_eval_terms(const unsigned char* t, int n_terms, const double* L) {
    // products in batches of 8 (independent table reads, leaf reads and multiplies in flight together), then the
    // batch is folded into the sum in table order: the only serial chain is one add per term
    constexpr int B = 8;
    double acc = 0;
    for (int i = 0; i < n_terms; i += B) {
        double p[B];
#pragma unroll
        for (int u = 0; u < B; u++) {
            const int ii = i + u < n_terms ? i + u : n_terms - 1;
            int f0, f1, f2, f3 = 0;
            if (NF == 4) {
                const unsigned int w = reinterpret_cast<const unsigned int*>(t)[ii];  // terms are 4-byte aligned
                f0 = w & 255; f1 = (w >> 8) & 255; f2 = (w >> 16) & 255; f3 = w >> 24;
            } else {
                f0 = t[3 * ii]; f1 = t[3 * ii + 1]; f2 = t[3 * ii + 2];
            }
            double q = (L[f0 & 127] * L[f1]) * L[f2];
            if (NF == 4) q = q * L[f3];
            p[u] = (f0 & 128) ? -q : q;
        }
#pragma unroll
        for (int u = 0; u < B; u++)
            if (i + u < n_terms) acc = (i + u == 0) ? p[u] : acc + p[u];
    }
    return acc;
}

// stages 1-4: q1, q2 = 5 normalised correspondences (x, y).  Returns false when the constraint matrix is singular
// (cv::solve returns false and zero-fills -> no usable model).
CVM_HD bool five_point_setup(const double (*q1)[2], const double (*q2)[2], FPSetup& fs) {
    FP_TICK(0);
    // 1-2. Q and its completed right singular basis
    double At[9 * 9], W[5], Vt[25];
    for (int i = 0; i < 5; i++) {
        const double x1 = q1[i][0], y1 = q1[i][1], x2 = q2[i][0], y2 = q2[i][1];
        double* r = At + 9 * i;
        r[0] = x2 * x1; r[1] = x2 * y1; r[2] = x2;
        r[3] = y2 * x1; r[4] = y2 * y1; r[5] = y2;
        r[6] = x1;      r[7] = y1;      r[8] = 1.0;
    }
    jacobi_svd<9, 5, 9>(At, 9, W, Vt);
    const double* EE = At + 9 * 5;  // EE[a*9 + 3i + j]: basis matrix a, entry (i, j)
    for (int k = 0; k < 36; k++) fs.EE[k] = EE[k];
    FP_TICK(1);

    // 3. getCoeffMat: the 10 cubic constraints (det E = 0, 2 E E^T E - tr(E E^T) E = 0) expanded over
    //    e[], e[]^2, e[]^3 exactly as OpenCV's generated code sums them
    double A[10][20];
    {
        double L[108 + 5];
        for (int k = 0; k < 36; k++) {
            const double e = EE[k], e2 = e * e;
            L[k] = e;
            L[36 + k] = e2;
            L[72 + k] = e2 * e;
        }
        for (int k = 0; k < 5; k++) L[108 + k] = FP_TAB(FP_COEF_CONST)[k];
        // getCoeffMat ends by permuting its columns (left block of the system = computed columns
        // 6 8 18 15 12 5 14 7 4 11, right block = 19 13 1 16 17 3 10 9 2 0)
        for (int r = 0; r < 10; r++)
            for (int c = 0; c < 20; c++) {
                const int k = 20 * r + FP_TAB(FP_COEF_PERM)[c];
                const int s = FP_TAB(FP_COEF_START)[k], e = FP_TAB(FP_COEF_START)[k + 1];
                A[r][c] = fp_eval_terms<4>(FP_TAB(FP_COEF_TERMS) + 4 * s, e - s, L);
            }
    }

    FP_TICK(2);
    // A[:, :10].inv() * A[:, 10:] is folded by cv::MatExpr into cv::solve(A[:, :10], A[:, 10:], DECOMP_LU) =
    // hal::LU64f: elimination with partial pivoting applied to the right-hand side, then back-substitution.
    // Left block: A[r][0..9]; right block (the result): A[r][10..19].
    for (int i = 0; i < 10; i++) {
        int k = i;
        for (int j = i + 1; j < 10; j++)
            if (fabs(A[j][i]) > fabs(A[k][i])) k = j;
        if (fabs(A[k][i]) < DBL_EPSILON * 100) return false;
        if (k != i) {
            for (int j = i; j < 20; j++) {
                const double t = A[i][j];
                A[i][j] = A[k][j];
                A[k][j] = t;
            }
        }
        const double d = -1 / A[i][i];
        for (int j = i + 1; j < 10; j++) {
            const double alpha = A[j][i] * d;
            for (int c = i + 1; c < 10; c++) A[j][c] += alpha * A[i][c];
            for (int c = 10; c < 20; c++) A[j][c] += alpha * A[i][c];
        }
    }
    for (int i = 9; i >= 0; i--)
        for (int j = 10; j < 20; j++) {
            double s = A[i][j];
            for (int k = i + 1; k < 10; k++) s -= A[i][k] * A[k][j];
            A[i][j] = s / A[i][i];
        }

    FP_TICK(3);
    // 4. B(z): rows (4,5), (6,7), (8,9) -> [cubic | cubic | quartic], highest power first (OpenCV layout b[13]);
    //    row1 - row2 with the structural zeros of OpenCV's shifted rows (0 - x = -x and x - 0 = x exactly)
    double (*B)[13] = fs.B;
    for (int i = 0; i < 3; i++) {
        const double* r1 = &A[4 + 2 * i][10];
        const double* r2 = &A[5 + 2 * i][10];
        double* b = B[i];
        b[0] = 0.0 - r2[0];   b[1] = r1[0] - r2[1]; b[2] = r1[1] - r2[2]; b[3] = r1[2];
        b[4] = 0.0 - r2[3];   b[5] = r1[3] - r2[4]; b[6] = r1[4] - r2[5]; b[7] = r1[5];
        b[8] = 0.0 - r2[6];   b[9] = r1[6] - r2[7]; b[10] = r1[7] - r2[8]; b[11] = r1[8] - r2[9]; b[12] = r1[9];
    }
    // det B(z): c[0..10], each a sum of signed triple products of b[] in OpenCV's order
    {
        const double* bf = &B[0][0];
        for (int k = 0; k < 11; k++) {
            const int s = FP_TAB(FP_POLY_START)[k], e = FP_TAB(FP_POLY_START)[k + 1];
            fs.coeffs[k] = fp_eval_terms<3>(FP_TAB(FP_POLY_TERMS) + 3 * s, e - s, bf);
        }
    }
    FP_TICK(4);
    return true;
}

// stage 6: candidates in root-slot order; models: up to 10 matrices, 9 doubles each, row-major
CVM_HD int five_point_models(const FPSetup& fs, const Cplx* roots, int n, double* models) {
    const double* EE = fs.EE;
    const double (*B)[13] = fs.B;
    int count = 0;
    for (int i = 0; i < n; i++) {
        if (fabs(roots[i].im) > 1e-10) continue;
        const double z1 = roots[i].re, z2 = z1 * z1, z3 = z2 * z1, z4 = z3 * z1;
        double bz[9];
        for (int j = 0; j < 3; j++) {
            const double* br = B[j];
            bz[3 * j + 0] = br[0] * z3 + br[1] * z2 + br[2] * z1 + br[3];
            bz[3 * j + 1] = br[4] * z3 + br[5] * z2 + br[6] * z1 + br[7];
            bz[3 * j + 2] = br[8] * z4 + br[9] * z3 + br[10] * z2 + br[11] * z1 + br[12];
        }
        double w[3], vt[9];
        svd_tall<3, 3>(bz, w, nullptr, vt);  // SVD::solveZ: last row of vt
        const double* xy1 = vt + 6;
        if (fabs(xy1[2]) < 1e-10) continue;
        const double x = xy1[0] / xy1[2], y = xy1[1] / xy1[2];
        // Evec = EE.col(0)*x + EE.col(1)*y + EE.col(2)*z + EE.col(3); Evec /= norm(Evec), as the cv2 binary runs it on
        // an FMA-capable x86 host: addWeighted = fma(a, alpha, b*beta) + 0, scaleAdd = fma(c, z, .), add; norm =
        // sequential sum of squares with the 9th element folded by fma (tail of the 4-wide loop); scale by 1/norm.
        double* e = models + 9 * count;
        double acc = 0;
        for (int k = 0; k < 9; k++) {
            double v = fma(EE[k], x, EE[9 + k] * y) + 0.0;
            v = fma(EE[18 + k], z1, v);
            v = v + EE[27 + k];
            e[k] = v;
            acc = k < 8 ? acc + v * v : fma(v, v, acc);
        }
        const double inv = 1.0 / sqrt(acc);
        for (int k = 0; k < 9; k++) e[k] = e[k] * inv;
        count++;
    }
    FP_TICK(6);
    return count;
}

// q1, q2: 5 normalised correspondences (x, y); models: up to 10 matrices, 9 doubles each, row-major.
CVM_HD int five_point(const double (*q1)[2], const double (*q2)[2], double* models) {
    FPSetup fs;
    if (!five_point_setup(q1, q2, fs)) return 0;
    Cplx roots[10];
    const int n = dk_solve(fs.coeffs, roots);
    FP_TICK(5);
    return five_point_models(fs, roots, n, models);
}

// EMEstimatorCallback::computeError for one correspondence (double arithmetic in OpenCV's operation order,
// float32 result) — the residual behind the inlier mask of cv2.findEssentialMat (mono_slam.py:111)
CVM_HD float sampson_f32(const double* E, double x1, double y1, double x2, double y2) {
    const double Ex1_0 = (E[0] * x1 + E[1] * y1) + E[2];
    const double Ex1_1 = (E[3] * x1 + E[4] * y1) + E[5];
    const double Ex1_2 = (E[6] * x1 + E[7] * y1) + E[8];
    const double Etx2_0 = (E[0] * x2 + E[3] * y2) + E[6];
    const double Etx2_1 = (E[1] * x2 + E[4] * y2) + E[7];
    const double x2tEx1 = (x2 * Ex1_0 + y2 * Ex1_1) + Ex1_2;
    const double a = Ex1_0 * Ex1_0, b = Ex1_1 * Ex1_1, c = Etx2_0 * Etx2_0, d = Etx2_1 * Etx2_1;
    return (float)(x2tEx1 * x2tEx1 / (((a + b) + c) + d));
}

}  // namespace cvm
