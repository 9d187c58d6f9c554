// fivepoint.cuh — Nister's five-point essential-matrix solver, as OpenCV's EMEstimatorCallback::runKernel
// (calib3d/five-point.cpp) executes it; one thread per 5-subset.  Replaces the minimal solver inside
//   cv2.findEssentialMat(..., method=cv2.RANSAC)          ros_ws/src/mono_slam.py:111
//
// Stage-by-stage fidelity (pinned by oracle/opencv_replay.py against cv2 4.13):
//   1. Q (5x9) with e = vec(E) row-major, e[3i+j] multiplying x2_i * x1_j;
//   2. null space = rows 5..8 of Vt of SVD::compute(Q, FULL_UV): OpenCV's Jacobi SVD completes the basis
//      with fixed pseudo-random sign vectors, so the basis (hence the hidden variable z and the order in
//      which Durand-Kerner returns the roots) is a deterministic function of the row space of Q;
//   3. 10 cubic constraints (det E = 0, 2 E E^T E - tr(E E^T) E = 0) in Nister's monomial order, reduced by
//      Gauss-Jordan with partial pivoting (OpenCV: A[:, :10].inv() * A[:, 10:]);
//   4. det of the 3x3 polynomial matrix B(z) -> degree-10 polynomial;
//   5. cv::solvePoly's Durand-Kerner (Gauss-Seidel order, start (1+i)^k).  OpenCV always runs 300 sweeps;
//      the iterates reach their rounding-noise floor after ~20 and then only jitter in the last bits, so the
//      sweep loop stops once the largest update stalls (see dk_solve);
//   6. real roots in slot order -> null vector of B(z) (SVD::solveZ) -> E = x E0 + y E1 + z E2 + E3, unit norm.
#pragma once
#include "cvmath.cuh"

namespace cvm {

// quadratic monomial index of (a, b) and cubic monomial index of (quad m, c); variables 0=x 1=y 2=z 3=1.
// Cubic order = the column order of OpenCV's constraint matrix:
// x3 y3 x2y xy2 x2z x2 y2z y2 xyz xy | xz2 xz x yz2 yz y z3 z2 z 1
CVM_HD int t2(int a, int b) {
    constexpr int T[4][4] = {{0, 1, 2, 3}, {1, 4, 5, 6}, {2, 5, 7, 8}, {3, 6, 8, 9}};
    return T[a][b];
}
CVM_HD int t3(int m, int c) {
    constexpr int T[10][4] = {{0, 2, 4, 5},    {2, 3, 8, 9},     {4, 8, 10, 11},   {5, 9, 11, 12},  {3, 1, 6, 7},
                              {8, 6, 13, 14},  {9, 7, 14, 15},   {10, 13, 16, 17}, {11, 14, 17, 18}, {12, 15, 18, 19}};
    return T[m][c];
}

struct Cplx {
    double re, im;
};
CVM_HD Cplx cmul(Cplx a, Cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
CVM_HD Cplx cdiv(Cplx a, Cplx b) {
    const double t = 1. / (b.re * b.re + b.im * b.im);
    return {(a.re * b.re + a.im * b.im) * t, (-a.re * b.im + a.im * b.re) * t};
}

// One Durand-Kerner run for a fixed degree N (cv::solvePoly's loop): Gauss-Seidel updates in slot order, start
// (1+i)^k.  Fully unrolled for N known at compile time so that the roots live in registers and the independent
// numerator (Horner) chains of all slots can overlap the serial denominator chain; the arithmetic of every update is
// performed in exactly OpenCV's order.  OpenCV always runs 300 sweeps; the iterates reach their rounding floor long
// before and then only jitter in the last bits, so the sweep loop stops once the largest update stalls.
template <int N>
CVM_HD void dk_run(const double* coeffs, Cplx* roots, int max_iters) {
    Cplx r[N];
    {
        Cplx p = {1, 0};
        const Cplx q = {1, 1};
#pragma unroll
        for (int i = 0; i < N; i++) {
            r[i] = p;
            p = cmul(p, q);
        }
    }
    double prev = DBL_MAX;
    int stalled = 0;
    for (int iter = 0; iter < max_iters; iter++) {
        double max_diff = 0;
#pragma unroll
        for (int i = 0; i < N; i++) {
            const Cplx p = r[i];
            Cplx num = {coeffs[N], 0}, denom = {coeffs[N], 0};
#pragma unroll
            for (int j = 0; j < N; j++) {
                num = cmul(num, p);
                num.re += coeffs[N - j - 1];
                if (j != i) {
                    const Cplx d = {p.re - r[j].re, p.im - r[j].im};
                    // coincident iterates (OpenCV's multiple-root branch) do not occur for generic data; the factor
                    // is skipped like OpenCV does before its special-case correction
                    const Cplx nd = cmul(denom, d);
                    const bool nz = d.re != 0 || d.im != 0;
                    denom.re = nz ? nd.re : denom.re;
                    denom.im = nz ? nd.im : denom.im;
                }
            }
            num = cdiv(num, denom);
            r[i].re = p.re - num.re;
            r[i].im = p.im - num.im;
            max_diff = fmax(max_diff, sqrt(num.re * num.re + num.im * num.im));
        }
        if (max_diff <= 0) break;
        // stop after two stalled sweeps below 1e-8 (quadratic convergence has hit the rounding floor)
        if (max_diff < 1e-8 && max_diff >= 0.5 * prev) {
            if (++stalled >= 2) break;
        } else {
            stalled = 0;
        }
        prev = max_diff;
    }
#pragma unroll
    for (int i = 0; i < N; i++) {
        if (fabs(r[i].im) < 1e-100) r[i].im = 0;
        roots[i] = r[i];
    }
}

// cv::solvePoly for a real polynomial of degree <= 10, coeffs[k] multiplies z^k.  Returns the degree n
// actually solved (leading coefficients <= DBL_EPSILON are dropped as OpenCV does); roots[0..n-1].
CVM_HD int dk_solve(const double* coeffs, Cplx* roots, int max_iters = 300) {
    int n = 10;
    for (; n > 1; n--)
        if (fabs(coeffs[n]) > DBL_EPSILON) break;
    switch (n) {
        case 10: dk_run<10>(coeffs, roots, max_iters); break;
        case 9: dk_run<9>(coeffs, roots, max_iters); break;
        case 8: dk_run<8>(coeffs, roots, max_iters); break;
        case 7: dk_run<7>(coeffs, roots, max_iters); break;
        case 6: dk_run<6>(coeffs, roots, max_iters); break;
        case 5: dk_run<5>(coeffs, roots, max_iters); break;
        case 4: dk_run<4>(coeffs, roots, max_iters); break;
        case 3: dk_run<3>(coeffs, roots, max_iters); break;
        case 2: dk_run<2>(coeffs, roots, max_iters); break;
        default: dk_run<1>(coeffs, roots, max_iters); break;
    }
    return n;
}

// polynomial helpers, coefficient k multiplies z^k
template <int NA, int NB>
CVM_HD void pmul_acc(const double* a, const double* b, double* out, double sign) {
    for (int i = 0; i < NA; i++)
        for (int j = 0; j < NB; j++) out[i + j] += sign * (a[i] * b[j]);
}

// q1, q2: 5 normalised correspondences (x, y); models: up to 10 matrices, 9 doubles each, row-major.
CVM_HD int five_point(const double (*q1)[2], const double (*q2)[2], double* models) {
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

    // 3. constraints.  Lq[s][m]: quadratic form of (E E^T - 1/2 tr(E E^T) I), symmetric index s of (i, k)
    double A[10][20];
    for (int r = 0; r < 10; r++)
        for (int c = 0; c < 20; c++) A[r][c] = 0;
    {
        double Lq[6][10];
        const int si[6] = {0, 0, 0, 1, 1, 2}, sk[6] = {0, 1, 2, 1, 2, 2};
        for (int s = 0; s < 6; s++) {
            for (int m = 0; m < 10; m++) Lq[s][m] = 0;
            for (int l = 0; l < 3; l++)
                for (int a = 0; a < 4; a++)
                    for (int b = 0; b < 4; b++)
                        Lq[s][t2(a, b)] += EE[a * 9 + 3 * si[s] + l] * EE[b * 9 + 3 * sk[s] + l];
        }
        for (int m = 0; m < 10; m++) {
            const double tr = 0.5 * (Lq[0][m] + Lq[3][m] + Lq[5][m]);
            Lq[0][m] -= tr;
            Lq[3][m] -= tr;
            Lq[5][m] -= tr;
        }
        const int sym[3][3] = {{0, 1, 2}, {1, 3, 4}, {2, 4, 5}};
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) {
                double* row = A[3 * i + j];
                for (int k = 0; k < 3; k++) {
                    const double* L = Lq[sym[i][k]];
                    for (int m = 0; m < 10; m++)
                        for (int c = 0; c < 4; c++) row[t3(m, c)] += L[m] * EE[c * 9 + 3 * k + j];
                }
            }
        // det(E) = E[0,:] . (E[1,:] x E[2,:])
        double cr[3][10];
        for (int d = 0; d < 3; d++) {
            for (int m = 0; m < 10; m++) cr[d][m] = 0;
            const int u = (d + 1) % 3, v = (d + 2) % 3;
            for (int a = 0; a < 4; a++)
                for (int b = 0; b < 4; b++)
                    cr[d][t2(a, b)] += EE[a * 9 + 3 + u] * EE[b * 9 + 6 + v] - EE[a * 9 + 3 + v] * EE[b * 9 + 6 + u];
        }
        for (int d = 0; d < 3; d++)
            for (int m = 0; m < 10; m++)
                for (int c = 0; c < 4; c++) A[9][t3(m, c)] += cr[d][m] * EE[c * 9 + d];
    }

    // Gauss-Jordan with partial pivoting on [A(:, 0:10) | A(:, 10:20)]
    for (int c = 0; c < 10; c++) {
        int piv = c;
        double best = fabs(A[c][c]);
        for (int r = c + 1; r < 10; r++)
            if (fabs(A[r][c]) > best) {
                best = fabs(A[r][c]);
                piv = r;
            }
        if (best < DBL_MIN) return 0;  // singular: OpenCV's inv() returns zeros -> no usable model
        if (piv != c)
            for (int k = c; k < 20; k++) {
                const double t = A[c][k];
                A[c][k] = A[piv][k];
                A[piv][k] = t;
            }
        const double inv = 1. / A[c][c];
        for (int k = c; k < 20; k++) A[c][k] *= inv;
        for (int r = 0; r < 10; r++) {
            if (r == c) continue;
            const double f = A[r][c];
            if (f == 0) continue;
            for (int k = c; k < 20; k++) A[r][k] -= f * A[c][k];
        }
    }

    // 4. B(z): rows (4,5), (6,7), (8,9) -> [cubic | cubic | quartic], highest power first (OpenCV layout b[13])
    double B[3][13];
    for (int i = 0; i < 3; i++) {
        const double* r1 = &A[4 + 2 * i][10];
        const double* r2 = &A[5 + 2 * i][10];
        double* b = B[i];
        b[0] = -r2[0];        b[1] = r1[0] - r2[1]; b[2] = r1[1] - r2[2]; b[3] = r1[2];
        b[4] = -r2[3];        b[5] = r1[3] - r2[4]; b[6] = r1[4] - r2[5]; b[7] = r1[5];
        b[8] = -r2[6];        b[9] = r1[6] - r2[7]; b[10] = r1[7] - r2[8]; b[11] = r1[8] - r2[9]; b[12] = r1[9];
    }
    double coeffs[11];
    {
        // ascending-power copies of the blocks
        double P[3][3][5];
        for (int i = 0; i < 3; i++) {
            for (int k = 0; k < 4; k++) {
                P[i][0][k] = B[i][3 - k];
                P[i][1][k] = B[i][7 - k];
            }
            P[i][0][4] = P[i][1][4] = 0;
            for (int k = 0; k < 5; k++) P[i][2][k] = B[i][12 - k];
        }
        for (int k = 0; k < 11; k++) coeffs[k] = 0;
        // det = sum_j B0j * cofactor_j ; minors of rows 1,2
        double m0[8], m1[8], m2[7];
        for (int k = 0; k < 8; k++) m0[k] = m1[k] = 0;
        for (int k = 0; k < 7; k++) m2[k] = 0;
        pmul_acc<4, 5>(P[1][1], P[2][2], m0, 1.);   // B11*B22
        pmul_acc<5, 4>(P[1][2], P[2][1], m0, -1.);  // - B12*B21
        pmul_acc<4, 5>(P[1][0], P[2][2], m1, 1.);   // B10*B22
        pmul_acc<5, 4>(P[1][2], P[2][0], m1, -1.);  // - B12*B20
        pmul_acc<4, 4>(P[1][0], P[2][1], m2, 1.);   // B10*B21
        pmul_acc<4, 4>(P[1][1], P[2][0], m2, -1.);  // - B11*B20
        pmul_acc<4, 8>(P[0][0], m0, coeffs, 1.);
        pmul_acc<4, 8>(P[0][1], m1, coeffs, -1.);
        pmul_acc<5, 7>(P[0][2], m2, coeffs, 1.);
    }

    // 5. roots
    Cplx roots[10];
    const int n = dk_solve(coeffs, roots);

    // 6. candidates in slot order
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
        double* e = models + 9 * count;
        double nrm = 0;
        for (int k = 0; k < 9; k++) {
            e[k] = EE[k] * x + EE[9 + k] * y + EE[18 + k] * z1 + EE[27 + k];
            nrm += e[k] * e[k];
        }
        nrm = sqrt(nrm);
        for (int k = 0; k < 9; k++) e[k] /= nrm;
        count++;
    }
    return count;
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
