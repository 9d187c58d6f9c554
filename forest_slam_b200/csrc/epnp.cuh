// epnp.cuh — host+device restatement of the minimal solver inside
// (a restatement of OpenCV arithmetic — OpenCV is Apache-2.0 licensed, (C) its contributors; nothing here is taken from /root/reference)
//   cv2.solvePnPRansac(points3D, mkpts1_l, K0, dist_coeffs_l, reprojectionError=1.0, confidence=0.99,
//                      iterationsCount=1000, flags=cv2.SOLVEPNP_ITERATIVE)              stereo_slam.py:294-295
// OpenCV runs its RANSAC kernel with SOLVEPNP_EPNP on 5-point subsets (calib3d/solvepnp.cpp): the image points are
// undistorted to normalised float32 coordinates (cvUndistortPointsInternal, 5 fixed-point iterations), re-scaled by
// the camera matrix, and handed to EPnP (calib3d/epnp.cpp, Lepetit/Moreno-Noguer/Fua): PCA control points, barycentric
// coordinates, the 2n x 12 system M, eigen-decomposition of M^T M (OpenCV: Jacobi SVD), three beta approximations each
// refined by 5 Gauss-Newton steps (Householder QR), absolute orientation (Horn, SVD of a 3x3), best reprojection
// error wins; then cv::Rodrigues(R) gives the model (rvec, tvec).
//
// With n = 5 the system M^T M has an (at least) two-dimensional exact null space, so the basis OpenCV's Jacobi SVD
// returns for it is decided by rounding noise: the restatement therefore follows OpenCV's arithmetic OPERATION BY
// OPERATION (summation orders, the SVD-based cvInvert / cvSolve back-substitution, Householder QR with its quirks),
// compiled without FMA contraction; the Jacobi rotations use core/lapack.cpp's own plain-arithmetic hypot (cvm::hypot_cv).
// OpenCV is not vendored under /root/reference (Dockerfile:62-67): pinned against the cv2 4.13 binary by
// tests/test_oracle_cpu.py (building blocks against cv2.invert / cv2.solve / cv2.SVDecomp / cv2.undistortPoints, the
// whole solver against cv2.solvePnP(flags=SOLVEPNP_EPNP) on 5 points).
#pragma once
#include "cvmath.cuh"
#include "pnpmath.cuh"
#include "posemath.cuh"

#ifndef EPNP_WAVEFRONT
#define EPNP_WAVEFRONT false
#endif

namespace cvm {

// ---------------------------------------------------------------------------------------- building blocks
// cvUndistortPointsInternal for one CV_32FC2 point, no R / P, TermCriteria(MAX_ITER, 5): float32 normalised result.
CVM_HD void undistort_point_f32(float u, float v, const CamD& cam, float& xo, float& yo) {
    const double ifx = 1. / cam.fx, ify = 1. / cam.fy;
    double x = u, y = v;
    x = (x - cam.cx) * ifx;
    y = (y - cam.cy) * ify;
    const double x0 = x, y0 = y;
    const double k1 = cam.k[0], k2 = cam.k[1], p1 = cam.k[2], p2 = cam.k[3], k3 = cam.k[4];
    for (int j = 0; j < 5; j++) {
        const double r2 = x * x + y * y;
        const double icdist = (1 + ((0. * r2 + 0.) * r2 + 0.) * r2) / (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
        if (icdist < 0) {
            x = ((double)u - cam.cx) * ifx;
            y = ((double)v - cam.cy) * ify;
            break;
        }
        const double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x) + 0. * r2 + 0. * r2 * r2;
        const double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y + 0. * r2 + 0. * r2 * r2;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
    }
    xo = (float)x;
    yo = (float)y;
}

// cv::solve(A[M x N], b[M], x[N], DECOMP_SVD): JacobiSVD on A^T, then SVBkSb with one right-hand side.
template <int M, int N>
CVM_HD void svd_solve(const double* A, const double* b, double* x) {
    double At[N * M], W[N], Vt[N * N];
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++) At[j * M + i] = A[i * N + j];
    jacobi_svd<M, N, N>(At, M, W, Vt);
    double threshold = 0;
    for (int j = 0; j < N; j++) x[j] = 0;
    for (int i = 0; i < N; i++) threshold += W[i];
    threshold *= DBL_EPSILON * 2;
    for (int i = 0; i < N; i++) {
        double wi = W[i];
        if (fabs(wi) <= threshold) continue;
        wi = 1 / wi;
        double s = 0;
        for (int j = 0; j < M; j++) s += At[i * M + j] * b[j];
        s *= wi;
        for (int j = 0; j < N; j++) x[j] = x[j] + s * Vt[i * N + j];
    }
}

// cv::invert(A[3x3], DECOMP_SVD): SVD::compute + SVD::backSubst with an empty right-hand side (pseudo-inverse).
CVM_HD void svd_invert3(const double* A, double* inv) {
    double w[3], u[9], vt[9];
    svd_tall<3, 3>(A, w, u, vt);
    double threshold = 0, buffer[3];
    for (int i = 0; i < 9; i++) inv[i] = 0;
    for (int i = 0; i < 3; i++) threshold += w[i];
    threshold *= DBL_EPSILON * 2;
    for (int i = 0; i < 3; i++) {
        double wi = w[i];
        if (fabs(wi) <= threshold) continue;
        wi = 1 / wi;
        for (int j = 0; j < 3; j++) buffer[j] = u[j * 3 + i] * wi;
        for (int r = 0; r < 3; r++) {
            const double s = vt[i * 3 + r];
            for (int j = 0; j < 3; j++) inv[r * 3 + j] = inv[r * 3 + j] + s * buffer[j];
        }
    }
}

// epnp::qr_solve (Householder QR of A[NR x NC], in place, with the reference's pivot-scan quirk: the scan for
// eta never looks at the last row).  Returns false when OpenCV returns early (singular column): x untouched.
template <int NR, int NC>
CVM_HD bool epnp_qr_solve(double* A, double* b, double* X) {
    double A1[NC], A2[NC];
    for (int k = 0; k < NC; k++) {
        double* ppAkk = A + k * NC + k;
        double* ppAik1 = ppAkk;
        double eta = fabs(*ppAik1);
        for (int i = k + 1; i < NR; i++) {
            const double elt = fabs(*ppAik1);
            if (eta < elt) eta = elt;
            ppAik1 += NC;
        }
        if (eta == 0) return false;
        double* ppAik2 = ppAkk;
        double sum2 = 0.0;
        const double inv_eta = 1. / eta;
        for (int i = k; i < NR; i++) {
            *ppAik2 *= inv_eta;
            sum2 += *ppAik2 * *ppAik2;
            ppAik2 += NC;
        }
        double sigma = sqrt(sum2);
        if (*ppAkk < 0) sigma = -sigma;
        *ppAkk += sigma;
        A1[k] = sigma * *ppAkk;
        A2[k] = -eta * sigma;
        for (int j = k + 1; j < NC; j++) {
            double* ppAik = ppAkk;
            double sum = 0;
            for (int i = k; i < NR; i++) {
                sum += *ppAik * ppAik[j - k];
                ppAik += NC;
            }
            const double tau = sum / A1[k];
            ppAik = ppAkk;
            for (int i = k; i < NR; i++) {
                ppAik[j - k] -= tau * *ppAik;
                ppAik += NC;
            }
        }
    }
    // b <- Qt b
    for (int j = 0; j < NC; j++) {
        double* ppAjj = A + j * NC + j;
        double* ppAij = ppAjj;
        double tau = 0;
        for (int i = j; i < NR; i++) {
            tau += *ppAij * b[i];
            ppAij += NC;
        }
        tau /= A1[j];
        ppAij = ppAjj;
        for (int i = j; i < NR; i++) {
            b[i] -= tau * *ppAij;
            ppAij += NC;
        }
    }
    // X = R^-1 b
    X[NC - 1] = b[NC - 1] / A2[NC - 1];
    for (int i = NC - 2; i >= 0; i--) {
        const double* ppAij = A + i * NC + (i + 1);
        double sum = 0;
        for (int j = i + 1; j < NC; j++) {
            sum += *ppAij * X[j];
            ppAij++;
        }
        X[i] = (b[i] - sum) / A2[i];
    }
    return true;
}

CVM_HD double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
CVM_HD double dist2_3(const double* p1, const double* p2) {
    return (p1[0] - p2[0]) * (p1[0] - p2[0]) + (p1[1] - p2[1]) * (p1[1] - p2[1]) +
           (p1[2] - p2[2]) * (p1[2] - p2[2]);
}

// ---------------------------------------------------------------------------------------- EPnP proper
template <int NP>
struct EPnP {
    double fu, fv, uc, vc;
    double pws[NP * 3], us[NP * 2], alphas[NP * 4], pcs[NP * 3];
    double cws[4][3], ccs[4][3];

    // epnp::epnp + init_points: object points as double, image points = normalised float32 * f + c
    CVM_HD void init(const CamD& cam, const double (*obj)[3], const float (*nrm)[2]) {
        fu = cam.fx; fv = cam.fy; uc = cam.cx; vc = cam.cy;
        for (int i = 0; i < NP; i++) {
            pws[3 * i] = obj[i][0];
            pws[3 * i + 1] = obj[i][1];
            pws[3 * i + 2] = obj[i][2];
            us[2 * i] = nrm[i][0] * fu + uc;
            us[2 * i + 1] = nrm[i][1] * fv + vc;
        }
    }

    CVM_HD void choose_control_points() {
        cws[0][0] = cws[0][1] = cws[0][2] = 0;
        for (int i = 0; i < NP; i++)
            for (int j = 0; j < 3; j++) cws[0][j] += pws[3 * i + j];
        for (int j = 0; j < 3; j++) cws[0][j] /= NP;
        double PW0[NP * 3], pw0tpw0[9], dc[3], u[9], vt[9];
        for (int i = 0; i < NP; i++)
            for (int j = 0; j < 3; j++) PW0[3 * i + j] = pws[3 * i + j] - cws[0][j];
        // mulTransposed(PW0, PW0tPW0, 1): upper triangle, k ascending, then completeSymm
        for (int i = 0; i < 3; i++)
            for (int j = i; j < 3; j++) {
                double s = 0;
                for (int k = 0; k < NP; k++) s += PW0[3 * k + i] * PW0[3 * k + j];
                pw0tpw0[3 * i + j] = s;
                pw0tpw0[3 * j + i] = s;
            }
        svd_tall<3, 3>(pw0tpw0, dc, u, vt);  // UCt = U^T: uct[3*(i-1)+j] = u[j][i-1]
        for (int i = 1; i < 4; i++) {
            const double k = sqrt(dc[i - 1] / NP);
            for (int j = 0; j < 3; j++) cws[i][j] = cws[0][j] + k * u[3 * j + (i - 1)];
        }
    }

    CVM_HD void compute_barycentric_coordinates() {
        double cc[9], ci[9];
        for (int i = 0; i < 3; i++)
            for (int j = 1; j < 4; j++) cc[3 * i + j - 1] = cws[j][i] - cws[0][i];
        svd_invert3(cc, ci);
        for (int i = 0; i < NP; i++) {
            const double* pi = pws + 3 * i;
            double* a = alphas + 4 * i;
            for (int j = 0; j < 3; j++)
                a[1 + j] = ci[3 * j] * (pi[0] - cws[0][0]) + ci[3 * j + 1] * (pi[1] - cws[0][1]) +
                           ci[3 * j + 2] * (pi[2] - cws[0][2]);
            a[0] = 1.0f - a[1] - a[2] - a[3];
        }
    }

    // `ut8` = rows 8..11 of U^T (the four right-most singular vectors): the only rows EPnP reads
    CVM_HD void compute_ccs(const double* betas, const double* ut8) {
        for (int i = 0; i < 4; i++) ccs[i][0] = ccs[i][1] = ccs[i][2] = 0.0f;
        for (int i = 0; i < 4; i++) {
            const double* v = ut8 + 12 * (3 - i);
            for (int j = 0; j < 4; j++)
                for (int k = 0; k < 3; k++) ccs[j][k] += betas[i] * v[3 * j + k];
        }
    }

    CVM_HD void compute_pcs() {
        for (int i = 0; i < NP; i++) {
            const double* a = alphas + 4 * i;
            double* pc = pcs + 3 * i;
            for (int j = 0; j < 3; j++)
                pc[j] = a[0] * ccs[0][j] + a[1] * ccs[1][j] + a[2] * ccs[2][j] + a[3] * ccs[3][j];
        }
    }

    CVM_HD void solve_for_sign() {
        if (pcs[2] < 0.0) {
            for (int i = 0; i < 4; i++)
                for (int j = 0; j < 3; j++) ccs[i][j] = -ccs[i][j];
            for (int i = 0; i < NP; i++) {
                pcs[3 * i] = -pcs[3 * i];
                pcs[3 * i + 1] = -pcs[3 * i + 1];
                pcs[3 * i + 2] = -pcs[3 * i + 2];
            }
        }
    }

    CVM_HD void estimate_R_and_t(double R[3][3], double t[3]) {
        double pc0[3] = {0, 0, 0}, pw0[3] = {0, 0, 0};
        for (int i = 0; i < NP; i++)
            for (int j = 0; j < 3; j++) {
                pc0[j] += pcs[3 * i + j];
                pw0[j] += pws[3 * i + j];
            }
        for (int j = 0; j < 3; j++) {
            pc0[j] /= NP;
            pw0[j] /= NP;
        }
        double abt[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, abt_d[3], abt_u[9], abt_vt[9];
        for (int i = 0; i < NP; i++) {
            const double* pc = pcs + 3 * i;
            const double* pw = pws + 3 * i;
            for (int j = 0; j < 3; j++) {
                abt[3 * j] += (pc[j] - pc0[j]) * (pw[0] - pw0[0]);
                abt[3 * j + 1] += (pc[j] - pc0[j]) * (pw[1] - pw0[1]);
                abt[3 * j + 2] += (pc[j] - pc0[j]) * (pw[2] - pw0[2]);
            }
        }
        svd_tall<3, 3>(abt, abt_d, abt_u, abt_vt);
        // R = U V^T with V = (Vt)^T: R[i][j] = sum_k U[i][k] * V[j][k] = sum_k U[i][k] * Vt[k][j]
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++)
                R[i][j] = abt_u[3 * i] * abt_vt[j] + abt_u[3 * i + 1] * abt_vt[3 + j] + abt_u[3 * i + 2] * abt_vt[6 + j];
        const double det = R[0][0] * R[1][1] * R[2][2] + R[0][1] * R[1][2] * R[2][0] + R[0][2] * R[1][0] * R[2][1] -
                           R[0][2] * R[1][1] * R[2][0] - R[0][1] * R[1][0] * R[2][2] - R[0][0] * R[1][2] * R[2][1];
        if (det < 0) {
            R[2][0] = -R[2][0];
            R[2][1] = -R[2][1];
            R[2][2] = -R[2][2];
        }
        t[0] = pc0[0] - dot3(R[0], pw0);
        t[1] = pc0[1] - dot3(R[1], pw0);
        t[2] = pc0[2] - dot3(R[2], pw0);
    }

    CVM_HD double reprojection_error(const double R[3][3], const double t[3]) {
        double sum2 = 0.0;
        for (int i = 0; i < NP; i++) {
            const double* pw = pws + 3 * i;
            const double Xc = dot3(R[0], pw) + t[0];
            const double Yc = dot3(R[1], pw) + t[1];
            const double inv_Zc = 1.0 / (dot3(R[2], pw) + t[2]);
            const double ue = uc + fu * Xc * inv_Zc;
            const double ve = vc + fv * Yc * inv_Zc;
            const double u = us[2 * i], v = us[2 * i + 1];
            sum2 += sqrt((u - ue) * (u - ue) + (v - ve) * (v - ve));
        }
        return sum2 / NP;
    }

    CVM_HD double compute_R_and_t(const double* ut8, const double* betas, double R[3][3], double t[3]) {
        compute_ccs(betas, ut8);
        compute_pcs();
        solve_for_sign();
        estimate_R_and_t(R, t);
        return reprojection_error(R, t);
    }

    static CVM_HD void compute_L_6x10(const double* ut8, double* l_6x10) {
        const double* v[4] = {ut8 + 12 * 3, ut8 + 12 * 2, ut8 + 12 * 1, ut8};
        double dv[4][6][3];
        for (int i = 0; i < 4; i++) {
            int a = 0, b = 1;
            for (int j = 0; j < 6; j++) {
                dv[i][j][0] = v[i][3 * a] - v[i][3 * b];
                dv[i][j][1] = v[i][3 * a + 1] - v[i][3 * b + 1];
                dv[i][j][2] = v[i][3 * a + 2] - v[i][3 * b + 2];
                b++;
                if (b > 3) {
                    a++;
                    b = a + 1;
                }
            }
        }
        for (int i = 0; i < 6; i++) {
            double* row = l_6x10 + 10 * i;
            row[0] = dot3(dv[0][i], dv[0][i]);
            row[1] = 2.0f * dot3(dv[0][i], dv[1][i]);
            row[2] = dot3(dv[1][i], dv[1][i]);
            row[3] = 2.0f * dot3(dv[0][i], dv[2][i]);
            row[4] = 2.0f * dot3(dv[1][i], dv[2][i]);
            row[5] = dot3(dv[2][i], dv[2][i]);
            row[6] = 2.0f * dot3(dv[0][i], dv[3][i]);
            row[7] = 2.0f * dot3(dv[1][i], dv[3][i]);
            row[8] = 2.0f * dot3(dv[2][i], dv[3][i]);
            row[9] = dot3(dv[3][i], dv[3][i]);
        }
    }

    CVM_HD void compute_rho(double* rho) {
        rho[0] = dist2_3(cws[0], cws[1]);
        rho[1] = dist2_3(cws[0], cws[2]);
        rho[2] = dist2_3(cws[0], cws[3]);
        rho[3] = dist2_3(cws[1], cws[2]);
        rho[4] = dist2_3(cws[1], cws[3]);
        rho[5] = dist2_3(cws[2], cws[3]);
    }

    // betas10 = [B11 B12 B22 B13 B23 B33 B14 B24 B34 B44]
    static CVM_HD void find_betas_approx_1(const double* L, const double* rho, double* betas) {
        double l[6 * 4], b4[4];
        for (int i = 0; i < 6; i++) {
            l[4 * i] = L[10 * i];
            l[4 * i + 1] = L[10 * i + 1];
            l[4 * i + 2] = L[10 * i + 3];
            l[4 * i + 3] = L[10 * i + 6];
        }
        svd_solve<6, 4>(l, rho, b4);
        if (b4[0] < 0) {
            betas[0] = sqrt(-b4[0]);
            betas[1] = -b4[1] / betas[0];
            betas[2] = -b4[2] / betas[0];
            betas[3] = -b4[3] / betas[0];
        } else {
            betas[0] = sqrt(b4[0]);
            betas[1] = b4[1] / betas[0];
            betas[2] = b4[2] / betas[0];
            betas[3] = b4[3] / betas[0];
        }
    }
    static CVM_HD void find_betas_approx_2(const double* L, const double* rho, double* betas) {
        double l[6 * 3], b3[3];
        for (int i = 0; i < 6; i++) {
            l[3 * i] = L[10 * i];
            l[3 * i + 1] = L[10 * i + 1];
            l[3 * i + 2] = L[10 * i + 2];
        }
        svd_solve<6, 3>(l, rho, b3);
        if (b3[0] < 0) {
            betas[0] = sqrt(-b3[0]);
            betas[1] = (b3[2] < 0) ? sqrt(-b3[2]) : 0.0;
        } else {
            betas[0] = sqrt(b3[0]);
            betas[1] = (b3[2] > 0) ? sqrt(b3[2]) : 0.0;
        }
        if (b3[1] < 0) betas[0] = -betas[0];
        betas[2] = 0.0;
        betas[3] = 0.0;
    }
    static CVM_HD void find_betas_approx_3(const double* L, const double* rho, double* betas) {
        double l[6 * 5], b5[5];
        for (int i = 0; i < 6; i++)
            for (int j = 0; j < 5; j++) l[5 * i + j] = L[10 * i + j];
        svd_solve<6, 5>(l, rho, b5);
        if (b5[0] < 0) {
            betas[0] = sqrt(-b5[0]);
            betas[1] = (b5[2] < 0) ? sqrt(-b5[2]) : 0.0;
        } else {
            betas[0] = sqrt(b5[0]);
            betas[1] = (b5[2] > 0) ? sqrt(b5[2]) : 0.0;
        }
        if (b5[1] < 0) betas[0] = -betas[0];
        betas[2] = b5[3] / betas[0];
        betas[3] = 0.0;
    }

    static CVM_HD void gauss_newton(const double* L, const double* rho, double* betas) {
        double a[6 * 4], b[6], x[4] = {0, 0, 0, 0};
        for (int k = 0; k < 5; k++) {
            for (int i = 0; i < 6; i++) {
                const double* rowL = L + i * 10;
                double* rowA = a + i * 4;
                rowA[0] = 2 * rowL[0] * betas[0] + rowL[1] * betas[1] + rowL[3] * betas[2] + rowL[6] * betas[3];
                rowA[1] = rowL[1] * betas[0] + 2 * rowL[2] * betas[1] + rowL[4] * betas[2] + rowL[7] * betas[3];
                rowA[2] = rowL[3] * betas[0] + rowL[4] * betas[1] + 2 * rowL[5] * betas[2] + rowL[8] * betas[3];
                rowA[3] = rowL[6] * betas[0] + rowL[7] * betas[1] + rowL[8] * betas[2] + 2 * rowL[9] * betas[3];
                b[i] = rho[i] - (rowL[0] * betas[0] * betas[0] + rowL[1] * betas[0] * betas[1] +
                                 rowL[2] * betas[1] * betas[1] + rowL[3] * betas[0] * betas[2] +
                                 rowL[4] * betas[1] * betas[2] + rowL[5] * betas[2] * betas[2] +
                                 rowL[6] * betas[0] * betas[3] + rowL[7] * betas[1] * betas[3] +
                                 rowL[8] * betas[2] * betas[3] + rowL[9] * betas[3] * betas[3]);
            }
            epnp_qr_solve<6, 4>(a, b, x);
            for (int i = 0; i < 4; i++) betas[i] += x[i];
        }
    }

    // epnp::compute_pose -> R (row-major 3x3), t.  `ut` (144 elements) and `d` (12) are the workspace of the 12x12
    // decomposition: plain local arrays on the host, per-thread strided shared memory in the RANSAC solver kernel
    // (local memory there spilled the whole batch's matrices through L1/L2 to DRAM: ransac_pnp.cu).
    CVM_HD void compute_pose(double* R_out, double* t_out) {
        double ut[12 * 12], d[12];
        compute_pose_ws<double*, double*>(R_out, t_out, ut, d);
    }
    template <class PA, class PW>
    CVM_HD void compute_pose_ws(double* R_out, double* t_out, PA ut, PW d) {
        double ut8[4 * 12];
        compute_basis_ws<PA, PW>(ut, d, ut8);
        compute_pose_from_basis(ut8, R_out, t_out);
    }

    // First half of epnp::compute_pose: control points, barycentric coordinates, M^T M and its 12x12 decomposition.
    // ut8 = rows 8..11 of U^T, the only part the second half reads.  (The RANSAC solver kernel runs the two halves as two
    // launches: only this one needs the 1.25 KB per-thread matrix workspace that limits it to four warps per SM.)
    template <class PA, class PW>
    CVM_HD void compute_basis_ws(PA ut, PW d, double* ut8) {
        choose_control_points();
        compute_barycentric_coordinates();
        // M (2 NP x 12): rows 2i / 2i+1 = alphas_i[k] * (fu, 0, uc - u_i) / (0, fv, vc - v_i) for the four control points k.
        // Its entries are recomputed where MtM needs them (one multiply each, the same bits every time) instead of being
        // kept as a 960-byte local array: the solver kernel is bound by what it keeps in local memory.
        auto Mkc = [&](int k, int c) -> double {
            const double a = alphas[4 * (k >> 1) + c / 3];
            const int comp = c % 3;
            if ((k & 1) == 0) return comp == 0 ? a * fu : comp == 1 ? 0.0 : a * (uc - us[2 * (k >> 1)]);
            return comp == 0 ? 0.0 : comp == 1 ? a * fv : a * (vc - us[2 * (k >> 1) + 1]);
        };
        // MtM = M^T M (mulTransposed order: upper triangle, k ascending, completeSymm); it is symmetric, so the
        // transpose OpenCV hands to JacobiSVD is MtM itself: rows of At end as U^T
        for (int i = 0; i < 12; i++)
            for (int j = i; j < 12; j++) {
                double s = 0;
                for (int k = 0; k < 2 * NP; k++) s += Mkc(k, i) * Mkc(k, j);
                ut[i * 12 + j] = s;
                ut[j * 12 + i] = s;
            }
        // V is not an output of epnp's cvSVD call
        // (WAVEFRONT = true is bit-identical — 104/104 parity tests — but measured SLOWER on B200, twice: PnP stage 8.0 ms
        // against 6.4 ms per 256 pairs in round 1, and 42.9k against 52.2k stereo pairs/s with the two-launch solver (build
        // with -DEPNP_WAVEFRONT=true): the six predicated chains cost more issue slots and address arithmetic than they hide)
        jacobi_svd<12, 12, 12, /*KEEP_V*/ false, PA, PW, /*WAVEFRONT*/ EPNP_WAVEFRONT>(ut, 12, d, nullptr);
        for (int i = 0; i < 48; i++) ut8[i] = ut[96 + i];
    }

    // Second half: the three beta approximations, Gauss-Newton, absolute orientation, best reprojection error.
    CVM_HD void compute_pose_from_basis(const double* ut8, double* R_out, double* t_out) {
        double l_6x10[60], rho[6];
        compute_L_6x10(ut8, l_6x10);
        compute_rho(rho);
        // the three approximations share ONE copy of gauss_newton / compute_R_and_t (rolled loop: the second half of the
        // solver stalled on instruction fetch as much as on its dependencies); the best is kept with OpenCV's strict
        // comparisons: N = 1; if (err2 < err1) N = 2; if (err3 < err_N) N = 3
        double best_err = 0;
#pragma unroll 1
        for (int a = 1; a <= 3; a++) {
            double betas[4], R[3][3], t[3];
            if (a == 1) find_betas_approx_1(l_6x10, rho, betas);
            else if (a == 2) find_betas_approx_2(l_6x10, rho, betas);
            else find_betas_approx_3(l_6x10, rho, betas);
            gauss_newton(l_6x10, rho, betas);
            const double err = compute_R_and_t(ut8, betas, R, t);
            if (a == 1 || err < best_err) {
                best_err = err;
                for (int i = 0; i < 3; i++) {
                    t_out[i] = t[i];
                    for (int j = 0; j < 3; j++) R_out[3 * i + j] = R[i][j];
                }
            }
        }
    }

    // What the second half needs of the first, as a flat record (RANSAC solver, two launches): 15 + 10 + 20 + 12 doubles of
    // the problem (STATE_HEAD), then ut8 (48).
    static constexpr int STATE_HEAD = NP * 3 + NP * 2 + NP * 4 + 12;
    static constexpr int STATE_DOUBLES = STATE_HEAD + 48;
    CVM_HD void save_state_head(double* st) const {
        int o = 0;
        for (int i = 0; i < NP * 3; i++) st[o++] = pws[i];
        for (int i = 0; i < NP * 2; i++) st[o++] = us[i];
        for (int i = 0; i < NP * 4; i++) st[o++] = alphas[i];
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 3; j++) st[o++] = cws[i][j];
    }
    CVM_HD void load_state_head(const CamD& cam, const double* st) {
        fu = cam.fx; fv = cam.fy; uc = cam.cx; vc = cam.cy;
        int o = 0;
        for (int i = 0; i < NP * 3; i++) pws[i] = st[o++];
        for (int i = 0; i < NP * 2; i++) us[i] = st[o++];
        for (int i = 0; i < NP * 4; i++) alphas[i] = st[o++];
        for (int i = 0; i < 4; i++)
            for (int j = 0; j < 3; j++) cws[i][j] = st[o++];
    }
};

// cv::Rodrigues, matrix -> vector (calib3d/calibration.cpp cvRodrigues2): R is re-orthogonalised by an SVD first
CVM_HD void rodrigues_mat2vec(const double* Rin, double* r) {
    double w[3], U[9], Vt[9], R[9];
    svd_tall<3, 3>(Rin, w, U, Vt);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) R[3 * i + j] = U[3 * i] * Vt[j] + U[3 * i + 1] * Vt[3 + j] + U[3 * i + 2] * Vt[6 + j];
    double rx = R[7] - R[5], ry = R[2] - R[6], rz = R[3] - R[1];
    const double s = sqrt((rx * rx + ry * ry + rz * rz) * 0.25);
    double c = (R[0] + R[4] + R[8] - 1) * 0.5;
    c = c > 1. ? 1. : c < -1. ? -1. : c;
    double theta = acos(c);
    if (s < 1e-5) {
        if (c > 0) {
            rx = ry = rz = 0;
        } else {
            double t;
            t = (R[0] + 1) * 0.5;
            rx = sqrt(fmax(t, 0.));
            t = (R[4] + 1) * 0.5;
            ry = sqrt(fmax(t, 0.)) * (R[1] < 0 ? -1. : 1.);
            t = (R[8] + 1) * 0.5;
            rz = sqrt(fmax(t, 0.)) * (R[2] < 0 ? -1. : 1.);
            if (fabs(rx) < fabs(ry) && fabs(rx) < fabs(rz) && (R[5] > 0) != (ry * rz > 0)) rz = -rz;
            theta /= sqrt(rx * rx + ry * ry + rz * rz);
            rx *= theta;
            ry *= theta;
            rz *= theta;
        }
    } else {
        double vth = 1 / (2 * s);
        vth *= theta;
        rx *= vth;
        ry *= vth;
        rz *= vth;
    }
    r[0] = rx;
    r[1] = ry;
    r[2] = rz;
}

// PnPRansacCallback::runKernel for one 5-subset: undistort -> EPnP -> Rodrigues.  model = rvec | tvec.
// `ut` / `d`: the 144 + 12 element workspace of the 12x12 decomposition (see EPnP::compute_pose_ws)
template <class PA, class PW>
CVM_HD void pnp_minimal_epnp5_ws(const CamD& cam, const double (*obj)[3], const float (*img)[2], double* model, PA ut,
                                 PW d) {
    float nrm[5][2];
    for (int i = 0; i < 5; i++) undistort_point_f32(img[i][0], img[i][1], cam, nrm[i][0], nrm[i][1]);
    EPnP<5> e;
    e.init(cam, obj, nrm);
    double R[9], t[3];
    e.template compute_pose_ws<PA, PW>(R, t, ut, d);
    rodrigues_mat2vec(R, model);
    model[3] = t[0];
    model[4] = t[1];
    model[5] = t[2];
}

// the same in two halves (ransac_pnp.cu): `state` = EPnP<5>::STATE_DOUBLES doubles per solve, written / read in place
// (global memory in the solver kernels: no local copy)
template <class PA, class PW>
CVM_HD void pnp_minimal_epnp5_basis(const CamD& cam, const double (*obj)[3], const float (*img)[2], double* state, PA ut,
                                    PW d) {
    float nrm[5][2];
    for (int i = 0; i < 5; i++) undistort_point_f32(img[i][0], img[i][1], cam, nrm[i][0], nrm[i][1]);
    EPnP<5> e;
    e.init(cam, obj, nrm);
    e.template compute_basis_ws<PA, PW>(ut, d, state + EPnP<5>::STATE_HEAD);
    e.save_state_head(state);
}
CVM_HD void pnp_minimal_epnp5_finish(const CamD& cam, const double* state, double* model) {
    EPnP<5> e;
    double R[9], t[3];
    e.load_state_head(cam, state);
    e.compute_pose_from_basis(state + EPnP<5>::STATE_HEAD, R, t);
    rodrigues_mat2vec(R, model);
    model[3] = t[0];
    model[4] = t[1];
    model[5] = t[2];
}

CVM_HD void pnp_minimal_epnp5(const CamD& cam, const double (*obj)[3], const float (*img)[2], double* model) {
    float nrm[5][2];
    for (int i = 0; i < 5; i++) undistort_point_f32(img[i][0], img[i][1], cam, nrm[i][0], nrm[i][1]);
    EPnP<5> e;
    e.init(cam, obj, nrm);
    double R[9], t[3];
    e.compute_pose(R, t);
    rodrigues_mat2vec(R, model);
    model[3] = t[0];
    model[4] = t[1];
    model[5] = t[2];
}

}  // namespace cvm
