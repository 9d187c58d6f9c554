// posemath.cuh — host+device restatement of cv::decomposeEssentialMat and of the per-point work of
// cv::recoverPose (calib3d/five-point.cpp) + cv::triangulatePoints (calib3d/triangulate.cpp):
// a 4x4 DLT system per point and view pair, solved by OpenCV's Jacobi SVD (last row of Vt), followed by
// the cheirality / distance gates.  Replaces cv2.recoverPose(E, p0, p1, focal=, pp=)  mono_slam.py:112.
#pragma once
#include "cvmath.cuh"

namespace cvm {

CVM_HD double det3(const double* m) {
    return m[0] * (m[4] * m[8] - m[5] * m[7]) - m[1] * (m[3] * m[8] - m[5] * m[6]) +
           m[2] * (m[3] * m[7] - m[4] * m[6]);
}

CVM_HD void matmul33(const double* a, const double* b, double* c) {
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += a[3 * i + k] * b[3 * k + j];
            c[3 * i + j] = s;
        }
}

// decomposeEssentialMat: SVD(E) = U D Vt, det fixes, R1 = U W Vt, R2 = U W^T Vt, t = U[:, 2]
CVM_HD void decompose_essential(const double* E, double* R1, double* R2, double* t) {
    double w[3], U[9], Vt[9];
    svd_tall<3, 3>(E, w, U, Vt);
    if (det3(U) < 0)
        for (int i = 0; i < 9; i++) U[i] *= -1.;
    if (det3(Vt) < 0)
        for (int i = 0; i < 9; i++) Vt[i] *= -1.;
    const double W[9] = {0, 1, 0, -1, 0, 0, 0, 0, 1};
    const double Wt[9] = {0, -1, 0, 1, 0, 0, 0, 0, 1};
    double UW[9];
    matmul33(U, W, UW);
    matmul33(UW, Vt, R1);
    matmul33(U, Wt, UW);
    matmul33(UW, Vt, R2);
    t[0] = U[2];
    t[1] = U[5];
    t[2] = U[8];
}

// Last row of Vt of cv::SVD::compute(A) for a 4x4 A — the same arithmetic as jacobi_svd<4, 4, 4> (cvmath.cuh), operation
// for operation, but with every index a compile-time constant so that the 16 + 16 + 4 doubles live in registers.  The
// generic routine's dynamic row pointers put them in local memory: ncu showed 382 MB of DRAM writes per launch of
// k_rp_cheirality and lg_throttle / long_scoreboard as its top stalls.  Only what triangulatePoints reads is produced
// (the row of Vt that the descending sort puts last); the completion of degenerate left vectors never touches Vt.
CVM_HD void svd4_last_vt_row(const double* A, double* out) {
    const double eps = DBL_EPSILON * 10;
    double a[4][4], v[4][4], w[4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int k = 0; k < 4; k++) {
            a[i][k] = A[k * 4 + i];  // OpenCV feeds JacobiSVD with the transpose
            v[i][k] = (i == k) ? 1. : 0.;
        }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) sd += a[i][k] * a[i][k];
        w[i] = sd;
    }
    for (int iter = 0; iter < 30; iter++) {
        bool changed = false;
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = i + 1; j < 4; j++) {
                double aa = w[i], p = 0, bb = w[j];
#pragma unroll
                for (int k = 0; k < 4; k++) p += a[i][k] * a[j][k];
                if (!(fabs(p) <= eps * sqrt(aa * bb))) {
                    p *= 2;
                    const double beta = aa - bb, gamma = hypot_cv(p, beta);
                    double c, sn;
                    if (beta < 0) {
                        const double delta = (gamma - beta) * 0.5;
                        sn = sqrt(delta / gamma);
                        c = p / (gamma * sn * 2);
                    } else {
                        c = sqrt((gamma + beta) / (gamma * 2));
                        sn = p / (gamma * c * 2);
                    }
                    aa = bb = 0;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const double t0 = c * a[i][k] + sn * a[j][k];
                        const double t1 = -sn * a[i][k] + c * a[j][k];
                        a[i][k] = t0;
                        a[j][k] = t1;
                        aa += t0 * t0;
                        bb += t1 * t1;
                    }
                    w[i] = aa;
                    w[j] = bb;
                    changed = true;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const double t0 = c * v[i][k] + sn * v[j][k];
                        const double t1 = -sn * v[i][k] + c * v[j][k];
                        v[i][k] = t0;
                        v[j][k] = t1;
                    }
                }
            }
        if (!changed) break;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        double sd = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) sd += a[i][k] * a[i][k];
        w[i] = sqrt(sd);
    }
    // OpenCV's selection sort (descending, strict <): only w and the rows of Vt are needed here
#pragma unroll
    for (int i = 0; i < 3; i++) {
        int j = i;
        double wj = w[i];
#pragma unroll
        for (int k = i + 1; k < 4; k++)
            if (wj < w[k]) {
                wj = w[k];
                j = k;
            }
#pragma unroll
        for (int jj = i + 1; jj < 4; jj++)
            if (j == jj) {
                const double t = w[i];
                w[i] = w[jj];
                w[jj] = t;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const double u = v[i][k];
                    v[i][k] = v[jj][k];
                    v[jj][k] = u;
                }
            }
    }
#pragma unroll
    for (int k = 0; k < 4; k++) out[k] = v[3][k];
}

// One point, one candidate second camera P = [R | t] (first camera [I | 0]); (x1, y1), (x2, y2) normalised.
// Returns the recoverPose mask bit: z*w > 0, z/w < dist, z' > 0, z' < dist.
CVM_HD bool cheirality(const double* R, const double* t, double x1, double y1, double x2, double y2,
                       double dist) {
    double A[16];
    // view 0: x * P0[2,:] - P0[0,:], y * P0[2,:] - P0[1,:] with P0 = [I | 0]
    A[0] = x1 * 0. - 1.; A[1] = x1 * 0. - 0.; A[2] = x1 * 1. - 0.; A[3] = x1 * 0. - 0.;
    A[4] = y1 * 0. - 0.; A[5] = y1 * 0. - 1.; A[6] = y1 * 1. - 0.; A[7] = y1 * 0. - 0.;
    for (int k = 0; k < 4; k++) {
        const double p0 = k < 3 ? R[k] : t[0], p1 = k < 3 ? R[3 + k] : t[1], p2 = k < 3 ? R[6 + k] : t[2];
        A[8 + k] = x2 * p2 - p0;
        A[12 + k] = y2 * p2 - p1;
    }
    double v3[4];
    svd4_last_vt_row(A, v3);
    double X = v3[0], Y = v3[1], Z = v3[2];
    const double Wh = v3[3];
    bool ok = Z * Wh > 0;
    X /= Wh;
    Y /= Wh;
    Z /= Wh;
    const double one = Wh / Wh;
    ok = ok && (Z < dist);
    const double z2 = ((R[6] * X + R[7] * Y) + R[8] * Z) + t[2] * one;
    ok = ok && (z2 > 0) && (z2 < dist);
    return ok;
}

}  // namespace cvm
