// pnpmath.cuh — host+device restatement of the OpenCV pieces behind
//   cv2.solvePnPRansac(points3D, mkpts1_l, K0, dist_coeffs_l, reprojectionError=1.0, confidence=0.99,
//                      iterationsCount=1000, flags=cv2.SOLVEPNP_ITERATIVE)            stereo_slam.py:294-295
// cv::projectPoints (calib3d/calibration.cpp, cvProjectPoints2Internal) and the float32 reprojection error of
// PnPRansacCallback::computeError (calib3d/solvepnp.cpp).
#pragma once
#include "cvmath.cuh"

namespace cvm {

struct CamD {
    double fx, fy, cx, cy;
    double k[5];  // k1 k2 p1 p2 k3
};

// cvProjectPoints2Internal for one point, R already expanded from rvec; result as the float32 pair OpenCV
// stores into the CV_32F projpoints matrix.
CVM_HD void project_point_f32(const double* R, const double* t, const CamD& cam, double X, double Y, double Z,
                              float& u, float& v) {
    double x = R[0] * X + R[1] * Y + R[2] * Z + t[0];
    double y = R[3] * X + R[4] * Y + R[5] * Z + t[1];
    double z = R[6] * X + R[7] * Y + R[8] * Z + t[2];
    z = z ? 1. / z : 1;
    x *= z;
    y *= z;
    const double r2 = x * x + y * y, r4 = r2 * r2, r6 = r4 * r2;
    const double a1 = 2 * x * y, a2 = r2 + 2 * x * x, a3 = r2 + 2 * y * y;
    const double cdist = 1 + cam.k[0] * r2 + cam.k[1] * r4 + cam.k[4] * r6;
    // OpenCV also forms icdist2 = 1 / (1 + k4 r2 + k5 r4 + k6 r6) and adds s1 r2 + s2 r4 (+ s3, s4): with the reference's
    // five coefficients (k4..k6 = s1..s4 = 0) that is 1 / 1 and + 0 for every finite point, so x * cdist * icdist2 + ... is
    // x * cdist + ... bit for bit; a non-finite point projects to inf / NaN either way and fails the <= test either way.
    // (The literal restatement cost a second fp64 division per evaluation: the scoring kernel's largest item.)
    const double xd0 = x * cdist + cam.k[2] * a1 + cam.k[3] * a2;
    const double yd0 = y * cdist + cam.k[2] * a3 + cam.k[3] * a1;
    u = (float)(xd0 * cam.fx + cam.cx);
    v = (float)(yd0 * cam.fy + cam.cy);
}

// PnPRansacCallback::computeError for one point: float32 difference, float32 squared norm
CVM_HD float reproj_err_f32(float iu, float iv, float pu, float pv) {
#ifdef __CUDA_ARCH__
    const float dx = __fsub_rn(iu, pu), dy = __fsub_rn(iv, pv);
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
#else
    const float dx = iu - pu, dy = iv - pv;
    volatile float a = dx * dx, b = dy * dy;  // no contraction on the host either
    return a + b;
#endif
}

}  // namespace cvm
