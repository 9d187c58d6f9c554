// host_harness.cu — TEST INFRASTRUCTURE.  Compiles the __host__ __device__ numerical headers of the product
// (csrc/cvmath.cuh, fivepoint.cuh, ...) for the HOST so that the CPU test-suite can pin the exact code the
// kernels execute against cv2 without a GPU.  Never loaded by the product path.
#include "../../forest_slam_b200/csrc/fivepoint.cuh"
#include "../../forest_slam_b200/csrc/posemath.cuh"
#include "../../forest_slam_b200/csrc/epnp.cuh"
#include <vector>

extern "C" {

int hh_five_point(const double* q1, const double* q2, double* models) {
    return cvm::five_point((const double(*)[2])q1, (const double(*)[2])q2, models);
}

// stages 1-4 alone: null-space basis EE[36], B(z) [39], polynomial c[11] (pinned against the traced cv2 values)
int hh_five_point_setup(const double* q1, const double* q2, double* EE, double* B, double* c) {
    cvm::FPSetup fs;
    const bool ok = cvm::five_point_setup((const double(*)[2])q1, (const double(*)[2])q2, fs);
    for (int k = 0; k < 36; k++) EE[k] = fs.EE[k];
    for (int k = 0; k < 39; k++) B[k] = (&fs.B[0][0])[k];
    for (int k = 0; k < 11; k++) c[k] = fs.coeffs[k];
    return ok ? 1 : 0;
}

// cv::solvePoly restatement alone: roots (re, im) in slot order; max_iters = 300 reproduces OpenCV's fixed sweep count
int hh_solve_poly(const double* coeffs, double* roots, int max_iters, int early_stop) {
    cvm::Cplx r[10];
    int n;
    if (early_stop) n = cvm::dk_solve(coeffs, r, max_iters);
    else {
        cvm::DKState st;
        cvm::dk_init(coeffs, st);
        // plain 300 sweeps: advance one sweep at a time and ignore the stop rule
        for (int it = 1; it <= max_iters; it++) {
            st.stalled = 0; st.floor_md = DBL_MAX;
            cvm::dk_advance(coeffs, st, it);
        }
        n = cvm::dk_finish(st, r);
    }
    for (int i = 0; i < n; i++) { roots[2 * i] = r[i].re; roots[2 * i + 1] = r[i].im; }
    return n;
}

// exact-cycle statistics of the Durand-Kerner iteration: first sweep t at which the full state equals the state
// p sweeps earlier (p <= 16); returns t (or -1), *period = p
int hh_dk_cycle(const double* coeffs, int* period) {
    cvm::DKState st;
    cvm::dk_init(coeffs, st);
    cvm::Cplx hist[17][10];
    for (int it = 1; it <= 300; it++) {
        st.stalled = 0; st.floor_md = DBL_MAX;
        cvm::dk_advance(coeffs, st, it);
        for (int i = 0; i < 10; i++) hist[it % 17][i] = st.r[i];
        for (int p = 1; p <= 16 && p < it; p++) {
            bool same = true;
            for (int i = 0; i < st.n; i++)
                same &= hist[(it - p) % 17][i].re == st.r[i].re && hist[(it - p) % 17][i].im == st.r[i].im;
            if (same) { *period = p; return it; }
        }
    }
    *period = 0;
    return -1;
}

// every sweep's state: states[(it-1)*20 + 2*i + {0,1}], it = 1..300; returns the degree solved
int hh_dk_states(const double* coeffs, double* states) {
    cvm::DKState st;
    cvm::dk_init(coeffs, st);
    for (int it = 1; it <= 300; it++) {
        st.stalled = 0; st.floor_md = DBL_MAX;
        cvm::dk_advance(coeffs, st, it);
        for (int i = 0; i < 10; i++) { states[(it - 1) * 20 + 2 * i] = st.r[i].re; states[(it - 1) * 20 + 2 * i + 1] = st.r[i].im; }
    }
    return st.n;
}

void hh_ransac_subsets(int count, int n_iters, int* out) {
    cvm::RNG rng(0xFFFFFFFFFFFFFFFFull);
    for (int it = 0; it < n_iters; it++) {
        int* idx = out + 5 * it;
        for (int i = 0; i < 5; i++) {
            int v;
            bool dup;
            do {
                v = rng.uniform(0, count);
                dup = false;
                for (int k = 0; k < i; k++) dup |= idx[k] == v;
            } while (dup);
            idx[i] = v;
        }
    }
}

int hh_update_num_iters(double p, double ep, int model_points, int max_iters) {
    return cvm::ransac_update_num_iters(p, ep, model_points, max_iters);
}

void hh_svd33(const double* A, double* w, double* u, double* vt) { cvm::svd_tall<3, 3>(A, w, u, vt); }
void hh_svd44(const double* A, double* w, double* u, double* vt) { cvm::svd_tall<4, 4>(A, w, u, vt); }

// serial findEssentialMat(RANSAC) built from the SAME header code the kernel k_essential_ransac runs
// (sampler, five_point, sampson_f32, adaptive cap): returns the number of inliers, fills E[9], mask[n], iters.
int hh_find_essential(const float* p0, const float* p1, int n, double focal, double ppx, double ppy, double prob,
                      double threshold, int max_iters, double* E_out, unsigned char* mask, int* iters) {
    if (n < 5) return -1;
    std::vector<double> x1(n), y1(n), x2(n), y2(n);
    for (int i = 0; i < n; i++) {
        const cvm::CvNormalise nrm(focal, ppx, ppy);
        x1[i] = nrm.x(p0[2 * i]);
        y1[i] = nrm.y(p0[2 * i + 1]);
        x2[i] = nrm.x(p1[2 * i]);
        y2[i] = nrm.y(p1[2 * i + 1]);
    }
    double thr = threshold / ((focal + focal) / 2);
    const float thr2 = (float)(thr * thr);
    cvm::RNG rng(0xFFFFFFFFFFFFFFFFull);
    int niters = max_iters > 1 ? max_iters : 1, best = 0, it = 0;
    double bestE[9] = {0};
    for (it = 0; it < niters; it++) {
        int idx[5];
        if (n == 5) {
            for (int i = 0; i < 5; i++) idx[i] = i;
        } else {
            for (int i = 0; i < 5; i++) {
                int v;
                bool dup;
                do {
                    v = rng.uniform(0, n);
                    dup = false;
                    for (int k = 0; k < i; k++) dup |= idx[k] == v;
                } while (dup);
                idx[i] = v;
            }
        }
        double q1[5][2], q2[5][2], models[90];
        for (int i = 0; i < 5; i++) {
            q1[i][0] = x1[idx[i]]; q1[i][1] = y1[idx[i]];
            q2[i][0] = x2[idx[i]]; q2[i][1] = y2[idx[i]];
        }
        const int nm = cvm::five_point(q1, q2, models);
        for (int k = 0; k < nm; k++) {
            int c = 0;
            for (int i = 0; i < n; i++) c += cvm::sampson_f32(models + 9 * k, x1[i], y1[i], x2[i], y2[i]) <= thr2;
            if (c > (best > 4 ? best : 4)) {
                best = c;
                for (int i = 0; i < 9; i++) bestE[i] = models[9 * k + i];
                niters = cvm::ransac_update_num_iters(prob, (double)(n - c) / n, 5, niters);
            }
        }
        if (n == 5) { it = 1; break; }
    }
    *iters = it;
    for (int i = 0; i < 9; i++) E_out[i] = bestE[i];
    for (int i = 0; i < n; i++)
        mask[i] = best ? (n == 5 ? 1 : (cvm::sampson_f32(bestE, x1[i], y1[i], x2[i], y2[i]) <= thr2)) : 0;
    return n == 5 && best ? 5 : best;
}

// recoverPose from the same header code as k_recover_pose
int hh_recover_pose(const double* E, const float* p0, const float* p1, int n, double focal, double ppx, double ppy,
                    double dist, double* R_out, double* t_out, unsigned char* mask) {
    double R1[9], R2[9], tp[3], tm[3];
    cvm::decompose_essential(E, R1, R2, tp);
    for (int i = 0; i < 3; i++) tm[i] = -tp[i];
    int cnt[4] = {0, 0, 0, 0};
    std::vector<unsigned char> bits(n);
    for (int i = 0; i < n; i++) {
        const cvm::CvNormalise nrm(focal, ppx, ppy);
        const double x1 = nrm.x(p0[2 * i]), y1 = nrm.y(p0[2 * i + 1]);
        const double x2 = nrm.x(p1[2 * i]), y2 = nrm.y(p1[2 * i + 1]);
        const bool m0 = cvm::cheirality(R1, tp, x1, y1, x2, y2, dist);
        const bool m1 = cvm::cheirality(R2, tp, x1, y1, x2, y2, dist);
        const bool m2 = cvm::cheirality(R1, tm, x1, y1, x2, y2, dist);
        const bool m3 = cvm::cheirality(R2, tm, x1, y1, x2, y2, dist);
        cnt[0] += m0; cnt[1] += m1; cnt[2] += m2; cnt[3] += m3;
        bits[i] = (unsigned char)(m0 | (m1 << 1) | (m2 << 2) | (m3 << 3));
    }
    int best;
    if (cnt[0] >= cnt[1] && cnt[0] >= cnt[2] && cnt[0] >= cnt[3]) best = 0;
    else if (cnt[1] >= cnt[0] && cnt[1] >= cnt[2] && cnt[1] >= cnt[3]) best = 1;
    else if (cnt[2] >= cnt[0] && cnt[2] >= cnt[1] && cnt[2] >= cnt[3]) best = 2;
    else best = 3;
    for (int i = 0; i < 9; i++) R_out[i] = (best & 1) ? R2[i] : R1[i];
    for (int i = 0; i < 3; i++) t_out[i] = (best & 2) ? tm[i] : tp[i];
    for (int i = 0; i < n; i++) mask[i] = ((bits[i] >> best) & 1) ? 255 : 0;
    return cnt[best];
}

static cvm::CamD hh_cam(const double* cam9) {
    cvm::CamD c;
    c.fx = cam9[0]; c.fy = cam9[1]; c.cx = cam9[2]; c.cy = cam9[3];
    for (int i = 0; i < 5; i++) c.k[i] = cam9[4 + i];
    return c;
}
// cam9 = fx fy cx cy k1 k2 p1 p2 k3
void hh_undistort(const float* uv, int n, const double* cam9, float* out) {
    const cvm::CamD c = hh_cam(cam9);
    for (int i = 0; i < n; i++) cvm::undistort_point_f32(uv[2 * i], uv[2 * i + 1], c, out[2 * i], out[2 * i + 1]);
}
void hh_epnp5(const float* obj, const float* img, const double* cam9, double* model) {
    const cvm::CamD c = hh_cam(cam9);
    double o[5][3];
    float im[5][2];
    for (int i = 0; i < 5; i++) {
        for (int j = 0; j < 3; j++) o[i][j] = (double)obj[3 * i + j];
        im[i][0] = img[2 * i];
        im[i][1] = img[2 * i + 1];
    }
    cvm::pnp_minimal_epnp5(c, o, im, model);
}
void hh_svd_solve64(const double* A, const double* b, double* x) { cvm::svd_solve<6, 4>(A, b, x); }
void hh_svd_invert3(const double* A, double* inv) { cvm::svd_invert3(A, inv); }
void hh_svd12(const double* A, double* w, double* ut) {
    double At[144], vt[144];
    for (int i = 0; i < 12; i++) for (int j = 0; j < 12; j++) At[j * 12 + i] = A[i * 12 + j];
    cvm::jacobi_svd<12, 12, 12>(At, 12, w, vt);
    for (int i = 0; i < 144; i++) ut[i] = At[i];
}
void hh_rodrigues_mat2vec(const double* R, double* r) { cvm::rodrigues_mat2vec(R, r); }

// serial solvePnPRansac (RANSAC stage only: EPnP minimal sets, float32 reprojection test, adaptive cap) from the SAME
// header code k_pnp_ransac runs.  model_out = rvec|tvec of the best minimal-sample model; returns the inlier count.
int hh_pnp_ransac(const float* p3, const float* p2, int n, const double* cam9, double reproj, double conf, int max_iters,
                  double* model_out, unsigned char* mask, int* iters) {
    const cvm::CamD cam = hh_cam(cam9);
    const float thr2 = (float)(reproj * reproj);
    cvm::RNG rng(0xFFFFFFFFFFFFFFFFull);
    int niters = max_iters > 1 ? max_iters : 1, best = 0, it = 0;
    double bm[6] = {0};
    for (it = 0; it < niters; it++) {
        int idx[5];
        for (int i = 0; i < 5; i++) {
            int v;
            bool dup;
            do {
                v = rng.uniform(0, n);
                dup = false;
                for (int k = 0; k < i; k++) dup |= idx[k] == v;
            } while (dup);
            idx[i] = v;
        }
        double obj[5][3], m[6], R[9];
        float img[5][2];
        for (int i = 0; i < 5; i++) {
            for (int j = 0; j < 3; j++) obj[i][j] = (double)p3[3 * idx[i] + j];
            img[i][0] = p2[2 * idx[i]];
            img[i][1] = p2[2 * idx[i] + 1];
        }
        cvm::pnp_minimal_epnp5(cam, obj, img, m);
        cvm::rodrigues_vec2mat(m, R);
        int c = 0;
        for (int i = 0; i < n; i++) {
            float pu, pv;
            cvm::project_point_f32(R, m + 3, cam, (double)p3[3 * i], (double)p3[3 * i + 1], (double)p3[3 * i + 2], pu, pv);
            c += cvm::reproj_err_f32(p2[2 * i], p2[2 * i + 1], pu, pv) <= thr2;
        }
        if (c > (best > 4 ? best : 4)) {
            best = c;
            for (int i = 0; i < 6; i++) bm[i] = m[i];
            niters = cvm::ransac_update_num_iters(conf, (double)(n - c) / n, 5, niters);
        }
    }
    *iters = it;
    double R[9];
    cvm::rodrigues_vec2mat(bm, R);
    for (int i = 0; i < 6; i++) model_out[i] = bm[i];
    for (int i = 0; i < n; i++) {
        float pu, pv;
        cvm::project_point_f32(R, bm + 3, cam, (double)p3[3 * i], (double)p3[3 * i + 1], (double)p3[3 * i + 2], pu, pv);
        mask[i] = best ? (cvm::reproj_err_f32(p2[2 * i], p2[2 * i + 1], pu, pv) <= thr2) : 0;
    }
    return best;
}
}
