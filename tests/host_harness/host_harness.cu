// host_harness.cu — TEST INFRASTRUCTURE.  Compiles the __host__ __device__ numerical headers of the product
// (csrc/cvmath.cuh, fivepoint.cuh, ...) for the HOST so that the CPU test-suite can pin the exact code the
// kernels execute against cv2 without a GPU.  Never loaded by the product path.
#include "../../forest_slam_b200/csrc/fivepoint.cuh"

extern "C" {

int hh_five_point(const double* q1, const double* q2, double* models) {
    return cvm::five_point((const double(*)[2])q1, (const double(*)[2])q2, models);
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
}
