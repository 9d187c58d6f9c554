// host-side trace of slow Durand-Kerner cases (scratch)
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
static int g_trace = 0; static int g_last_iter = 0; static double g_md[400];
#define FP_DK_TRACE(iter, md, r) do { g_last_iter = iter; if (iter < 400) g_md[iter] = md; } while (0)
#include "../forest_slam_b200/csrc/fivepoint.cuh"
int main() {
    srand(7);
    auto rnd = []() { return rand() / (double)RAND_MAX; };
    auto gauss = [&]() { return sqrt(-2 * log(rnd() + 1e-12)) * cos(6.283185307 * rnd()); };
    const int K = 1665, NS = 4096;
    std::vector<double> pts(K * 4);
    const double f = 642.9;
    for (int i = 0; i < K; i++) {
        double X = rnd() * 30 - 15, Y = rnd() * 6 - 3, Z = 2 + rnd() * 40;
        double th = 0.01, tx = 0.01, ty = 0.002, tz = 0.13;
        double X2 = cos(th) * X + sin(th) * Z + tx, Y2 = Y + ty, Z2 = -sin(th) * X + cos(th) * Z + tz;
        double x1 = X / Z + gauss() * 0.5 / f, y1 = Y / Z + gauss() * 0.5 / f, x2 = X2 / Z2 + gauss() * 0.5 / f, y2 = Y2 / Z2 + gauss() * 0.5 / f;
        if (rnd() < 0.27) { x2 = rnd() * 1.4 - 0.7; y2 = rnd() * 0.9 - 0.45; }
        pts[4 * i] = x1; pts[4 * i + 1] = y1; pts[4 * i + 2] = x2; pts[4 * i + 3] = y2;
    }
    int shown = 0;
    for (int s = 0; s < NS; s++) {
        double q1[5][2], q2[5][2];
        int ks[5];
        for (int i = 0; i < 5; i++) { int k = rand() % K; ks[i] = k; q1[i][0] = pts[4 * k]; q1[i][1] = pts[4 * k + 1]; q2[i][0] = pts[4 * k + 2]; q2[i][1] = pts[4 * k + 3]; }
        double models[90];
        int c = cvm::five_point(q1, q2, models);
        if (g_last_iter > 100 && shown < 6) {
            shown++;
            printf("subset %d idx %d %d %d %d %d: sweeps %d nmodels %d; max_diff at 5,10,20,30,50,100,200,299: %.3g %.3g %.3g %.3g %.3g %.3g %.3g %.3g\n", s, ks[0], ks[1], ks[2], ks[3], ks[4], g_last_iter + 1, c,
                   g_md[5], g_md[10], g_md[20], g_md[30], g_md[50], g_md[100], g_md[200], g_md[299]);
        }
    }
    return 0;
}
