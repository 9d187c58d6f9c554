// host-side statistics: at which Durand-Kerner sweep does the state become a bitwise fixed point / enter a cycle of
// period <= 4?  (scratch; FP_DK_TRACE sees the roots after every sweep)
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <vector>
struct Snap { double v[20]; };
static Snap g_hist[301]; static int g_n = 0;
#define FP_DK_TRACE(iter, md, r) do { if (iter < 301) { for (int _i = 0; _i < 10; _i++) { g_hist[iter].v[2*_i] = (_i < N) ? r[_i].re : 0; g_hist[iter].v[2*_i+1] = (_i < N) ? r[_i].im : 0; } g_n = iter + 1; } } while (0)
#include "../forest_slam_b200/csrc/fivepoint.cuh"
int main() {
    srand(7);
    auto rnd = []() { return rand() / (double)RAND_MAX; };
    auto gauss = [&]() { return sqrt(-2 * log(rnd() + 1e-12)) * cos(6.283185307 * rnd()); };
    const int K = 1665, NS = 4000;
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
    int hist_fix[302] = {0}, hist_cyc[302] = {0}, percnt[129] = {0};
    int never = 0, total = 0;
    for (int s = 0; s < NS; s++) {
        double q1[5][2], q2[5][2];
        for (int i = 0; i < 5; i++) { int k = rand() % K; q1[i][0] = pts[4 * k]; q1[i][1] = pts[4 * k + 1]; q2[i][0] = pts[4 * k + 2]; q2[i][1] = pts[4 * k + 3]; }
        double models[90];
        g_n = 0;
        cvm::five_point(q1, q2, models);
        if (g_n == 0) continue;
        total++;
        // first sweep t (0-based) whose state equals the state `per` sweeps earlier
        static int first[129]; for (int i = 0; i < 129; i++) first[i] = -1;
        for (int per = 1; per <= 128; per++)
            for (int t = per; t < g_n; t++)
                if (!memcmp(&g_hist[t], &g_hist[t - per], sizeof(Snap))) { first[per] = t; break; }
        if (first[1] >= 0) hist_fix[first[1]]++;
        int best = -1, bp = 0;
        for (int per = 1; per <= 128; per++) if (first[per] >= 0 && (best < 0 || first[per] < best)) { best = first[per]; bp = per; }
        if (best >= 0) { hist_cyc[best]++; percnt[bp]++; } else never++;
    }
    { int le8=0, le32=0, le128=0; for (int q=1;q<=128;q++){ if(q<=8) le8+=percnt[q]; if(q<=32) le32+=percnt[q]; le128+=percnt[q]; } printf("period<=8: %d, <=32: %d, <=128: %d\n", le8, le32, le128); }
    printf("solves %d; never periodic (<=4) within 300: %d (%.1f%%); by period: 1:%d 2:%d 3:%d 4:%d\n", total, never, 100.0 * never / total, percnt[1], percnt[2], percnt[3], percnt[4]);
    int cf = 0, cc = 0;
    for (int t = 0; t <= 300; t++) {
        cf += hist_fix[t]; cc += hist_cyc[t];
        if (t % 8 == 7 || t == 300) printf("by sweep %3d: fixed point %.1f%%, cycle<=4 %.1f%%\n", t + 1, 100.0 * cf / total, 100.0 * cc / total);
    }
    return 0;
}
