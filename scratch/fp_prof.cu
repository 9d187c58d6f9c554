// scratch micro-benchmark: latency distribution of cvm::five_point over realistic RANSAC subsets (not part of the product)
#define FS_PROFILE_FIVEPOINT
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <algorithm>
#include "../forest_slam_b200/csrc/fivepoint.cuh"

__global__ void k(const double* q, int nsub, double* models, int* cnt, long long* cyc, int lanes) {
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (lane >= lanes) return;
    const int s = w * lanes + lane;
    if (s >= nsub) return;
    double q1[5][2], q2[5][2];
    for (int i = 0; i < 5; i++) { q1[i][0] = q[s * 20 + 4 * i]; q1[i][1] = q[s * 20 + 4 * i + 1]; q2[i][0] = q[s * 20 + 4 * i + 2]; q2[i][1] = q[s * 20 + 4 * i + 3]; }
    long long t0 = clock64();
    int c = cvm::five_point(q1, q2, models + (size_t)s * 90);
    long long t1 = clock64();
    cnt[s] = c;
    cyc[s] = t1 - t0;
}

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
    std::vector<double> h(NS * 20);
    for (int s = 0; s < NS; s++)
        for (int i = 0; i < 5; i++) { int k = rand() % K; for (int c = 0; c < 4; c++) h[s * 20 + 4 * i + c] = pts[4 * k + c]; }
    double *q, *models; int* cnt; long long* cyc;
    cudaMalloc(&q, h.size() * 8); cudaMemcpy(q, h.data(), h.size() * 8, cudaMemcpyHostToDevice);
    cudaMalloc(&models, (size_t)NS * 90 * 8); cudaMalloc(&cnt, NS * 4); cudaMalloc(&cyc, NS * 8);
    for (int lanes_ : {1000, 8, 32}) {
        int lanes = lanes_;
        const bool single = lanes == 1000;
        if (single) lanes = 1;
        const int warps = single ? 1 : (NS + lanes - 1) / lanes;
        const int blocks = single ? 1 : (warps * 32 + 63) / 64;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        k<<<blocks, single ? 32 : 64>>>(q, single ? 1 : NS, models, cnt, cyc, lanes);
        int zero[512] = {0}; cudaMemcpyToSymbol(cvm::fp_sweeps, zero, sizeof(zero));
        cudaEventRecord(e0);
        k<<<blocks, single ? 32 : 64>>>(q, single ? 1 : NS, models, cnt, cyc, lanes);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        std::vector<long long> c(NS); cudaMemcpy(c.data(), cyc, NS * 8, cudaMemcpyDeviceToHost);
        std::sort(c.begin(), c.end());
        int hist[512]; cudaMemcpyFromSymbol(hist, cvm::fp_sweeps, sizeof(hist));
        printf("lanes %d: kernel %.3f ms; solve cycles p50 %lld p90 %lld p99 %lld max %lld\n", lanes, ms, c[NS / 2], c[NS * 9 / 10], c[NS * 99 / 100], c[NS - 1]);
        long long pr[8]; cudaMemcpyFromSymbol(pr, cvm::fp_prof, sizeof(pr));
        printf("one thread's stages: svd %lld build %lld gj %lld det %lld dk %lld cand %lld\n", pr[1]-pr[0], pr[2]-pr[1], pr[3]-pr[2], pr[4]-pr[3], pr[5]-pr[4], pr[6]-pr[5]);
        printf("sweep histogram (bucket of 10): ");
        for (int b = 0; b < 52; b++) { int t = 0; for (int j = 0; j < 10 && b * 10 + j < 512; j++) t += hist[b * 10 + j]; if (t) printf("[%d-%d]:%d ", b * 10, b * 10 + 9, t); }
        printf("\n");
    }
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
