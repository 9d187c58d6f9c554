// scratch: fp64 latency / issue rate on this GPU
#include <cstdio>
template <int CH, int OP>
__global__ void k(double* out, long long* cyc, int iters, double seed) {
    double a[CH];
    for (int c = 0; c < CH; c++) a[c] = seed + c * 1e-3 + threadIdx.x * 1e-6;
    const double m = 1.0000001, ad = 1e-9;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CH; c++) {
            if (OP == 0) a[c] = fma(a[c], m, ad);
            else if (OP == 1) a[c] = __dmul_rn(a[c], m);
            else if (OP == 2) a[c] = __dadd_rn(a[c], ad);
            else if (OP == 3) a[c] = 1.0 / a[c];
            else a[c] = sqrt(a[c]);
        }
    }
    long long t1 = clock64();
    double s = 0;
    for (int c = 0; c < CH; c++) s += a[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int CH, int OP>
void run(const char* name, int blocks, int threads) {
    double* out; long long* cyc;
    cudaMalloc(&out, (size_t)blocks * threads * 8); cudaMalloc(&cyc, 8);
    const int iters = 4096;
    k<CH, OP><<<blocks, threads>>>(out, cyc, iters, 1.5);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<CH, OP><<<blocks, threads>>>(out, cyc, iters, 1.5);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double ops = (double)blocks * threads * iters * CH;
    printf("%-6s chains %2d grid %4dx%4d: %8.1f cyc/iter (%.2f cyc/op/warp) | %.3f ms, %.2f Tops/s\n", name, CH, blocks, threads, (double)c / iters, (double)c / iters / CH, ms, ops / ms / 1e9);
    cudaFree(out); cudaFree(cyc);
}
int main() {
    run<1, 0>("dfma", 1, 32); run<8, 0>("dfma", 1, 32); run<16, 0>("dfma", 1, 32);
    run<1, 1>("dmul", 1, 32); run<1, 2>("dadd", 1, 32); run<8, 2>("dadd", 1, 32);
    run<1, 3>("ddiv", 1, 32); run<4, 3>("ddiv", 1, 32); run<1, 4>("dsqrt", 1, 32); run<4, 4>("dsqrt", 1, 32);
    run<8, 0>("dfma", 1, 128); run<8, 0>("dfma", 1, 256); run<8, 0>("dfma", 1, 1024);
    run<8, 0>("dfma", 148, 1024); run<8, 0>("dfma", 148 * 2, 1024);
    run<4, 3>("ddiv", 148 * 2, 1024); run<4, 4>("dsqrt", 148 * 2, 1024);
    return 0;
}
