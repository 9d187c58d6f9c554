// pipe_peaks.cu — measured CUDA-core pipe rates of this B200 ("measure with a micro-benchmark and record next
// to MEASURED_PEAKS.json"): fp64 DFMA, fp64 DADD+DMUL (the solvers run without FMA contraction), popc32, MUFU.EX2,
// fp32 FFMA.  Every kernel: 148 x 8 CTAs x 256 threads, 8 independent chains per thread, CUDA events, best of 5.
// Output: one JSON object (profiles/pipe_peaks.json).   nvcc -arch=sm_100a -O3 -o pipe_peaks pipe_peaks.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096, CH = 8;

__global__ void k_dfma(double* out, double a, double b) {
    double x[CH];
    for (int i = 0; i < CH; i++) x[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; it++)
#pragma unroll
        for (int i = 0; i < CH; i++) x[i] = fma(x[i], a, b);
    double s = 0;
    for (int i = 0; i < CH; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_dmuladd(double* out, double a, double b) {
    double x[CH];
    for (int i = 0; i < CH; i++) x[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; it++)
#pragma unroll
        for (int i = 0; i < CH; i++) x[i] = __dadd_rn(__dmul_rn(x[i], a), b);   // two instructions, no contraction
    double s = 0;
    for (int i = 0; i < CH; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma(float* out, float a, float b) {
    float x[CH];
    for (int i = 0; i < CH; i++) x[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; it++)
#pragma unroll
        for (int i = 0; i < CH; i++) x[i] = fmaf(x[i], a, b);
    float s = 0;
    for (int i = 0; i < CH; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_popc(unsigned* out, unsigned a) {
    unsigned x[CH];
    for (int i = 0; i < CH; i++) x[i] = threadIdx.x * 2654435761u + i;
    for (int it = 0; it < ITERS; it++)
#pragma unroll
        for (int i = 0; i < CH; i++) x[i] = __popc(x[i] ^ a) + (x[i] << 3);     // 1 popc + 1 lop + 1 shift/add
    unsigned s = 0;
    for (int i = 0; i < CH; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ex2(float* out, float a) {
    float x[CH];
    for (int i = 0; i < CH; i++) x[i] = (threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < ITERS; it++)
#pragma unroll
        for (int i = 0; i < CH; i++) x[i] = exp2f(x[i] * a);                    // ex2.approx.ftz + 1 fmul
    float s = 0;
    for (int i = 0; i < CH; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
double best_ms(F launch) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    launch();
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        launch();
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int blocks = p.multiProcessorCount * 8, threads = 256;
    void* buf;
    cudaMalloc(&buf, (size_t)blocks * threads * 8);
    const double n = (double)blocks * threads * ITERS * CH;
    const double t_dfma = best_ms([&] { k_dfma<<<blocks, threads>>>((double*)buf, 1.0000001, 1e-9); });
    const double t_dma = best_ms([&] { k_dmuladd<<<blocks, threads>>>((double*)buf, 1.0000001, 1e-9); });
    const double t_ffma = best_ms([&] { k_ffma<<<blocks, threads>>>((float*)buf, 1.0000001f, 1e-9f); });
    const double t_popc = best_ms([&] { k_popc<<<blocks, threads>>>((unsigned*)buf, 0x5a5a5a5au); });
    const double t_ex2 = best_ms([&] { k_ex2<<<blocks, threads>>>((float*)buf, 0.999f); });
    int clk = 0;
    cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"sm_clock_mhz_max\": %.0f, "
           "\"fp64_tflops_fma\": %.2f, \"fp64_tflops_no_fma\": %.2f, \"fp64_ginstr_no_fma\": %.1f, "
           "\"fp32_tflops_fma\": %.2f, \"popc_gops\": %.1f, \"ex2_gops\": %.1f, "
           "\"how\": \"scratch/pipe_peaks.cu: %d CTAs x %d threads, %d independent chains x %d iterations per thread, "
           "CUDA events, best of 5; fp64_tflops_no_fma = DMUL + DADD pairs counted as 2 flop (the rate of the solvers, "
           "which restate OpenCV without FMA contraction); popc/ex2 = instructions/s of that pipe\"}\n",
           p.name, p.multiProcessorCount, clk / 1e3, 2 * n / t_dfma / 1e9, 2 * n / t_dma / 1e9, 2 * n / t_dma / 1e6,
           2 * n / t_ffma / 1e9, n / t_popc / 1e6, n / t_ex2 / 1e6, blocks, threads, CH, ITERS);
    return 0;
}
