// micro-benchmark: issue cost (clk per warp-instruction per SM sub-partition) of redux.sync / shfl / vimnmx on sm_100a
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(unsigned* out, long long* clk, int iters) {
    unsigned a[8];
    for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 2654435761u + i;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) a[i] = __reduce_max_sync(0xffffffffu, a[i] + it);
            if (OP == 1) a[i] = __shfl_xor_sync(0xffffffffu, a[i] + it, 16);
            if (OP == 2) a[i] = max(a[i] + it, a[(i + 1) & 7]);
            if (OP == 3) a[i] = __vimax3_u32(a[i] + it, a[(i + 1) & 7], a[(i + 2) & 7]);
            if (OP == 4) a[i] = a[i] + it;
        }
    }
    long long t1 = clock64();
    unsigned s = 0;
    for (int i = 0; i < 8; i++) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) clk[0] = t1 - t0;
}
template <int OP>
void run(const char* name, int threads) {
    unsigned* out; long long* clk;
    cudaMalloc(&out, 1 << 22); cudaMalloc(&clk, 8);
    const int iters = 4096;
    k<OP><<<148, threads>>>(out, clk, iters);
    k<OP><<<148, threads>>>(out, clk, iters);
    long long c; cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    const double per = (double)c / (iters * 8.0) / (threads / 128.0);
    printf("%-10s warps/SMSP=%d : %.2f clk per warp-instr per SMSP (loop body includes one IADD per op)\n", name, threads / 128, per);
    cudaFree(out); cudaFree(clk);
}
int main() {
    for (int th : {128, 256, 512}) {
        run<4>("iadd", th); run<0>("redux+iadd", th); run<1>("shfl+iadd", th); run<2>("vimnmx+iadd", th); run<3>("vimnmx3+iadd", th);
    }
    return 0;
}
