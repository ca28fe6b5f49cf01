// tools/ubench_dfma_warps.cu -- how many warps of ONE CTA does it take to fill the FP64 pipe of an SM?  (not product code)
// Every thread runs 36 independent DFMA chains (the register block of k_chol_band_reg); prints cycles per warp-DFMA of the CTA.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_dfma_warps tools/ubench_dfma_warps.cu
#include <cuda_runtime.h>
#include <cstdio>
template <int ILP>
__global__ void k(double* out, long long* cyc, int iters) {
    double a[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) a[i] = threadIdx.x * 1e-3 + i;
    const double b = 1.0000001 + threadIdx.x * 1e-9, c = 1e-9;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) a[i] = fma(a[i], b, c);
    }
    __syncthreads();
    const long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += a[i];
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int ILP>
void run(const char* name) {
    double* out; long long* cyc;
    cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 8);
    const int iters = 200;
    for (int warps : {1, 2, 3, 4, 6, 8, 12, 16, 24, 32}) {
        k<ILP><<<1, warps * 32>>>(out, cyc, iters);
        k<ILP><<<1, warps * 32>>>(out, cyc, iters);
        long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        const double dfma = (double)iters * ILP * warps;
        printf("%s: %2d warps: %8lld cycles, %.2f cycles per warp-DFMA of the SM (%.1f DFMA lanes / clk)\n", name, warps, h, h / dfma, 32.0 * dfma / h);
    }
}
int main() { run<36>("ILP 36"); run<8>("ILP 8"); return 0; }
