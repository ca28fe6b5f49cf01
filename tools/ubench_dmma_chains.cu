// tools/ubench_dmma_chains.cu -- DMMA m8n8k4 rate vs independent accumulator chains per warp and warps per SM sub-partition (not product code).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_dmma_chains tools/ubench_dmma_chains.cu
#include <cuda_runtime.h>
#include <cstdio>
template <int NC>
__global__ void k_dmma(double* out, int iters) {
    double c[NC][2];
    for (int i = 0; i < NC; ++i) { c[i][0] = 0; c[i][1] = 0; }
    const double a = 1.0 + threadIdx.x * 1e-6, b = 1.0 - threadIdx.x * 1e-6;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NC; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0;
    for (int i = 0; i < NC; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NC>
void run(double* out, int sms, int threads) {
    const int iters = 20000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_dmma<NC><<<sms, threads>>>(out, iters); cudaDeviceSynchronize();
    cudaEventRecord(e0); k_dmma<NC><<<sms, threads>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double tf = 2.0 * 256 * NC * iters * (double)sms * (threads / 32) / (ms * 1e-3) / 1e12;
    // cycles per DMMA per warp at 1.9 GHz nominal
    printf("chains %2d warps/SMSP %d: %.2f TFLOP/s, %.1f ns per DMMA per warp\n", NC, threads / 128, tf, ms * 1e6 / ((double)NC * iters));
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    double* out; cudaMalloc(&out, 8 * 1024 * 1024);
    for (int t : {128, 256, 512}) {
        run<1>(out, p.multiProcessorCount, t); run<2>(out, p.multiProcessorCount, t); run<3>(out, p.multiProcessorCount, t); run<4>(out, p.multiProcessorCount, t);
        run<6>(out, p.multiProcessorCount, t); run<8>(out, p.multiProcessorCount, t); run<12>(out, p.multiProcessorCount, t); run<16>(out, p.multiProcessorCount, t);
    }
    return 0;
}
