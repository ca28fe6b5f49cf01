// tools/ubench_fp64_mix.cu -- one DMMA warp + three DFMA warps per SM sub-partition: shows that the FP64 pipe is shared and that a
// saturating DMMA stream starves the DFMA warps (total time = sum of the two alone).  Not product code.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_fp64_mix tools/ubench_fp64_mix.cu
#include <cuda_runtime.h>
#include <cstdio>
// per SMSP: 1 DMMA warp (4 chains) + 3 DFMA warps (ILP chains each). n1 / n2 = iteration counts (0 = role idle)
template <int ILP>
__global__ void __launch_bounds__(512, 1) k_mix(double* out, long long* tt, int n1, int n2) {
    const int warp = threadIdx.x >> 5;
    const long long t0 = clock64();
    double s = 0;
    if (warp < 4) {
        double c[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
        const double a = 1.0 + threadIdx.x * 1e-6, b = 1.0 - threadIdx.x * 1e-6;
        for (int it = 0; it < n1; ++it) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                             : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
        }
        for (int i = 0; i < 4; ++i) s += c[i][0] + c[i][1];
    } else {
        double a[ILP];
        for (int i = 0; i < ILP; ++i) a[i] = threadIdx.x * 1e-3 + i;
        const double b = 1.0000001, c = 1e-9;
        for (int it = 0; it < n2; ++it) {
#pragma unroll
            for (int i = 0; i < ILP; ++i) a[i] = fma(a[i], b, c);
        }
        for (int i = 0; i < ILP; ++i) s += a[i];
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) tt[warp] = t1 - t0;
}
template <int ILP>
void run(double* out, long long* tt, int sms, int n1, int n2) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_mix<ILP><<<sms, 512>>>(out, tt, n1, n2); cudaDeviceSynchronize();
    cudaEventRecord(e0); k_mix<ILP><<<sms, 512>>>(out, tt, n1, n2); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[16]; cudaMemcpy(h, tt, sizeof(h), cudaMemcpyDeviceToHost);
    printf("ILP %d n1 %6d n2 %6d: %.3f ms; dmma warp %lld cyc (%.1f/DMMA), dfma warp %lld cyc (%.2f/DFMA)\n", ILP, n1, n2, ms, h[0],
           n1 ? (double)h[0] / (4.0 * n1) : 0.0, h[4], n2 ? (double)h[4] / ((double)ILP * n2) : 0.0);
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    double* out; cudaMalloc(&out, 8 * 1024 * 1024);
    long long* tt; cudaMalloc(&tt, 256);
    const int sms = p.multiProcessorCount;
    // DMMA alone: 10000 iters x 4 = 40000 DMMAs x 16 cyc = 640k cycles
    run<1>(out, tt, sms, 10000, 0);
    // DFMA alone, 3 warps/SMSP
    run<1>(out, tt, sms, 0, 40000); run<2>(out, tt, sms, 0, 20000); run<4>(out, tt, sms, 0, 10000); run<8>(out, tt, sms, 0, 5000);
    // both
    run<1>(out, tt, sms, 10000, 40000); run<2>(out, tt, sms, 10000, 20000); run<4>(out, tt, sms, 10000, 10000); run<8>(out, tt, sms, 10000, 5000);
    run<8>(out, tt, sms, 10000, 20000); run<1>(out, tt, sms, 10000, 10000);
    return 0;
}
