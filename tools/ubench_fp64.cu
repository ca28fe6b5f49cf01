// tools/ubench_fp64.cu -- FP64 pipe micro-benchmarks on the B200 box (not product code).
// Measures: dependent-chain DFMA peak, DMMA m8n8k4 (f64 tensor) peak, and the per-call
// cost of the double-precision atan2 / asin / sqrt / division used by the BA factor.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_fp64 tools/ubench_fp64.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cmath>

__global__ void k_dfma(double* out, int iters) {
    double a[8];
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-3 + i;
    const double b = 1.0000001, c = 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = fma(a[i], b, c);
    }
    double s = 0;
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_dmma(double* out, int iters) {
    double c[8][2];
    for (int i = 0; i < 8; ++i) { c[i][0] = 0; c[i][1] = 0; }
    const double a = 1.0 + threadIdx.x * 1e-6, b = 1.0 - threadIdx.x * 1e-6;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0;
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>
__global__ void k_func(double* out, int iters) {
    double x = 0.3 + threadIdx.x * 1e-3, y = 0.7 + blockIdx.x * 1e-4, acc = 0;
    for (int it = 0; it < iters; ++it) {
        double r;
        if (OP == 0) r = atan2(x, y);
        else if (OP == 1) r = asin(x * 0.9);
        else if (OP == 2) r = sqrt(x + y);
        else if (OP == 3) r = x / y;
        else r = rsqrt(x + y);
        acc += r;
        x = x + 1e-9 * r; y = y - 1e-9 * r;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <typename F>
static float time_ms(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f();
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    f();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms;
}

int main() {
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    const int sms = p.multiProcessorCount, blocks = sms * 8, threads = 256;
    double* out;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    const int iters = 20000;
    float ms = time_ms([&] { k_dfma<<<blocks, threads>>>(out, iters); });
    const double dfma = 2.0 * 8 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
    printf("{\"sms\": %d, \"dfma_tflops\": %.3f", sms, dfma);
    ms = time_ms([&] { k_dmma<<<blocks, threads>>>(out, iters); });
    const double dmma = 2.0 * 256 * 8 * iters * (double)blocks * (threads / 32) / (ms * 1e-3) / 1e12;
    printf(", \"dmma_m8n8k4_tflops\": %.3f", dmma);
    const char* names[5] = {"atan2", "asin", "sqrt", "div", "rsqrt"};
    const int fi = 2000;
    float t[5];
    t[0] = time_ms([&] { k_func<0><<<blocks, threads>>>(out, fi); });
    t[1] = time_ms([&] { k_func<1><<<blocks, threads>>>(out, fi); });
    t[2] = time_ms([&] { k_func<2><<<blocks, threads>>>(out, fi); });
    t[3] = time_ms([&] { k_func<3><<<blocks, threads>>>(out, fi); });
    t[4] = time_ms([&] { k_func<4><<<blocks, threads>>>(out, fi); });
    for (int i = 0; i < 5; ++i) {
        // cost in DFMA-equivalents: time per call-thread relative to one DFMA-thread at the measured peak
        const double calls = (double)fi * blocks * threads;
        const double dfma_equiv = (t[i] * 1e-3) / calls * (dfma * 1e12 / 2.0);
        printf(", \"%s_gcalls_s\": %.2f, \"%s_dfma_equiv\": %.1f", names[i], calls / (t[i] * 1e-3) / 1e9, names[i], dfma_equiv);
    }
    printf("}\n");
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { fprintf(stderr, "cuda error %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
