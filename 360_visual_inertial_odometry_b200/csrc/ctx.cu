// ctx.cu -- context lifetime, stream, launch accounting and the FMA-peak probe.
#include <cstdlib>

#include "common.cuh"

static std::string g_create_error;

extern "C" int vio360_version(void) { return 100; }

extern "C" const char* vio360_last_error(const vio360_ctx* ctx) {
    return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

extern "C" int vio360_create(int device, vio360_ctx** out) {
    if (!out) { g_create_error = "vio360_create: out is NULL"; return VIO360_ERR_INVALID; }
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) {
        g_create_error = std::string("vio360_create: no CUDA device (") + cudaGetErrorString(e) +
                         "); this library has no CPU fallback";
        return VIO360_ERR_CUDA;
    }
    if (device < 0 || device >= count) { g_create_error = "vio360_create: bad device index"; return VIO360_ERR_INVALID; }
    vio360_ctx* ctx = new vio360_ctx();
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaGetDeviceProperties(&ctx->prop, device)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) {
        g_create_error = std::string("vio360_create: ") + cudaGetErrorString(e);
        delete ctx;
        return VIO360_ERR_CUDA;
    }
    if (ctx->prop.major < 10) {
        g_create_error = "vio360_create: device is not sm_100-class (kernels are built for sm_100a only)";
        cudaStreamDestroy(ctx->stream);
        delete ctx;
        return VIO360_ERR_CUDA;
    }
    ctx->num_sms = ctx->prop.multiProcessorCount;
    { const char* e = getenv("VIO360_DEBUG_SYNC"); ctx->debug_sync = e && e[0] == '1'; }
    *out = ctx;
    return VIO360_OK;
}

extern "C" void vio360_destroy(vio360_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    vio360_fe_release(ctx);
    for (auto& t : ctx->timed) { cudaEventDestroy(t.e0); cudaEventDestroy(t.e1); }
    for (auto e : ctx->event_pool) cudaEventDestroy(e);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    if (ctx->dscratch) cudaFree(ctx->dscratch);
    for (auto& r : ctx->registered) cudaHostUnregister(reinterpret_cast<void*>(r.first));   // ranges the caller forgot
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int vio360_synchronize(vio360_ctx* ctx) {
    if (!ctx) return VIO360_ERR_INVALID;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VIO360_OK;
}

extern "C" int vio360_enable_kernel_timing(vio360_ctx* ctx, int on) {
    if (!ctx) return VIO360_ERR_INVALID;
    ctx->timing = on != 0;
    return VIO360_OK;
}

// folds finished launches into the per-name totals
static int drain_timed(vio360_ctx* ctx) {
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (auto& t : ctx->timed) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, t.e0, t.e1) == cudaSuccess) {
            auto& acc = ctx->kernel_ms[t.name];
            acc.first += ms; acc.second += 1;
        }
        ctx->event_pool.push_back(t.e0); ctx->event_pool.push_back(t.e1);
    }
    ctx->timed.clear();
    return VIO360_OK;
}

extern "C" int vio360_kernel_time(vio360_ctx* ctx, const char* name, double* total_ms, long long* launches) {
    if (!ctx || !name) return VIO360_ERR_INVALID;
    int rc = drain_timed(ctx);
    if (rc) return rc;
    auto it = ctx->kernel_ms.find(name);
    if (total_ms) *total_ms = it == ctx->kernel_ms.end() ? 0.0 : it->second.first;
    if (launches) *launches = it == ctx->kernel_ms.end() ? 0 : it->second.second;
    return VIO360_OK;
}

extern "C" int vio360_reset_kernel_times(vio360_ctx* ctx) {
    if (!ctx) return VIO360_ERR_INVALID;
    int rc = drain_timed(ctx);
    if (rc) return rc;
    ctx->kernel_ms.clear();
    return VIO360_OK;
}

// Page-locks the whole pages INSIDE a caller-owned host range, so that later copies from it run at the PCIe rate.
// The range is not widened: its first and last page may be shared with neighbouring heap allocations, and a page can
// be registered only once.  The frame copies are cut at the registration's boundaries (vio360_h2d_linear).
extern "C" int vio360_host_register(vio360_ctx* ctx, const void* ptr, size_t bytes) {
    if (!ctx || !ptr || bytes == 0) return ctx ? vio360_fail(ctx, VIO360_ERR_INVALID, "host_register: bad arguments") : VIO360_ERR_INVALID;
    const uintptr_t key = reinterpret_cast<uintptr_t>(ptr);
    if (ctx->registered_by_ptr.count(key)) return VIO360_OK;   // already registered through this context
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const uintptr_t page = 4096, a = (key + page - 1) & ~(page - 1), e = (key + bytes) & ~(page - 1);
    if (e <= a) return vio360_fail(ctx, VIO360_ERR_INVALID, "host_register(%p, %zu): the range holds no whole page", ptr, bytes);
    cudaError_t rc = cudaHostRegister(reinterpret_cast<void*>(a), (size_t)(e - a), cudaHostRegisterPortable);
    if (rc != cudaSuccess) {
        cudaGetLastError();
        return vio360_fail(ctx, VIO360_ERR_CUDA, "host_register(%p, %zu): %s%s", ptr, bytes, cudaGetErrorString(rc),
                           rc == cudaErrorHostMemoryAlreadyRegistered ? " (the range overlaps memory that is already page-locked)" : "");
    }
    ctx->registered[a] = e;
    ctx->registered_by_ptr[key] = a;
    return VIO360_OK;
}
extern "C" int vio360_host_unregister(vio360_ctx* ctx, const void* ptr) {
    if (!ctx || !ptr) return VIO360_ERR_INVALID;
    auto it = ctx->registered_by_ptr.find(reinterpret_cast<uintptr_t>(ptr));
    if (it == ctx->registered_by_ptr.end()) return vio360_fail(ctx, VIO360_ERR_INVALID, "host_unregister(%p): not registered through this context", ptr);
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    VIO_CUDA(ctx, cudaDeviceSynchronize());   // no copy from the range may still be in flight (frames travel on a second stream)
    const uintptr_t a = it->second;
    cudaError_t rc = cudaHostUnregister(reinterpret_cast<void*>(a));
    ctx->registered.erase(a);
    ctx->registered_by_ptr.erase(it);
    if (rc != cudaSuccess) { cudaGetLastError(); return vio360_fail(ctx, VIO360_ERR_CUDA, "host_unregister: %s", cudaGetErrorString(rc)); }
    return VIO360_OK;
}

extern "C" void* vio360_stream(vio360_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" long long vio360_launch_count(const vio360_ctx* ctx) { return ctx ? ctx->launches : 0; }

// ---- FMA pipe probe: 8 independent dependent-FMA chains per thread, all SMs busy
template <typename T>
__global__ void __launch_bounds__(256) k_fma_probe(T* out, int iters, T a, T b) {
    T x0 = a + threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
        x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

template <typename T>
static int probe(vio360_ctx* ctx, double* tflops) {
    const int blocks = ctx->num_sms * 8, threads = 256, iters = 4096;
    T* d = nullptr;
    VIO_CUDA(ctx, cudaMalloc(&d, sizeof(T) * blocks * threads));
    cudaEvent_t e0, e1;
    VIO_CUDA(ctx, cudaEventCreate(&e0)); VIO_CUDA(ctx, cudaEventCreate(&e1));
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        VIO_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
        k_fma_probe<T><<<blocks, threads, 0, ctx->stream>>>(d, iters, (T)0.999, (T)0.001);
        ctx->launches++;
        VIO_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
        VIO_CUDA(ctx, cudaEventSynchronize(e1));
        float ms = 0;
        VIO_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
        const double flops = 2.0 * 8.0 * iters * (double)blocks * threads;
        if (rep > 0 && ms > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    *tflops = best;
    return VIO360_OK;
}

extern "C" int vio360_probe_fma_peak(vio360_ctx* ctx, double* fp64_tflops, double* fp32_tflops) {
    if (!ctx) return VIO360_ERR_INVALID;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc;
    if (fp64_tflops && (rc = probe<double>(ctx, fp64_tflops))) return rc;
    if (fp32_tflops && (rc = probe<float>(ctx, fp32_tflops))) return rc;
    return VIO360_OK;
}
