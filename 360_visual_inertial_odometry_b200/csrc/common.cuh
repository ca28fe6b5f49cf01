// common.cuh -- context object and error plumbing shared by all translation units of
// libvio360_cuda.so.  No torch types, no CPU fallbacks.
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "../../include/vio360.h"

struct vio360_fe_state;  // fe.cu

struct vio360_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaDeviceProp prop{};
    int num_sms = 148;
    long long launches = 0;
    std::string err;
    vio360_fe_state* fe = nullptr;
    // small pinned + device scratch used to ship descriptor tables
    void* pinned = nullptr;
    size_t pinned_bytes = 0;
    // device scratch of the stateless entry points (vio360_pnp_evaluate, vio360_triangulate): grown on demand, kept
    // between calls (those calls end with a stream synchronise, so it is never in use when the next one starts)
    void* dscratch = nullptr;
    size_t dscratch_bytes = 0;
    // optional per-kernel CUDA-event timing (bench.py's roofline numbers)
    bool timing = false;
    bool debug_sync = false;     // VIO360_DEBUG_SYNC=1: synchronise after every launch so a fault names its kernel
    struct TimedLaunch { const char* name; cudaEvent_t e0, e1; };
    std::vector<TimedLaunch> timed;
    std::vector<cudaEvent_t> event_pool;
    std::unordered_map<std::string, std::pair<double, long long>> kernel_ms;   // name -> (total ms, launches)
    // Host ranges page-locked through vio360_host_register: begin -> end of the WHOLE pages inside the caller's range
    // (never widened: the pages at both ends may be shared with a neighbouring allocation).  A copy must lie inside
    // one registration or outside all of them (the runtime rejects a copy that straddles a boundary), so host frame
    // copies are cut at these boundaries (vio360_h2d_linear).
    std::map<uintptr_t, uintptr_t> registered;
    std::unordered_map<uintptr_t, uintptr_t> registered_by_ptr;   // caller's pointer -> begin of its registration
};

// Brackets one kernel launch with events on the context's stream when timing is on.
struct KernelTimer {
    vio360_ctx* ctx; size_t idx; bool on;
    KernelTimer(vio360_ctx* c, const char* name) : ctx(c), idx(0), on(c->timing) {
        if (!on) return;
        auto get = [&]() { cudaEvent_t e; if (!ctx->event_pool.empty()) { e = ctx->event_pool.back(); ctx->event_pool.pop_back(); } else cudaEventCreate(&e); return e; };
        vio360_ctx::TimedLaunch t{name, get(), get()};
        cudaEventRecord(t.e0, ctx->stream);
        ctx->timed.push_back(t); idx = ctx->timed.size() - 1;
    }
    ~KernelTimer() { if (on) cudaEventRecord(ctx->timed[idx].e1, ctx->stream); }
};

inline int vio360_fail(vio360_ctx* ctx, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    return code;
}

#define VIO_CUDA(ctx, call)                                                              \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return vio360_fail((ctx), VIO360_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, \
                               #call, cudaGetErrorString(e_));                           \
    } while (0)

#define VIO_LAUNCH_CHECK(ctx)                                                            \
    do {                                                                                 \
        (ctx)->launches++;                                                               \
        cudaError_t e_ = cudaGetLastError();                                             \
        if (e_ == cudaSuccess && (ctx)->debug_sync) e_ = cudaStreamSynchronize((ctx)->stream); \
        if (e_ != cudaSuccess)                                                           \
            return vio360_fail((ctx), VIO360_ERR_CUDA, "%s:%d kernel launch -> %s", __FILE__, \
                               __LINE__, cudaGetErrorString(e_));                        \
    } while (0)

// RAII-less device buffer helper (explicit free; sized lazily)
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t reserve(size_t n) {
        if (n <= bytes) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};
struct PinBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t reserve(size_t n) {
        if (n <= bytes) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; bytes = 0;
        cudaError_t e = cudaMallocHost(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

void vio360_fe_release(vio360_ctx* ctx);  // fe.cu

inline cudaError_t vio360_pinned(vio360_ctx* ctx, size_t bytes, void** out) {   // the context's pinned staging buffer, grown on demand
    if (bytes > ctx->pinned_bytes) {
        if (ctx->pinned) cudaFreeHost(ctx->pinned);
        ctx->pinned = nullptr; ctx->pinned_bytes = 0;
        const size_t want = (bytes + (1u << 16) - 1) & ~(size_t)((1u << 16) - 1);
        cudaError_t e = cudaMallocHost(&ctx->pinned, want);
        if (e != cudaSuccess) return e;
        ctx->pinned_bytes = want;
    }
    *out = ctx->pinned;
    return cudaSuccess;
}
inline cudaError_t vio360_scratch(vio360_ctx* ctx, size_t bytes, void** out) {
    if (bytes > ctx->dscratch_bytes) {
        if (ctx->dscratch) cudaFree(ctx->dscratch);
        ctx->dscratch = nullptr; ctx->dscratch_bytes = 0;
        const size_t want = (bytes + (1u << 20) - 1) & ~(size_t)((1u << 20) - 1);
        cudaError_t e = cudaMalloc(&ctx->dscratch, want);
        if (e != cudaSuccess) return e;
        ctx->dscratch_bytes = want;
    }
    *out = ctx->dscratch;
    return cudaSuccess;
}

// Launch with programmatic dependent launch allowed: the kernel may be scheduled while the previous kernel of the
// stream is still draining (once all of that kernel's CTAs have executed griddepcontrol.launch_dependents or exited);
// it must execute griddepcontrol.wait before it touches anything the previous kernel wrote.  Hides the launch gap
// between the short kernels of one BA step (tile -> reduce -> pull).
template <typename... KArgs, typename... Args>
inline cudaError_t vio360_launch_pdl(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Host-to-device copy of `bytes` contiguous bytes, cut at the boundaries of the registered host ranges so that every
// piece lies inside one registration (DMA straight from the caller's pages) or in pageable memory (the short fringes).
inline cudaError_t vio360_h2d_linear(vio360_ctx* ctx, void* dst, const void* src, size_t bytes, cudaStream_t st) {
    uintptr_t p = reinterpret_cast<uintptr_t>(src);
    const uintptr_t end = p + bytes;
    uint8_t* d = static_cast<uint8_t*>(dst);
    if (!ctx->registered.empty()) {
        auto it = ctx->registered.upper_bound(p);
        if (it != ctx->registered.begin()) { auto pr = std::prev(it); if (pr->second > p) it = pr; }
        for (; it != ctx->registered.end() && it->first < end && p < end; ++it) {
            const uintptr_t a = it->first > p ? it->first : p, e = it->second < end ? it->second : end;
            if (a > p) {   // pageable piece in front of this registration
                cudaError_t rc = cudaMemcpyAsync(d, reinterpret_cast<const void*>(p), a - p, cudaMemcpyHostToDevice, st);
                if (rc != cudaSuccess) return rc;
                d += a - p; p = a;
            }
            if (e > p) {
                cudaError_t rc = cudaMemcpyAsync(d, reinterpret_cast<const void*>(p), e - p, cudaMemcpyHostToDevice, st);
                if (rc != cudaSuccess) return rc;
                d += e - p; p = e;
            }
        }
    }
    if (p < end) return cudaMemcpyAsync(d, reinterpret_cast<const void*>(p), end - p, cudaMemcpyHostToDevice, st);
    return cudaSuccess;
}
