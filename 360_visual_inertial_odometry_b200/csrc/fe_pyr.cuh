// fe_pyr.cuh -- K1: integer image pyramid (cv::pyrDown semantics, bit exact).
//
// level l+1 (x,y) = ( sum_{i,j} g_i g_j  L_l(reflect101(2x-2+i), reflect101(2y-2+j)) + 128 ) >> 8
// with g = [1 4 6 4 1]  (OpenCV modules/imgproc/src/pyramids.cpp, 8U path, called from
// buildOpticalFlowPyramid inside cv::calcOpticalFlowPyrLK, reference call site
// src/processing/FeatureTracker.cpp:238).
//
// One CTA produces a 128 x 16 output tile: the (2*16+3) x (2*128+3) input footprint is
// read once from global memory with aligned 32-bit loads, reduced horizontally into
// shared memory (u16), then vertically, and written back as coalesced uchar4 stores.
// HBM bound: algorithmic traffic is 1.25 bytes per input pixel.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace vio360 {

constexpr int kMaxLevels = 8;

struct PyrJob {
    const uint8_t* src; uint8_t* dst;
    int spitch, dpitch;
};

__host__ __device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * len - 2 - p;
    return p;
}

constexpr int kPyrTW = 128;  // output tile width
constexpr int kPyrTH = 16;   // output tile height
constexpr int kPyrRows = 2 * kPyrTH + 3;

__global__ void __launch_bounds__(256) k_pyr_down(const PyrJob* __restrict__ jobs, int w, int h, int dw, int dh) {
    __shared__ uint16_t hs[kPyrRows][kPyrTW + 8];
    const PyrJob job = jobs[blockIdx.z];
    const int ox0 = blockIdx.x * kPyrTW, oy0 = blockIdx.y * kPyrTH;
    const int iy0 = 2 * oy0 - 2;
    // fast path: every input byte the tile touches is inside the image, rows 4-byte aligned
    const bool interior = (2 * ox0 - 4 >= 0) && (2 * (ox0 + kPyrTW) + 4 <= w) && (iy0 >= 0) &&
                          (iy0 + kPyrRows <= h) && ((job.spitch & 3) == 0) &&
                          ((reinterpret_cast<uintptr_t>(job.src) & 3) == 0);
    constexpr int kGroups = kPyrTW / 4;
    for (int it = threadIdx.x; it < kPyrRows * kGroups; it += blockDim.x) {
        const int r = it / kGroups, g = it - r * kGroups;
        const int ox = ox0 + 4 * g;      // first of 4 output columns
        uint16_t o[4];
        if (interior) {
            const uint32_t* row = reinterpret_cast<const uint32_t*>(job.src + (size_t)(iy0 + r) * job.spitch + (2 * ox - 4));
            const uint32_t w0 = __ldg(row), w1 = __ldg(row + 1), w2 = __ldg(row + 2), w3 = __ldg(row + 3);
            // bytes: w0 = in[2ox-4..2ox-1], w1 = in[2ox..2ox+3], w2 = in[2ox+4..2ox+7], w3 = in[2ox+8..2ox+11]
            int b[12];
            b[0] = (w0 >> 16) & 255; b[1] = w0 >> 24;
            b[2] = w1 & 255; b[3] = (w1 >> 8) & 255; b[4] = (w1 >> 16) & 255; b[5] = w1 >> 24;
            b[6] = w2 & 255; b[7] = (w2 >> 8) & 255; b[8] = (w2 >> 16) & 255; b[9] = w2 >> 24;
            b[10] = w3 & 255; b[11] = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k)   // output ox+k is centred on input 2(ox+k) = b[2k+2]
                o[k] = (uint16_t)(b[2 * k] + 4 * b[2 * k + 1] + 6 * b[2 * k + 2] + 4 * b[2 * k + 3] + b[2 * k + 4]);
        } else {
            const uint8_t* row = job.src + (size_t)reflect101(iy0 + r, h) * job.spitch;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int cx = 2 * (ox + k);
                int v = 0;
                if (ox + k < dw)
                    v = row[reflect101(cx - 2, w)] + 4 * row[reflect101(cx - 1, w)] + 6 * row[reflect101(cx, w)] +
                        4 * row[reflect101(cx + 1, w)] + row[reflect101(cx + 2, w)];
                o[k] = (uint16_t)v;
            }
        }
        *reinterpret_cast<uint2*>(&hs[r][4 * g]) = make_uint2(o[0] | ((uint32_t)o[1] << 16), o[2] | ((uint32_t)o[3] << 16));
    }
    __syncthreads();
    for (int it = threadIdx.x; it < kPyrTH * kGroups; it += blockDim.x) {
        const int ry = it / kGroups, g = it - ry * kGroups;
        const int oy = oy0 + ry, ox = ox0 + 4 * g;
        if (oy >= dh || ox >= dw) continue;
        uint32_t out = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int v = hs[2 * ry][4 * g + k] + 4 * hs[2 * ry + 1][4 * g + k] + 6 * hs[2 * ry + 2][4 * g + k] +
                          4 * hs[2 * ry + 3][4 * g + k] + hs[2 * ry + 4][4 * g + k];
            out |= (uint32_t)((v + 128) >> 8) << (8 * k);
        }
        uint8_t* dp = job.dst + (size_t)oy * job.dpitch + ox;
        if (ox + 3 < dw && (job.dpitch & 3) == 0) {
            *reinterpret_cast<uint32_t*>(dp) = out;
        } else {
            for (int k = 0; k < 4 && ox + k < dw; ++k) dp[k] = (uint8_t)(out >> (8 * k));
        }
    }
}


// ---------------------------------------------------------------------------------------
// Fast path (w % 8 == 0, 8-byte aligned rows): a warp marches down a 240-column strip.
// Each lane owns 8 input columns (one 64-bit load per row), keeps the vertical 5-tap
// window of the last five rows as four packed 2 x u16 registers per row, and the
// horizontal 5-tap pass takes its two left / one right neighbour columns from the
// adjacent lanes by shuffle -- no shared memory, no barriers.  Lanes 0 and 31 only supply
// halos, so a warp writes 120 output bytes per row (one 32-bit store per lane).
// All arithmetic is exact: vertical sums <= 16*255, full sums <= 256*255 + 128 < 2^16, so
// two columns share one 32-bit register without carries between the halves.
constexpr int kPyrStripRows = 32;          // output rows per warp (upper bound; halved while the grid underfills the GPU)
constexpr int kPyrWarpCols = 240;          // input columns whose outputs one warp writes
constexpr int kPyrFastThreads = 256;

struct PyrRow { uint32_t e0, o0, e1, o1; };   // (c0,c2) (c1,c3) (c4,c6) (c5,c7) as 2 x u16

__device__ __forceinline__ PyrRow pyr_load_row(const uint8_t* __restrict__ p) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(p));
    PyrRow r;
    r.e0 = __byte_perm(v.x, 0u, 0x4240); r.o0 = __byte_perm(v.x, 0u, 0x4341);
    r.e1 = __byte_perm(v.y, 0u, 0x4240); r.o1 = __byte_perm(v.y, 0u, 0x4341);
    return r;
}

__device__ __forceinline__ uint32_t pyr_v5(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e) {
    return (a + e) + 4u * (b + d) + 6u * c;
}

__device__ __forceinline__ PyrRow pyr_unpack(uint2 v) {
    PyrRow r;
    r.e0 = __byte_perm(v.x, 0u, 0x4240); r.o0 = __byte_perm(v.x, 0u, 0x4341);
    r.e1 = __byte_perm(v.y, 0u, 0x4240); r.o1 = __byte_perm(v.y, 0u, 0x4341);
    return r;
}

// horizontal 5-tap + rounding of one output row from the vertical sums of this lane and its neighbours
__device__ __forceinline__ uint32_t pyr_hrow(const PyrRow& r0, const PyrRow& r1, const PyrRow& r2, const PyrRow& r3,
                                             const PyrRow& r4, bool left_edge, bool right_edge) {
    const uint32_t ve0 = pyr_v5(r0.e0, r1.e0, r2.e0, r3.e0, r4.e0);   // (V0,V2)
    const uint32_t vo0 = pyr_v5(r0.o0, r1.o0, r2.o0, r3.o0, r4.o0);   // (V1,V3)
    const uint32_t ve1 = pyr_v5(r0.e1, r1.e1, r2.e1, r3.e1, r4.e1);   // (V4,V6)
    const uint32_t vo1 = pyr_v5(r0.o1, r1.o1, r2.o1, r3.o1, r4.o1);   // (V5,V7)
    // neighbours: (V6,V7) of the lane to the left, V0 of the lane to the right
    uint32_t lft = __shfl_up_sync(0xffffffffu, __byte_perm(ve1, vo1, 0x7632), 1);
    uint32_t rgt = __shfl_down_sync(0xffffffffu, ve0, 1);
    if (left_edge) lft = __byte_perm(ve0, vo0, 0x5432);    // reflect: V[-2] = V2, V[-1] = V1
    if (right_edge) rgt = ve1 >> 16;                        // reflect: V[8] = V6
    const uint32_t a01 = __byte_perm(lft, ve0, 0x5410);     // (V-2, V0)
    const uint32_t b01 = __byte_perm(lft, vo0, 0x5432);     // (V-1, V1)
    const uint32_t c01 = __byte_perm(ve0, ve1, 0x5432);     // (V2, V4)
    const uint32_t b23 = __byte_perm(vo0, vo1, 0x5432);     // (V3, V5)
    const uint32_t c23 = __byte_perm(ve1, rgt, 0x5432);     // (V6, V8)
    const uint32_t p01 = (a01 + c01) + 4u * (b01 + vo0) + 6u * ve0 + 0x00800080u;
    const uint32_t p23 = (c01 + c23) + 4u * (b23 + vo1) + 6u * ve1 + 0x00800080u;
    return __byte_perm(p01, p23, 0x7531);
}

// strip_rows (a multiple of 4) output rows per warp; four output rows per loop trip with the eight input
// rows of the next trip already in flight (2 KB of loads per warp outstanding: at ~24 resident warps per SM
// that is the ~35 KB per SM HBM3e needs in flight; with 1 KB the kernel stalled at 4.0 TB/s).
__global__ void __launch_bounds__(kPyrFastThreads)
k_pyr_down8(const PyrJob* __restrict__ jobs, int w, int h, int dw, int dh, int nwx, int nsy, int total_warps,
            int strip_rows) {
    const int gw = blockIdx.x * (kPyrFastThreads / 32) + (threadIdx.x >> 5);
    if (gw >= total_warps) return;
    const int lane = threadIdx.x & 31;
    const int wx = gw % nwx, t = gw / nwx, sy = t % nsy, jb = t / nsy;
    const PyrJob job = jobs[jb];
    const int col0 = wx * kPyrWarpCols - 8 + 8 * lane;           // first input column of this lane
    const int lcol = min(max(col0, 0), w - 8);                    // clamped load column
    const bool left_edge = col0 == 0, right_edge = col0 + 8 == w;
    const bool writer = lane >= 1 && lane <= 30 && col0 >= 0 && col0 < w;
    const uint8_t* src = job.src + lcol;
    const int y0 = sy * strip_rows, y1 = min(y0 + strip_rows, dh);
    auto load = [&](int r) { return __ldg(reinterpret_cast<const uint2*>(src + (size_t)reflect101(r, h) * job.spitch)); };
    PyrRow r0 = pyr_unpack(load(2 * y0 - 2)), r1 = pyr_unpack(load(2 * y0 - 1)), r2 = pyr_unpack(load(2 * y0));
    uint2 n[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) n[i] = load(2 * y0 + 1 + i);
    uint8_t* dst = job.dst + (size_t)y0 * job.dpitch + (col0 >> 1);
    for (int y = y0; y < y1; y += 4) {
        PyrRow q[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = pyr_unpack(n[i]);
        if (y + 4 < y1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) n[i] = load(2 * y + 9 + i);
        }
        const uint32_t o0 = pyr_hrow(r0, r1, r2, q[0], q[1], left_edge, right_edge);
        const uint32_t o1 = pyr_hrow(r2, q[0], q[1], q[2], q[3], left_edge, right_edge);
        const uint32_t o2 = pyr_hrow(q[1], q[2], q[3], q[4], q[5], left_edge, right_edge);
        const uint32_t o3 = pyr_hrow(q[3], q[4], q[5], q[6], q[7], left_edge, right_edge);
        if (writer) {
            *reinterpret_cast<uint32_t*>(dst) = o0;
            if (y + 1 < y1) *reinterpret_cast<uint32_t*>(dst + job.dpitch) = o1;
            if (y + 2 < y1) *reinterpret_cast<uint32_t*>(dst + 2 * (size_t)job.dpitch) = o2;
            if (y + 3 < y1) *reinterpret_cast<uint32_t*>(dst + 3 * (size_t)job.dpitch) = o3;
        }
        dst += 4 * (size_t)job.dpitch;
        r0 = q[5]; r1 = q[6]; r2 = q[7];
    }
}

}  // namespace vio360
