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

}  // namespace vio360
