// fe_lk.cuh -- K3 (+K2 fused): pyramidal Lucas-Kanade tracker, one warp per feature.
//
// Restates cv::calcOpticalFlowPyrLK's LKTrackerInvoker (OpenCV 4.x
// modules/video/src/lkpyramid.cpp) as called by FeatureTracker::TrackOpticalFlow
// (src/processing/FeatureTracker.cpp:238-240): fixed-point bilinear sampling with 14-bit
// weights, 5 fractional bits on the patch, Scharr derivatives (computed here from a
// 24x24 shared-memory tile instead of a materialised derivative image), REFLECT_101
// image apron, zero derivative apron, float 2x2 solve, the eps^2 / oscillation exits and
// the level-0 error epilogue.  The 2x2 normal-equation sums are exact integer warp
// reductions (OpenCV accumulates the same integers in float lanes; the difference is
// below 2^-23 relative and is what the 0.05 px position tolerance absorbs).
// This translation unit is compiled with --fmad=false so float expressions round like
// the non-FMA baseline build of lkpyramid.cpp.
#pragma once

#include "fe_pyr.cuh"

namespace vio360 {

struct FrameLevels {
    const uint8_t* p[kMaxLevels];
    int pitch[kMaxLevels];
};

struct TrackJobDev {
    FrameLevels prev, cur;
    const float* prev_xy;
    float* cur_xy;
    uint8_t* status;
    float* err;    // may be null
    uint8_t* keep; // may be null
    int n;
};

struct LkParams {
    int nlevels;
    int w[kMaxLevels], h[kMaxLevels];
    int max_iters;
    double eps2, min_eig;
    float polar_thr, margin;
};

constexpr int kWin = 21;
constexpr int kNpx = kWin * kWin;       // 441
constexpr int kPpl = (kNpx + 31) / 32;  // 14 window pixels per lane
constexpr int kTRows = 24, kTPitch = 32;
constexpr int kDDim = 22;
constexpr int kJRows = 32, kJPitch = 36;
constexpr int kLkWarps = 4;

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Fill a (rows x pitch) shared tile with image bytes starting at (x0, y0); x0 % 4 == 0.
// Interior tiles use aligned 32-bit loads; tiles touching the border index with
// REFLECT_101 byte by byte (the pyramid apron of buildOpticalFlowPyramid).
template <int ROWS, int PITCH, int WORDS>
__device__ __forceinline__ void stage_tile(uint8_t* s, const uint8_t* __restrict__ img, int pitch, int w, int h,
                                           int x0, int y0, int lane) {
    const bool interior = x0 >= 0 && y0 >= 0 && (x0 + 4 * WORDS) <= w && (y0 + ROWS) <= h && ((pitch & 3) == 0) &&
                          ((reinterpret_cast<uintptr_t>(img) & 3) == 0);
    if (interior) {
        for (int it = lane; it < ROWS * WORDS; it += 32) {
            const int r = it / WORDS, c = it - r * WORDS;
            const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(img + (size_t)(y0 + r) * pitch + x0) + c);
            *reinterpret_cast<uint32_t*>(s + r * PITCH + 4 * c) = v;
        }
    } else {
        for (int it = lane; it < ROWS * WORDS * 4; it += 32) {
            const int r = it / (WORDS * 4), c = it - r * (WORDS * 4);
            s[r * PITCH + c] = img[(size_t)reflect101(y0 + r, h) * pitch + reflect101(x0 + c, w)];
        }
    }
}

__device__ __forceinline__ void bilinear_weights(float a, float b, int& w00, int& w01, int& w10, int& w11) {
    w00 = __float2int_rn((1.f - a) * (1.f - b) * 16384.f);
    w01 = __float2int_rn(a * (1.f - b) * 16384.f);
    w10 = __float2int_rn((1.f - a) * b * 16384.f);
    w11 = 16384 - w00 - w01 - w10;
}

__global__ void __launch_bounds__(kLkWarps * 32) k_lk_track(const TrackJobDev* __restrict__ jobs, LkParams P) {
    __shared__ __align__(16) uint8_t sT_all[kLkWarps][kTRows * kTPitch];
    __shared__ __align__(16) uint32_t sD_all[kLkWarps][kDDim * kDDim];
    __shared__ __align__(16) uint8_t sJ_all[kLkWarps][kJRows * kJPitch];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const TrackJobDev& job = jobs[blockIdx.y];
    const int pt = blockIdx.x * kLkWarps + warp;
    if (pt >= job.n) return;   // warp-uniform; only __syncwarp is used below
    uint8_t* sT = sT_all[warp];
    uint32_t* sD = sD_all[warp];
    uint8_t* sJ = sJ_all[warp];

    // window pixel owned by this lane in round k: idx = lane + 32k
    int wofs[kPpl];   // (wy << 8) | wx, or -1
#pragma unroll
    for (int k = 0; k < kPpl; ++k) {
        const int idx = lane + 32 * k;
        const int wy = idx / kWin, wx = idx - wy * kWin;
        wofs[k] = idx < kNpx ? ((wy << 8) | wx) : -1;
    }

    const float px0 = job.prev_xy[2 * pt], py0 = job.prev_xy[2 * pt + 1];
    const float half = (kWin - 1) * 0.5f;
    float outx = 0.f, outy = 0.f;   // nextPts[ptidx]
    int status = 1;
    float errv = 0.f;

    for (int level = P.nlevels - 1; level >= 0; --level) {
        const int w = P.w[level], h = P.h[level];
        const float sc = (float)(1. / (1 << level));
        float px = px0 * sc, py = py0 * sc;
        float nx, ny;
        if (level == P.nlevels - 1) { nx = px; ny = py; }
        else { nx = outx * 2.f; ny = outy * 2.f; }
        outx = nx; outy = ny;
        px -= half; py -= half;
        const int ipx = (int)floorf(px), ipy = (int)floorf(py);
        if (ipx < -kWin || ipx >= w || ipy < -kWin || ipy >= h) {
            if (level == 0) { status = 0; errv = 0.f; }
            continue;
        }
        // ---- template: raw 24x24 tile -> Scharr 22x22 -> 21x21 fixed-point patch in registers
        const uint8_t* imgI = job.prev.p[level];
        const int pitchI = job.prev.pitch[level];
        const int xT0 = (ipx - 1) & ~3, sh = (ipx - 1) - xT0;
        __syncwarp();
        stage_tile<kTRows, kTPitch, 7>(sT, imgI, pitchI, w, h, xT0, ipy - 1, lane);
        __syncwarp();
        for (int it = lane; it < kDDim * kDDim; it += 32) {
            const int dy = it / kDDim, dx = it - dy * kDDim;
            const int gx = ipx + dx, gy = ipy + dy;
            uint32_t packed = 0;
            if (gx >= 0 && gx < w && gy >= 0 && gy < h) {   // derivative apron is zero
                const uint8_t* t = sT + (dy + 1) * kTPitch + sh + dx + 1;   // centre
                const int a00 = t[-kTPitch - 1], a01 = t[-kTPitch], a02 = t[-kTPitch + 1];
                const int a10 = t[-1], a12 = t[1];
                const int a20 = t[kTPitch - 1], a21 = t[kTPitch], a22 = t[kTPitch + 1];
                const int ix = (3 * (a02 + a22) + 10 * a12) - (3 * (a00 + a20) + 10 * a10);
                const int iy = 3 * ((a20 - a00) + (a22 - a02)) + 10 * (a21 - a01);
                packed = ((uint32_t)ix & 0xffffu) | ((uint32_t)iy << 16);
            }
            sD[it] = packed;
        }
        __syncwarp();
        int w00, w01, w10, w11;
        bilinear_weights(px - ipx, py - ipy, w00, w01, w10, w11);
        int Iv[kPpl], Ixy[kPpl];   // patch value (5 frac bits); packed (Ix | Iy << 16)
        int sA11 = 0, sA12 = 0, sA22 = 0;
#pragma unroll
        for (int k = 0; k < kPpl; ++k) {
            Iv[k] = 0; Ixy[k] = 0;
            if (wofs[k] >= 0) {
                const int wy = wofs[k] >> 8, wx = wofs[k] & 255;
                const uint8_t* t = sT + (wy + 1) * kTPitch + sh + wx + 1;
                const int iv = (t[0] * w00 + t[1] * w01 + t[kTPitch] * w10 + t[kTPitch + 1] * w11 + (1 << 8)) >> 9;
                const uint32_t* dp = sD + wy * kDDim + wx;
                const uint32_t d00 = dp[0], d01 = dp[1], d10 = dp[kDDim], d11 = dp[kDDim + 1];
                const int ixv = ((int)(short)(d00 & 0xffff) * w00 + (int)(short)(d01 & 0xffff) * w01 +
                                 (int)(short)(d10 & 0xffff) * w10 + (int)(short)(d11 & 0xffff) * w11 + (1 << 13)) >> 14;
                const int iyv = (((int)d00 >> 16) * w00 + ((int)d01 >> 16) * w01 + ((int)d10 >> 16) * w10 +
                                 ((int)d11 >> 16) * w11 + (1 << 13)) >> 14;
                Iv[k] = iv;
                Ixy[k] = (ixv & 0xffff) | (iyv << 16);
                sA11 += ixv * ixv; sA12 += ixv * iyv; sA22 += iyv * iyv;
            }
        }
        const float FLT_SCALE = 1.f / (1 << 20);
        const float A11 = (float)warp_sum_ll(sA11) * FLT_SCALE;
        const float A12 = (float)warp_sum_ll(sA12) * FLT_SCALE;
        const float A22 = (float)warp_sum_ll(sA22) * FLT_SCALE;
        float D = A11 * A22 - A12 * A12;
        const float tt = A11 - A22;
        const float minEig = (A22 + A11 - sqrtf(tt * tt + 4.f * A12 * A12)) / (float)(2 * kWin * kWin);
        if ((double)minEig < P.min_eig || D < 1.1920928955078125e-07f) {
            if (level == 0) status = 0;
            continue;
        }
        D = 1.f / D;
        nx -= half; ny -= half;
        float pdx = 0.f, pdy = 0.f;
        const uint8_t* imgJ = job.cur.p[level];
        const int pitchJ = job.cur.pitch[level];
        int tx0 = 0, ty0 = 0;
        bool tile_ok = false;
        for (int j = 0; j < P.max_iters; ++j) {
            const int inx = (int)floorf(nx), iny = (int)floorf(ny);
            if (inx < -kWin || inx >= w || iny < -kWin || iny >= h) {
                if (level == 0) status = 0;
                break;
            }
            if (!tile_ok || inx < tx0 || inx + 22 > tx0 + kJPitch || iny < ty0 || iny + 22 > ty0 + kJRows) {
                tx0 = (inx - 5) & ~3; ty0 = iny - 5;
                __syncwarp();
                stage_tile<kJRows, kJPitch, 9>(sJ, imgJ, pitchJ, w, h, tx0, ty0, lane);
                __syncwarp();
                tile_ok = true;
            }
            bilinear_weights(nx - inx, ny - iny, w00, w01, w10, w11);
            const uint8_t* base = sJ + (iny - ty0) * kJPitch + (inx - tx0);
            int sb1 = 0, sb2 = 0;
#pragma unroll
            for (int k = 0; k < kPpl; ++k) {
                if (wofs[k] >= 0) {
                    const uint8_t* q = base + (wofs[k] >> 8) * kJPitch + (wofs[k] & 255);
                    const int diff = ((q[0] * w00 + q[1] * w01 + q[kJPitch] * w10 + q[kJPitch + 1] * w11 + (1 << 8)) >> 9) - Iv[k];
                    sb1 += diff * (int)(short)(Ixy[k] & 0xffff);
                    sb2 += diff * (Ixy[k] >> 16);
                }
            }
            const float b1 = (float)warp_sum_ll(sb1) * FLT_SCALE;
            const float b2 = (float)warp_sum_ll(sb2) * FLT_SCALE;
            const float dx = (A12 * b2 - A22 * b1) * D;
            const float dy = (A12 * b1 - A11 * b2) * D;
            nx += dx; ny += dy;
            outx = nx + half; outy = ny + half;
            if ((double)dx * (double)dx + (double)dy * (double)dy <= P.eps2) break;
            if (j > 0 && fabs((double)(dx + pdx)) < 0.01 && fabs((double)(dy + pdy)) < 0.01) {
                outx -= dx * 0.5f; outy -= dy * 0.5f;
                break;
            }
            pdx = dx; pdy = dy;
        }
        if (status && level == 0) {   // error epilogue (err is always requested by the reference)
            const float qx = outx - half, qy = outy - half;
            const int inx = (int)floorf(qx), iny = (int)floorf(qy);
            if (inx < -kWin || inx >= w || iny < -kWin || iny >= h) {
                status = 0;
            } else if (job.err) {
                if (!tile_ok || inx < tx0 || inx + 22 > tx0 + kJPitch || iny < ty0 || iny + 22 > ty0 + kJRows) {
                    tx0 = (inx - 5) & ~3; ty0 = iny - 5;
                    __syncwarp();
                    stage_tile<kJRows, kJPitch, 9>(sJ, imgJ, pitchJ, w, h, tx0, ty0, lane);
                    __syncwarp();
                }
                bilinear_weights(qx - inx, qy - iny, w00, w01, w10, w11);
                const uint8_t* base = sJ + (iny - ty0) * kJPitch + (inx - tx0);
                int se = 0;
#pragma unroll
                for (int k = 0; k < kPpl; ++k) {
                    if (wofs[k] >= 0) {
                        const uint8_t* q = base + (wofs[k] >> 8) * kJPitch + (wofs[k] & 255);
                        const int diff = ((q[0] * w00 + q[1] * w01 + q[kJPitch] * w10 + q[kJPitch + 1] * w11 + (1 << 8)) >> 9) - Iv[k];
                        se += diff < 0 ? -diff : diff;
                    }
                }
                errv = (float)warp_sum_ll(se) / (float)(32 * kWin * kWin);
            }
        }
    }
    if (lane == 0) {
        job.cur_xy[2 * pt] = outx; job.cur_xy[2 * pt + 1] = outy;
        job.status[pt] = (uint8_t)status;
        if (job.err) job.err[pt] = errv;
        if (job.keep) {   // FeatureTracker.cpp:118-126 ; Camera.cpp:120-123, 134-139
            const float W0 = (float)P.w[0], H0 = (float)P.h[0];
            const float vr = outy / H0;
            const bool polar = (vr < P.polar_thr) || (vr > (1.0f - P.polar_thr));
            const bool nearb = (outx < P.margin) || (outx > W0 - P.margin) || (outy < P.margin) || (outy > H0 - P.margin);
            job.keep[pt] = (uint8_t)(status && !polar && !nearb);
        }
    }
}

}  // namespace vio360
