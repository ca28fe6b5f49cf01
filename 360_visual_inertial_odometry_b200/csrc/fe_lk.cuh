// fe_lk.cuh -- K3 (+K2 fused): pyramidal Lucas-Kanade tracker, one warp per feature.
//
// Restates cv::calcOpticalFlowPyrLK's LKTrackerInvoker (OpenCV 4.x
// modules/video/src/lkpyramid.cpp) as called by FeatureTracker::TrackOpticalFlow
// (src/processing/FeatureTracker.cpp:238-240): fixed-point bilinear sampling with 14-bit
// weights, 5 fractional bits on the patch, Scharr derivatives (computed here from a
// 24-row shared-memory tile instead of a materialised derivative image), REFLECT_101
// image apron, zero derivative apron, float 2x2 solve, the eps^2 / oscillation exits and
// the level-0 error epilogue.
//
// Mapping.  The 21x21 window is covered by 126 horizontal runs of 4 pixels (6 runs per
// row, the last three columns of a row are padding with zero derivatives); lane l owns
// runs l, l+32, l+64, l+96.  A run reads two 32-bit words per tile row and realigns them
// with funnel shifts; one dp2a (16-bit weight pair x 8-bit pixel pair) evaluates half of
// a bilinear tap, so a pixel costs two dp2a instead of four multiply-adds.  The
// derivative samples of the template stay in registers for all iterations.  The
// mismatch vector is accumulated as sum(J*Ix) - sum(I*Ix) in exact integers (identical to
// OpenCV's sum((J-I)*Ix)); the 2x2 normal-equation sums are exact 64-bit warp reductions
// built from two 32-bit REDUX per sum (OpenCV accumulates the same integers in float
// lanes; the difference is below 2^-23 relative and is what the 0.05 px position
// tolerance absorbs).  This translation unit is compiled with --fmad=false so float
// expressions round like the non-FMA baseline build of lkpyramid.cpp.
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
    const int* n_ptr;   // when set, the point count is read from device memory (per-frame chain, fe_seq.cuh) and n is ignored
};

struct LkParams {
    int nlevels;
    int w[kMaxLevels], h[kMaxLevels];
    int max_iters;
    double eps2, min_eig;
    float polar_thr, margin;
};

constexpr int kWin = 21;
constexpr int kRunsPerRow = 6;                    // 6 x 4 = 24 >= 21 columns
constexpr int kRuns = kWin * kRunsPerRow;         // 126
constexpr int kRpl = 4;                           // runs per lane
constexpr int kTRows = 24, kTWords = 8;           // template tile: 24 rows x 32 bytes
constexpr int kDRows = 22, kDWords = 24;          // derivative grid: two planes (Ix, Iy) of 22 (+2 zero) rows x 24 ints; a run's
                                                  // four values sit at word 4 * run, so 128-bit accesses are conflict free
constexpr int kDPlane = (kDRows + 2) * kDWords;
constexpr int kJRows = 32, kJWords = 9;           // search tile: 32 rows x 36 bytes
constexpr int kLkWarpWords = kTRows * kTWords + 2 * kDPlane + kJRows * kJWords;   // shared words per feature (warp)

// exact 64-bit warp sum of 32-bit lane partials from two hardware integer reductions
__device__ __forceinline__ long long warp_sum_i32(int v) {
    const int lo = __reduce_add_sync(0xffffffffu, v & 0xffff);
    const int hi = __reduce_add_sync(0xffffffffu, v >> 16);
    return (long long)hi * 65536ll + (long long)lo;
}

// d = c + a.lo16 * b.byte[0|2] + a.hi16 * b.byte[1|3]   (a signed 16-bit pair, b unsigned bytes)
__device__ __forceinline__ int dp2a_lo_su(int a, unsigned b, int c) {
    int d;
    asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ int dp2a_hi_su(int a, unsigned b, int c) {
    int d;
    asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// d = c + sum_i a.byte[i] (unsigned) * b.byte[i] (signed)
__device__ __forceinline__ int dp4a_us(unsigned a, int b, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Tile staging.  A (ROWS x WORDS) tile whose top-left byte is image pixel (x0, y0), x0 % 4 == 0.
// Interior tiles use aligned 32-bit loads held in registers until the warp may overwrite the
// shared tile; tiles touching the border are rare and go through one out-of-line routine that
// indexes with REFLECT_101 byte by byte (the pyramid apron of buildOpticalFlowPyramid).
template <int ROWS, int WORDS>
__device__ __forceinline__ bool tile_interior(const uint8_t* img, int pitch, int w, int h, int x0, int y0) {
    return x0 >= 0 && y0 >= 0 && (x0 + 4 * WORDS) <= w && (y0 + ROWS) <= h && ((pitch & 3) == 0) &&
           ((reinterpret_cast<uintptr_t>(img) & 3) == 0);
}
template <int WORDS>
__device__ __forceinline__ uint32_t tile_word_fast(const uint8_t* __restrict__ img, int pitch, int x0, int y0, int it) {
    const int r = it / WORDS, c = it - r * WORDS;
    return __ldg(reinterpret_cast<const uint32_t*>(img + (size_t)(y0 + r) * pitch + x0) + c);
}
// Asynchronous (LDGSTS) fill of an interior tile: no registers are held while the loads fly.
template <int ROWS, int WORDS>
__device__ __forceinline__ void tile_fill_async(uint32_t* s, const uint8_t* __restrict__ img, int pitch, int x0, int y0, int lane) {
#pragma unroll
    for (int i = 0; i < (ROWS * WORDS) / 32; ++i) {
        const int it = lane + 32 * i, r = it / WORDS, c = it - r * WORDS;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s + it);
        const uint8_t* src = img + (size_t)(y0 + r) * pitch + x0 + 4 * c;
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __noinline__ void tile_fill_border(uint32_t* s, const uint8_t* __restrict__ img, int pitch, int w, int h,
                                              int x0, int y0, int words, int total, int lane) {
#pragma unroll 1
    for (int it = lane; it < total; it += 32) {
        const int r = it / words, c = it - r * words;
        const uint8_t* row = img + (size_t)reflect101(y0 + r, h) * pitch;
        const int x = x0 + 4 * c;
        s[it] = (uint32_t)row[reflect101(x, w)] | ((uint32_t)row[reflect101(x + 1, w)] << 8) |
                ((uint32_t)row[reflect101(x + 2, w)] << 16) | ((uint32_t)row[reflect101(x + 3, w)] << 24);
    }
}

__device__ __forceinline__ void bilinear_weights(float a, float b, int& wt, int& wb) {
    const int w00 = __float2int_rn((1.f - a) * (1.f - b) * 16384.f);
    const int w01 = __float2int_rn(a * (1.f - b) * 16384.f);
    const int w10 = __float2int_rn((1.f - a) * b * 16384.f);
    const int w11 = 16384 - w00 - w01 - w10;
    wt = (w00 & 0xffff) | (w01 << 16);   // (w00, w01) as a signed 16-bit pair
    wb = (w10 & 0xffff) | (w11 << 16);
}

// The four bilinear patch values (5 fractional bits) of a run whose first tap is byte
// (4 * k + s) of tile row `top`; `bot` is the row below.
__device__ __forceinline__ void run_sample(const uint32_t* __restrict__ top, const uint32_t* __restrict__ bot, int s8,
                                           int wt, int wb, int* v) {
    const uint32_t t0 = top[0], t1 = top[1], b0 = bot[0], b1 = bot[1];
    const uint32_t ta = __funnelshift_rc(t0, t1, s8), tb = __funnelshift_rc(t0, t1, s8 + 8);
    const uint32_t ba = __funnelshift_rc(b0, b1, s8), bb = __funnelshift_rc(b0, b1, s8 + 8);
    v[0] = dp2a_lo_su(wt, ta, dp2a_lo_su(wb, ba, 256)) >> 9;
    v[1] = dp2a_lo_su(wt, tb, dp2a_lo_su(wb, bb, 256)) >> 9;
    v[2] = dp2a_hi_su(wt, ta, dp2a_hi_su(wb, ba, 256)) >> 9;
    v[3] = dp2a_hi_su(wt, tb, dp2a_hi_su(wb, bb, 256)) >> 9;
}

// Derivatives outside the image are zero (the zero apron of calcScharrDeriv's buffer); only
// windows that leave the image need this pass.
__device__ __noinline__ void deriv_zero_outside(int* sDx, int* sDy, int ipx, int ipy, int w, int h, int lane) {
#pragma unroll 1
    for (int i = lane; i < kDRows * kDRows; i += 32) {
        const int dy = i / kDRows, dx = i - dy * kDRows;
        const unsigned gx = (unsigned)(ipx + dx), gy = (unsigned)(ipy + dy);
        if (gx >= (unsigned)w || gy >= (unsigned)h) { sDx[dy * kDWords + dx] = 0; sDy[dy * kDWords + dx] = 0; }
    }
}

// kLkWarps warps per CTA; REGS = register cap (8 warps x 128 regs -> 16 warps / SM, 7 x 96 -> 21, 6 x 80 -> 24).
// PERSIST = false: warp w of CTA (x, y) tracks feature x * kLkWarps + w of job y and exits.  The iteration count of
// a feature is data dependent (13 on average, 30 at most per level), and a CTA keeps its registers and shared memory
// until its slowest warp is done: with 8-warp CTAs a quarter of the warp slots of an SM idle (ncu: 11.9 of 16 warps
// active).  PERSIST = true: the grid is exactly the warps that fit the device, and every warp takes the next feature
// of the batch from a device counter until none is left, so a slot is never empty before the batch ends.  Work item
// i = (job i / max_n, feature i % max_n); consecutive items are consecutive features of one frame (L2 locality).
// `queue`: [0] next item, [1] warps that have finished; the last warp to finish resets both for the next launch.
template <int kLkWarps, int REGS, bool PERSIST>
__global__ void __launch_bounds__(kLkWarps * 32) __maxnreg__(REGS) k_lk_track(const TrackJobDev* __restrict__ jobs, LkParams P, int max_n, int n_items,
                                                                            unsigned* __restrict__ queue) {
    extern __shared__ __align__(16) uint32_t lk_smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t* sT = lk_smem + warp * kLkWarpWords;
    int* sDx = reinterpret_cast<int*>(sT + kTRows * kTWords);
    int* sDy = sDx + kDPlane;
    uint32_t* sJ = reinterpret_cast<uint32_t*>(sDy + kDPlane);
    // runs owned by this lane: window row wy, column segment seg (columns 4*seg .. 4*seg+3);
    // offJ = wy * kJWords + seg is the only per-run register kept across the level loop
    int offJ[kRpl];
    unsigned colmask = 0;   // bit (4k + j): pixel j of run k is a real window pixel
#pragma unroll
    for (int k = 0; k < kRpl; ++k) {
        const int r = lane + 32 * k;
        const bool real = r < kRuns;
        const int rr = real ? r : kRuns - 1;
        const int wy = rr / kRunsPerRow, seg = rr - wy * kRunsPerRow;
        offJ[k] = wy * kJWords + seg;
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (real && 4 * seg + j < kWin) colmask |= 1u << (4 * k + j);
    }
  for (;;) {
    int job_idx, pt;
    if (PERSIST) {
        unsigned item = 0;
        if (lane == 0) item = atomicAdd(queue, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= (unsigned)n_items) break;
        job_idx = (int)(item / (unsigned)max_n); pt = (int)(item - (unsigned)job_idx * (unsigned)max_n);
    } else {
        job_idx = blockIdx.y; pt = blockIdx.x * (int)(blockDim.x >> 5) + warp;   // launched with <= kLkWarps warps per CTA
    }
    const TrackJobDev& job = jobs[job_idx];
    if (pt >= (job.n_ptr ? *job.n_ptr : job.n)) {   // warp-uniform; only __syncwarp is used below
        if (PERSIST) continue;
        return;
    }

#define LK_OFF_T(k) ((offJ[k] / kJWords + 1) * kTWords + offJ[k] % kJWords)
// padding runs (k == 3 on lanes 30, 31) read the two zero rows below the derivative grid
#define LK_OFF_D(k) (((k) == 3 && lane >= 30) ? kDRows * kDWords : (offJ[k] / kJWords) * kDWords + 4 * (offJ[k] % kJWords))

    const float px0 = job.prev_xy[2 * pt], py0 = job.prev_xy[2 * pt + 1];
    const float half = (kWin - 1) * 0.5f;
    float outx = 0.f, outy = 0.f;   // nextPts[ptidx]
    int status = 1;
    float errv = 0.f;
    bool pfT = false;   // the next level's template tile is already in flight into sT

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
        nx -= half; ny -= half;
        const uint8_t* imgI = job.prev.p[level];
        const int pitchI = job.prev.pitch[level];
        const uint8_t* imgJ = job.cur.p[level];
        const int pitchJ = job.cur.pitch[level];
        // ---- one global round trip: template tile and the first search tile together
        const int xT0 = (ipx - 1) & ~3, yT0 = ipy - 1;
        const int oT = ipx - xT0;                  // byte column of window pixel 0 in the tile (1..4)
        int tx0 = 0, ty0 = 0;
        bool tile_ok = false;
        {
            // template tile: prefetched during the previous level's iterations when interior
            const bool haveT = pfT;
            pfT = false;
            const int inx = (int)floorf(nx), iny = (int)floorf(ny);
            const bool jin = !(inx < -kWin || inx >= w || iny < -kWin || iny >= h);
            __syncwarp();   // every lane is done with the previous level's tiles
            if (!haveT) {
                if (tile_interior<kTRows, kTWords>(imgI, pitchI, w, h, xT0, yT0)) tile_fill_async<kTRows, kTWords>(sT, imgI, pitchI, xT0, yT0, lane);
                else tile_fill_border(sT, imgI, pitchI, w, h, xT0, yT0, kTWords, kTRows * kTWords, lane);
            }
            // search tile: issued now, needed only after the Scharr / template phase
            bool jAsync = false;
            if (jin) {
                tx0 = (inx - 4) & ~3; ty0 = iny - 5;
                jAsync = tile_interior<kJRows, kJWords>(imgJ, pitchJ, w, h, tx0, ty0);
                if (jAsync) tile_fill_async<kJRows, kJWords>(sJ, imgJ, pitchJ, tx0, ty0, lane);
                else tile_fill_border(sJ, imgJ, pitchJ, w, h, tx0, ty0, kJWords, kJRows * kJWords, lane);
            }
            if (jAsync) cp_async_wait<1>();   // everything but the search tile (committed last) has landed
            else cp_async_wait<0>();
            tile_ok = jin;
            __syncwarp();
        }
        int wtI, wbI;   // template weights, kept for the error epilogue
        bilinear_weights(px - ipx, py - ipy, wtI, wbI);
        int Ix[kRpl * 4], Iy[kRpl * 4];
        int sA11 = 0, sA12 = 0, sA22 = 0, cI1 = 0, cI2 = 0;
        // A window whose 22 x 22 derivative taps all lie inside the image (the usual case) needs no zero
        // apron, and then interpolation and Scharr commute exactly: sum_k iw_k * Scharr(I)(p_k) ==
        // Scharr(sum_k iw_k * I(. + k)) in integers (no intermediate rounding on either side).  So the
        // bilinear sums Iw (23 bits) are formed once on a 23 x 23 grid, the template value is
        // (Iw + 256) >> 9 and the derivatives are (Scharr(Iw) + 2^13) >> 14 -- one pass instead of a
        // derivative grid plus a four-tap interpolation of each derivative.
        const bool border_window = ipx < 0 || ipy < 0 || ipx + kDRows > w || ipy + kDRows > h;   // warp-uniform, rare
        if (!border_window) {
            int* G = sDx;   // [24 rows][24 cols]: G[r][c] = Iw at window offset (c - 1, r - 1)
            {
                const int cr = lane % 6, g = lane / 6;        // g 0..4 work, lanes 30, 31 idle
                if (g < 5) {
                    const bool full = g < 4;                  // grid rows 5g .. 5g+4; g == 4 has rows 20, 21, 22 only
                    const int r0 = g * 5;
                    const int b0 = oT - 1 + 4 * cr;           // tile byte of grid column 4 * cr
                    const uint32_t* q = sT + r0 * kTWords + (b0 >> 2);
                    const int s8 = 8 * (b0 & 3);
                    uint32_t ta = __funnelshift_rc(q[0], q[1], s8), tb = __funnelshift_rc(q[0], q[1], s8 + 8);
                    int* o = G + r0 * kDWords + 4 * cr;
#pragma unroll
                    for (int i = 0; i < 5; ++i) {
                        if (i < 3 || full) {
                            q += kTWords;
                            const uint32_t ba = __funnelshift_rc(q[0], q[1], s8), bb = __funnelshift_rc(q[0], q[1], s8 + 8);
                            *reinterpret_cast<int4*>(o) = make_int4(dp2a_lo_su(wtI, ta, dp2a_lo_su(wbI, ba, 0)), dp2a_lo_su(wtI, tb, dp2a_lo_su(wbI, bb, 0)),
                                                                    dp2a_hi_su(wtI, ta, dp2a_hi_su(wbI, ba, 0)), dp2a_hi_su(wtI, tb, dp2a_hi_su(wbI, bb, 0)));
                            o += kDWords;
                            ta = ba; tb = bb;
                        }
                    }
                }
            }
            __syncwarp();
#pragma unroll
            for (int k = 0; k < kRpl; ++k) {
                // rows y-1, y, y+1 and columns x0-1 .. x0+4 of the window run (x0 = 4 * seg, y = wy)
                const int* gp = G + (((k) == 3 && lane >= 30) ? 0 : (offJ[k] / kJWords) * kDWords + 4 * (offJ[k] % kJWords));
                const int4 a4 = *reinterpret_cast<const int4*>(gp), b4 = *reinterpret_cast<const int4*>(gp + kDWords),
                           c4 = *reinterpret_cast<const int4*>(gp + 2 * kDWords);
                const int2 a2 = *reinterpret_cast<const int2*>(gp + 4), b2 = *reinterpret_cast<const int2*>(gp + kDWords + 4),
                           c2 = *reinterpret_cast<const int2*>(gp + 2 * kDWords + 4);
                const int A[6] = {a4.x, a4.y, a4.z, a4.w, a2.x, a2.y}, B[6] = {b4.x, b4.y, b4.z, b4.w, b2.x, b2.y},
                          Cc[6] = {c4.x, c4.y, c4.z, c4.w, c2.x, c2.y};
                int sv[6], dv[6];
#pragma unroll
                for (int c = 0; c < 6; ++c) { sv[c] = 3 * (A[c] + Cc[c]) + 10 * B[c]; dv[c] = Cc[c] - A[c]; }
                const bool pad_run = k == 3 && lane >= 30, seg5 = offJ[k] % kJWords == 5;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    int ixv = (sv[j + 2] - sv[j] + (1 << 13)) >> 14;
                    int iyv = (3 * (dv[j] + dv[j + 2]) + 10 * dv[j + 1] + (1 << 13)) >> 14;
                    const int ival = (B[j + 1] + 256) >> 9;
                    if (pad_run || (seg5 && j >= 1)) { ixv = 0; iyv = 0; }   // padding pixels (window columns 21..23)
                    Ix[4 * k + j] = ixv; Iy[4 * k + j] = iyv;
                    sA11 += ixv * ixv; sA12 += ixv * iyv; sA22 += iyv * iyv;
                    cI1 += ival * ixv; cI2 += ival * iyv;
                }
            }
        } else {
            // Zero what the Scharr stage below never writes: columns 22, 23 of every row (the taps right of the
            // window) and rows 22, 23 (read by the two padding runs of lanes 30 / 31).  Done per border window
            // (rare) and not once per warp: the interior path uses the same words as its 24 x 24 grid.
            if (lane < kDRows) {
                *reinterpret_cast<int2*>(sDx + lane * kDWords + kDRows) = make_int2(0, 0);
                *reinterpret_cast<int2*>(sDy + lane * kDWords + kDRows) = make_int2(0, 0);
            }
            for (int i = lane; i < 2 * kDWords; i += 32) { sDx[kDRows * kDWords + i] = 0; sDy[kDRows * kDWords + i] = 0; }
            // ---- Scharr on the 22 x 22 grid (ipx + dx, ipy + dy): lane = (4-column run cr, row group g).
            // Per tile row one dp4a gives the horizontal smoothing 3a+10b+3c and one the horizontal
            // difference c-a of a column; the vertical halves are combined from a 3-row register window.
            {
                const int cr = lane % 6, g = lane / 6;            // g 0..4 work, lanes 30, 31 idle
                if (g < 5) {
                    const bool full = g < 4;                      // rows 0-4, 5-9, 10-14, 15-19; g == 4 has rows 20, 21 only
                    const int r0 = g * 5;
                    const int b0 = oT - 1 + 4 * cr;               // tile byte of column (dx0 - 1), dx0 = 4 * cr
                    const uint32_t* q = sT + r0 * kTWords + (b0 >> 2);
                    const int s8 = 8 * (b0 & 3);
                    int hA[4], hB[4], hC[4], dA[4], dB[4], dC[4];
    #define LK_ROW_HD(H, Dd)                                                                   \
                    {                                                                          \
                        const uint32_t a0 = __funnelshift_rc(q[0], q[1], s8);                   \
                        const uint32_t a1 = __funnelshift_rc(q[1], q[2], s8);                   \
                        const uint32_t x1 = __funnelshift_r(a0, a1, 8), x2 = __funnelshift_r(a0, a1, 16), \
                                       x3 = __funnelshift_r(a0, a1, 24);                        \
                        H[0] = dp4a_us(a0, 0x00030a03, 0); Dd[0] = dp4a_us(a0, 0x000100ff, 0);  \
                        H[1] = dp4a_us(x1, 0x00030a03, 0); Dd[1] = dp4a_us(x1, 0x000100ff, 0);  \
                        H[2] = dp4a_us(x2, 0x00030a03, 0); Dd[2] = dp4a_us(x2, 0x000100ff, 0);  \
                        H[3] = dp4a_us(x3, 0x00030a03, 0); Dd[3] = dp4a_us(x3, 0x000100ff, 0);  \
                        q += kTWords;                                                           \
                    }
                    // output row (r0 + I) from the window (H0, D0), (H1, D1) and the freshly read row (H2, D2)
    #define LK_ROW_OUT(I, H0, D0, H1, D1, H2, D2)                                              \
                    if ((I) < 2 || full) {                                                     \
                        LK_ROW_HD(H2, D2)                                                      \
                        int ix[4], iy[4];                                                      \
                        _Pragma("unroll") for (int j = 0; j < 4; ++j) {                        \
                            ix[j] = 3 * (D0[j] + D2[j]) + 10 * D1[j];                           \
                            iy[j] = H2[j] - H0[j];                                             \
                        }                                                                      \
                        int* ox = sDx + (r0 + (I)) * kDWords + 4 * cr;                         \
                        int* oy = sDy + (r0 + (I)) * kDWords + 4 * cr;                         \
                        if (cr < 5) {                                                          \
                            *reinterpret_cast<int4*>(ox) = make_int4(ix[0], ix[1], ix[2], ix[3]); \
                            *reinterpret_cast<int4*>(oy) = make_int4(iy[0], iy[1], iy[2], iy[3]); \
                        } else {   /* columns 20, 21; 22 and 23 stay zero */                   \
                            *reinterpret_cast<int2*>(ox) = make_int2(ix[0], ix[1]);            \
                            *reinterpret_cast<int2*>(oy) = make_int2(iy[0], iy[1]);            \
                        }                                                                      \
                    }
                    LK_ROW_HD(hA, dA)
                    LK_ROW_HD(hB, dB)
                    LK_ROW_OUT(0, hA, dA, hB, dB, hC, dC)
                    LK_ROW_OUT(1, hB, dB, hC, dC, hA, dA)
                    LK_ROW_OUT(2, hC, dC, hA, dA, hB, dB)
                    LK_ROW_OUT(3, hA, dA, hB, dB, hC, dC)
                    LK_ROW_OUT(4, hB, dB, hC, dC, hA, dA)
    #undef LK_ROW_OUT
    #undef LK_ROW_HD
                }
            }
            __syncwarp();
            if (ipx < 0 || ipy < 0 || ipx + kDRows > w || ipy + kDRows > h) {   // warp-uniform, rare
                deriv_zero_outside(sDx, sDy, ipx, ipy, w, h, lane);
                __syncwarp();
            }
            // ---- template samples: I (only for sum I*Ix, sum I*Iy), Ix, Iy per owned pixel
            const int w00 = (int)(short)(wtI & 0xffff), w01 = wtI >> 16, w10 = (int)(short)(wbI & 0xffff), w11 = wbI >> 16;
            {
                const int kT = oT >> 2, sT8 = 8 * (oT & 3);
    #pragma unroll
                for (int k = 0; k < kRpl; ++k) {
                    int iv[4];
                    run_sample(sT + LK_OFF_T(k) + kT, sT + LK_OFF_T(k) + kT + kTWords, sT8, wtI, wbI, iv);
                    const int oD = LK_OFF_D(k);
                    const int4 xa = *reinterpret_cast<const int4*>(sDx + oD), xb = *reinterpret_cast<const int4*>(sDx + oD + kDWords);
                    const int4 ya = *reinterpret_cast<const int4*>(sDy + oD), yb = *reinterpret_cast<const int4*>(sDy + oD + kDWords);
                    // the fifth tap of a run in the last column segment would be column 24 = the next row's
                    // column 0; that tap belongs to padding pixel 23 and must read zero
                    const bool seg5 = offJ[k] % kJWords == 5;
                    const int tx[5] = {xa.x, xa.y, xa.z, xa.w, seg5 ? 0 : sDx[oD + 4]}, bx[5] = {xb.x, xb.y, xb.z, xb.w, seg5 ? 0 : sDx[oD + kDWords + 4]};
                    const int ty[5] = {ya.x, ya.y, ya.z, ya.w, seg5 ? 0 : sDy[oD + 4]}, by[5] = {yb.x, yb.y, yb.z, yb.w, seg5 ? 0 : sDy[oD + kDWords + 4]};
    #pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        int ixv = (tx[j] * w00 + tx[j + 1] * w01 + bx[j] * w10 + bx[j + 1] * w11 + (1 << 13)) >> 14;
                        int iyv = (ty[j] * w00 + ty[j + 1] * w01 + by[j] * w10 + by[j + 1] * w11 + (1 << 13)) >> 14;
                        // window column 21 (padding) still sees derivative column 21 on its left tap;
                        // columns 22, 23 and the padding runs only see zeros
                        if (j == 1 && offJ[k] % kJWords == 5) { ixv = 0; iyv = 0; }
                        Ix[4 * k + j] = ixv; Iy[4 * k + j] = iyv;
                        sA11 += ixv * ixv; sA12 += ixv * iyv; sA22 += iyv * iyv;
                        cI1 += iv[j] * ixv; cI2 += iv[j] * iyv;
                    }
                }
            }
        }
        const float FLT_SCALE = 1.f / (1 << 20);
        const float A11 = (float)warp_sum_i32(sA11) * FLT_SCALE;
        const float A12 = (float)warp_sum_i32(sA12) * FLT_SCALE;
        const float A22 = (float)warp_sum_i32(sA22) * FLT_SCALE;
        // The template tile is dead from here on (the error epilogue re-reads it at level 0 only):
        // start the next level's template tile, whose position is known, behind the iterations.
        if (level > 0) {
            const int ln = level - 1, wn = P.w[ln], hn = P.h[ln];
            const float scn = (float)(1. / (1 << ln));
            const int ipxn = (int)floorf(px0 * scn - half), ipyn = (int)floorf(py0 * scn - half);
            if (!(ipxn < -kWin || ipxn >= wn || ipyn < -kWin || ipyn >= hn)) {
                const int xTn = (ipxn - 1) & ~3, yTn = ipyn - 1;
                if (tile_interior<kTRows, kTWords>(job.prev.p[ln], job.prev.pitch[ln], wn, hn, xTn, yTn)) {
                    tile_fill_async<kTRows, kTWords>(sT, job.prev.p[ln], job.prev.pitch[ln], xTn, yTn, lane);
                    pfT = true;
                }
            }
        }
        if (pfT) cp_async_wait<1>(); else cp_async_wait<0>();   // the search tile of this level has landed
        __syncwarp();
        float D = A11 * A22 - A12 * A12;
        const float tt = A11 - A22;
        const float minEig = (A22 + A11 - sqrtf(tt * tt + 4.f * A12 * A12)) / (float)(2 * kWin * kWin);
        if ((double)minEig < P.min_eig || D < 1.1920928955078125e-07f) {
            if (level == 0) status = 0;
            continue;
        }
        D = 1.f / D;
        float pdx = 0.f, pdy = 0.f;
        // (re)load the search tile around (inx, iny) when the window leaves the staged one
        auto ensure_tile = [&](int inx, int iny) {
            if (!tile_ok || inx < tx0 || inx + 25 > tx0 + 4 * kJWords || iny < ty0 || iny + 22 > ty0 + kJRows) {
                tx0 = (inx - 4) & ~3; ty0 = iny - 5;
                if (tile_interior<kJRows, kJWords>(imgJ, pitchJ, w, h, tx0, ty0)) {
                    uint32_t jw[9];
#pragma unroll
                    for (int i = 0; i < 9; ++i) jw[i] = tile_word_fast<kJWords>(imgJ, pitchJ, tx0, ty0, lane + 32 * i);
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 9; ++i) sJ[lane + 32 * i] = jw[i];
                } else {
                    __syncwarp();
                    tile_fill_border(sJ, imgJ, pitchJ, w, h, tx0, ty0, kJWords, kJRows * kJWords, lane);
                }
                __syncwarp();
                tile_ok = true;
            }
        };
        for (int j = 0; j < P.max_iters; ++j) {
            const int inx = (int)floorf(nx), iny = (int)floorf(ny);
            if (inx < -kWin || inx >= w || iny < -kWin || iny >= h) {
                if (level == 0) status = 0;
                break;
            }
            ensure_tile(inx, iny);
            int wt, wb;
            bilinear_weights(nx - inx, ny - iny, wt, wb);
            const int ox = inx - tx0;
            const uint32_t* base = sJ + (iny - ty0) * kJWords + (ox >> 2);
            const int s8 = 8 * (ox & 3);
            int sb1 = -cI1, sb2 = -cI2;
#pragma unroll
            for (int k = 0; k < kRpl; ++k) {
                int jv[4];
                run_sample(base + offJ[k], base + offJ[k] + kJWords, s8, wt, wb, jv);
#pragma unroll
                for (int q = 0; q < 4; ++q) { sb1 += jv[q] * Ix[4 * k + q]; sb2 += jv[q] * Iy[4 * k + q]; }
            }
            const float b1 = (float)warp_sum_i32(sb1) * FLT_SCALE;
            const float b2 = (float)warp_sum_i32(sb2) * FLT_SCALE;
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
                ensure_tile(inx, iny);
                int wtJ, wbJ;
                bilinear_weights(qx - inx, qy - iny, wtJ, wbJ);
                const int ox = inx - tx0;
                const uint32_t* base = sJ + (iny - ty0) * kJWords + (ox >> 2);
                const int s8 = 8 * (ox & 3);
                const int kT = oT >> 2, sT8 = 8 * (oT & 3);
                int se = 0;
#pragma unroll
                for (int k = 0; k < kRpl; ++k) {
                    int jv[4], iv[4];
                    run_sample(base + offJ[k], base + offJ[k] + kJWords, s8, wtJ, wbJ, jv);
                    run_sample(sT + LK_OFF_T(k) + kT, sT + LK_OFF_T(k) + kT + kTWords, sT8, wtI, wbI, iv);
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if ((colmask >> (4 * k + q)) & 1u) { const int d = jv[q] - iv[q]; se += d < 0 ? -d : d; }
                }
                errv = (float)warp_sum_i32(se) / (float)(32 * kWin * kWin);
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
    if (!PERSIST) break;
    __syncwarp();   // the next feature's tiles overwrite this one's
  }
    if (PERSIST && lane == 0) {
        __threadfence();
        const unsigned total = gridDim.x * kLkWarps;
        if (atomicAdd(queue + 1, 1u) == total - 1) { queue[0] = 0; queue[1] = 0; __threadfence(); }
    }
}

}  // namespace vio360
