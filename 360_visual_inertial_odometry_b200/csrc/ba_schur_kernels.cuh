// ba_schur_kernels.cuh -- included by ba.cu after BaDev / warp_sum / block_sum_to.
// See ba_schur.cuh for the algorithm and the reference code this replaces.
#pragma once

#include "ba_schur.cuh"

namespace vio360 {

// upper-triangle index t (0..20) of a symmetric 6x6 -> (a, b), a <= b, row-major
__device__ __forceinline__ void tri6(int t, int& a, int& b) {
    a = 0;
    int rem = t;
    while (rem >= 6 - a) { rem -= 6 - a; ++a; }
    b = a + rem;
}

__device__ __forceinline__ void named_bar_sync(int id, int count) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int count) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}
constexpr int kBarProd = 1, kBarFull = 2, kBarEmpty = 4;   // ids: producers-only, full[2], empty[2]

// PT = double: the parity path (1e-9).  PT = float: the north star's optional fp32 mode -- phase A (projection, atan2 /
// asin, Jacobians, Huber) runs in float on the FP32 pipe, which the DMMA bursts of the consumers do not occupy; its
// results are widened and phases B / C / D (sums, 3x3 Cholesky, F'F, Z, the SYRK) stay in double.
template <typename PT>
__global__ void __launch_bounds__(kSThreads, 1)
k_schur_tile(BaDev d, const SchurChunk* __restrict__ chunks, const int* __restrict__ slots, const int* __restrict__ batch_rng,
             const int* __restrict__ batch_lm, const double* __restrict__ wide_inv, bool apply_loss,
             double* __restrict__ partS,
             double* __restrict__ partRhs, double* __restrict__ partCost) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* Zst = reinterpret_cast<double*>(smem_raw);          // [2][kSK][kSZPitch] operand stages, Z^T: column-major in k
    double* St = Zst + 2 * kSK * kSZPitch;                      // [kSTiles][64] S tile, DMMA fragment order
    double* sPose = St + kSTiles * 64;                          // [kSWin][kPoseStride]
    double* sLmk = sPose + kSWin * kPoseStride;                 // [kSMaxLm][9]: L^-1 (6), (E'E)^-1 g (3)
    double* sBlk = sLmk + kSMaxLm * 9;                          // [32] block-sum scratch
    float* sPoseF = reinterpret_cast<float*>(reinterpret_cast<int*>(sBlk + 32) + 32);   // [kSWin][kPoseStride] (PT = float only)
    pdl_launch_dependents();   // the slab reduce may be scheduled as soon as SMs free up; it waits for this grid before reading
    const SchurChunk ch = chunks[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kSTiles * 64; i += kSThreads) St[i] = 0.0;
    for (int i = tid; i < ch.span * kPoseStride; i += kSThreads) {
        const double v = d.posebuf[(size_t)ch.pbase * kPoseStride + i];
        sPose[i] = v;
        if (sizeof(PT) == 4) sPoseF[i] = (float)v;
    }
    for (int i = tid; i < ch.spanB * kPoseStride; i += kSThreads) {   // second pose group of a cross chunk: slots kSGrp ..
        const double v = d.posebuf[(size_t)ch.pbaseB * kPoseStride + i];
        sPose[kSGrp * kPoseStride + i] = v;
        if (sizeof(PT) == 4) sPoseF[kSGrp * kPoseStride + i] = (float)v;
    }
    const bool wide = (ch.flags & kChunkWide) != 0, cross = (ch.flags & kChunkCross) != 0;
    __syncthreads();
    const int nbatch = (ch.lm_end - ch.lm_begin + kSMaxLm - 1) / kSMaxLm;
    double pacc0 = 0.0, pacc1 = 0.0;   // producer lane j of pose slot w: entries 2j, 2j+1 of [F'F tri (21) | F'(r - E ig) (6)]
    double cost = 0.0;

    if (tid < kSProd) {
        // ================= producers: thread = (pose slot w, landmark slot ls) =================
        const int w = tid >> 4, ls = tid & 15;
        const int* myslot = slots + (size_t)ch.batch0 * kSProd + tid;
        const int* mylm = batch_lm + (size_t)ch.batch0 * kSMaxLm;    // landmark ids of the chunk's batches
        int oi_next = nbatch > 0 ? __ldg(myslot) : -1;   // observation of slot (w, ls) in batch 0
        bool pose_const = true;
        if (w < ch.span) pose_const = d.free_map[ch.pbase + w] < 0;
        else if (w >= kSGrp && w - kSGrp < ch.spanB) pose_const = d.free_map[ch.pbaseB + w - kSGrp] < 0;
        // observation data of the batch about to be evaluated: fetched one batch ahead (after
        // phase A of the previous batch), so that no global-memory latency sits in phase A
        double nu = 0.0, nv = 0.0, nP0 = 0.0, nP1 = 0.0, nP2 = 0.0;
        bool nout = false;
        if (oi_next >= 0) {
            const size_t l = (size_t)__ldg(mylm + ls);
            nu = __ldg(d.u + oi_next); nv = __ldg(d.v + oi_next); nout = __ldg(d.outlier + oi_next) != 0;
            nP0 = __ldg(d.points + 3 * l); nP1 = __ldg(d.points + 3 * l + 1); nP2 = __ldg(d.points + 3 * l + 2);
        }
        for (int b = 0; b < nbatch; ++b) {
            const int cur = ch.lm_begin + b * kSMaxLm;
            const int NB = min(kSMaxLm, ch.lm_end - cur);
            double* Zs = Zst + (b & 1) * kSK * kSZPitch;
            double* sRed = Zs;                                  // alias: [kSProd][kSRed]
            const int oi = oi_next;
            if (b + 1 < nbatch) oi_next = __ldg(myslot + (size_t)(b + 1) * kSProd);
            // phase B inputs, one value per lane (e = 9: constant flag, 10..12: LM diagonal of the
            // landmark), requested here and first used after phase A
            double pre = 0.0;
            if (tid < kSMaxLm * 16) {
                const int bl = tid >> 4, e = tid & 15;
                if (bl < NB) {
                    const int lid = __ldg(mylm + (size_t)b * kSMaxLm + bl);
                    if (wide) { if (e < 9) pre = __ldg(wide_inv + 9 * (size_t)lid + e); }   // L^-1 (6), (E'E + D^2)^-1 g (3) of the WHOLE landmark
                    else if (e == 9) pre = __ldg(d.point_fixed + lid) != 0 ? 1.0 : 0.0;
                    else if (e >= 10 && e < 13 && d.D) pre = __ldg(d.D + 6 * (size_t)d.K + 3 * (size_t)lid + (e - 10));
                } else if (e == 9) pre = 1.0;
            }
            if (b >= 2) named_bar_sync(kBarEmpty + (b & 1), kSThreads);   // consumers are done with this stage
            // ---- phase A
            double r[2] = {0.0, 0.0}, F[12], E[6];
#pragma unroll
            for (int j = 0; j < 12; ++j) F[j] = 0.0;
#pragma unroll
            for (int j = 0; j < 6; ++j) E[j] = 0.0;
            if (oi >= 0) {
                if (sizeof(PT) == 4) {
                    const float Pwf[3] = {(float)nP0, (float)nP1, (float)nP2};
                    float rf[2], Ff[12], Ef[6];
                    cost += (double)obs_evaluate<true>(sPoseF + kPoseStride * w, Pwf, (float)nu, (float)nv, nout, d.cam, apply_loss, rf, Ff, Ef);
                    r[0] = rf[0]; r[1] = rf[1];
#pragma unroll
                    for (int j = 0; j < 12; ++j) F[j] = Ff[j];
#pragma unroll
                    for (int j = 0; j < 6; ++j) E[j] = Ef[j];
                } else {
                    const double Pw[3] = {nP0, nP1, nP2};
                    cost += obs_evaluate<true>(sPose + kPoseStride * w, Pw, nu, nv, nout, d.cam, apply_loss, r, F, E);
                }
                if (pose_const) {   // constant pose: no F block (its E block still enters E'E)
#pragma unroll
                    for (int j = 0; j < 12; ++j) F[j] = 0.0;
                }
            }
            if (b + 1 < nbatch && oi_next >= 0) {
                const size_t l = (size_t)__ldg(mylm + (size_t)(b + 1) * kSMaxLm + ls);
                nu = __ldg(d.u + oi_next); nv = __ldg(d.v + oi_next); nout = __ldg(d.outlier + oi_next) != 0;
                nP0 = __ldg(d.points + 3 * l); nP1 = __ldg(d.points + 3 * l + 1); nP2 = __ldg(d.points + 3 * l + 2);
            }
            {
                double* q = sRed + tid * kSRed;
                q[0] = E[0] * E[0] + E[3] * E[3]; q[1] = E[0] * E[1] + E[3] * E[4]; q[2] = E[0] * E[2] + E[3] * E[5];
                q[3] = E[1] * E[1] + E[4] * E[4]; q[4] = E[1] * E[2] + E[4] * E[5]; q[5] = E[2] * E[2] + E[5] * E[5];
                q[6] = E[0] * r[0] + E[3] * r[1]; q[7] = E[1] * r[0] + E[4] * r[1]; q[8] = E[2] * r[0] + E[5] * r[1];
            }
            named_bar_sync(kBarProd, kSProd);
            // ---- phase B: per-landmark sums over the pose slots, Cholesky, L^-1, (E'E)^-1 g.
            // thread (bl, e) = tid>>4, tid&15 sums entry e < 9 of landmark bl; the nine sums of a
            // landmark sit in one half warp and are gathered by shuffles, lane e == 0 factorises.
            if (wide) {
                // pass 1 (k_wide_prepare) factorised E'E + D^2 over ALL observations of the landmark: lanes e < 9 hold it
                if (tid < kSMaxLm * 16 && (tid & 15) < 9) sLmk[(tid >> 4) * 9 + (tid & 15)] = pre;
            } else if (tid < kSMaxLm * 16) {   // 8 whole warps
                const int bl = tid >> 4, e = tid & 15;
                double sum = 0.0;
                if (e < 9) {   // four independent partial sums: a 6-deep instead of a 24-deep add chain
                    double s4[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
                    for (int pw = 0; pw < kSWin; ++pw) s4[pw & 3] += sRed[(pw * kSMaxLm + bl) * kSRed + e];
                    sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
                }
                const int base = lane & 16;
                double hq[9];
#pragma unroll
                for (int j = 0; j < 9; ++j) hq[j] = __shfl_sync(0xffffffffu, sum, base + j);
                const double lm_const = __shfl_sync(0xffffffffu, pre, base + 9);
                const double Dl0 = __shfl_sync(0xffffffffu, pre, base + 10), Dl1 = __shfl_sync(0xffffffffu, pre, base + 11),
                             Dl2 = __shfl_sync(0xffffffffu, pre, base + 12);
                if (e == 0) {
                    double* q = sLmk + bl * 9;
                    double li[6] = {0, 0, 0, 0, 0, 0}, ig[3] = {0, 0, 0};
                    if (lm_const == 0.0) {
                        hq[0] += Dl0 * Dl0; hq[3] += Dl1 * Dl1; hq[5] += Dl2 * Dl2;   // zeros without an LM diagonal
                        // E'E = L L': only L^-1 is needed (reciprocal square roots, no divisions)
                        const double i00 = vio_rsqrt(hq[0]);
                        const double l10 = hq[1] * i00, l20 = hq[2] * i00;
                        const double i11 = vio_rsqrt(hq[3] - l10 * l10);
                        const double l21 = (hq[4] - l20 * l10) * i11;
                        const double i22 = vio_rsqrt(hq[5] - l20 * l20 - l21 * l21);
                        const double i10 = -l10 * i00 * i11;
                        const double i21 = -l21 * i11 * i22;
                        const double i20 = -(l20 * i00 + l21 * i10) * i22;
                        li[0] = i00; li[1] = i10; li[2] = i11; li[3] = i20; li[4] = i21; li[5] = i22;
                        bool ok = true;
#pragma unroll
                        for (int j = 0; j < 6; ++j) ok = ok && isfinite(li[j]);
                        if (!ok) {   // rank-deficient landmark: not eliminated (as the general kernel)
#pragma unroll
                            for (int j = 0; j < 6; ++j) li[j] = 0.0;
                        }
                        // (E'E)^-1 g = L^-T (L^-1 g)
                        const double g0 = hq[6], g1 = hq[7], g2 = hq[8];
                        const double y0 = li[0] * g0, y1 = li[1] * g0 + li[2] * g1, y2 = li[3] * g0 + li[4] * g1 + li[5] * g2;
                        ig[0] = li[0] * y0 + li[1] * y1 + li[3] * y2;
                        ig[1] = li[2] * y1 + li[4] * y2;
                        ig[2] = li[5] * y2;
                    }
#pragma unroll
                    for (int j = 0; j < 6; ++j) q[j] = li[j];
                    q[6] = ig[0]; q[7] = ig[1]; q[8] = ig[2];
                }
            }
            named_bar_sync(kBarProd, kSProd);
            // ---- phase C
            const double* q = sLmk + ls * 9;
            const double s0 = r[0] - (E[0] * q[6] + E[1] * q[7] + E[2] * q[8]);
            const double s1 = r[1] - (E[3] * q[6] + E[4] * q[7] + E[5] * q[8]);
            // per-pose sums over the 16 landmark lanes of this half warp: recursive-halving
            // reduce-scatter of the 32 (27 used) entries; lane j ends with entries 2j, 2j+1
            {
                constexpr int TA[32] = {0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 4, 4, 5, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
                constexpr int TB[32] = {0, 1, 2, 3, 4, 5, 1, 2, 3, 4, 5, 2, 3, 4, 5, 3, 4, 5, 4, 5, 5, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
                double x[16];
                const bool up8 = (lane & 8) != 0;
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    double lo, hi;   // entries i and i + 16 of this thread (compile-time indices)
                    lo = F[TA[i]] * F[TB[i]] + F[6 + TA[i]] * F[6 + TB[i]];
                    if (i + 16 < 21) {
                        hi = F[TA[i + 16]] * F[TB[i + 16]] + F[6 + TA[i + 16]] * F[6 + TB[i + 16]];
                    } else if (i + 16 < 27) {
                        hi = F[i + 16 - 21] * s0 + F[6 + i + 16 - 21] * s1;
                    } else {
                        hi = 0.0;
                    }
                    const double keep = up8 ? hi : lo, send = up8 ? lo : hi;
                    x[i] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
                }
#define SCHUR_RS_STEP(M, HALF)                                                                    \
                {                                                                                 \
                    const bool up = (lane & (M)) != 0;                                            \
                    _Pragma("unroll") for (int i = 0; i < (HALF); ++i) {                          \
                        const double keep = up ? x[i + (HALF)] : x[i], send = up ? x[i] : x[i + (HALF)]; \
                        x[i] = keep + __shfl_xor_sync(0xffffffffu, send, (M));                    \
                    }                                                                             \
                }
                SCHUR_RS_STEP(4, 8)
                SCHUR_RS_STEP(2, 4)
                SCHUR_RS_STEP(1, 2)
#undef SCHUR_RS_STEP
                if (!cross) { pacc0 += x[0]; pacc1 += x[1]; }   // a cross chunk owns no diagonal block, rhs entry or cost
            }
            // ---- operand stage: Z = (F'E) L^-T (zero for empty slots, constant poses and constant landmarks)
            {
                double z[3][6];
#pragma unroll
                for (int a = 0; a < 6; ++a) {
                    const double w0 = F[a] * E[0] + F[6 + a] * E[3];
                    const double w1 = F[a] * E[1] + F[6 + a] * E[4];
                    const double w2 = F[a] * E[2] + F[6 + a] * E[5];
                    z[0][a] = w0 * q[0];
                    z[1][a] = w0 * q[1] + w1 * q[2];
                    z[2][a] = w0 * q[3] + w1 * q[4] + w2 * q[5];
                }
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    double2* dst = reinterpret_cast<double2*>(Zs + (3 * ls + c) * kSZPitch + 6 * w);
                    dst[0] = make_double2(z[c][0], z[c][1]);
                    dst[1] = make_double2(z[c][2], z[c][3]);
                    dst[2] = make_double2(z[c][4], z[c][5]);
                }
            }
            __threadfence_block();
            named_bar_arrive(kBarFull + (b & 1), kSThreads);
        }
    } else {
        // ================= consumers: S_tile -= Z Z' over the stage's 48 columns =================
        const int cw = warp - kSProd / 32;
        const int fr = lane >> 2, fc = lane & 3;
        int rng_next = nbatch > 0 ? __ldg(batch_rng + ch.batch0) : 0;
        for (int b = 0; b < nbatch; ++b) {
            const double* Zs = Zst + (b & 1) * kSK * kSZPitch;
            // pose slots [wlo, whi] used by this batch: Z is zero outside, so only the 8x8 tiles
            // tlo..thi of the window change (a band of ~degree poses inside the 24-pose window)
            const int rng = rng_next;
            if (b + 1 < nbatch) rng_next = __ldg(batch_rng + ch.batch0 + b + 1);
            const int wlo = rng & 0xff, whi = (rng >> 8) & 0xff;
            const int tlo = (6 * wlo) >> 3, thi = (6 * whi + 5) >> 3;
            const int blo = tlo >> 1, nb2 = whi >= wlo ? (thi >> 1) - blo + 1 : 0;
            const int nblk = nb2 * (nb2 + 1) / 2;
            named_bar_sync(kBarFull + (b & 1), kSThreads);
            for (int blk = cw; blk < nblk; blk += kSCons / 32) {
                int bi = 0, rem = blk;
                while (rem >= nb2 - bi) { rem -= nb2 - bi; ++bi; }
                const int bj = blo + bi + rem;
                bi += blo;
                // a cross chunk owns only the blocks between its two pose groups: 16-row blocks 0..4 hold group A
                // (rows < 72), blocks 4..8 group B
                if (cross && (bi > (6 * kSGrp - 1) / 16 || bj < (6 * kSGrp) / 16)) continue;
                const bool diag = bi == bj;
                double2* c00 = reinterpret_cast<double2*>(St + schur_tile_index(2 * bi, 2 * bj) * 64) + lane;
                double2* c01 = reinterpret_cast<double2*>(St + schur_tile_index(2 * bi, 2 * bj + 1) * 64) + lane;
                double2* c11 = reinterpret_cast<double2*>(St + schur_tile_index(2 * bi + 1, 2 * bj + 1) * 64) + lane;
                double2* c10 = diag ? c00 : reinterpret_cast<double2*>(St + schur_tile_index(2 * bi + 1, 2 * bj) * 64) + lane;
                double2 a00 = *c00, a01 = *c01, a11 = *c11, a10 = make_double2(0.0, 0.0);
                if (!diag) a10 = *c10;
                const double* za = Zs + fc * kSZPitch + 16 * bi + fr;
                const double* zb = Zs + fc * kSZPitch + 16 * bj + fr;
                // two copies of the k loop: a predicated mma.sync costs a WARPSYNC + NOP pair each
                if (diag) {
#pragma unroll 4
                    for (int ks = 0; ks < kSK / 4; ++ks) {
                        const double x0 = -za[4 * ks * kSZPitch], x1 = -za[4 * ks * kSZPitch + 8];
                        const double y0 = zb[4 * ks * kSZPitch], y1 = zb[4 * ks * kSZPitch + 8];
                        dmma884(a00.x, a00.y, x0, y0);
                        dmma884(a01.x, a01.y, x0, y1);
                        dmma884(a11.x, a11.y, x1, y1);
                    }
                } else {
#pragma unroll 4
                    for (int ks = 0; ks < kSK / 4; ++ks) {
                        const double x0 = -za[4 * ks * kSZPitch], x1 = -za[4 * ks * kSZPitch + 8];
                        const double y0 = zb[4 * ks * kSZPitch], y1 = zb[4 * ks * kSZPitch + 8];
                        dmma884(a00.x, a00.y, x0, y0);
                        dmma884(a01.x, a01.y, x0, y1);
                        dmma884(a11.x, a11.y, x1, y1);
                        dmma884(a10.x, a10.y, x1, y0);
                    }
                }
                *c00 = a00; *c01 = a01; *c11 = a11;
                if (!diag) *c10 = a10;
            }
            if (b + 2 < nbatch) {
                __threadfence_block();
                named_bar_arrive(kBarEmpty + (b & 1), kSThreads);
            }
        }
    }
    __syncthreads();
    // ---- fold the per-pose sums into the tile (diagonal pose blocks) and the chunk's rhs slab
    double* outRhs = partRhs + (size_t)blockIdx.x * kST;
    if (tid < kSProd) {
        const int pw = tid >> 4, j = tid & 15;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int t = 2 * j + h;
            const double v = h ? pacc1 : pacc0;
            if (t < 21) {
                int a, b;
                tri6(t, a, b);
                const int R = 6 * pw + a, Cc = 6 * pw + b;
                St[schur_elem_index(R, Cc)] += v;                                            // R <= Cc: always a stored tile
                if (a != b && (Cc >> 3) == (R >> 3)) St[schur_elem_index(Cc, R)] += v;       // mirror inside a diagonal tile
            } else if (t < 27) {
                outRhs[6 * pw + (t - 21)] = v;
            }
        }
    }
    __syncthreads();
    // ---- tile -> slab, upper 8x8 tiles of the chunk's span only.  The slab is pose-block major
    // [pose slot i][pose slot j][6 x 6] so that k_schur_reduce reads a block as 288 contiguous bytes; a lane's
    // two columns (even, odd) always fall into the same 6-wide block.
    {
        double* outS = partS + (size_t)blockIdx.x * kST * kST;
        const int nt = cross ? kSNT : 2 * ((6 * ch.span + 15) >> 4);
        const int fr = lane >> 2, fc = lane & 3;
        for (int ti = 0; ti < nt; ++ti) {
            const int R = 8 * ti + fr, rb = R / 6, ra = R - 6 * rb;
            for (int tj = ti + warp; tj < nt; tj += kSThreads / 32) {
                const double2 v = *(reinterpret_cast<const double2*>(St + schur_tile_index(ti, tj) * 64) + lane);
                const int C = 8 * tj + 2 * fc, cb = C / 6, ca = C - 6 * cb;
                *reinterpret_cast<double2*>(outS + (size_t)(rb * kSWin + cb) * 36 + ra * 6 + ca) = v;
            }
        }
    }
    // ---- cost of the chunk
    if (cross) cost = 0.0;
    cost = warp_sum(cost);
    if (lane == 0) sBlk[warp] = cost;
    __syncthreads();
    if (warp == 0) {
        double t = lane < kSThreads / 32 ? sBlk[lane] : 0.0;
        t = warp_sum(t);
        if (lane == 0) partCost[blockIdx.x] = t;
    }
}

// Dense S (both triangles) + rhs + cost from the compact band buffer [n_free][band_w][36 | rhs n | cost]
__global__ void __launch_bounds__(256)
k_expand_band(const double* __restrict__ band, int n_free, int band_w, double* __restrict__ packed) {
    const int n = 6 * n_free;
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nn = (size_t)n * n;
    if (idx < nn) {
        const int i = (int)(idx / n), j = (int)(idx - (size_t)i * n);
        const int bi = i / 6, bj = j / 6, a = i - 6 * bi, b = j - 6 * bj;
        double v = 0.0;
        if (bj >= bi) { if (bj - bi < band_w) v = band[((size_t)bi * band_w + (bj - bi)) * 36 + 6 * a + b]; }
        else if (bi - bj < band_w) v = band[((size_t)bj * band_w + (bi - bj)) * 36 + 6 * b + a];
        packed[idx] = v;
    } else if (idx < nn + n + 1) {
        packed[idx] = band[(size_t)n_free * band_w * 36 + (idx - nn)];
    }
}

// Sums the chunk slabs into S (and rhs, cost) in a fixed order: one CTA per pair of free poses
// (bi <= bj), thread = (entry of the 6x6 block, one of kRedLanes strided walkers of the list of
// chunks covering pose bi); the walkers' partial sums are combined in a fixed order.
// cov_ptr / cov_idx: for every pose, the chunks whose window contains it; reach[p] = last pose
// that shares a chunk with p.  FULL: also write the mirrored block, add the pose part of the LM
// diagonal (levenberg_marquardt_strategy.cc:78-89) and sum the chunk costs, so that no further
// pass over S is needed; !FULL: upper block triangle only (k_schur_warp then adds the landmarks
// that fit no window and k_finish_S mirrors).
// One block pair (bi <= bj) of the slab sum; called by every thread of the CTA (threads beyond 36 * kRedLanes idle
// through the loads).  first: this CTA also sums the chunk costs.  part / partR: shared scratch.
template <bool FULL, int kRedLanes>
__device__ __forceinline__ void schur_reduce_block(const int bi, const int bj, const bool first, double* __restrict__ S, int n,
                                                   double* __restrict__ rhs, double* __restrict__ cost_out,
                                                   const double* __restrict__ partS, const double* __restrict__ partRhs,
                                                   const double* __restrict__ partCost, int nchunks,
                                                   const int* __restrict__ cov_ptr, const int4* __restrict__ cov_idx,
                                                   const int* __restrict__ reach, const int* __restrict__ free_pose,
                                                   const double* __restrict__ D, bool add_diag, int band_w,
                                                   double (*part)[36], double (*partR)[6]) {
    const int t = threadIdx.x, e = t % 36, g = t / 36;
    const int a = e / 6, b = e - 6 * a;
    const int pi = free_pose[bi], pj = free_pose[bj];
    const int q0 = cov_ptr[pi], q1 = cov_ptr[pi + 1];
    const bool walker = g < kRedLanes;
    const bool in_env = walker && pj <= reach[pi] && (bi < bj || a <= b);
    const bool do_rhs = walker && bi == bj && e < 6;   // the walkers' first six threads also sum the pose's rhs entry e
    double s = 0.0, sr = 0.0;
    if (in_env || do_rhs) {
        // four chunks per round: the (chunk, window) entries and the slab loads of a round are
        // independent, so a walker pays two memory latencies per four chunks instead of per chunk
        for (int q = q0 + g; q < q1; q += 4 * kRedLanes) {
            int4 ent[4];   // {chunk, slot of pose pi, first column pose | columns << 16, slot of the first column}
            double v[4], vr[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) ent[u] = q + u * kRedLanes < q1 ? __ldg(cov_idx + q + u * kRedLanes) : make_int4(-1, 0, 0, 0);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int first = ent[u].z & 0xffff, cols = ent[u].z >> 16;
                v[u] = (in_env && ent[u].x >= 0 && pj >= first && pj < first + cols)
                           ? partS[(size_t)ent[u].x * kST * kST + (size_t)(ent[u].y * kSWin + (ent[u].w + pj - first)) * 36 + e]   // pose-block major slab
                           : 0.0;
                vr[u] = (do_rhs && ent[u].x >= 0) ? partRhs[(size_t)ent[u].x * kST + 6 * ent[u].y + e] : 0.0;
            }
            s += (v[0] + v[1]) + (v[2] + v[3]);
            sr += (vr[0] + vr[1]) + (vr[2] + vr[3]);
        }
    }
    if (walker) {
        part[g][e] = s;
        if (e < 6) partR[g][e] = sr;
    }
    __syncthreads();
    if (t < 36 && (bi < bj || a <= b)) {
        double tot = part[0][e];
#pragma unroll
        for (int k = 1; k < kRedLanes; ++k) tot += part[k][e];
        if (FULL && bi == bj && a == b && add_diag && D) { const double dd = D[6 * (size_t)pi + a]; tot += dd * dd; }
        const size_t ri = 6 * (size_t)bi + a, rj = 6 * (size_t)bj + b;
        if (FULL && band_w > 0) {
            // compact exchange layout [block row bi][bj - bi < band_w][6 x 6] (vio360_ba_reduced_band_dev)
            double* blk = S + ((size_t)bi * band_w + (bj - bi)) * 36;
            blk[6 * a + b] = tot;
            if (bi == bj && a != b) blk[6 * b + a] = tot;
        } else {
            S[ri * n + rj] = tot;
            if (FULL && ri != rj) S[rj * n + ri] = tot;
        }
    }
    if (bi == bj && t >= 36 && t < 42) {   // rhs of this pose
        const int ar = t - 36;
        double tot = partR[0][ar];
#pragma unroll
        for (int k = 1; k < kRedLanes; ++k) tot += partR[k][ar];
        rhs[6 * bi + ar] = tot;
    }
    if (FULL && first && t >= 64 && t < 96) {   // cost: fixed-order sum of the chunk partials
        double tot = 0.0;
        for (int i = t - 64; i < nchunks; i += 32) tot += partCost[i];
        tot = warp_sum(tot);
        if (t == 64) *cost_out = tot;
    }
}

// Pass 1 for WIDE landmarks (observing poses span more than kSWin): one warp per landmark sums E'E (+ D^2) and E'r over
// ALL its observations, factorises E'E = L L' and leaves L^-1 (6) and (E'E)^-1 g (3) for the chunk kernel, whose
// jobs each see only two pose groups of the landmark.  Same arithmetic as phase B of k_schur_tile.
template <typename PT>
__global__ void __launch_bounds__(128) k_wide_prepare(BaDev d, const int* __restrict__ wide, int n_wide, bool apply_loss, double* __restrict__ wide_inv) {
    const int wi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (wi >= n_wide) return;
    const int l = wide[wi];
    const double Pw[3] = {d.points[3 * (size_t)l], d.points[3 * (size_t)l + 1], d.points[3 * (size_t)l + 2]};
    double hq[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int q = d.lm_ptr[l] + lane; q < d.lm_ptr[l + 1]; q += 32) {
        double r[2], F[12], E[6], pose[kPoseStride];
        const int p = d.pose[q];
#pragma unroll
        for (int j = 0; j < kPoseStride; ++j) pose[j] = __ldg(d.posebuf + kPoseStride * (size_t)p + j);
        eval_obs_as<PT>(pose, Pw, d.u[q], d.v[q], d.outlier[q] != 0, d.cam, apply_loss, r, F, E);
        hq[0] += E[0] * E[0] + E[3] * E[3]; hq[1] += E[0] * E[1] + E[3] * E[4]; hq[2] += E[0] * E[2] + E[3] * E[5];
        hq[3] += E[1] * E[1] + E[4] * E[4]; hq[4] += E[1] * E[2] + E[4] * E[5]; hq[5] += E[2] * E[2] + E[5] * E[5];
        hq[6] += E[0] * r[0] + E[3] * r[1]; hq[7] += E[1] * r[0] + E[4] * r[1]; hq[8] += E[2] * r[0] + E[5] * r[1];
    }
#pragma unroll
    for (int j = 0; j < 9; ++j) hq[j] = warp_sum(hq[j]);
    if (lane != 0) return;
    double li[6] = {0, 0, 0, 0, 0, 0}, ig[3] = {0, 0, 0};
    if (d.point_fixed[l] == 0) {
        if (d.D) {
            const double* Dl = d.D + 6 * (size_t)d.K + 3 * (size_t)l;
            hq[0] += Dl[0] * Dl[0]; hq[3] += Dl[1] * Dl[1]; hq[5] += Dl[2] * Dl[2];
        }
        const double i00 = vio_rsqrt(hq[0]);
        const double l10 = hq[1] * i00, l20 = hq[2] * i00;
        const double i11 = vio_rsqrt(hq[3] - l10 * l10);
        const double l21 = (hq[4] - l20 * l10) * i11;
        const double i22 = vio_rsqrt(hq[5] - l20 * l20 - l21 * l21);
        const double i10 = -l10 * i00 * i11;
        const double i21 = -l21 * i11 * i22;
        const double i20 = -(l20 * i00 + l21 * i10) * i22;
        li[0] = i00; li[1] = i10; li[2] = i11; li[3] = i20; li[4] = i21; li[5] = i22;
        bool ok = true;
#pragma unroll
        for (int j = 0; j < 6; ++j) ok = ok && isfinite(li[j]);
        if (!ok) {
#pragma unroll
            for (int j = 0; j < 6; ++j) li[j] = 0.0;
        }
        const double y0 = li[0] * hq[6], y1 = li[1] * hq[6] + li[2] * hq[7], y2 = li[3] * hq[6] + li[4] * hq[7] + li[5] * hq[8];
        ig[0] = li[0] * y0 + li[1] * y1 + li[3] * y2;
        ig[1] = li[2] * y1 + li[4] * y2;
        ig[2] = li[5] * y2;
    }
    double* q = wide_inv + 9 * (size_t)l;
#pragma unroll
    for (int j = 0; j < 6; ++j) q[j] = li[j];
    q[6] = ig[0]; q[7] = ig[1]; q[8] = ig[2];
}

template <bool FULL, int kRedLanes>
__global__ void __launch_bounds__(36 * kRedLanes)
k_schur_reduce(double* __restrict__ S, int n, double* __restrict__ rhs, double* __restrict__ cost_out,
               const double* __restrict__ partS, const double* __restrict__ partRhs, const double* __restrict__ partCost,
               int nchunks, const SchurChunk* __restrict__ chunks, const int* __restrict__ cov_ptr,
               const int4* __restrict__ cov_idx, const int* __restrict__ reach, const int* __restrict__ free_pose,
               const double* __restrict__ D, bool add_diag, bool band_grid, int band_w, int row0, int n_free, const PeerPublish pub) {
    // band_grid: blockIdx.x counts block columns from the diagonal (the host sized the grid to the
    // widest row of the envelope and zeroed S outside it once per structure).  row0: first block row
    // of the grid (a rank of a peer-memory exchange sums only the rows its landmarks touch).
    pdl_wait();                // launched with programmatic dependent launch behind k_schur_tile: its slabs are complete from here on
    pdl_launch_dependents();   // (multi-GPU) k_band_pull may be scheduled behind this grid
    const int bi = blockIdx.y + row0, bj = band_grid ? bi + blockIdx.x : blockIdx.x;
    if (bi > bj || bj >= n_free) return;
    __shared__ double part[kRedLanes][36];
    __shared__ double partR[kRedLanes][6];
    schur_reduce_block<FULL, kRedLanes>(bi, bj, blockIdx.y == 0 && bi == bj, S, n, rhs, cost_out, partS, partRhs, partCost, nchunks,
                                        cov_ptr, cov_idx, reach, free_pose, D, add_diag, band_w, part, partR);
    peer_publish_block(pub, bi);   // multi-GPU: release the block row to the other ranks once its last block is stored (ba_peer.cuh)
}

}  // namespace vio360
