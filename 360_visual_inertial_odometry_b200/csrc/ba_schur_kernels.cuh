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

__global__ void __launch_bounds__(kSThreads, 1)
k_schur_tile(BaDev d, const SchurChunk* __restrict__ chunks, bool apply_loss, double* __restrict__ partS,
             double* __restrict__ partRhs, double* __restrict__ partCost) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* Z = reinterpret_cast<double*>(smem_raw);            // [kST][kSZPitch] operand stage
    double* sRed = Z;                                           // alias: [kSThreads][kSRed] reduction scratch
    double* St = Z + kST * kSZPitch;                            // [kSTiles][64] S tile, DMMA fragment order
    double* sPose = St + kSTiles * 64;                          // [kSWin][kPoseStride]
    double* sLmk = sPose + kSWin * kPoseStride;                 // [kSMaxLm][9]: L^-1 (6), (E'E)^-1 g (3)
    double* sBlk = sLmk + kSMaxLm * 9;                          // [32] block-sum scratch
    short* sOst = reinterpret_cast<short*>(sBlk + 32);          // [kSMaxLm][kSWin] thread that holds (landmark, pose)
    int* sLmp = reinterpret_cast<int*>(sOst + kSMaxLm * kSWin); // [kSMaxLm + 1] lm_ptr of the batch

    const SchurChunk ch = chunks[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kST * kSZPitch; i += kSThreads) Z[i] = 0.0;
    for (int i = tid; i < kSTiles * 64; i += kSThreads) St[i] = 0.0;
    for (int i = tid; i < ch.span * kPoseStride; i += kSThreads) sPose[i] = d.posebuf[(size_t)ch.pbase * kPoseStride + i];

    // per-pose sums: thread (w, e) with w = tid / 9, e = tid % 9 owns entries 9*round + e of
    // [F'F upper triangle (21) | F'(r - E ig) (6)] of pose slot w
    double pacc[3] = {0.0, 0.0, 0.0};
    double cost = 0.0;
    int cur = ch.lm_begin;
    while (cur < ch.lm_end) {
        __syncthreads();   // stage / tile free: previous batch's DMMA (or the initial zeroing) is done
        if (tid <= kSMaxLm) sLmp[tid] = d.lm_ptr[min(cur + tid, ch.lm_end)];
        for (int i = tid; i < kSMaxLm * kSWin; i += kSThreads) sOst[i] = -1;
        __syncthreads();
        const int o0 = sLmp[0];
        int NB = 0;
        while (NB < kSMaxLm && cur + NB < ch.lm_end && sLmp[NB + 1] - o0 <= kSThreads) ++NB;
        const int nobs = sLmp[NB] - o0;
        const int NBp = (NB + 3) & ~3;
        // ---- phase A: evaluate one observation per thread
        const bool act = tid < nobs;
        int ls = 0, w = 0;
        bool emit = false;     // free pose and free landmark: contributes to Z
        bool pfree = false;
        double r[2], F[12], E[6];
        if (act) {
            const int i = o0 + tid;
            const int p = d.pose[i], l = d.point[i];
            ls = l - cur; w = p - ch.pbase;
            const double Pw[3] = {__ldg(d.points + 3 * (size_t)l), __ldg(d.points + 3 * (size_t)l + 1),
                                  __ldg(d.points + 3 * (size_t)l + 2)};
            cost += obs_evaluate<true>(sPose + kPoseStride * w, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
            pfree = d.free_map[p] >= 0;
            emit = pfree && d.point_fixed[l] == 0;
            sOst[ls * kSWin + w] = (short)tid;
            double* q = sRed + tid * kSRed;
            q[0] = E[0] * E[0] + E[3] * E[3]; q[1] = E[0] * E[1] + E[3] * E[4]; q[2] = E[0] * E[2] + E[3] * E[5];
            q[3] = E[1] * E[1] + E[4] * E[4]; q[4] = E[1] * E[2] + E[4] * E[5]; q[5] = E[2] * E[2] + E[5] * E[5];
            q[6] = E[0] * r[0] + E[3] * r[1]; q[7] = E[1] * r[0] + E[4] * r[1]; q[8] = E[2] * r[0] + E[5] * r[1];
        }
        __syncthreads();
        // ---- phase B: per-landmark sums, Cholesky, L^-1, (E'E)^-1 g
        if (tid < NB * 9) {
            const int bl = tid / 9, e = tid - bl * 9;
            double s = 0.0;
            for (int j = sLmp[bl] - o0; j < sLmp[bl + 1] - o0; ++j) s += sRed[j * kSRed + e];
            sLmk[tid] = s;
        }
        __syncthreads();
        if (tid < NB) {
            const int l = cur + tid;
            double* q = sLmk + tid * 9;
            double h[6] = {q[0], q[1], q[2], q[3], q[4], q[5]};
            const double g0 = q[6], g1 = q[7], g2 = q[8];
            double li[6] = {0, 0, 0, 0, 0, 0}, ig[3] = {0, 0, 0};
            if (d.point_fixed[l] == 0) {
                if (d.D) {
                    const double* Dl = d.D + 6 * (size_t)d.K + 3 * (size_t)l;
                    h[0] += Dl[0] * Dl[0]; h[3] += Dl[1] * Dl[1]; h[5] += Dl[2] * Dl[2];
                }
                const double l00 = sqrt(h[0]);
                const double l10 = h[1] / l00, l20 = h[2] / l00;
                const double l11 = sqrt(h[3] - l10 * l10);
                const double l21 = (h[4] - l20 * l10) / l11;
                const double l22 = sqrt(h[5] - l20 * l20 - l21 * l21);
                const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
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
                const double y0 = li[0] * g0, y1 = li[1] * g0 + li[2] * g1, y2 = li[3] * g0 + li[4] * g1 + li[5] * g2;
                ig[0] = li[0] * y0 + li[1] * y1 + li[3] * y2;
                ig[1] = li[2] * y1 + li[4] * y2;
                ig[2] = li[5] * y2;
            }
#pragma unroll
            for (int j = 0; j < 6; ++j) q[j] = li[j];
            q[6] = ig[0]; q[7] = ig[1]; q[8] = ig[2];
        }
        __syncthreads();
        // ---- phase C: per-pose sums of F'F and F'(r - E ig) in three rounds of nine entries
        double li[6] = {0, 0, 0, 0, 0, 0};
        double s0 = 0.0, s1 = 0.0;
        if (act) {
            const double* q = sLmk + ls * 9;
#pragma unroll
            for (int j = 0; j < 6; ++j) li[j] = q[j];
            s0 = r[0] - (E[0] * q[6] + E[1] * q[7] + E[2] * q[8]);
            s1 = r[1] - (E[3] * q[6] + E[4] * q[7] + E[5] * q[8]);
        }
#pragma unroll
        for (int round = 0; round < 3; ++round) {
            if (act) {
                double* q = sRed + tid * kSRed;
#pragma unroll
                for (int e = 0; e < 9; ++e) {
                    const int t = 9 * round + e;   // compile-time after unrolling
                    double v;
                    if (t < 21) {
                        int a = 0, rem = t;
                        while (rem >= 6 - a) { rem -= 6 - a; ++a; }
                        const int b = a + rem;
                        v = F[a] * F[b] + F[6 + a] * F[6 + b];
                    } else {
                        v = F[t - 21] * s0 + F[6 + t - 21] * s1;
                    }
                    q[e] = pfree ? v : 0.0;
                }
            }
            __syncthreads();
            if (tid < kSWin * 9) {
                const int pw = tid / 9, e = tid - pw * 9;
                double s = 0.0;
                for (int bl = 0; bl < NB; ++bl) {
                    const int src = sOst[bl * kSWin + pw];
                    if (src >= 0) s += sRed[src * kSRed + e];
                }
                pacc[round] += s;
            }
            __syncthreads();
        }
        // ---- operand stage: Z = (F'E) L^-T for every (pose slot, landmark) of the batch, zeros elsewhere
        if (act) {
            double* zr = Z + (6 * w) * kSZPitch + 3 * ls;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double w0 = F[a] * E[0] + F[6 + a] * E[3];
                const double w1 = F[a] * E[1] + F[6 + a] * E[4];
                const double w2 = F[a] * E[2] + F[6 + a] * E[5];
                zr[a * kSZPitch + 0] = emit ? w0 * li[0] : 0.0;
                zr[a * kSZPitch + 1] = emit ? w0 * li[1] + w1 * li[2] : 0.0;
                zr[a * kSZPitch + 2] = emit ? w0 * li[3] + w1 * li[4] + w2 * li[5] : 0.0;
            }
        }
        for (int t = tid; t < NBp * kSWin; t += kSThreads) {
            const int bl = t / kSWin, pw = t - bl * kSWin;
            if (bl >= NB || sOst[t] < 0) {
                double* zr = Z + (6 * pw) * kSZPitch + 3 * bl;
#pragma unroll
                for (int a = 0; a < 6; ++a) { zr[a * kSZPitch] = 0.0; zr[a * kSZPitch + 1] = 0.0; zr[a * kSZPitch + 2] = 0.0; }
            }
        }
        __syncthreads();
        // ---- phase D: S_tile -= Z Z' over the batch's 3*NBp columns, 16x16 blocks of the upper triangle
        {
            const int nb2 = (6 * ch.span + 15) >> 4;
            const int nblk = nb2 * (nb2 + 1) / 2;
            const int ksteps = (3 * NBp) >> 2;
            const int fr = lane >> 2, fc = lane & 3;
            for (int blk = warp; blk < nblk; blk += kSWarps) {
                int bi = 0, rem = blk;
                while (rem >= nb2 - bi) { rem -= nb2 - bi; ++bi; }
                const int bj = bi + rem;
                const bool diag = bi == bj;
                double2* c00 = reinterpret_cast<double2*>(St + schur_tile_index(2 * bi, 2 * bj) * 64) + lane;
                double2* c01 = reinterpret_cast<double2*>(St + schur_tile_index(2 * bi, 2 * bj + 1) * 64) + lane;
                double2* c11 = reinterpret_cast<double2*>(St + schur_tile_index(2 * bi + 1, 2 * bj + 1) * 64) + lane;
                double2* c10 = diag ? c00 : reinterpret_cast<double2*>(St + schur_tile_index(2 * bi + 1, 2 * bj) * 64) + lane;
                double2 a00 = *c00, a01 = *c01, a11 = *c11, a10 = make_double2(0.0, 0.0);
                if (!diag) a10 = *c10;
                const double* za = Z + (16 * bi + fr) * kSZPitch + fc;
                const double* zb = Z + (16 * bj + fr) * kSZPitch + fc;
#pragma unroll 3
                for (int ks = 0; ks < ksteps; ++ks) {
                    const double x0 = -za[4 * ks], x1 = -za[8 * kSZPitch + 4 * ks];
                    const double y0 = zb[4 * ks], y1 = zb[8 * kSZPitch + 4 * ks];
                    dmma884(a00.x, a00.y, x0, y0);
                    dmma884(a01.x, a01.y, x0, y1);
                    dmma884(a11.x, a11.y, x1, y1);
                    if (!diag) dmma884(a10.x, a10.y, x1, y0);
                }
                *c00 = a00; *c01 = a01; *c11 = a11;
                if (!diag) *c10 = a10;
            }
        }
        cur += NB;
    }
    __syncthreads();
    // ---- fold the per-pose sums into the tile (diagonal pose blocks) and the chunk's rhs slab
    double* outRhs = partRhs + (size_t)blockIdx.x * kST;
    if (tid < kSWin * 9) {
        const int pw = tid / 9, e = tid - pw * 9;
#pragma unroll
        for (int round = 0; round < 3; ++round) {
            const int t = 9 * round + e;
            if (t < 21) {
                int a, b;
                tri6(t, a, b);
                const int R = 6 * pw + a, Cc = 6 * pw + b;
                St[schur_elem_index(R, Cc)] += pacc[round];                       // R <= Cc: always a stored tile
                if (a != b && (Cc >> 3) == (R >> 3)) St[schur_elem_index(Cc, R)] += pacc[round];   // mirror inside a diagonal tile
            } else {
                outRhs[6 * pw + (t - 21)] = pacc[round];
            }
        }
    }
    __syncthreads();
    // ---- tile -> slab (dense kST x kST, upper 8x8 tiles of the chunk's span only)
    {
        double* outS = partS + (size_t)blockIdx.x * kST * kST;
        const int nt = 2 * ((6 * ch.span + 15) >> 4);
        const int fr = lane >> 2, fc = lane & 3;
        for (int ti = 0; ti < nt; ++ti)
            for (int tj = ti + warp; tj < nt; tj += kSWarps) {
                const double2 v = *(reinterpret_cast<const double2*>(St + schur_tile_index(ti, tj) * 64) + lane);
                *reinterpret_cast<double2*>(outS + (size_t)(8 * ti + fr) * kST + 8 * tj + 2 * fc) = v;
            }
    }
    // ---- cost of the chunk
    cost = warp_sum(cost);
    if (lane == 0) sBlk[warp] = cost;
    __syncthreads();
    if (warp == 0) {
        double t = lane < kSWarps ? sBlk[lane] : 0.0;
        t = warp_sum(t);
        if (lane == 0) partCost[blockIdx.x] = t;
    }
}

// Sums the chunk slabs into the upper block triangle of S (and rhs) in a fixed order.
// cov_ptr / cov_idx: for every pose, the chunks whose window contains it.
__global__ void __launch_bounds__(256)
k_schur_reduce(double* __restrict__ S, int n, double* __restrict__ rhs, const double* __restrict__ partS,
               const double* __restrict__ partRhs, const SchurChunk* __restrict__ chunks,
               const int* __restrict__ cov_ptr, const int* __restrict__ cov_idx, const int* __restrict__ free_pose) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < (long long)n * n) {
        const int ri = (int)(idx / n), rj = (int)(idx - (long long)ri * n);
        const int bi = ri / 6, bj = rj / 6;
        if (bi < bj || (bi == bj && ri <= rj)) {
            const int pi = free_pose[bi], pj = free_pose[bj];
            const int a = ri - 6 * bi, b = rj - 6 * bj;
            double s = 0.0;
            for (int q = cov_ptr[pi]; q < cov_ptr[pi + 1]; ++q) {
                const int c = cov_idx[q];
                const SchurChunk ch = chunks[c];
                if (pj >= ch.pbase && pj < ch.pbase + ch.span)
                    s += partS[(size_t)c * kST * kST + (size_t)(6 * (pi - ch.pbase) + a) * kST + 6 * (pj - ch.pbase) + b];
            }
            S[idx] = s;
        }
    } else if (idx < (long long)n * n + n) {
        const int ri = (int)(idx - (long long)n * n);
        const int bi = ri / 6, pi = free_pose[bi], a = ri - 6 * bi;
        double s = 0.0;
        for (int q = cov_ptr[pi]; q < cov_ptr[pi + 1]; ++q) {
            const int c = cov_idx[q];
            s += partRhs[(size_t)c * kST + 6 * (pi - chunks[c].pbase) + a];
        }
        rhs[ri] = s;
    }
}

}  // namespace vio360
