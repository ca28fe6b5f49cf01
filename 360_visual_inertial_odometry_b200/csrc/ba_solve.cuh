// ba_solve.cuh -- included by ba.cu.  SURVEY 8-F rank 3: consume the reduced camera system on the
// device instead of handing it back to Ceres:
//   k_chol_skyline  S x = rhs by a block (6x6) envelope Cholesky.  Replaces the dense LLT of
//                   SchurComplementSolver::SolveReducedLinearSystem (thirdparty/ceres-solver/
//                   internal/ceres/schur_complement_solver.cc, DenseSchurComplementSolver).  In a
//                   sliding window a pose only shares landmarks with its ~20 neighbours, so S is
//                   block banded; `reach[j]` (monotone, from the chunk table) bounds the non-zero
//                   block rows of column j and the factor stays inside that envelope.
//   k_back_substitute  per landmark y = (E'E + D^2)^-1 (E'r - E'F x)  = SchurEliminator::BackSubstitute
//                   (schur_eliminator_impl.h:310-380), one warp per landmark, factors re-evaluated.
// One CTA does the factorisation: it is a chain of nb dependent block columns; the banded case
// (the one the reference produces) needs ~nb * 3 barriers and ~n * bw^2 flops.  A fully dense S
// is handled by the same code, slowly -- documented limit.
#pragma once

namespace vio360 {

constexpr int kCholThreads = 1024;

// lower Cholesky of the 6x6 A (row-major, lower part read) and the inverse of the factor;
// division free: reciprocal square roots of the pivots
__device__ __forceinline__ bool chol6_inv(const double* A, double* L, double* Li) {
    bool ok = true;
    double inv[6];
#pragma unroll
    for (int i = 0; i < 36; ++i) { L[i] = 0.0; Li[i] = 0.0; }
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double d = A[j * 6 + j];
#pragma unroll
        for (int k = 0; k < j; ++k) d -= L[j * 6 + k] * L[j * 6 + k];
        if (!(d > 0.0)) { ok = false; d = 1.0; }
        inv[j] = vio_rsqrt(d);
        L[j * 6 + j] = d * inv[j];
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            double s = A[i * 6 + j];
#pragma unroll
            for (int k = 0; k < j; ++k) s -= L[i * 6 + k] * L[j * 6 + k];
            L[i * 6 + j] = s * inv[j];
        }
    }
#pragma unroll
    for (int c = 0; c < 6; ++c) {   // L * Li = I, column by column (forward substitution)
#pragma unroll
        for (int i = c; i < 6; ++i) {
            double s = (i == c) ? 1.0 : 0.0;
#pragma unroll
            for (int k = c; k < i; ++k) s -= L[i * 6 + k] * Li[k * 6 + c];
            Li[i * 6 + c] = s * inv[i];
        }
    }
    return ok;
}

// A: n x n row-major work copy of S (lower triangle used and overwritten by the factor);
// Linv: nb x 36 scratch; b: rhs in, solution out; status[0] = 0 ok / 1 not positive definite.
// The right-hand side rides along as one more row below the matrix (row index n of the
// augmented lower triangle), so the forward substitution happens inside the factorisation.
__global__ void __launch_bounds__(kCholThreads) k_chol_skyline(double* __restrict__ A, int nb, const int* __restrict__ reach,
                                                               double* __restrict__ Linv, double* __restrict__ b, int* __restrict__ status) {
    extern __shared__ double panel[];   // [(rows below the diagonal block, + the rhs row) x 6] of the current column
    __shared__ double sD[36], sL[36], sLi[36], sy[6];
    __shared__ int sbad;
    const int tid = threadIdx.x, n = 6 * nb;
    if (tid == 0) sbad = 0;
    __syncthreads();
    // ---- factorisation (+ forward substitution of the rhs row)
    for (int j = 0; j < nb; ++j) {
        const int rj = reach[j], r = rj - j;   // block rows j+1 .. rj are inside the envelope
        const int rows = 6 * r;
        if (tid < 36) sD[tid] = A[(size_t)(6 * j + tid / 6) * n + 6 * j + tid % 6];
        __syncthreads();
        if (tid == 0) { if (!chol6_inv(sD, sL, sLi)) sbad = 1; }
        __syncthreads();
        if (tid < 36) {
            A[(size_t)(6 * j + tid / 6) * n + 6 * j + tid % 6] = sL[tid];
            Linv[(size_t)j * 36 + tid] = sLi[tid];
        }
        // L_ij = A_ij Ljj^-T, one thread per row of the column panel; row `rows` is the rhs: y_j = Ljj^-1 b_j
        for (int row = tid; row <= rows; row += kCholThreads) {
            double* a = row < rows ? A + (size_t)(6 * (j + 1) + row) * n + 6 * j : b + 6 * j;
            double v[6], o[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) v[k] = a[k];
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k <= c; ++k) s += v[k] * sLi[c * 6 + k];
                o[c] = s;
            }
#pragma unroll
            for (int c = 0; c < 6; ++c) { a[c] = o[c]; panel[row * 6 + c] = o[c]; }
        }
        __syncthreads();
        // trailing update inside the envelope: A_ik -= L_ij L_kj^T for j < k <= i <= rj, and the rhs row
        for (int ck = tid & 127; ck < rows; ck += 128) {
            const double* lk = panel + ck * 6;
            const double k0 = lk[0], k1 = lk[1], k2 = lk[2], k3 = lk[3], k4 = lk[4], k5 = lk[5];
            const int first = (ck / 6) * 6;   // block lower part: rows of block ck/6 and below
            for (int ri = first + (tid >> 7); ri <= rows; ri += kCholThreads / 128) {
                const double* li = panel + ri * 6;
                const double s = li[0] * k0 + li[1] * k1 + li[2] * k2 + li[3] * k3 + li[4] * k4 + li[5] * k5;
                if (ri < rows) A[(size_t)(6 * (j + 1) + ri) * n + 6 * (j + 1) + ck] -= s;
                else b[6 * (j + 1) + ck] -= s;
            }
        }
        __syncthreads();
    }
    // ---- backward substitution L^T x = y
    for (int j = nb - 1; j >= 0; --j) {
        const int r = reach[j] - j;
        if (tid < 32) {   // warp 0: s_c = y_c - sum_rows L[row][c] x[row]
            double acc[6] = {0, 0, 0, 0, 0, 0};
            for (int row = tid; row < 6 * r; row += 32) {
                const double* a = A + (size_t)(6 * (j + 1) + row) * n + 6 * j;
                const double xv = b[6 * (j + 1) + row];
#pragma unroll
                for (int c = 0; c < 6; ++c) acc[c] += a[c] * xv;
            }
#pragma unroll
            for (int c = 0; c < 6; ++c)
#pragma unroll
                for (int sft = 16; sft > 0; sft >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], sft);
            if (tid < 6) {
                double s[6];
#pragma unroll
                for (int c = 0; c < 6; ++c) s[c] = b[6 * j + c] - acc[c];
                double x = 0.0;   // x = Ljj^-T s : x_t = sum_{k >= t} Linv[k][t] s_k
                for (int k = tid; k < 6; ++k) x += Linv[(size_t)j * 36 + k * 6 + tid] * s[k];
                sy[tid] = x;
            }
            __syncwarp();
            if (tid < 6) b[6 * j + tid] = sy[tid];
            __threadfence_block();
            __syncwarp();
        }
        // only warp 0 works in this phase; the other warps wait once at the end
    }
    __syncthreads();
    if (tid == 0) status[0] = sbad;
}

// Banded fast path of the same factorisation: when no column reaches further than kCholWb - 1 block
// rows (sliding windows: ~20), the active (kCholWb x kCholWb)-block part of the trailing matrix lives
// in shared memory as a circular buffer (block (i, k) at slot (i mod Wb, k mod Wb)); a finished
// column frees its slot for the block row that enters the window, fetched from the untouched S
// while the column is being processed.  Three block barriers per column, no global round trips
// on the dependent path.  L's columns are streamed out compactly for the backward substitution.
constexpr int kCholWb = 24;

__device__ __forceinline__ bool chol6_diag(const double* A, double* L, double* invd) {
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 36; ++i) L[i] = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double d = A[j * 6 + j];
#pragma unroll
        for (int k = 0; k < j; ++k) d -= L[j * 6 + k] * L[j * 6 + k];
        if (!(d > 0.0)) { ok = false; d = 1.0; }
        const double iv = vio_rsqrt(d);
        invd[j] = iv;
        L[j * 6 + j] = d * iv;
#pragma unroll
        for (int i = j + 1; i < 6; ++i) {
            double s = A[i * 6 + j];
#pragma unroll
            for (int k = 0; k < j; ++k) s -= L[i * 6 + k] * L[j * 6 + k];
            L[i * 6 + j] = s * iv;
        }
    }
    return ok;
}

// S: n x n row-major (read only); Lcol: nb x (Wb-1) x 36 (block rows below the diagonal of each
// column); Ldiag: nb x 42 (L_jj, 6 reciprocal pivots); b: rhs in, x out.
__global__ void __launch_bounds__(kCholThreads) k_chol_band(const double* __restrict__ S, int nb, const int* __restrict__ reach,
                                                            double* __restrict__ Lcol, double* __restrict__ Ldiag, double* __restrict__ b,
                                                            int* __restrict__ status) {
    extern __shared__ double sm[];
    double* Wsm = sm;                                         // [Wb][Wb][36]
    double* panel = Wsm + kCholWb * kCholWb * 36;             // [(Wb-1)*6 + 1][6]
    double* sb = panel + ((kCholWb - 1) * 6 + 1) * 6;         // [n]
    __shared__ double sL[36], sInv[6];
    __shared__ int sbad;
    const int tid = threadIdx.x, n = 6 * nb;
    auto slot = [&](int i, int k) { return Wsm + ((i % kCholWb) * kCholWb + (k % kCholWb)) * 36; };
    if (tid == 0) sbad = 0;
    for (int e = tid; e < n; e += kCholThreads) sb[e] = b[e];
    {   // first window: lower triangle of block rows 0 .. min(Wb, nb) - 1
        const int w0 = min(kCholWb, nb), wr = 6 * w0;
        for (int e = tid; e < wr * wr; e += kCholThreads) {
            const int row = e / wr, colm = e - row * wr;
            if (colm / 6 > row / 6) continue;
            slot(row / 6, colm / 6)[(row % 6) * 6 + colm % 6] = S[(size_t)row * n + colm];
        }
    }
    __syncthreads();
    // the block row that enters the window when column j leaves (row j + Wb, columns j+1 .. j+Wb) is
    // fetched three columns ahead: one element per thread, three loads in flight
    const int pa = tid / (6 * kCholWb), pc = tid - pa * (6 * kCholWb);   // row in block, column offset
    auto fetch_row = [&](int j) -> double {
        const int inew = j + kCholWb;
        return (inew < nb && pa < 6) ? S[(size_t)(6 * inew + pa) * n + 6 * (j + 1) + pc] : 0.0;
    };
    double preA = fetch_row(0), preB = fetch_row(1), preC = fetch_row(2);
    // Column j: [panel phase, all threads] barrier [trailing update on warps 1..31 | warp 0 updates the
    // next diagonal block and factorises it] barrier -- the serial 6x6 pivot hides behind the update.
    if (tid == 0) { if (!chol6_diag(slot(0, 0), sL, sInv)) sbad = 1; }
    __syncthreads();
#define CHOL_STEP(J, PRE)                                                                            \
    if ((J) < nb) {                                                                                \
        const int j = (J);                                                                         \
        const int r = reach[j] - j, rows = 6 * r;                                                  \
        if (tid < 36) Ldiag[(size_t)j * 42 + tid] = sL[tid];                                       \
        if (tid < 6) Ldiag[(size_t)j * 42 + 36 + tid] = sInv[tid];                                 \
        if (tid <= rows) {   /* panel row: x Ljj^T = v by substitution; row `rows` is the rhs */   \
            const int bi = tid / 6, a = tid - bi * 6;                                              \
            double* src = tid < rows ? slot(j + 1 + bi, j) + a * 6 : sb + 6 * j;                   \
            double o[6];                                                                           \
            _Pragma("unroll") for (int c = 0; c < 6; ++c) {                                        \
                double v = src[c];                                                                 \
                _Pragma("unroll") for (int k = 0; k < c; ++k) v -= o[k] * sL[c * 6 + k];           \
                o[c] = v * sInv[c];                                                                \
            }                                                                                      \
            _Pragma("unroll") for (int c = 0; c < 6; ++c) panel[tid * 6 + c] = o[c];               \
            if (tid < rows) {                                                                      \
                _Pragma("unroll") for (int c = 0; c < 6; ++c) Lcol[((size_t)j * (kCholWb - 1) + bi) * 36 + a * 6 + c] = o[c]; \
            } else {                                                                               \
                _Pragma("unroll") for (int c = 0; c < 6; ++c) sb[6 * j + c] = o[c];                \
            }                                                                                      \
        }                                                                                          \
        __syncthreads();                                                                           \
        if (tid < 32) {   /* pivot warp: block (j+1, j+1) -= L L^T of the first panel block, then factorise */ \
            if (j + 1 < nb) {                                                                      \
                double* dblk = slot(j + 1, j + 1);                                                 \
                if (rows >= 6) {                                                                   \
                    for (int e = tid; e < 36; e += 32) {                                           \
                        const double* li = panel + (e / 6) * 6;                                    \
                        const double* lk = panel + (e % 6) * 6;                                    \
                        dblk[e] -= li[0] * lk[0] + li[1] * lk[1] + li[2] * lk[2] + li[3] * lk[3] + li[4] * lk[4] + li[5] * lk[5]; \
                    }                                                                              \
                }                                                                                  \
                __syncwarp();                                                                      \
                if (tid == 0) { if (!chol6_diag(dblk, sL, sInv)) sbad = 1; }                       \
            }                                                                                      \
        } else {                                                                                   \
            const int t2 = tid - 32;   /* 992 threads: column ck = t2 % 124, row phase t2 / 124 (8) */ \
            for (int ck = t2 % 124; ck < rows; ck += 124) {                                        \
                const double* lk = panel + ck * 6;                                                 \
                const double k0 = lk[0], k1 = lk[1], k2 = lk[2], k3 = lk[3], k4 = lk[4], k5 = lk[5]; \
                const int kb = ck / 6, kc = ck - kb * 6;                                           \
                for (int ri = kb * 6 + t2 / 124; ri <= rows; ri += 8) {                            \
                    if (ri < 6) continue;   /* block (j+1, j+1) belongs to the pivot warp */       \
                    const double* li = panel + ri * 6;                                             \
                    const double sdot = li[0] * k0 + li[1] * k1 + li[2] * k2 + li[3] * k3 + li[4] * k4 + li[5] * k5; \
                    if (ri < rows) slot(j + 1 + ri / 6, j + 1 + kb)[(ri % 6) * 6 + kc] -= sdot;    \
                    else sb[6 * (j + 1) + ck] -= sdot;                                             \
                }                                                                                  \
            }                                                                                      \
        }                                                                                          \
        if (j + kCholWb < nb && pa < 6) slot(j + kCholWb, j + 1 + pc / 6)[pa * 6 + pc % 6] = PRE;  /* row slot j mod Wb is free now */ \
        PRE = fetch_row(j + 3);                                                                    \
        __syncthreads();                                                                           \
    }
    for (int j3 = 0; j3 < nb; j3 += 3) {
        CHOL_STEP(j3, preA)
        CHOL_STEP(j3 + 1, preB)
        CHOL_STEP(j3 + 2, preC)
    }
#undef CHOL_STEP
    __syncthreads();
    for (int e = tid; e < n; e += kCholThreads) b[e] = sb[e];
    if (tid == 0) status[0] = sbad;
}

// Register-window form of the banded factorisation (round 2).  k_chol_band keeps the active window in shared memory
// and its trailing update issues one broadcast load per multiply-add: the shared-memory pipe, not the FP64 pipe, sets
// its 4 us per column.  Here every 6x6 block of the window lives in the REGISTERS of one thread for as long as it is
// in the window:
//   * an owner thread stands for an unordered pair {a, b} of residues mod Wb = 24.  At column j the window holds block
//     rows / columns j .. j + 23, one per residue; the pair's block is (i, k) = the two window indices with those
//     residues, larger first.  When column k leaves the window its block (i, k) is finished, and the block that enters
//     with the new row, (k + 24, i), has the SAME residue pair: the thread moves straight on to it (prefetched with
//     cp.async from the untouched S into the thread's shared-memory slot when the previous block was taken up), so no
//     register block ever idles;
//   * twelve warps own the 276 off-diagonal pairs as 8 x 4 tiles of (a, b): a warp's loads of P touch at most 12
//     distinct rows (one or two shared-memory wavefronts per 128-bit load) instead of 32 unrelated ones;
//   * warp 0 owns the 24 diagonal blocks (lower triangles) and has its scheduler (and FP64 pipe quarter) to itself: the
//     tile warps are the warps 1-3, 5-7, 9-11, 13-15.  In the update phase of column j its lane of residue j + 1 updates
//     and then factorises the next pivot -- the only serial part of the whole factorisation -- while the tile warps are
//     busy with their 216 DFMA per thread (measured: one warp alone saturates its scheduler's FP64 pipe,
//     tools/ubench_dfma_warps.cu, so the pivot chain would otherwise queue behind them);
//   * per column: [panel] 138 threads solve one row each of the column's blocks against L_jj (handed over through
//     shared memory by their owners), results go to shared memory transposed ([m][residue][6]) and to Lcol; one more
//     thread finishes the right-hand side of block j.  barrier.  [update] as above; the owners of column j + 1 hand
//     their blocks over.  barrier.
// Output layout identical to k_chol_band (Lcol, Ldiag, b), so k_chol_band_back serves both.
constexpr int kChol2Threads = 384;        // 12 warps at <= 168 registers: warp 0 = diagonal blocks, nine tile warps, two that only help with the panel
constexpr int kChol2PanelT0 = 192;        // the panel rows are solved by threads 192 .. 330

// block (i, k) of S (zeros outside the matrix) -> a thread's prefetch slot, asynchronously
__device__ __forceinline__ void chol2_prefetch(const double* __restrict__ S, int n, int nb, int i, int k, double* slot) {
    if (i < nb && k < nb) {
        const double* src = S + (size_t)(6 * i) * n + 6 * k;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(slot);
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int h = 0; h < 3; ++h)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst + (r * 6 + 2 * h) * 8), "l"(src + (size_t)r * n + 2 * h) : "memory");
    } else {
#pragma unroll
        for (int e = 0; e < 18; ++e) reinterpret_cast<double2*>(slot)[e] = make_double2(0.0, 0.0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}
// slot (row-major block) -> registers
__device__ __forceinline__ void chol2_take(double (&blk)[36], const double* slot) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
    for (int e = 0; e < 18; ++e) { const double2 v = reinterpret_cast<const double2*>(slot)[e]; blk[2 * e] = v.x; blk[2 * e + 1] = v.y; }
}
__device__ __forceinline__ void chol2_store(const double (&blk)[36], double* dst) {
#pragma unroll
    for (int e = 0; e < 18; ++e) reinterpret_cast<double2*>(dst)[e] = make_double2(blk[2 * e], blk[2 * e + 1]);
}
// dg = lower triangle (packed, dg[i * (i + 1) / 2 + k]) of the updated diagonal block (j, j): right-looking in-place
// Cholesky (the dependent chain per step is rsqrt, one product, one update); L_jj and the reciprocal pivots go to
// shared memory for the panel and to Ldiag; false if not positive definite.  (The inverse of L_jj, which would turn the
// panel rows into independent dot products, costs this chain more than it saves the panel: measured, removed.)
__device__ __forceinline__ bool chol2_factor(double (&dg)[21], double* sL, double* sInv, double* __restrict__ Ldiag_j) {
    bool ok = true;
    double inv[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
        double d = dg[j * (j + 1) / 2 + j];
        if (!(d > 0.0)) { ok = false; d = 1.0; }
        const double iv = vio_rsqrt(d);
        inv[j] = iv;
        dg[j * (j + 1) / 2 + j] = d * iv;
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (i > j) dg[i * (i + 1) / 2 + j] *= iv;
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
            for (int k = 0; k < 6; ++k)
                if (i > j && k > j && k <= i) dg[i * (i + 1) / 2 + k] -= dg[i * (i + 1) / 2 + j] * dg[k * (k + 1) / 2 + j];
    }
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int k = 0; k < 6; ++k) { const double v = k <= i ? dg[i * (i + 1) / 2 + k] : 0.0; sL[i * 6 + k] = v; Ldiag_j[i * 6 + k] = v; }
#pragma unroll
    for (int e = 0; e < 6; ++e) { sInv[e] = inv[e]; Ldiag_j[36 + e] = inv[e]; }
    return ok;
}

__global__ void __maxnreg__(168) k_chol_band_reg(const double* __restrict__ S, int nb, double* __restrict__ Lcol,
                                                  double* __restrict__ Ldiag, double* __restrict__ b, int* __restrict__ status) {
    extern __shared__ double sm2[];
    double* panelT = sm2;                               // [6][Wb][6]: panelT[m][res][r] = P_res[r][m]
    double* Acol = panelT + 6 * kCholWb * 6;            // [Wb][36]: blocks (i, j + 1), i by residue, for the next panel phase
    double* pref = Acol + kCholWb * 36;                 // [threads][36]: every owner's NEXT block, prefetched with cp.async
    double* sb = pref + kChol2Threads * 36;             // [n]
    __shared__ __align__(16) double sL[36], sInv[6], sY[6];   // L_jj and its reciprocal pivots, y_j
    __shared__ int sbad;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, n = 6 * nb;
    const bool diag = warp == 0 && lane < kCholWb;
    // Residue pair of this thread.  The 24 residues form three groups of 8; nine warps hold the 276 off-diagonal pairs:
    // three the 28 pairs inside one group (8 distinct rows of P per load), six an 8 x 4 tile between two groups (12 rows).
    // Warp w runs on scheduler w % 4: the diagonal warp shares scheduler 0 with two tile warps only, the others get
    // 3 + 2 + 2, so that the serial pivot chain never queues behind three streams of DFMA.
    int a = lane, bb = lane;
    bool bulk = false;
    {
        const int tile_of_warp[12] = {-1, 2, 5, 7, 0, 3, 6, 8, 1, 4, -1, -1};
        const int t = tile_of_warp[warp];
        if (t >= 0 && t < 3) {   // pairs inside group t: lane -> (x, y), x > y
            int x = 1;
            while ((x + 1) * x / 2 <= lane) ++x;
            const int y = lane - x * (x - 1) / 2;
            a = 8 * t + x; bb = 8 * t + y;
            bulk = lane < 28;
        } else if (t >= 3) {     // tiles 3, 4: groups 1 x 0; 5, 6: 2 x 0; 7, 8: 2 x 1 (two halves of the lower group each)
            const int q = (t - 3) >> 1, ga = q == 0 ? 1 : 2, gb = q == 2 ? 1 : 0;
            a = 8 * ga + (lane >> 2); bb = 8 * gb + 4 * ((t - 3) & 1) + (lane & 3);
            bulk = true;
        }
    }
    const int prow = tid - kChol2PanelT0;               // panel row of this thread (0 .. 138 take part)
    double* mine = pref + tid * 36;
    if (tid == 0) sbad = 0;
    for (int e = tid; e < n; e += kChol2Threads) sb[e] = b[e];
    double blk[36];   // tile warps: the block, row-major; diagonal lanes: dg = the packed lower triangle in blk[0..20]
    double (&dg)[21] = reinterpret_cast<double (&)[21]>(blk);
    // column 0: block (a, b), the block after it is (b + 24, a); a diagonal owner moves down the diagonal
    if (diag) {
        chol2_prefetch(S, n, nb, a, a, mine);
        asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
            for (int k = 0; k <= i; ++k) dg[i * (i + 1) / 2 + k] = mine[i * 6 + k];
        chol2_prefetch(S, n, nb, a + kCholWb, a + kCholWb, mine);
        if (a == 0 && !chol2_factor(dg, sL, sInv, Ldiag)) sbad = 1;
    } else if (bulk) {
        chol2_prefetch(S, n, nb, a, bb, mine);
        chol2_take(blk, mine);
        chol2_prefetch(S, n, nb, bb + kCholWb, a, mine);
        if (bb == 0) chol2_store(blk, Acol + a * 36);
    }
    __syncthreads();
    for (int j = 0; j < nb; ++j) {
        const int jm = j % kCholWb;
        // ---- panel: one row of the column's blocks per thread, X L_jj' = A by substitution; row 138 finishes rhs block j
        if (prow >= 0 && prow <= (kCholWb - 1) * 6) {
            const bool rhs = prow == (kCholWb - 1) * 6;
            const int bi = prow / 6, ar = prow - bi * 6;
            const int res = (jm + 1 + bi) % kCholWb;
            const double* src = rhs ? sb + 6 * j : Acol + res * 36 + ar * 6;
            double o[6];
#pragma unroll
            for (int c = 0; c < 6; ++c) {   // o L_jj' = v by substitution
                double v = src[c];
#pragma unroll
                for (int k = 0; k < c; ++k) v -= o[k] * sL[c * 6 + k];
                o[c] = v * sInv[c];
            }
            if (rhs) {
#pragma unroll
                for (int c = 0; c < 6; ++c) { sb[6 * j + c] = o[c]; sY[c] = o[c]; }
            } else {
#pragma unroll
                for (int c = 0; c < 6; ++c) {
                    panelT[(c * kCholWb + res) * 6 + ar] = o[c];
                    Lcol[((size_t)j * (kCholWb - 1) + bi) * 36 + ar * 6 + c] = o[c];
                }
            }
        }
        __syncthreads();
        // ---- update
        if (diag) {
            const int dd = (a - jm + kCholWb) % kCholWb;   // this lane's block is (j + dd, j + dd)
            if (dd >= 1) {
                const double* pd = panelT + a * 6;
#pragma unroll
                for (int m = 0; m < 6; ++m) {
                    const double2 r0 = *reinterpret_cast<const double2*>(pd + m * kCholWb * 6), r1 = *reinterpret_cast<const double2*>(pd + m * kCholWb * 6 + 2),
                                  r2 = *reinterpret_cast<const double2*>(pd + m * kCholWb * 6 + 4);
                    const double rv[6] = {r0.x, r0.y, r1.x, r1.y, r2.x, r2.y};
#pragma unroll
                    for (int r = 0; r < 6; ++r)
#pragma unroll
                        for (int c = 0; c <= r; ++c) dg[r * (r + 1) / 2 + c] -= rv[r] * rv[c];
                }
                if (dd == 1 && j + 1 < nb) {   // the next pivot: the serial part, overlapped with the other warps' updates
                                if (!chol2_factor(dg, sL, sInv, Ldiag + (size_t)(j + 1) * 42)) sbad = 1;
                            }
            } else {   // (j, j) is finished: (j + 24, j + 24) enters, (j + 48, j + 48) is prefetched
                asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
                for (int i = 0; i < 6; ++i)
#pragma unroll
                    for (int k = 0; k <= i; ++k) dg[i * (i + 1) / 2 + k] = mine[i * 6 + k];
                chol2_prefetch(S, n, nb, j + 2 * kCholWb, j + 2 * kCholWb, mine);
            }
        } else {
            if (prow >= 0 && prow < (kCholWb - 1) * 6) {   // right-hand side rows below block j: b -= P y_j
                const int row = 6 * (j + 1) + prow;
                if (row < n) {
                    const int res = (jm + 1 + prow / 6) % kCholWb, ar = prow % 6;
                    double acc = 0.0;
#pragma unroll
                    for (int m = 0; m < 6; ++m) acc += panelT[(m * kCholWb + res) * 6 + ar] * sY[m];
                    sb[row] -= acc;
                }
            }
            if (bulk) {
                const int da = (a - jm + kCholWb) % kCholWb, db = (bb - jm + kCholWb) % kCholWb;
                const int di = max(da, db), dk = min(da, db);
                if (dk != 0) {
                    // block (i, k) = (j + di, j + dk): rows of P_i times rows of P_k; the residue of i is a's when da > db
                    const double* pr = panelT + (da > db ? a : bb) * 6;
                    const double* pc = panelT + (da > db ? bb : a) * 6;
#pragma unroll
                    for (int m = 0; m < 6; ++m) {
                        const double2 r0 = *reinterpret_cast<const double2*>(pr + m * kCholWb * 6), r1 = *reinterpret_cast<const double2*>(pr + m * kCholWb * 6 + 2),
                                      r2 = *reinterpret_cast<const double2*>(pr + m * kCholWb * 6 + 4);
                        const double2 c0 = *reinterpret_cast<const double2*>(pc + m * kCholWb * 6), c1 = *reinterpret_cast<const double2*>(pc + m * kCholWb * 6 + 2),
                                      c2 = *reinterpret_cast<const double2*>(pc + m * kCholWb * 6 + 4);
                        const double rv[6] = {r0.x, r0.y, r1.x, r1.y, r2.x, r2.y}, cv[6] = {c0.x, c0.y, c1.x, c1.y, c2.x, c2.y};
#pragma unroll
                        for (int r = 0; r < 6; ++r)
#pragma unroll
                            for (int c = 0; c < 6; ++c) blk[r * 6 + c] -= rv[r] * cv[c];
                    }
                } else {
                    // block (j + di, j) is finished: take up (j + 24, j + di), prefetch the one after it, (j + di + 24, j + 24)
                    chol2_take(blk, mine);
                    chol2_prefetch(S, n, nb, j + di + kCholWb, j + kCholWb, mine);
                }
                // blocks of column j + 1 join the next panel: (j + di, j + 1) after its update, (j + 24, j + 1) as taken up
                if (dk == 1 || (dk == 0 && di == 1)) chol2_store(blk, Acol + (dk == 1 ? (da > db ? a : bb) : jm) * 36);
            }
        }
        __syncthreads();
    }
    for (int e = tid; e < n; e += kChol2Threads) b[e] = sb[e];
    if (tid == 0) status[0] = sbad;
}

// Column-oriented backward substitution L^T x = y for the factor of k_chol_band_reg (which writes all 23 block rows of a
// column, zeros beyond its reach).  k_chol_band_back reduces 138 rows per
// column on one warp (0.7 us per column).  Here the sums are kept per block and updated in parallel the moment an x_j is
// known: thread (q, c) adds (L(j, k)' x_j)[c] to the pending sum of block k = j - 1 - q, a 6-term dot product; the sum
// of block j - 1 is then complete, and the six lanes that hold it go straight on to x_{j-1} = L^-T (y - sum) with the
// stored inverse.  One barrier per column; the next column's 6 + 6 factor entries are in flight meanwhile
// (two columns ahead measured the same: the chain, not the loads, sets the 0.44 us per column).
constexpr int kCholBackThreads = 160;   // 23 x 6 = 138 workers
__global__ void __launch_bounds__(kCholBackThreads) k_chol_band_back_col(int nb, const double* __restrict__ Lcol, const double* __restrict__ Ldiag,
                                                                         double* __restrict__ Linv, double* __restrict__ b) {
    extern __shared__ double sbk[];
    double* sx = sbk;            // [n]: y in, x out
    double* sacc = sbk + 6 * nb; // [n]: pending sums
    const int tid = threadIdx.x, n = 6 * nb;
    const int q = tid / 6, c = tid - 6 * q;
    const bool worker = q < kCholWb - 1;
    for (int e = tid; e < n; e += kCholBackThreads) { sx[e] = b[e]; sacc[e] = 0.0; }
    // the inverses of all diagonal factors, in parallel (one thread per block column): off the factorisation's chain
    for (int j = tid; j < nb; j += kCholBackThreads) {
        const double* L = Ldiag + (size_t)j * 42;
        double li[36];
#pragma unroll
        for (int e = 0; e < 36; ++e) li[e] = 0.0;
#pragma unroll
        for (int cc = 0; cc < 6; ++cc)
#pragma unroll
            for (int i = 0; i < 6; ++i)
                if (i >= cc) {
                    double sacc2 = (i == cc) ? 1.0 : 0.0;
#pragma unroll
                    for (int k = 0; k < 6; ++k)
                        if (k >= cc && k < i) sacc2 -= L[i * 6 + k] * li[k * 6 + cc];
                    li[i * 6 + cc] = sacc2 * L[36 + i];
                }
#pragma unroll
        for (int e = 0; e < 36; ++e) Linv[(size_t)j * 36 + e] = li[e];
    }
    __syncthreads();   // Linv is read back through global memory by other threads of this CTA
    // column c of the block L(j, j - 1 - q) = Lcol[(j - 1 - q) * 23 + q][a][c], a = 0 .. 5
    auto fetch_l = [&](int j, double* dst) {
        const int k = j - 1 - q;
#pragma unroll
        for (int a = 0; a < 6; ++a) dst[a] = (worker && k >= 0) ? Lcol[((size_t)k * (kCholWb - 1) + q) * 36 + a * 6 + c] : 0.0;
    };
    // column t = tid of L_jj^-1 (rows k >= t), for the six solving lanes
    auto fetch_inv = [&](int j, double* dst) {
#pragma unroll
        for (int k = 0; k < 6; ++k) dst[k] = (tid < 6 && j >= 0) ? Linv[(size_t)j * 36 + k * 6 + tid] : 0.0;
    };
    double lcur[6], lnxt[6], icur[6], inxt[6];
    fetch_l(nb - 1, lcur);
    fetch_inv(nb - 1, icur);
    __syncthreads();
    if (tid < 6) {   // x_{nb-1} = L^-T y
        double xv = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k) xv += icur[k] * sx[6 * (nb - 1) + k];
        __syncwarp(0x3f);
        sx[6 * (nb - 1) + tid] = xv;
    }
    fetch_inv(nb - 2, icur);
    __syncthreads();
    for (int j = nb - 1; j > 0; --j) {
        fetch_l(j - 1, lnxt);
        fetch_inv(j - 2, inxt);
        // pending sums of the blocks j - 23 .. j - 1 take x_j in
        const int k = j - 1 - q;
        double s0 = 0.0;
        if (worker && k >= 0) {
            const double* xj = sx + 6 * j;
            double s1 = 0.0;
            s0 = lcur[0] * xj[0]; s1 = lcur[1] * xj[1];
            s0 += lcur[2] * xj[2]; s1 += lcur[3] * xj[3];
            s0 += lcur[4] * xj[4]; s1 += lcur[5] * xj[5];
            s0 = sacc[6 * k + c] + (s0 + s1);
            if (q > 0) sacc[6 * k + c] = s0;
        }
        if (tid < 6) {   // q == 0: block j - 1 is complete; x_{j-1} = L^-T (y - sum), the six components across six lanes
            const double sv = sx[6 * (j - 1) + tid] - s0;
            double xv = 0.0;
#pragma unroll
            for (int kk = 0; kk < 6; ++kk) xv += icur[kk] * __shfl_sync(0x3fu, sv, kk);
            sx[6 * (j - 1) + tid] = xv;
        }
#pragma unroll
        for (int a = 0; a < 6; ++a) { lcur[a] = lnxt[a]; icur[a] = inxt[a]; }
        __syncthreads();
    }
    for (int e = tid; e < n; e += kCholBackThreads) b[e] = sx[e];
}

// Backward substitution L^T x = y for the banded factor: one warp, the next column's data is
// fetched while the current one is reduced (the chain is nb dependent 6x6 solves).
__global__ void __launch_bounds__(32) k_chol_band_back(int nb, const int* __restrict__ reach, const double* __restrict__ Lcol,
                                                       const double* __restrict__ Ldiag, double* __restrict__ b) {
    extern __shared__ double sx[];   // [n]
    const int lane = threadIdx.x, n = 6 * nb;
    for (int e = lane; e < n; e += 32) sx[e] = b[e];
    __syncwarp();
    constexpr int kRowsPerLane = ((kCholWb - 1) * 6 + 31) / 32;   // 5
    double cur[kRowsPerLane][6], nxt[kRowsPerLane][6];
    double dcur[2], dnxt[2];   // lanes 0..20 hold L_jj (36) and the reciprocal pivots (6): two doubles each
    auto fetch = [&](int j, double (*dst)[6], double* dd) {
        const int rows = 6 * (reach[j] - j);
        const double* col = Lcol + (size_t)j * (kCholWb - 1) * 36;
#pragma unroll
        for (int q = 0; q < kRowsPerLane; ++q) {
            const int row = lane + 32 * q;
#pragma unroll
            for (int c = 0; c < 6; ++c) dst[q][c] = row < rows ? col[row * 6 + c] : 0.0;
        }
        dd[0] = lane < 21 ? Ldiag[(size_t)j * 42 + 2 * lane] : 0.0;
        dd[1] = lane < 21 ? Ldiag[(size_t)j * 42 + 2 * lane + 1] : 0.0;
    };
    if (nb > 0) fetch(nb - 1, cur, dcur);
    for (int j = nb - 1; j >= 0; --j) {
        if (j > 0) fetch(j - 1, nxt, dnxt);
        const int rows = 6 * (reach[j] - j);
        double acc[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int q = 0; q < kRowsPerLane; ++q) {
            const int row = lane + 32 * q;
            const double xv = row < rows ? sx[6 * (j + 1) + row] : 0.0;
#pragma unroll
            for (int c = 0; c < 6; ++c) acc[c] += cur[q][c] * xv;
        }
#pragma unroll
        for (int c = 0; c < 6; ++c)
#pragma unroll
            for (int sft = 16; sft > 0; sft >>= 1) acc[c] += __shfl_xor_sync(0xffffffffu, acc[c], sft);
        // x_c = (y_c - acc_c - sum_{k > c} L[k][c] x_k) / L[c][c], every lane redundantly (values via shuffles)
        double x[6];
#pragma unroll
        for (int c = 5; c >= 0; --c) {
            double v = sx[6 * j + c] - acc[c];
#pragma unroll
            for (int k = c + 1; k < 6; ++k) {
                const int e = k * 6 + c;   // L[k][c] sits in lane e / 2, element e % 2
                v -= __shfl_sync(0xffffffffu, dcur[e & 1], e >> 1) * x[k];
            }
            x[c] = v * __shfl_sync(0xffffffffu, dcur[(36 + c) & 1], (36 + c) >> 1);
        }
        __syncwarp();
        if (lane < 6) sx[6 * j + lane] = x[lane];
        __syncwarp();
#pragma unroll
        for (int q = 0; q < kRowsPerLane; ++q)
#pragma unroll
            for (int c = 0; c < 6; ++c) cur[q][c] = nxt[q][c];
        dcur[0] = dnxt[0]; dcur[1] = dnxt[1];
    }
    __syncwarp();
    for (int e = lane; e < n; e += 32) b[e] = sx[e];
}

// one warp per landmark of [lm_begin, lm_end): dl = (E'E + D^2)^-1 sum_i E_i'(r_i - F_i x_p(i))
__global__ void __launch_bounds__(128) k_back_substitute(BaDev d, int lm_begin, int lm_end, bool apply_loss, const double* __restrict__ x,
                                                         double* __restrict__ dl) {
    const int l = lm_begin + (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (l >= lm_end) return;
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};   // E'E upper (6), E's (3)
    const double Pw[3] = {d.points[3 * (size_t)l], d.points[3 * (size_t)l + 1], d.points[3 * (size_t)l + 2]};
    for (int q = d.lm_ptr[l] + lane; q < d.lm_ptr[l + 1]; q += 32) {
        const int p = d.pose[q];
        double r[2], F[12], E[6];
        obs_evaluate<true>(d.posebuf + (size_t)kPoseStride * p, Pw, d.u[q], d.v[q], d.outlier[q] != 0, d.cam, apply_loss, r, F, E);
        const int fm = d.free_map[p];
        double s0 = r[0], s1 = r[1];
        if (fm >= 0) {
#pragma unroll
            for (int k = 0; k < 6; ++k) { s0 -= F[k] * x[6 * fm + k]; s1 -= F[6 + k] * x[6 * fm + k]; }
        }
        acc[0] += E[0] * E[0] + E[3] * E[3]; acc[1] += E[0] * E[1] + E[3] * E[4]; acc[2] += E[0] * E[2] + E[3] * E[5];
        acc[3] += E[1] * E[1] + E[4] * E[4]; acc[4] += E[1] * E[2] + E[4] * E[5]; acc[5] += E[2] * E[2] + E[5] * E[5];
        acc[6] += E[0] * s0 + E[3] * s1; acc[7] += E[1] * s0 + E[4] * s1; acc[8] += E[2] * s0 + E[5] * s1;
    }
#pragma unroll
    for (int j = 0; j < 9; ++j) acc[j] = warp_sum(acc[j]);
    if (lane != 0) return;
    double out[3] = {0.0, 0.0, 0.0};
    if (d.point_fixed[l] == 0) {
        if (d.D) {
            const double* Dl = d.D + 6 * (size_t)d.K + 3 * (size_t)l;
            acc[0] += Dl[0] * Dl[0]; acc[3] += Dl[1] * Dl[1]; acc[5] += Dl[2] * Dl[2];
        }
        const double i00 = vio_rsqrt(acc[0]);
        const double l10 = acc[1] * i00, l20 = acc[2] * i00;
        const double i11 = vio_rsqrt(acc[3] - l10 * l10);
        const double l21 = (acc[4] - l20 * l10) * i11;
        const double i22 = vio_rsqrt(acc[5] - l20 * l20 - l21 * l21);
        const double i10 = -l10 * i00 * i11, i21 = -l21 * i11 * i22, i20 = -(l20 * i00 + l21 * i10) * i22;
        if (isfinite(i00) && isfinite(i11) && isfinite(i22) && isfinite(i10) && isfinite(i21) && isfinite(i20)) {
            const double y0 = i00 * acc[6], y1 = i10 * acc[6] + i11 * acc[7], y2 = i20 * acc[6] + i21 * acc[7] + i22 * acc[8];
            out[0] = i00 * y0 + i10 * y1 + i20 * y2;
            out[1] = i11 * y1 + i21 * y2;
            out[2] = i22 * y2;
        }
    }
    dl[3 * (size_t)l] = out[0]; dl[3 * (size_t)l + 1] = out[1]; dl[3 * (size_t)l + 2] = out[2];
}

}  // namespace vio360
