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
