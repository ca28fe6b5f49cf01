// ba_schur.cuh -- K7+K8+K9 fused without global atomics: factor evaluation, Gauss-Newton block
// accumulation and per-landmark Schur elimination for a *chunk* of landmarks whose
// observing poses fit a window of kSWin consecutive poses.
//
// Replaces the ParallelFor over ResidualBlock::Evaluate -> BAFactor::Evaluate
// (thirdparty/ceres-solver/internal/ceres/program_evaluator.h:187-258; src/optimization/
// Factors.cpp:327-542) fused with SchurEliminator<2,3,6>::Eliminate
// (schur_eliminator_impl.h:179-308).  Where Ceres guards every cell of the reduced camera
// system with a std::mutex (:211,414,551), a CTA here owns a private 144 x 144 tile of S in
// shared memory for its pose window:
//
//   phase A  one thread per observation: residual, Jacobians F (2x6), E (2x3), Huber
//   phase B  per landmark: E'E (+D^2), g = E'r by a segmented sum; Cholesky E'E = L L';
//            L^-1 and (E'E)^-1 g
//   phase C  per observation: F'F and F'(r - E (E'E)^-1 g) summed per pose through a slot
//            table (no atomics), Z = (F'E) L^-T written to the operand stage
//   phase D  S_tile -= Z Z'  on the FP64 tensor pipe (mma.sync m8n8k4.f64 = DMMA): the
//            per-landmark 6x3 * 3x3 * 3x6 products of all pose pairs are one SYRK over the
//            concatenated landmark columns
//
// Every chunk writes its tile to a private slab; k_schur_reduce sums the slabs that cover a
// cell in a fixed order, so the result is deterministic.  Landmarks that do not fit a window
// (span > kSWin poses, or two observations from one pose) take k_schur_warp in ba.cu.
#pragma once

#include "ba_math.cuh"

namespace vio360 {

constexpr int kSWin = 24;                          // poses per chunk window
constexpr int kST = 6 * kSWin;                     // 144 tile rows
constexpr int kSNT = kST / 8;                      // 18 DMMA tiles per dimension
constexpr int kSTiles = kSNT * (kSNT + 1) / 2;     // 171 stored (upper) tiles
constexpr int kSThreads = 512;                     // observations per batch
constexpr int kSWarps = kSThreads / 32;
constexpr int kSMaxLm = 24;                        // landmarks per batch
constexpr int kSZPitch = 84;                       // doubles per stage row: 3*kSMaxLm padded, == 4 (mod 16)
constexpr int kSRed = 9;                           // doubles per observation in the reduction scratch

struct SchurChunk { int lm_begin, lm_end, pbase, span; };

constexpr size_t kSchurSmemBytes =
    sizeof(double) * ((size_t)kST * kSZPitch + (size_t)kSTiles * 64 + (size_t)kSWin * kPoseStride + (size_t)kSMaxLm * 9 + 32) +
    sizeof(short) * (size_t)kSMaxLm * kSWin + sizeof(int) * 32;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__device__ __forceinline__ int schur_tile_index(int ti, int tj) {   // ti <= tj
    return ti * kSNT - (ti * (ti - 1)) / 2 + (tj - ti);
}
// address (in doubles) of element (row, col) of the tile store; requires row/8 <= col/8
__device__ __forceinline__ int schur_elem_index(int row, int col) {
    return schur_tile_index(row >> 3, col >> 3) * 64 + (((row & 7) << 2) + ((col & 7) >> 1)) * 2 + (col & 1);
}

}  // namespace vio360
