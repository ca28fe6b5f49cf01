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
// The CTA is warp-specialised.  Twelve producer warps own a dense (pose slot, landmark slot)
// grid of 24 x 16 threads per batch of 16 landmarks:
//   phase A  one thread per observation: residual, Jacobians F (2x6), E (2x3), Huber
//   phase B  per landmark: E'E (+D^2), g = E'r summed over the pose slots; Cholesky
//            E'E = L L'; L^-1 and (E'E)^-1 g
//   phase C  per observation: F'F and F'(r - E (E'E)^-1 g); the 16 landmark lanes of a pose
//            slot are one half warp, so the per-pose sums are a shuffle reduce-scatter (no
//            atomics, no shared memory); Z = (F'E) L^-T goes to one of two operand stages
// while four consumer warps run
//   phase D  S_tile -= Z Z'  on the FP64 tensor pipe (mma.sync m8n8k4.f64 = DMMA): the
//            per-landmark 6x3 * 3x3 * 3x6 products of all pose pairs are one SYRK over the
//            concatenated landmark columns of the other stage.
// Stages are handed over with named barriers (bar.arrive / bar.sync), so the latency-bound
// evaluation and the pipe-bound SYRK overlap.
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
constexpr int kSMaxLm = 16;                        // landmarks per batch
constexpr int kSProd = kSWin * kSMaxLm;            // 384 producer threads: thread = (pose slot, landmark slot)
constexpr int kSCons = 128;                        // 4 consumer warps issue the DMMA
constexpr int kSThreads = kSProd + kSCons;         // 512
constexpr int kSK = 3 * kSMaxLm;                   // 48 stage columns per batch (12 DMMA k-steps)
constexpr int kSZPitch = 148;                      // doubles per stage column (>= kST, == 4 mod 16: conflict-free fragments)
constexpr int kSRed = 9;                           // doubles per observation in the reduction scratch

constexpr int kSGrp = kSWin / 2;                   // poses per group of a wide landmark's jobs
constexpr int kChunkWide = 1;                      // flags: (E'E + D^2)^-1 of the landmarks comes from k_wide_prepare
constexpr int kChunkCross = 2;                     //        two pose groups; only the blocks between them are this chunk's

// lm_begin .. lm_end: positions in the chunk's landmark list (batch_lm[(batch0 + b) * kSMaxLm + ls] = landmark id, -1
// = empty slot); batch0: first row of the slot table that belongs to this chunk (one row of kSProd observation
// indices, -1 = empty, per batch of kSMaxLm landmarks; built on the host in ba.cu:build_chunks).  pose slots
// 0 .. span-1 hold poses pbase ..; for a cross chunk slots kSGrp .. kSGrp+spanB-1 hold poses pbaseB ..
struct SchurChunk { int lm_begin, lm_end, pbase, span, batch0, pbaseB, spanB, flags; };

constexpr size_t kSchurSmemBytes =
    sizeof(double) * (2 * (size_t)kSK * kSZPitch + (size_t)kSTiles * 64 + (size_t)kSWin * kPoseStride + (size_t)kSMaxLm * 9 + 32) +
    sizeof(int) * 32 + sizeof(float) * ((size_t)kSWin * kPoseStride + 8);   // + the float copy of the window's poses (fp32 evaluation mode)

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
