// ba_peer.cuh -- reduced-camera-system exchange over NVLink / NVSwitch peer memory.
//
// Every rank of a sharded window eliminates its own landmark range and leaves a partial block band
// [block row][block column - row < w][6 x 6] | rhs | cost in a buffer the other ranks can map
// (cudaIpc handle across processes, plain peer access inside one process).  The collective has a PRODUCER and a
// CONSUMER side, and neither contains a grid-wide wait:
//   publish  k_schur_reduce, the kernel that sums a rank's chunk slabs into its band, counts the finished blocks of
//            every block row; the CTA that completes a row releases it to all ranks (fence.sys + st.release.sys of the
//            epoch into each rank's row-flag array, a few remote 4-byte stores per row).  It never waits.
//   pull     k_band_pull (one launch per rank, right behind its own reduce): a CTA waits for the row flags of the rows
//            its elements belong to -- local memory polls, normally already set, since the peers run in lock step --
//            then every thread PULLS its element from the ranks whose row range covers it, sums in rank order (all
//            ranks hold bit-identical results, independent of arrival order), adds the pose part of the
//            Levenberg-Marquardt diagonal and writes the dense mirrored [S | rhs | cost] the solver reads: the
//            allreduce, the mirror and the expansion in one pass, no staging copy.
// A rank's pull can run while a slower peer is still reducing its last rows, and no CTA ever waits for a CTA of its
// own grid or for a kernel that itself waits (no residency dead-lock, whatever else occupies the SMs).
//
// A rank's landmarks only touch a contiguous range of block rows (landmark-range shards of a
// co-visibility band), so a block row is read only from the ranks whose range covers it: the bytes
// crossing NVLink stay ~ one band in total instead of (world - 1) bands per rank.
//
// Buffer discipline: two band buffers per rank, used alternately (epoch parity).  Rank A overwrites
// the buffer of epoch e when it builds epoch e + 2, i.e. after its pull e + 1 returned, which waited for every
// covering peer's row flags e + 1, which a peer releases only in its reduce e + 1, i.e. after its own pull e (the
// last reader of A's buffer e) has completed in stream order.  No trailing barrier is needed.
#pragma once

#include <cstdint>

namespace vio360 {

constexpr int kMaxPeers = 8;
constexpr size_t kPeerHeaderBytes = 1024;               // trace words; the row flags [src rank][block row] follow it
constexpr long long kPeerTimeoutCycles = 4000000000ll;  // ~2 s at 1.9 GHz: a missing peer fails loudly, it never hangs the box

struct PeerView {
    const double* band[kMaxPeers];   // this epoch's partial band of every rank (own pointer for own rank)
    const unsigned* row_flags;       // THIS rank's row flags [src rank][n_free]: last epoch rank src released row r for
    int row_lo[kMaxPeers], row_hi[kMaxPeers];   // block rows (free-pose index) a rank's band covers
    int world, rank;
    unsigned* announce[kMaxPeers];   // mode 3: every rank's flag word of THIS rank (row_flags[r] + rank * n_free + row_lo[rank]); null otherwise
};

// what the producer side (k_schur_reduce) needs to release finished block rows to every rank
struct PeerPublish {
    unsigned* row_flags[kMaxPeers];  // every rank's row-flag array (remote pointers; own for own rank)
    unsigned* row_count;             // local: blocks of a row summed so far, over all epochs
    int world, rank, n_free, band_w;
    unsigned epoch;                  // 0: not a peer build
    unsigned builds;                 // peer builds of this rank so far, this one included (row_count counts over all of them)
    int mode;                        // 0: st.release.sys per rank and row; 1: one fence.sys, then relaxed stores (default);
                                     // 2: one release per BUILD (the CTA that completes the last row releases every row)
    int row_lo, row_hi;              // this rank's block rows
};

__device__ __forceinline__ double ld_peer_f64(const double* p) {
    double v;
    asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys_u32(unsigned* p, unsigned v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// called by every thread of a reduce CTA after it stored block (bi, .) of the band
__device__ __forceinline__ void peer_publish_block(const PeerPublish& pub, int bi) {
    if (pub.epoch == 0u) return;
    __syncthreads();                                       // the CTA's stores are done
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned per_row = (unsigned)min(pub.band_w, pub.n_free - bi);
        const unsigned old = atomicAdd(pub.row_count + bi, 1u);
        if (old + 1u != pub.builds * per_row) return;      // not the CTA that completed the row
        if (pub.mode == 2) {
            // rank-level release: count completed rows in the slot behind the row counters
            const unsigned nrows = (unsigned)(pub.row_hi - pub.row_lo);
            __threadfence();
            const unsigned done = atomicAdd(pub.row_count + pub.n_free, 1u);
            if (done + 1u != pub.builds * nrows) return;
            __threadfence_system();
            for (int r = 0; r < pub.world; ++r)
                for (int row = pub.row_lo; row < pub.row_hi; ++row)
                    st_relaxed_sys_u32(pub.row_flags[r] + (size_t)pub.rank * pub.n_free + row, pub.epoch);
            return;
        }
        if (pub.mode == 1) {
            __threadfence_system();                        // the row's blocks (all CTAs: their fences precede their counts) are visible system wide
            for (int r = 0; r < pub.world; ++r) st_relaxed_sys_u32(pub.row_flags[r] + (size_t)pub.rank * pub.n_free + bi, pub.epoch);
            return;
        }
        __threadfence_system();
        for (int r = 0; r < pub.world; ++r)
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pub.row_flags[r] + (size_t)pub.rank * pub.n_free + bi), "r"(pub.epoch) : "memory");
    }
}

// Every thread pulls its elements of the band from the ranks whose rows cover them, sums in rank order, adds the
// LM diagonal and writes the dense mirrored [S | rhs | cost] (grid-stride over the band).
__device__ __forceinline__ void band_pull_sum_expand(const PeerView& pv, const int n_free, const int band_w,
                                                     const int* __restrict__ free_pose, const double* __restrict__ D,
                                                     double* __restrict__ packed) {
    const int n = 6 * n_free;
    const size_t nS = (size_t)n_free * band_w * 36, total = nS + n + 1;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        int row, bi = 0, bj = 0, e = 0;
        if (idx < nS) {
            const int blk = (int)(idx / 36);
            e = (int)(idx - (size_t)blk * 36);
            bi = blk / band_w; bj = bi + (blk - bi * band_w);
            if (bj >= n_free) continue;        // slots past the last block column
            row = bi;
        } else {
            row = idx < nS + n ? (int)(idx - nS) / 6 : -1;   // rhs of a pose : the cost (every rank)
        }
        double v[kMaxPeers];
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)
            v[p] = (p < pv.world && (row < 0 || (row >= pv.row_lo[p] && row < pv.row_hi[p]))) ? ld_peer_f64(pv.band[p] + idx) : 0.0;
        double s = v[0];
#pragma unroll
        for (int p = 1; p < kMaxPeers; ++p) s += v[p];   // rank order on every rank
        if (idx < nS) {
            const int a = e / 6, b = e - 6 * a;
            const size_t ri = 6 * (size_t)bi + a, rj = 6 * (size_t)bj + b;
            if (bi == bj) {
                if (a == b && D) { const double dd = D[6 * (size_t)free_pose[bi] + a]; s += dd * dd; }
                packed[ri * n + rj] = s;       // the band holds both halves of a diagonal block
            } else {
                packed[ri * n + rj] = s;
                packed[rj * n + ri] = s;
            }
        } else {
            packed[(size_t)n * n + (idx - nS)] = s;
        }
    }
}

__global__ void __launch_bounds__(256)
k_band_pull(const PeerView pv, const unsigned epoch, const int n_free, const int band_w,
            const int* __restrict__ free_pose, const double* __restrict__ D,
            double* __restrict__ packed, int* __restrict__ err_flag, unsigned long long* __restrict__ trace) {
    pdl_wait();   // launched with programmatic dependent launch behind this rank's k_schur_reduce: its band is complete from here on
    unsigned long long t_start = 0;
    if (trace && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
    // rows touched by this CTA's elements (grid-stride: the same relative rows every round would need per-round
    // waits; the grid covers the band in one round whenever it can, otherwise wait for everything up front)
    const int n = 6 * n_free;
    const size_t nS = (size_t)n_free * band_w * 36, total = nS + n + 1;
    const bool one_round = (size_t)gridDim.x * blockDim.x >= total;
    int r_first = 0, r_last = n_free - 1;
    bool need_all = !one_round;
    if (one_round) {
        const size_t e0 = (size_t)blockIdx.x * blockDim.x, e1 = min(e0 + blockDim.x, total) - 1;
        if (e0 >= total) return;
        auto row_of = [&](size_t idx) -> int { return idx < nS ? (int)(idx / ((size_t)band_w * 36)) : (idx < nS + n ? (int)((idx - nS) / 6) : -1); };
        const int ra = row_of(e0), rb = row_of(e1);
        if (ra < 0 || rb < 0 || rb < ra) need_all = true;          // the cost element (or the S / rhs boundary): every rank's first row
        else { r_first = ra; r_last = rb; }
    }
    // mode 3 (default): the band of this rank was completed by the previous kernel on this stream (k_schur_reduce runs
    // without any publishing code); CTA 0, the first to be scheduled, announces it to every rank, and a rank's rows are
    // all released by that one flag.  Only CTA 0 of a rank has to run for the others to proceed.
    const bool rank_level = pv.announce[0] != nullptr;
    if (rank_level && blockIdx.x == 0 && threadIdx.x < pv.world) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pv.announce[threadIdx.x]), "r"(epoch) : "memory");
    }
    // thread (p, k): rank p, k-th row of this CTA's row range (ranges are short: 256 elements are < 1 row)
    {
        const long long t0 = clock64();
        const int nrows = rank_level ? 1 : (need_all ? n_free : r_last - r_first + 1);
        for (int q = threadIdx.x; q < nrows * pv.world; q += blockDim.x) {
            const int p = q / nrows;
            int row = (need_all ? 0 : r_first) + (q - p * nrows);
            if (rank_level) {
                if (pv.row_hi[p] <= pv.row_lo[p] || (!need_all && (r_last < pv.row_lo[p] || r_first >= pv.row_hi[p]))) continue;   // this CTA reads nothing of rank p
                row = pv.row_lo[p];
            } else if (row < pv.row_lo[p] || row >= pv.row_hi[p]) continue;
            const unsigned* f = pv.row_flags + (size_t)p * n_free + row;
            while ((int)(ld_acquire_sys_u32(f) - epoch) < 0) {
                if (clock64() - t0 > kPeerTimeoutCycles) { *err_flag = 1; break; }
            }
        }
    }
    __syncthreads();
    if (trace && threadIdx.x == 0) {   // VIO360_PEER_TRACE: ns CTA 0 spent waiting / in total, summed over exchanges
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (blockIdx.x == 0) trace[0] += t - t_start;
    }
    band_pull_sum_expand(pv, n_free, band_w, free_pose, D, packed);
    if (trace) {
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (blockIdx.x == 0) trace[1] += t - t_start;
        }
    }
}

}  // namespace vio360
