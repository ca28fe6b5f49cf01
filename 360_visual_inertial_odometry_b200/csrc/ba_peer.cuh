// ba_peer.cuh -- reduced-camera-system exchange over NVLink / NVSwitch peer memory.
//
// Every rank of a sharded window eliminates its own landmark range and leaves a partial block band
// [block row][block column - row < w][6 x 6] | rhs | cost in a buffer the other ranks can map
// (cudaIpc handle across processes, plain peer access inside one process).  k_band_exchange is the
// whole collective: it announces the band, waits for the other ranks' announcements, then every rank
// PULLS the partial bands that cover a block over peer memory, sums them in rank order (so all ranks
// hold bit-identical results and the sum does not depend on arrival order), adds the pose part of the
// Levenberg-Marquardt diagonal and writes the dense [S | rhs | cost] the solver reads -- the allreduce,
// the mirror and the expansion in ONE launch, with no staging copy and no second pass.
//
// A rank's landmarks only touch a contiguous range of block rows (landmark-range shards of a
// co-visibility band), so a block row is read only from the ranks whose range covers it: the bytes
// crossing NVLink stay ~ one band in total instead of (world - 1) bands per rank.
//
// Buffer discipline: two band buffers per rank, used alternately (epoch parity).  Rank A overwrites
// the buffer of epoch e when it builds epoch e + 2, i.e. after its exchange e + 1 returned, which
// waited for every peer's announcement e + 1, which a peer posts only after its own exchange e (the
// last reader of A's buffer e) has completed in stream order.  No trailing barrier is needed.
#pragma once

#include <cstdint>

namespace vio360 {

constexpr int kMaxPeers = 8;
constexpr size_t kPeerHeaderBytes = 1024;               // arrival flags live at the head of the exchange buffer
constexpr long long kPeerTimeoutCycles = 4000000000ll;  // ~2 s at 1.9 GHz: a missing peer fails loudly, it never hangs the box

struct PeerView {
    const double* band[kMaxPeers];   // this epoch's partial band of every rank (own pointer for own rank)
    unsigned* flags[kMaxPeers];      // every rank's arrival flags: flags[r][s] = last epoch announced by rank s to rank r
    int row_lo[kMaxPeers], row_hi[kMaxPeers];   // block rows (free-pose index) a rank's band covers
    int world, rank;
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
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
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
k_band_exchange(const PeerView pv, const unsigned epoch, const int n_free, const int band_w,
                const int* __restrict__ free_pose, const double* __restrict__ D,
                double* __restrict__ packed, int* __restrict__ err_flag, unsigned long long* __restrict__ trace) {
    unsigned long long t_start = 0;
    if (trace && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
    // announce: this rank's band of `epoch` was completed by the previous kernel on this stream
    if (blockIdx.x == 0 && threadIdx.x < pv.world) {
        __threadfence_system();
        st_release_sys_u32(pv.flags[threadIdx.x] + pv.rank, epoch);
    }
    // wait for every rank's announcement (flags of this rank are local memory: the poll stays on this GPU)
    if (threadIdx.x < pv.world) {
        const unsigned* f = pv.flags[pv.rank] + threadIdx.x;
        const long long t0 = clock64();
        // (measured on 4 B200: acquire polls by one thread per peer beat relaxed polls + one fence per CTA, 13 vs 28 us)
        while ((int)(ld_acquire_sys_u32(f) - epoch) < 0) {
            if (clock64() - t0 > kPeerTimeoutCycles) { *err_flag = 1; break; }
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
