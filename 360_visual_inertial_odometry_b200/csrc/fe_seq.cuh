// fe_seq.cuh -- SURVEY 8-F rank 2: the per-frame chain of FeatureTracker::TrackFeatures
// (src/processing/FeatureTracker.cpp:61-206) with the feature list RESIDENT ON THE DEVICE:
//
//   track (:110, k_lk_track)  ->  post filter (:118-126)  ->  rotation RANSAC (:131, k_ransac_*)
//   ->  new Feature per inlier: id carried, track_count + 1 (:139-163)  ->  RemoveClusteredFeatures (:168)
//   ->  AssignFeaturesToGrid + LimitFeaturesPerGrid (:171-172, Frame.cpp:108-202)
//   ->  if count < max_features: DetectNewFeatures(image, CreateFeatureMask(features)) (:176-199), new ids,
//       grid cap again.
//
// The kernels here are the glue between the existing kernels: order-preserving stream compaction (the
// reference pushes survivors into fresh vectors in order), the two grid filters, the sample list of the
// RANSAC and the bookkeeping of ids / track counts.  One CTA per sequence: a frame has at most a few
// thousand features, and every step but the compaction scan is a per-feature map.  Counts never visit
// the host between the stages: downstream kernels read them from device memory (TrackJobDev::n_ptr,
// RansacJobDev::n, the occupancy count of k_occupancy_discs, DetectJobDev::max_corners_dev).
#pragma once

#include "fe_detect.cuh"
#include "fe_grid.cuh"
#include "fe_lk.cuh"
#include "fe_ransac.cuh"

namespace vio360 {

constexpr int kSeqThreads = 256;

struct SeqCounters {     // device-side scalars of one sequence (also copied back with the results)
    int n;               // features in the list (the previous frame's when a call starts, the new frame's when it ends)
    int next_id;         // FeatureTracker::m_next_feature_id
    int n_good;          // survivors of the post filter = RANSAC input
    int n_tracked;       // m_num_tracked: features created from RANSAC inliers (:165)
    int n_detected;      // m_num_detected (:197)
    int max_corners;     // corner limit of this frame's detection: max_features, or 0 when count >= max_features (:176)
    int n_occ;           // features masked out around (CreateFeatureMask)
    int overflow;        // the list would have exceeded its capacity (cannot happen for capacity >= 2 * max_features)
    int n_pre, next_id_pre;   // list length / id counter before this frame's detection (list 2 holds that list, so that
                              // k_seq_append can be repeated after the candidate buffers had to grow)
};

struct SeqDev {          // one sequence's device arrays, capacity `cap` features each
    SeqCounters* c;
    int* id[3]; float* xy[3]; int* tc[3];        // feature list: [0] between kernels / calls, [1] scratch, [2] pre-detection snapshot
    float* lk_xy; uint8_t* lk_status; uint8_t* lk_keep;
    float* good_prev; float* good_cur; int* good_src;
    uint8_t* inlier;     // RANSAC mask over the good list
    int* samples;        // ransac_iters x 3, built from the caller's random words
    const uint32_t* words;
    float* det_xy; int* det_n;                   // corners of this frame's detection
    int* out_id; float* out_xy; int* out_tc; int* out_counts;   // results (device staging or the caller's device arrays)
    RansacJobDev* rjob;  // this sequence's entry of the RANSAC job table (n is written here)
    int* occ_n_slot;     // this sequence's entry of the batch's occupancy-count array (k_occupancy_discs)
    int cap;
};

struct SeqParams {
    int W, H, max_features;
    int grid_cols, grid_rows, max_per_cell;
    int remove_clustered; float std_ratio;
    int ransac_iters;
    float polar_thr, margin;
};

// Order-preserving compaction of flags[0..n) by the whole CTA: s_idx[k] = index of the k-th survivor; returns
// the survivor count (in every thread).  s_scan: kSeqThreads ints of shared memory.
__device__ __forceinline__ int cta_compact(const uint8_t* __restrict__ flags, int n, int* __restrict__ s_idx, int* s_scan) {
    __shared__ int s_base;
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    for (int c0 = 0; c0 < n; c0 += kSeqThreads) {
        const int i = c0 + (int)threadIdx.x;
        const int f = (i < n && flags[i]) ? 1 : 0;
        // inclusive scan over the chunk (warp shuffles + one shared pass)
        int v = f;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, o); if ((threadIdx.x & 31) >= o) v += t; }
        if ((threadIdx.x & 31) == 31) s_scan[threadIdx.x >> 5] = v;
        __syncthreads();
        int woff = 0;
        for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) woff += s_scan[w];
        const int base = s_base;
        if (f) s_idx[base + woff + v - 1] = i;
        __syncthreads();
        if (threadIdx.x == kSeqThreads - 1) s_base = base + woff + v;
        __syncthreads();
    }
    return s_base;
}

// ---- after LK: post filter compaction (:118-126) + RANSAC sample list ---------------------------------
// Sample rule (the reference draws from std::random_device, :273-289, so any uniform draw of three distinct
// indices is a faithful run; this one is reproducible): word r -> index (r * n) >> 32; an index equal to an
// earlier one of the triple moves on to (index + 1) % n until distinct.
__global__ void __launch_bounds__(kSeqThreads) k_seq_post_lk(const SeqDev* __restrict__ seqs, SeqParams P) {
    extern __shared__ int seq_smem[];
    __shared__ int s_scan[kSeqThreads / 32];
    const SeqDev sd = seqs[blockIdx.x];
    const int n = sd.c->n;
    int* s_idx = seq_smem;
    const int ng = cta_compact(sd.lk_keep, n, s_idx, s_scan);
    for (int k = threadIdx.x; k < ng; k += kSeqThreads) {
        const int i = s_idx[k];
        sd.good_prev[2 * k] = sd.xy[0][2 * i]; sd.good_prev[2 * k + 1] = sd.xy[0][2 * i + 1];
        sd.good_cur[2 * k] = sd.lk_xy[2 * i]; sd.good_cur[2 * k + 1] = sd.lk_xy[2 * i + 1];
        sd.good_src[k] = i;
    }
    if (ng >= 3)
        for (int it = threadIdx.x; it < P.ransac_iters; it += kSeqThreads) {
            int idx[3];
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                int v = (int)(((unsigned long long)sd.words[3 * it + s] * (unsigned long long)ng) >> 32);
                for (int guard = 0; guard < 3; ++guard) {
                    bool dup = false;
                    for (int q = 0; q < s; ++q) dup = dup || idx[q] == v;
                    if (!dup) break;
                    v = (v + 1) % ng;
                }
                idx[s] = v;
                sd.samples[3 * it + s] = v;
            }
        }
    if (threadIdx.x == 0) { sd.c->n_good = ng; sd.rjob->n = ng; }
}

// shared-memory carve-up of the list kernels (cap = list capacity, ncell = grid cells)
struct SeqSmem {
    int* idx; int* track; int* cell_start; float* xy;              // 4-byte arrays: cap, cap, ncell + 2, 2 cap
    unsigned short* cell; unsigned short* member;                  // cap each
    unsigned char* clustered; uint8_t* keep;                       // ncell, cap
    __device__ SeqSmem(int* base, int cap, int ncell) {
        idx = base; track = base + cap; cell_start = base + 2 * cap;
        xy = reinterpret_cast<float*>(base + 2 * cap + ((ncell + 2 + 1) & ~1));
        cell = reinterpret_cast<unsigned short*>(xy + 2 * cap);
        member = cell + ((cap + 1) & ~1);
        clustered = reinterpret_cast<unsigned char*>(member + ((cap + 1) & ~1));
        keep = clustered + ((ncell + 3) & ~3);
    }
};
inline size_t seq_smem_bytes(int cap, int ncell) {
    return (size_t)cap * 16 + (size_t)((ncell + 3) & ~1) * 4 + (size_t)((cap + 1) & ~1) * 4 + (size_t)((ncell + 3) & ~3) + (size_t)cap + 16;
}

// Member lists of the grid cells: member[cell_start[c] .. cell_start[c + 1]) = the features of cell c in list order.
// One thread per cell scans the cell ids twice (count, fill): O(n) per thread instead of the O(n) scan per cell AND
// per statistic (or per feature, for the ranks) that the filters would otherwise need.  sm.cell must be filled.
__device__ __forceinline__ void cta_build_cells(int n, int ncell, const SeqSmem& sm) {
    for (int c = threadIdx.x; c < ncell; c += kSeqThreads) {
        int cnt = 0;
        for (int i = 0; i < n; ++i) cnt += sm.cell[i] == c;
        sm.cell_start[c + 1] = cnt;
    }
    __syncthreads();
    if (threadIdx.x < 32) {   // exclusive scan of the counts by one warp: a contiguous chunk per lane
        const int lane = threadIdx.x, per = (ncell + 31) / 32, c0 = min(lane * per, ncell), c1 = min(c0 + per, ncell);
        int sum = 0;
        for (int c = c0; c < c1; ++c) sum += sm.cell_start[c + 1];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        int run = incl - sum;
        if (lane == 0) sm.cell_start[0] = 0;
        for (int c = c0; c < c1; ++c) { run += sm.cell_start[c + 1]; sm.cell_start[c + 1] = run; }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < ncell; c += kSeqThreads) {
        int pos = sm.cell_start[c];
        for (int i = 0; i < n; ++i)
            if (sm.cell[i] == c) sm.member[pos++] = (unsigned short)i;
    }
    __syncthreads();
}

// grid filters on a list held in global memory (same arithmetic as fe_grid.cuh); positions, cell ids and the cells'
// member lists in shared memory
__device__ __forceinline__ void cta_cluster_keep(const float* __restrict__ xy, int n, const SeqParams& P, const SeqSmem& sm, uint8_t* __restrict__ keep) {
    const int ncell = P.grid_cols * P.grid_rows;
    const float cw = (float)P.W / P.grid_cols, chh = (float)P.H / P.grid_rows;
    const float diag = sqrtf(cw * cw + chh * chh);
    const float thr = diag * P.std_ratio;
    for (int i = threadIdx.x; i < n; i += kSeqThreads) {
        const float x = xy[2 * i], y = xy[2 * i + 1];
        sm.xy[2 * i] = x; sm.xy[2 * i + 1] = y;
        sm.cell[i] = (unsigned short)grid_cell_of(x, y, cw, chh, P.grid_cols, P.grid_rows);
    }
    __syncthreads();
    cta_build_cells(n, ncell, sm);
    for (int c = threadIdx.x; c < ncell; c += kSeqThreads) {
        const int k0 = sm.cell_start[c], k1 = sm.cell_start[c + 1], cnt = k1 - k0;
        bool clustered = false;
        if (n >= 4 && cnt >= 4) {   // sums in list order, as the reference's loops over the frame's features
            float mx = 0.f, my = 0.f;
            for (int k = k0; k < k1; ++k) { const int i = sm.member[k]; mx += sm.xy[2 * i]; my += sm.xy[2 * i + 1]; }
            mx /= (float)cnt; my /= (float)cnt;
            float var = 0.f;
            for (int k = k0; k < k1; ++k) { const int i = sm.member[k]; const float dx = sm.xy[2 * i] - mx, dy = sm.xy[2 * i + 1] - my; var += (dx * dx + dy * dy); }
            var /= (float)cnt;
            clustered = sqrtf(var) < thr;
        }
        sm.clustered[c] = clustered;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += kSeqThreads) keep[i] = !sm.clustered[sm.cell[i]];
    __syncthreads();
}

__device__ __forceinline__ void cta_grid_cap_keep(const float* __restrict__ xy, const int* __restrict__ tc, int n, const SeqParams& P,
                                                  const SeqSmem& sm, uint8_t* __restrict__ keep) {
    const float cw = (float)P.W / P.grid_cols, chh = (float)P.H / P.grid_rows;
    for (int i = threadIdx.x; i < n; i += kSeqThreads) {
        const float x = xy[2 * i], y = xy[2 * i + 1];
        const bool inside = !(x < 0 || x >= P.W || y < 0 || y >= P.H);
        sm.cell[i] = inside ? (unsigned short)grid_cell_of(x, y, cw, chh, P.grid_cols, P.grid_rows) : (unsigned short)0xffff;
        sm.track[i] = tc[i];
    }
    __syncthreads();
    cta_build_cells(n, P.grid_cols * P.grid_rows, sm);
    for (int i = threadIdx.x; i < n; i += kSeqThreads) {
        const unsigned short c = sm.cell[i];
        bool k = false;
        if (c != 0xffff) {   // rank among the cell's members by (track count desc, list index asc)
            const int t = sm.track[i];
            int rank = 0;
            for (int q = sm.cell_start[c]; q < sm.cell_start[c + 1]; ++q) {
                const int j = sm.member[q];
                rank += (sm.track[j] > t) || (sm.track[j] == t && j < i);
            }
            k = rank < P.max_per_cell;
        }
        keep[i] = k;
    }
    __syncthreads();
}

__device__ __forceinline__ void cta_gather(const SeqDev& sd, int from, int to, const int* s_idx, int m) {
    for (int k = threadIdx.x; k < m; k += kSeqThreads) {
        const int i = s_idx[k];
        sd.id[to][k] = sd.id[from][i]; sd.tc[to][k] = sd.tc[from][i];
        sd.xy[to][2 * k] = sd.xy[from][2 * i]; sd.xy[to][2 * k + 1] = sd.xy[from][2 * i + 1];
    }
    __syncthreads();
}

// ---- after RANSAC: inliers become this frame's features, cluster removal, grid cap ---------------------
__global__ void __launch_bounds__(kSeqThreads) k_seq_post_ransac(const SeqDev* __restrict__ seqs, SeqParams P) {
    extern __shared__ int seq_smem[];
    __shared__ int s_scan[kSeqThreads / 32];
    const SeqDev sd = seqs[blockIdx.x];
    SeqSmem sm(seq_smem, sd.cap, P.grid_cols * P.grid_rows);
    const int ng = sd.c->n_good;
    // inliers (:139-163): feature id carried, track count + 1, pixel = tracked position.  list 0 -> list 1
    int m = cta_compact(sd.inlier, ng, sm.idx, s_scan);
    for (int k = threadIdx.x; k < m; k += kSeqThreads) {
        const int g = sm.idx[k], i = sd.good_src[g];
        sd.id[1][k] = sd.id[0][i]; sd.tc[1][k] = sd.tc[0][i] + 1;
        sd.xy[1][2 * k] = sd.good_cur[2 * g]; sd.xy[1][2 * k + 1] = sd.good_cur[2 * g + 1];
    }
    __syncthreads();
    const int n_tracked = m;
    int cur = 1;
    if (P.remove_clustered && m >= 4) {                            // :168, :411
        cta_cluster_keep(sd.xy[cur], m, P, sm, sm.keep);
        const int m2 = cta_compact(sm.keep, m, sm.idx, s_scan);
        cta_gather(sd, cur, cur ^ 1, sm.idx, m2);
        cur ^= 1; m = m2;
    }
    cta_grid_cap_keep(sd.xy[cur], sd.tc[cur], m, P, sm, sm.keep);   // :171-172
    const int m3 = cta_compact(sm.keep, m, sm.idx, s_scan);
    cta_gather(sd, cur, cur ^ 1, sm.idx, m3);
    cur ^= 1; m = m3;
    // the list between kernels is list 0; list 2 keeps a copy for a repeated k_seq_append
    for (int k = threadIdx.x; k < m; k += kSeqThreads) {
        const int a = sd.id[cur][k], t = sd.tc[cur][k];
        const float x = sd.xy[cur][2 * k], y = sd.xy[cur][2 * k + 1];
        if (cur != 0) { sd.id[0][k] = a; sd.tc[0][k] = t; sd.xy[0][2 * k] = x; sd.xy[0][2 * k + 1] = y; }
        sd.id[2][k] = a; sd.tc[2][k] = t; sd.xy[2][2 * k] = x; sd.xy[2][2 * k + 1] = y;
    }
    if (threadIdx.x == 0) {
        SeqCounters& c = *sd.c;
        c.n = m; c.n_tracked = n_tracked; c.n_pre = m; c.next_id_pre = c.next_id;
        const bool detect = m < P.max_features;                    // :176
        c.max_corners = detect ? P.max_features : 0;
        c.n_occ = detect ? m : 0;                                  // every feature of the list is valid (Feature.cpp:25)
        *sd.occ_n_slot = c.n_occ;
        c.n_detected = 0;
    }
}

// ---- after detection: new features (:183-194), grid cap again (:199-200), results ----------------------
__global__ void __launch_bounds__(kSeqThreads) k_seq_append(const SeqDev* __restrict__ seqs, SeqParams P) {
    extern __shared__ int seq_smem[];
    __shared__ int s_scan[kSeqThreads / 32];
    const SeqDev sd = seqs[blockIdx.x];
    SeqSmem sm(seq_smem, sd.cap, P.grid_cols * P.grid_rows);
    SeqCounters& c = *sd.c;
    int m = c.n_pre;
    const bool detect = c.max_corners > 0;
    int nd = detect ? *sd.det_n : 0;
    int overflow = 0;
    if (m + nd > sd.cap) { nd = sd.cap - m; overflow = 1; }
    const int id0 = c.next_id_pre;
    for (int k = threadIdx.x; k < m; k += kSeqThreads) {           // start from the pre-detection list (idempotent)
        sd.id[0][k] = sd.id[2][k]; sd.tc[0][k] = sd.tc[2][k];
        sd.xy[0][2 * k] = sd.xy[2][2 * k]; sd.xy[0][2 * k + 1] = sd.xy[2][2 * k + 1];
    }
    __syncthreads();
    if (detect) {
        for (int k = threadIdx.x; k < nd; k += kSeqThreads) {
            sd.id[0][m + k] = id0 + k; sd.tc[0][m + k] = 0;        // Feature ctor: track count 0
            sd.xy[0][2 * (m + k)] = sd.det_xy[2 * k]; sd.xy[0][2 * (m + k) + 1] = sd.det_xy[2 * k + 1];
        }
        __syncthreads();
        m += nd;
        cta_grid_cap_keep(sd.xy[0], sd.tc[0], m, P, sm, sm.keep);
        const int m2 = cta_compact(sm.keep, m, sm.idx, s_scan);
        cta_gather(sd, 0, 1, sm.idx, m2);
        m = m2;
        for (int k = threadIdx.x; k < m; k += kSeqThreads) {
            sd.id[0][k] = sd.id[1][k]; sd.tc[0][k] = sd.tc[1][k];
            sd.xy[0][2 * k] = sd.xy[1][2 * k]; sd.xy[0][2 * k + 1] = sd.xy[1][2 * k + 1];
        }
        __syncthreads();
    }
    for (int k = threadIdx.x; k < m; k += kSeqThreads) {
        sd.out_id[k] = sd.id[0][k]; sd.out_tc[k] = sd.tc[0][k];
        sd.out_xy[2 * k] = sd.xy[0][2 * k]; sd.out_xy[2 * k + 1] = sd.xy[0][2 * k + 1];
    }
    if (threadIdx.x == 0) {
        c.n = m; c.next_id = id0 + nd; c.n_detected = nd; c.overflow = overflow;
        sd.out_counts[0] = m; sd.out_counts[1] = c.n_tracked; sd.out_counts[2] = nd; sd.out_counts[3] = overflow;
    }
}

}  // namespace vio360
