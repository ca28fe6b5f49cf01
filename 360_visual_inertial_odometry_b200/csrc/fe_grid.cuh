// fe_grid.cuh -- SURVEY 8-F rank 2: the per-frame feature-list filters of FeatureTracker::TrackFeatures that
// sit between tracking and detection, on the device so that the point list does not leave it:
//   k_cluster_filter  FeatureTracker::RemoveClusteredFeatures (src/processing/FeatureTracker.cpp:404-497)
//   k_grid_cap        Frame::AssignFeaturesToGrid + Frame::LimitFeaturesPerGrid (src/database/Frame.cpp:108-202)
// Both return a keep mask over the caller's feature order (the reference then compacts in order).
// One CTA per frame.  Float arithmetic follows the reference's expression order without contraction (this
// translation unit is compiled with --fmad=false).
#pragma once

#include <cstdint>

namespace vio360 {

constexpr int kGridMaxPoints = 16384;   // cell ids of one frame live in shared memory
constexpr int kGridThreads = 256;

struct GridJobDev {
    const float* xy;       // [n][2]
    const int* track;      // [n] track counts (grid cap only)
    uint8_t* keep;         // [n]
    int n;
};

// FeatureTracker.cpp:435-436 == Frame.cpp:131-132: float division, truncation, clamp to the last cell
// (negative coordinates, which the reference would index out of bounds with, go to cell 0)
__device__ __forceinline__ int grid_cell_of(float x, float y, float cw, float chh, int cols, int rows) {
    const int col = max(min((int)(x / cw), cols - 1), 0), row = max(min((int)(y / chh), rows - 1), 0);
    return row * cols + col;
}

__global__ void __launch_bounds__(kGridThreads)
k_cluster_filter(const GridJobDev* __restrict__ jobs, int width, int height, int cols, int rows, float std_ratio) {
    extern __shared__ unsigned short s_cell[];                    // [n] cell of every feature, then [cols * rows] flags
    const GridJobDev jb = jobs[blockIdx.x];
    const int n = jb.n, ncell = cols * rows;
    unsigned char* s_clustered = reinterpret_cast<unsigned char*>(s_cell + ((n + 1) & ~1));
    const float cw = (float)width / cols, chh = (float)height / rows;
    const float diag = sqrtf(cw * cw + chh * chh);
    const float thr = diag * std_ratio;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s_cell[i] = (unsigned short)grid_cell_of(jb.xy[2 * i], jb.xy[2 * i + 1], cw, chh, cols, rows);
    __syncthreads();
    // one thread per cell walks the list in feature order: the float sums must be sequential (:459-476)
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
        int cnt = 0;
        float mx = 0.f, my = 0.f;
        for (int i = 0; i < n; ++i)
            if (s_cell[i] == c) { mx += jb.xy[2 * i]; my += jb.xy[2 * i + 1]; ++cnt; }
        bool clustered = false;
        if (n >= 4 && cnt >= 4) {                       // :411 fewer than 4 features in the frame: nothing removed
            mx /= (float)cnt; my /= (float)cnt;
            float var = 0.f;
            for (int i = 0; i < n; ++i)
                if (s_cell[i] == c) { const float dx = jb.xy[2 * i] - mx, dy = jb.xy[2 * i + 1] - my; var += (dx * dx + dy * dy); }
            var /= (float)cnt;
            clustered = sqrtf(var) < thr;
        }
        s_clustered[c] = clustered;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) jb.keep[i] = !s_clustered[s_cell[i]];
}

__global__ void __launch_bounds__(kGridThreads)
k_grid_cap(const GridJobDev* __restrict__ jobs, int width, int height, int cols, int rows, int max_per_cell) {
    extern __shared__ unsigned short s_cell[];                    // [n] cell of every feature (0xffff: outside the image)
    const GridJobDev jb = jobs[blockIdx.x];
    const int n = jb.n;
    int* s_track = reinterpret_cast<int*>(s_cell + ((n + 1) & ~1));
    const float cw = (float)width / cols, chh = (float)height / rows;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float x = jb.xy[2 * i], y = jb.xy[2 * i + 1];
        const bool inside = !(x < 0 || x >= width || y < 0 || y >= height);      // Frame.cpp:125-128
        s_cell[i] = inside ? (unsigned short)grid_cell_of(x, y, cw, chh, cols, rows) : (unsigned short)0xffff;
        s_track[i] = jb.track[i];
    }
    __syncthreads();
    // A cell keeps its max_per_cell features with the highest track count (Frame.cpp:151-170).  std::sort on a
    // cell of at most 16 indices is libstdc++'s insertion sort, i.e. stable: ties keep feature order.  A
    // feature's rank in its cell is then the number of members that sort before it.
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const unsigned short c = s_cell[i];
        bool keep = false;
        if (c != 0xffff) {
            const int t = s_track[i];
            int rank = 0;
            for (int j = 0; j < n; ++j) rank += (s_cell[j] == c) && (s_track[j] > t || (s_track[j] == t && j < i));
            keep = rank < max_per_cell;
        }
        jb.keep[i] = keep;
    }
}

}  // namespace vio360
