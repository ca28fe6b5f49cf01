// fe_ransac.cuh -- rotation-only RANSAC on tracked ERP features.
// Replaces FeatureTracker::RejectOutliersRotationRANSAC / EstimateRotation /
// ComputeRotationInliers (src/processing/FeatureTracker.cpp:253-379) with
// Camera::PixelToBearing / AngularDistance (src/database/Camera.cpp:22-47, 89-98).
//
// The reference draws its 3-point samples from std::random_device, so the sample list is an
// input here (the host shim draws it the same way); everything after the draw is deterministic:
//   k_ransac_bearings    pixel -> unit bearing, fp32 with the reference's float/double mix
//   k_ransac_hypotheses  one warp per hypothesis: H = sum b2 b1^T of the 3 samples, R = U V^T of
//                        its SVD (det fixed to +1), skipped when |det R - 1| > 0.1, then the
//                        inlier count acos(clamp(R b1 . b2)) < threshold over all points
//   k_ransac_pick        first hypothesis with the largest count (strict > in the reference's
//                        loop), its mask; all ones when no hypothesis has an inlier
#pragma once

#include "common.cuh"

namespace vio360 {

struct RansacJobDev {
    const float* prev_xy; const float* cur_xy; int n;
    const int* samples;      // iters x 3 point indices
    float* b1; float* b2;    // n x 3 bearings (scratch)
    float* R;                // iters x 9 (scratch)
    int* counts;             // iters, -1 = hypothesis rejected by the determinant test
    uint8_t* mask;           // n
    int* best;               // [0] inlier count of the winner, [1] its index (-1: none)
};

__device__ __forceinline__ void pixel_to_bearing(float px, float py, int W, int H, float* b) {
    const float u_norm = px / (float)W, v_norm = py / (float)H;
    const float lon = (float)((double)((u_norm - 0.5f) * 2.0f) * 3.14159265358979323846);
    const float lat = (float)((double)(-(v_norm - 0.5f)) * 3.14159265358979323846);
    const float cos_lat = cosf(lat);
    float x = cos_lat * sinf(lon), y = -sinf(lat), z = cos_lat * cosf(lon);
    const float n2 = x * x + y * y + z * z;
    if (n2 > 0.f) { const float nn = sqrtf(n2); x /= nn; y /= nn; z /= nn; }
    b[0] = x; b[1] = y; b[2] = z;
}

__device__ __forceinline__ float det3f(const float* A) {
    return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) + A[2] * (A[3] * A[7] - A[4] * A[6]);
}

// R = U V^T of the SVD of M (row-major), with the smallest singular direction flipped when
// det < 0: one-sided Jacobi on the columns, singular values sorted descending.
__device__ void rotation_from_svd_f(const float* M, float* R) {
    float B[9], V[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { B[i] = M[i]; V[i] = (i % 4 == 0) ? 1.f : 0.f; }
    for (int sweep = 0; sweep < 30; ++sweep) {
        float off = 0.f;
#pragma unroll
        for (int pq = 0; pq < 3; ++pq) {
            const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
            float alpha = 0.f, beta = 0.f, gamma = 0.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                alpha += B[k * 3 + p] * B[k * 3 + p];
                beta += B[k * 3 + q] * B[k * 3 + q];
                gamma += B[k * 3 + p] * B[k * 3 + q];
            }
            const float lim = sqrtf(alpha * beta);
            if (gamma == 0.f || fabsf(gamma) <= 1e-9f * lim) continue;
            off = fmaxf(off, fabsf(gamma) / (lim > 0.f ? lim : 1.f));
            const float zeta = (beta - alpha) / (2.f * gamma);
            const float t = (zeta >= 0.f ? 1.f : -1.f) / (fabsf(zeta) + sqrtf(1.f + zeta * zeta));
            const float c = 1.f / sqrtf(1.f + t * t), s = c * t;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float bp = B[k * 3 + p], bq = B[k * 3 + q];
                B[k * 3 + p] = c * bp - s * bq; B[k * 3 + q] = s * bp + c * bq;
                const float vp = V[k * 3 + p], vq = V[k * 3 + q];
                V[k * 3 + p] = c * vp - s * vq; V[k * 3 + q] = s * vp + c * vq;
            }
        }
        if (off < 1e-7f) break;
    }
    float sv[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) sv[j] = sqrtf(B[j] * B[j] + B[3 + j] * B[3 + j] + B[6 + j] * B[6 + j]);
    int o0 = 0, o1 = 1, o2 = 2, tmp;
    if (sv[o1] > sv[o0]) { tmp = o0; o0 = o1; o1 = tmp; }
    if (sv[o2] > sv[o0]) { tmp = o0; o0 = o2; o2 = tmp; }
    if (sv[o2] > sv[o1]) { tmp = o1; o1 = o2; o2 = tmp; }
    const int ord[3] = {o0, o1, o2};
    float U[9], Vs[9];
    for (int j = 0; j < 3; ++j) {
        const int c = ord[j];
        for (int k = 0; k < 3; ++k) {
            U[k * 3 + j] = (sv[c] > 0.f) ? B[k * 3 + c] / sv[c] : 0.f;
            Vs[k * 3 + j] = V[k * 3 + c];
        }
    }
    if (!(sv[ord[2]] > 0.f) && sv[ord[1]] > 0.f) {   // rank 2: complete U with the cross product
        U[2] = U[3] * U[7] - U[6] * U[4];
        U[5] = U[6] * U[1] - U[0] * U[7];
        U[8] = U[0] * U[4] - U[3] * U[1];
    }
    auto mul_bt = [&]() {
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) R[i * 3 + j] = U[i * 3] * Vs[j * 3] + U[i * 3 + 1] * Vs[j * 3 + 1] + U[i * 3 + 2] * Vs[j * 3 + 2];
    };
    mul_bt();
    if (det3f(R) < 0.f) {
        Vs[2] = -Vs[2]; Vs[5] = -Vs[5]; Vs[8] = -Vs[8];   // V.col(2) *= -1 (FeatureTracker.cpp:346)
        mul_bt();
    }
}

__global__ void __launch_bounds__(256) k_ransac_bearings(const RansacJobDev* __restrict__ jobs, int W, int H) {
    const RansacJobDev& jb = jobs[blockIdx.y];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= jb.n) return;
    pixel_to_bearing(jb.prev_xy[2 * i], jb.prev_xy[2 * i + 1], W, H, jb.b1 + 3 * i);
    pixel_to_bearing(jb.cur_xy[2 * i], jb.cur_xy[2 * i + 1], W, H, jb.b2 + 3 * i);
}

__device__ __forceinline__ bool rotation_inlier(const float* R, const float* b1, const float* b2, float thr) {
    const float rx = R[0] * b1[0] + R[1] * b1[1] + R[2] * b1[2];
    const float ry = R[3] * b1[0] + R[4] * b1[1] + R[5] * b1[2];
    const float rz = R[6] * b1[0] + R[7] * b1[1] + R[8] * b1[2];
    float c = rx * b2[0] + ry * b2[1] + rz * b2[2];
    c = fmaxf(-1.0f, fminf(1.0f, c));
    return acosf(c) < thr;
}

constexpr int kRansacWarps = 4;

__global__ void __launch_bounds__(kRansacWarps * 32) k_ransac_hypotheses(const RansacJobDev* __restrict__ jobs, int iters, float thr) {
    const RansacJobDev& jb = jobs[blockIdx.y];
    const int it = blockIdx.x * kRansacWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (it >= iters || jb.n < 3) return;
    float Hm[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int s = 0; s < 3; ++s) {
        const int idx = jb.samples[3 * it + s];
        const float* a = jb.b1 + 3 * idx; const float* b = jb.b2 + 3 * idx;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) Hm[i * 3 + j] += b[i] * a[j];
    }
    float R[9];
    rotation_from_svd_f(Hm, R);   // same inputs on every lane: warp-uniform result
    if (lane < 9) jb.R[9 * it + lane] = R[lane];
    if (fabsf(det3f(R) - 1.0f) > 0.1f) {
        if (lane == 0) jb.counts[it] = -1;
        return;
    }
    int cnt = 0;
    for (int i = lane; i < jb.n; i += 32) cnt += rotation_inlier(R, jb.b1 + 3 * i, jb.b2 + 3 * i, thr) ? 1 : 0;
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if (lane == 0) jb.counts[it] = cnt;
}

__global__ void __launch_bounds__(256) k_ransac_pick(const RansacJobDev* __restrict__ jobs, int iters, float thr) {
    __shared__ int sbest[2];
    __shared__ unsigned long long skey[8];
    const RansacJobDev& jb = jobs[blockIdx.x];
    const int tid = threadIdx.x;
    if (jb.n < 3) {   // FeatureTracker.cpp:258-261
        for (int i = tid; i < jb.n; i += blockDim.x) jb.mask[i] = 1;
        if (tid == 0) { jb.best[0] = jb.n; jb.best[1] = -1; }
        return;
    }
    // largest count, earliest iteration: key = count << 32 | (0xffffffff - it)
    unsigned long long key = 0ull;
    for (int it = tid; it < iters; it += blockDim.x) {
        const int c = jb.counts[it];
        if (c > 0) key = max(key, ((unsigned long long)(unsigned int)c << 32) | (unsigned long long)(0xffffffffu - (unsigned int)it));
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) key = max(key, __shfl_xor_sync(0xffffffffu, key, s));
    if ((tid & 31) == 0) skey[tid >> 5] = key;
    __syncthreads();
    if (tid == 0) {
        unsigned long long k = 0ull;
        for (int w = 0; w < 8; ++w) k = max(k, skey[w]);
        sbest[0] = (int)(k >> 32);
        sbest[1] = k ? (int)(0xffffffffu - (unsigned int)(k & 0xffffffffull)) : -1;
        jb.best[0] = sbest[0]; jb.best[1] = sbest[1];
    }
    __syncthreads();
    const int bi = sbest[1];
    if (bi < 0) {   // no hypothesis found an inlier: keep everything (FeatureTracker.cpp:321-322)
        for (int i = tid; i < jb.n; i += blockDim.x) jb.mask[i] = 1;
        return;
    }
    float R[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) R[j] = jb.R[9 * bi + j];
    for (int i = tid; i < jb.n; i += blockDim.x) jb.mask[i] = rotation_inlier(R, jb.b1 + 3 * i, jb.b2 + 3 * i, thr) ? 1 : 0;
}

}  // namespace vio360
