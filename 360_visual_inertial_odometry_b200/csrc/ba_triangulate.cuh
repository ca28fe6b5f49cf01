// ba_triangulate.cuh -- included by ba.cu.  SURVEY 8-F rank 4: batched DLT triangulation of new map
// points, Estimator::TriangulateSinglePoint (src/processing/Estimator.cpp:1082-1139) for all matches of
// a keyframe pair (called per match at :1218).  fp32 like the reference; the smallest right singular
// vector of the 4x4 DLT matrix comes from a one-sided Jacobi SVD (Eigen::JacobiSVD<Matrix4f> is absent
// here; the vector is unique up to sign, which cancels in the dehomogenisation).
#pragma once

namespace vio360 {

struct Mat4f { float m[16]; };

__global__ void __launch_bounds__(128) k_triangulate(int n, const float* __restrict__ b1, const float* __restrict__ b2, Mat4f T1, Mat4f T2,
                                                     float* __restrict__ pts, unsigned char* __restrict__ valid) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float u0 = b1[3 * i], u1 = b1[3 * i + 1], u2 = b1[3 * i + 2];
    const float v0 = b2[3 * i], v1 = b2[3 * i + 1], v2 = b2[3 * i + 2];
    float B[16], V[16];
#pragma unroll
    for (int c = 0; c < 4; ++c) {   // Estimator.cpp:1098-1101
        B[0 * 4 + c] = u0 * T1.m[2 * 4 + c] - u2 * T1.m[0 * 4 + c];
        B[1 * 4 + c] = u1 * T1.m[2 * 4 + c] - u2 * T1.m[1 * 4 + c];
        B[2 * 4 + c] = v0 * T2.m[2 * 4 + c] - v2 * T2.m[0 * 4 + c];
        B[3 * 4 + c] = v1 * T2.m[2 * 4 + c] - v2 * T2.m[1 * 4 + c];
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) V[k] = (k % 5 == 0) ? 1.f : 0.f;
    for (int sweep = 0; sweep < 30; ++sweep) {
        float off = 0.f;
#pragma unroll
        for (int p = 0; p < 3; ++p)
#pragma unroll
            for (int q = p + 1; q < 4; ++q) {
                float alpha = 0.f, beta = 0.f, gamma = 0.f;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    alpha += B[k * 4 + p] * B[k * 4 + p];
                    beta += B[k * 4 + q] * B[k * 4 + q];
                    gamma += B[k * 4 + p] * B[k * 4 + q];
                }
                const float lim = sqrtf(alpha * beta);
                if (gamma == 0.f || fabsf(gamma) <= 1e-9f * lim) continue;
                off = fmaxf(off, fabsf(gamma) / (lim > 0.f ? lim : 1.f));
                const float zeta = (beta - alpha) / (2.f * gamma);
                const float t = (zeta >= 0.f ? 1.f : -1.f) / (fabsf(zeta) + sqrtf(1.f + zeta * zeta));
                const float c = 1.f / sqrtf(1.f + t * t), s = c * t;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float bp = B[k * 4 + p], bq = B[k * 4 + q];
                    B[k * 4 + p] = c * bp - s * bq; B[k * 4 + q] = s * bp + c * bq;
                    const float vp = V[k * 4 + p], vq = V[k * 4 + q];
                    V[k * 4 + p] = c * vp - s * vq; V[k * 4 + q] = s * vp + c * vq;
                }
            }
        if (off < 1e-7f) break;
    }
    int best = 0;
    float smin = 3.402823466e38f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float sv = B[j] * B[j] + B[4 + j] * B[4 + j] + B[8 + j] * B[8 + j] + B[12 + j] * B[12 + j];
        if (sv < smin) { smin = sv; best = j; }
    }
    float h[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {   // column `best` of V without dynamic register indexing
        float x = V[k * 4];
        x = best == 1 ? V[k * 4 + 1] : x; x = best == 2 ? V[k * 4 + 2] : x; x = best == 3 ? V[k * 4 + 3] : x;
        h[k] = x;
    }
    bool ok = !(fabsf(h[3]) < 1e-10f);                                   // :1108-1110
    const float x = h[0] / h[3], y = h[1] / h[3], z = h[2] / h[3];        // :1113
    ok = ok && isfinite(x) && isfinite(y) && isfinite(z);                 // :1119-1121
    pts[3 * i] = ok ? x : 0.f; pts[3 * i + 1] = ok ? y : 0.f; pts[3 * i + 2] = ok ? z : 0.f;
    valid[i] = ok ? 1 : 0;
}

}  // namespace vio360
