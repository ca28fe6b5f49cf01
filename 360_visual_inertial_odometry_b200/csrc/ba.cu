// ba.cu -- path B: bundle-adjustment factor evaluation on sm_100a.
//
// Kernels (SURVEY.md section 2.1 numbering):
//   K6  k_pose_prepare      per-pose chain (Factors.cpp:357-364), once per pose
//   K7  k_evaluate          residual + Jacobians (+Huber), one thread per observation
//   K8  k_blocks_landmark / k_blocks_pose   J^T J / J^T r blocks by segmented reductions
//   K9  k_schur_warp        factor evaluation fused with per-landmark Schur elimination
//   K10 k_chi2              chi-square classification
// Observations are re-ordered landmark-major once, in set_structure, so that every
// landmark is a contiguous segment; results are scattered back to the caller's order.
#include <unistd.h>

#include <algorithm>
#include <numeric>

#include "common.cuh"
#include "ba_math.cuh"
#include "ba_peer.cuh"
#include "ba_schur.cuh"

using namespace vio360;

struct BaDev {           // device view handed to kernels by value
    int K, M, O;
    const double* u;     // [O] landmark-major order
    const double* v;
    const int* pose;     // [O]
    const int* point;    // [O]
    const int* orig;     // [O] position in the caller's order
    const unsigned char* outlier;  // [O] landmark-major
    const int* lm_ptr;   // [M+1]
    const int* pose_ptr; // [K+1] segments of pose_order
    const int* pose_order; // [O] landmark-major indices grouped by pose
    const double* posebuf;   // [K*33]
    const double* points;    // [M*3]
    const int* free_map;     // [K]  reduced index or -1
    const unsigned char* point_fixed;  // [M]
    const double* D;         // [6K+3M] or nullptr
    CamConst cam;
};

struct vio360_ba {
    vio360_ctx* ctx = nullptr;
    int K = 0, M = 0, O = 0;
    bool has_structure = false, has_params = false, has_eval = false, has_eval_jac = false;
    bool has_reduced = false, reduced_apply_loss = false, solve_dirty = true, chol_attr_set = false;
    DevBuf d_chol, d_linv, d_x, d_dl, d_reachF;
    int max_reach = 0;   // widest column of the envelope, in blocks
    int red_band = 1;    // widest row of the chunk envelope in free-pose blocks (k_schur_reduce grid)
    int band_zero_n = -1;   // S outside the envelope is known to be zero for this n (-1: not known)
    DevBuf d_band;          // sharded builds: compact exchange buffer [n_free][band_w][36] | rhs | cost
    int band_w = 0;         // block columns per row in d_band (0: the last build was dense)
    bool dense_stale = false;   // d_packed has to be re-expanded from d_band before it is read
    PinBuf h_solve;
    int n_free = 0;
    int shard_rank = 0, shard_world = 1, lm_begin = 0, lm_end = 0;
    std::vector<int> lm_ptr, perm, free_map;
    std::vector<unsigned char> pose_fixed, point_fixed;
    double Tcb[16];
    CamConst cam;
    DevBuf d_u, d_v, d_pose, d_point, d_orig, d_outlier, d_lm_ptr, d_pose_ptr, d_pose_order;
    DevBuf d_Twb, d_Tcb, d_posebuf, d_pose_delta, d_points, d_free_map, d_point_fixed, d_D;
    DevBuf d_r, d_Jp, d_Jl, d_partials, d_packed, d_blocks, d_chi2, d_flag, d_blk_part;
    size_t blk_part_bytes = 0;
    PinBuf pin;
    bool use_D = false;
    bool eval_f32 = false;       // vio360_ba_set_eval_precision: per-observation math of the fused build / blocks in float
    // device LM loop (ba_lm.cuh): candidate point, step, per-parameter vectors, state
    DevBuf d_cand_pose_delta, d_cand_points, d_cand_posebuf, d_step, d_colsq, d_scale, d_lm_state, d_lm_part;
    PinBuf h_lm;
    // chunked (atomic-free) Schur path: see ba_schur.cuh
    std::vector<int> h_pose;                 // landmark-major pose index of every observation
    std::vector<SchurChunk> chunks;
    std::vector<int> irregular, wide, cov_ptr, cov_idx, free_pose, reach, slot_table, batch_rng, batch_lm;
    int n_regular_chunks = 0;
    int band_blocks = kSWin;   // block columns of S's band over ALL landmarks of the window (identical on every rank)
    DevBuf d_chunks, d_cov_ptr, d_cov_idx, d_free_pose, d_irregular, d_partS, d_partRhs, d_reach, d_slots, d_batch_rng, d_batch_lm, d_wide, d_wide_inv;
    bool chunks_dirty = true, smem_attr_set = false, chunk_error = false;
    // peer-memory exchange of the sharded reduced system (ba_peer.cuh)
    struct Peer {
        bool exported = false, attached = false;
        DevBuf local;                    // [flags | band parity 0 | band parity 1]
        void* base[kMaxPeers] = {};      // every rank's exchange buffer in this process' address space
        bool ipc_open[kMaxPeers] = {};
        int row_lo[kMaxPeers] = {}, row_hi[kMaxPeers] = {};
        size_t nd = 0, nd_pad = 0;       // doubles per band buffer (payload / padded)
        size_t header_bytes = 0;         // trace words + row flags [kMaxPeers][n_free] + row counters [n_free]
        int n_free = 0, band_w = 0;
        unsigned epoch = 0;              // exchanges completed (all ranks call in lock step)
        int* h_err = nullptr;            // pinned: set by the kernel when a peer never announced
        const void* zero_ptr = nullptr; int zero_n = -1;   // d_packed is known to be zero outside the band
        bool trace = false;                                 // VIO360_PEER_TRACE=1: device time stamps of the exchange kernel
        int world = 0;
        unsigned band_epoch = 0;         // epoch the last build wrote its band for
        unsigned builds = 0;             // sharded builds since the export (the reduce's row counters count over all of them)
        bool stale = false;              // set_structure / set_shard ran after the export: row ranges, band width and
                                         // chunk coverage the peers hold are those of the OLD structure
        int same_device = 1;             // ranks of the world that share this rank's device (tests, MPS)
    } peer;
    int peer_world() const { return peer.world; }
};

static double* peer_band(vio360_ba* ba, int r, unsigned epoch) {
    return reinterpret_cast<double*>(static_cast<char*>(ba->peer.base[r]) + ba->peer.header_bytes) + (size_t)(epoch & 1u) * ba->peer.nd_pad;
}
// How a rank's band reaches the peers (VIO360_PEER_PUBLISH).  0 / 1 / 2: k_schur_reduce counts finished blocks and
// releases rows (0, 1) or the whole band (2) as they complete -- measured on 2 B200 (profiles/r02_peer_publish_modes.txt):
// the per-CTA barrier + fence + atomic costs the reduce 13-40 us and buys nothing while the ranks run in lock step.
// 3 (default): the reduce is untouched and CTA 0 of the pull kernel announces.
static int peer_publish_mode() {
    static const int m = getenv("VIO360_PEER_PUBLISH") ? atoi(getenv("VIO360_PEER_PUBLISH")) : 3;
    return m;
}
static unsigned* peer_row_flags(vio360_ba* ba, int r) {     // rank r's row-flag array [src rank][n_free]
    return reinterpret_cast<unsigned*>(static_cast<char*>(ba->peer.base[r]) + kPeerHeaderBytes);
}
static void peer_release(vio360_ba* ba) {
    auto& P = ba->peer;
    for (int r = 0; r < kMaxPeers; ++r) { if (P.ipc_open[r] && P.base[r]) cudaIpcCloseMemHandle(P.base[r]); P.base[r] = nullptr; P.ipc_open[r] = false; }
    P.local.release();
    if (P.h_err) cudaFreeHost(P.h_err);
    P.h_err = nullptr; P.exported = P.attached = false; P.epoch = 0; P.zero_n = -1; P.stale = false; P.same_device = 1; P.builds = 0;
}

// ------------------------------------------------------------------------ kernels
__global__ void k_pose_prepare(int K, const double* __restrict__ Twb, const double* __restrict__ Tcb,
                               const double* __restrict__ delta, double* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= K) return;
    double T0[16], Tc[16], xi[6], o[kPoseStride];
    for (int i = 0; i < 16; ++i) { T0[i] = Twb[16 * (size_t)k + i]; Tc[i] = Tcb[i]; }
    for (int i = 0; i < 6; ++i) xi[i] = delta[6 * (size_t)k + i];
    pose_prepare(T0, Tc, xi, o);
    for (int i = 0; i < kPoseStride; ++i) out[kPoseStride * (size_t)k + i] = o[i];
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// deterministic block sum -> partials[blockIdx.x]
__device__ __forceinline__ void block_sum_to(double v, double* partials) {
    __shared__ double sm[32];
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) sm[w] = v;
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        double t = lane < nw ? sm[lane] : 0.0;
        t = warp_sum(t);
        if (lane == 0) partials[blockIdx.x] = t;
    }
}

__global__ void k_sum_partials(const double* __restrict__ partials, int n, double* __restrict__ out) {
    double t = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) t += partials[i];
    __shared__ double sm[32];
    t = warp_sum(t);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) sm[w] = t;
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        double s = lane < nw ? sm[lane] : 0.0;
        s = warp_sum(s);
        if (lane == 0) *out = s;
    }
}

// Per-observation evaluation in the precision PT (double: parity path; float: optional fp32 mode, results widened).
template <typename PT>
__device__ __forceinline__ void eval_obs_as(const double* pose, const double* Pw, double u, double v, bool outl, const CamConst& cam,
                                            bool apply_loss, double* r, double* F, double* E) {
    if (sizeof(PT) == 4) {
        float pf[kPoseStride], rf[2], Ff[12], Ef[6];
#pragma unroll
        for (int j = 0; j < kPoseStride; ++j) pf[j] = (float)pose[j];
        const float Pwf[3] = {(float)Pw[0], (float)Pw[1], (float)Pw[2]};
        obs_evaluate<true>(pf, Pwf, (float)u, (float)v, outl, cam, apply_loss, rf, Ff, Ef);
        r[0] = rf[0]; r[1] = rf[1];
#pragma unroll
        for (int j = 0; j < 12; ++j) F[j] = Ff[j];
#pragma unroll
        for (int j = 0; j < 6; ++j) E[j] = Ef[j];
    } else {
        obs_evaluate<true>(pose, Pw, u, v, outl, cam, apply_loss, r, F, E);
    }
}

#include "ba_schur_kernels.cuh"
#include "ba_solve.cuh"
#include "ba_triangulate.cuh"
#include "ba_lm.cuh"

template <bool WANT_JAC>
__global__ void __launch_bounds__(128) k_evaluate(BaDev d, bool apply_loss, double* __restrict__ r_out,
                                                  double* __restrict__ Jp_out, double* __restrict__ Jl_out,
                                                  double* __restrict__ partials) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double cost = 0.0;
    if (i < d.O) {
        const int p = d.pose[i], l = d.point[i];
        double pose[kPoseStride];
#pragma unroll
        for (int j = 0; j < kPoseStride; ++j) pose[j] = __ldg(d.posebuf + kPoseStride * (size_t)p + j);
        const double Pw[3] = {__ldg(d.points + 3 * (size_t)l), __ldg(d.points + 3 * (size_t)l + 1),
                              __ldg(d.points + 3 * (size_t)l + 2)};
        double r[2], F[12], E[6];
        cost = obs_evaluate<WANT_JAC>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
        const size_t o = (size_t)d.orig[i];
        r_out[2 * o] = r[0]; r_out[2 * o + 1] = r[1];
        if (WANT_JAC) {
#pragma unroll
            for (int j = 0; j < 12; ++j) Jp_out[12 * o + j] = F[j];
#pragma unroll
            for (int j = 0; j < 6; ++j) Jl_out[6 * o + j] = E[j];
        }
    }
    block_sum_to(cost, partials);
}

// Optional fp32 mode of the factor evaluation (north star: "1e-4 in an optional fp32 mode"): the same math with
// T = float on the FP32 pipe; prepared poses, points and observations are rounded to float on load, results are
// widened to the double buffers Ceres reads.  The cost partials are summed in double.
template <bool WANT_JAC>
__global__ void __launch_bounds__(128) k_evaluate_f32(BaDev d, bool apply_loss, double* __restrict__ r_out,
                                                      double* __restrict__ Jp_out, double* __restrict__ Jl_out,
                                                      double* __restrict__ partials) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double cost = 0.0;
    if (i < d.O) {
        const int p = d.pose[i], l = d.point[i];
        float pose[kPoseStride];
#pragma unroll
        for (int j = 0; j < kPoseStride; ++j) pose[j] = (float)__ldg(d.posebuf + kPoseStride * (size_t)p + j);
        const float Pw[3] = {(float)__ldg(d.points + 3 * (size_t)l), (float)__ldg(d.points + 3 * (size_t)l + 1),
                             (float)__ldg(d.points + 3 * (size_t)l + 2)};
        float r[2], F[12], E[6];
        cost = (double)obs_evaluate<WANT_JAC>(pose, Pw, (float)d.u[i], (float)d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
        const size_t o = (size_t)d.orig[i];
        r_out[2 * o] = r[0]; r_out[2 * o + 1] = r[1];
        if (WANT_JAC) {
#pragma unroll
            for (int j = 0; j < 12; ++j) Jp_out[12 * o + j] = F[j];
#pragma unroll
            for (int j = 0; j < 6; ++j) Jl_out[6 * o + j] = E[j];
        }
    }
    block_sum_to(cost, partials);
}

__global__ void __launch_bounds__(128) k_chi2(BaDev d, double thr, double* __restrict__ chi2,
                                              unsigned char* __restrict__ flag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.O) return;
    const int p = d.pose[i], l = d.point[i];
    double pose[12];
#pragma unroll
    for (int j = 0; j < 12; ++j) pose[j] = __ldg(d.posebuf + kPoseStride * (size_t)p + j);
    const double Pw[3] = {__ldg(d.points + 3 * (size_t)l), __ldg(d.points + 3 * (size_t)l + 1),
                          __ldg(d.points + 3 * (size_t)l + 2)};
    const double c = obs_chi2(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam);
    const size_t o = (size_t)d.orig[i];
    chi2[o] = c;
    flag[o] = (unsigned char)(c > thr);
}

__device__ __forceinline__ void load_pose(const BaDev& d, int p, double* pose) {
#pragma unroll
    for (int j = 0; j < kPoseStride; ++j) pose[j] = __ldg(d.posebuf + kPoseStride * (size_t)p + j);
}

// K8a: one warp per landmark: Hll (3x3), bl (3) by warp reduction; Hpl per observation.
template <typename PT>
__global__ void __launch_bounds__(128) k_blocks_landmark(BaDev d, bool apply_loss, double* __restrict__ Hll,
                                                         double* __restrict__ bl, double* __restrict__ Hpl) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= d.M) return;
    const int l = warp, lo = d.lm_ptr[l], hi = d.lm_ptr[l + 1];
    const bool lfix = d.point_fixed[l] != 0;
    const double Pw[3] = {d.points[3 * (size_t)l], d.points[3 * (size_t)l + 1], d.points[3 * (size_t)l + 2]};
    double h[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
    for (int base = lo; base < hi; base += 32) {
        const int i = base + lane;
        if (i < hi) {
            const int pk = d.pose[i];
            double pose[kPoseStride], r[2], F[12], E[6];
            load_pose(d, pk, pose);
            eval_obs_as<PT>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
            const bool pfix = d.free_map[pk] < 0;
            if (!lfix) {
                h[0] += E[0] * E[0] + E[3] * E[3]; h[1] += E[0] * E[1] + E[3] * E[4];
                h[2] += E[0] * E[2] + E[3] * E[5]; h[3] += E[1] * E[1] + E[4] * E[4];
                h[4] += E[1] * E[2] + E[4] * E[5]; h[5] += E[2] * E[2] + E[5] * E[5];
                g[0] += E[0] * r[0] + E[3] * r[1]; g[1] += E[1] * r[0] + E[4] * r[1];
                g[2] += E[2] * r[0] + E[5] * r[1];
            }
            if (Hpl) {
                const size_t o = 18 * (size_t)d.orig[i];
#pragma unroll
                for (int a = 0; a < 6; ++a)
#pragma unroll
                    for (int b = 0; b < 3; ++b)
                        Hpl[o + a * 3 + b] = (lfix || pfix) ? 0.0 : F[a] * E[b] + F[6 + a] * E[3 + b];
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) h[j] = warp_sum(h[j]);
#pragma unroll
    for (int j = 0; j < 3; ++j) g[j] = warp_sum(g[j]);
    if (lane == 0) {
        if (Hll) {
            double* H = Hll + 9 * (size_t)l;
            H[0] = h[0]; H[1] = h[1]; H[2] = h[2]; H[3] = h[1]; H[4] = h[3]; H[5] = h[4];
            H[6] = h[2]; H[7] = h[4]; H[8] = h[5];
        }
        if (bl) { bl[3 * (size_t)l] = g[0]; bl[3 * (size_t)l + 1] = g[1]; bl[3 * (size_t)l + 2] = g[2]; }
    }
}

// K8b: one block per pose: Hpp (6x6), bp (6) by a segmented block reduction over the
// pose's observations (no atomics, fixed summation order).
// gridDim.y > 1: the pose's observation list is cut into gridDim.y slices, one CTA each (a 10-keyframe window is 10
// poses of ~1900 observations: ten CTAs would leave the GPU empty); every CTA leaves its 27 sums in `part`, and the
// CTA that arrives last at the pose's counter adds the slices in slice order (fixed order, no atomics on the sums)
// and resets the counter for the next launch.
template <typename PT>
__global__ void __launch_bounds__(256) k_blocks_pose(BaDev d, bool apply_loss, bool skip_fixed_points_never,
                                                     double* __restrict__ Hpp, double* __restrict__ bp,
                                                     double* __restrict__ part, unsigned int* __restrict__ arrived) {
    const int k = blockIdx.x;
    const int nsl = gridDim.y, sl = blockIdx.y;
    const int lo0 = d.pose_ptr[k], hi0 = d.pose_ptr[k + 1];
    const int lo = lo0 + (int)((long long)(hi0 - lo0) * sl / nsl), hi = lo0 + (int)((long long)(hi0 - lo0) * (sl + 1) / nsl);
    double acc[27];
#pragma unroll
    for (int j = 0; j < 27; ++j) acc[j] = 0.0;
    if (d.free_map[k] >= 0) {
        double pose[kPoseStride];
        load_pose(d, k, pose);
        for (int q = lo + threadIdx.x; q < hi; q += blockDim.x) {
            const int i = d.pose_order[q];
            const int l = d.point[i];
            const double Pw[3] = {d.points[3 * (size_t)l], d.points[3 * (size_t)l + 1], d.points[3 * (size_t)l + 2]};
            double r[2], F[12], E[6];
            eval_obs_as<PT>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
            int t = 0;
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
                for (int b = a; b < 6; ++b) acc[t++] += F[a] * F[b] + F[6 + a] * F[6 + b];
#pragma unroll
            for (int a = 0; a < 6; ++a) acc[21 + a] += F[a] * r[0] + F[6 + a] * r[1];
        }
    }
    (void)skip_fixed_points_never;
    __shared__ double sm[8][27];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < 27; ++j) {
        const double s = warp_sum(acc[j]);
        if (lane == 0) sm[w][j] = s;
    }
    __syncthreads();
    if (threadIdx.x < 27) {
        double s = 0.0;
        for (int ww = 0; ww < (int)(blockDim.x >> 5); ++ww) s += sm[ww][threadIdx.x];
        sm[0][threadIdx.x] = s;
    }
    __syncthreads();
    if (nsl > 1) {
        __shared__ int s_last;
        if (threadIdx.x < 27) part[((size_t)k * nsl + sl) * 27 + threadIdx.x] = sm[0][threadIdx.x];
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_last = atomicAdd(arrived + k, 1u) == (unsigned int)(nsl - 1);
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        if (threadIdx.x < 27) {
            double s = 0.0;
            for (int j = 0; j < nsl; ++j) s += __ldcg(part + ((size_t)k * nsl + j) * 27 + threadIdx.x);
            sm[0][threadIdx.x] = s;
        }
        if (threadIdx.x == 0) arrived[k] = 0u;
        __syncthreads();
    }
    if (threadIdx.x < 36 && Hpp) {
        const int a = threadIdx.x / 6, b = threadIdx.x % 6;
        const int aa = a < b ? a : b, bb = a < b ? b : a;
        const int t = aa * 6 - aa * (aa - 1) / 2 + (bb - aa);
        Hpp[36 * (size_t)k + threadIdx.x] = sm[0][t];
    }
    if (threadIdx.x < 6 && bp) bp[6 * (size_t)k + threadIdx.x] = sm[0][21 + threadIdx.x];
}

// K9 (general form): one warp per landmark.  Lanes hold observations; E'E and E'r are
// reduced with shuffles; each lane then owns the row-block of its own pose and walks
// the landmark's observations (broadcast by shuffle) to emit -Y_i W_k^T into the upper
// block triangle of S.  Accumulation into S / rhs uses fp64 RED (order-insensitive to
// ~1e-16); the tiled kernel below removes the atomics for the regular case.
__global__ void __launch_bounds__(128) k_schur_warp(BaDev d, int lm_begin, int lm_end, const int* __restrict__ lm_list,
                                                    bool apply_loss, double* __restrict__ S, int n,
                                                    double* __restrict__ rhs, double* __restrict__ partials) {
    // lm_list != nullptr: process landmarks lm_list[0 .. lm_end) (the ones the chunked kernel cannot take)
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const bool valid = lm_list ? warp < lm_end : lm_begin + warp < lm_end;
    const int l = !valid ? 0 : (lm_list ? lm_list[warp] : lm_begin + warp);
    double cost = 0.0;
    if (valid) {
        const int lo = d.lm_ptr[l], hi = d.lm_ptr[l + 1];
        const bool lfix = d.point_fixed[l] != 0;
        const double Pw[3] = {d.points[3 * (size_t)l], d.points[3 * (size_t)l + 1], d.points[3 * (size_t)l + 2]};
        const bool single = (hi - lo) <= 32;
        double r[2], F[12], E[6];
        int pf = -1;
        // ---- pass 1: E'E (+D^2), g = E'r, cost
        double h[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
        for (int base = lo; base < hi; base += 32) {
            const int i = base + lane;
            if (i < hi) {
                const int pk = d.pose[i];
                double pose[kPoseStride];
                load_pose(d, pk, pose);
                cost += obs_evaluate<true>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
                pf = d.free_map[pk];
                h[0] += E[0] * E[0] + E[3] * E[3]; h[1] += E[0] * E[1] + E[3] * E[4];
                h[2] += E[0] * E[2] + E[3] * E[5]; h[3] += E[1] * E[1] + E[4] * E[4];
                h[4] += E[1] * E[2] + E[4] * E[5]; h[5] += E[2] * E[2] + E[5] * E[5];
                g[0] += E[0] * r[0] + E[3] * r[1]; g[1] += E[1] * r[0] + E[4] * r[1];
                g[2] += E[2] * r[0] + E[5] * r[1];
            }
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) h[j] = warp_sum(h[j]);
#pragma unroll
        for (int j = 0; j < 3; ++j) g[j] = warp_sum(g[j]);
        double inv[6] = {0, 0, 0, 0, 0, 0}, ig[3] = {0, 0, 0};
        if (!lfix) {
            if (d.D) {
                const double d0 = d.D[6 * (size_t)d.K + 3 * (size_t)l], d1 = d.D[6 * (size_t)d.K + 3 * (size_t)l + 1],
                             d2 = d.D[6 * (size_t)d.K + 3 * (size_t)l + 2];
                h[0] += d0 * d0; h[3] += d1 * d1; h[5] += d2 * d2;
            }
            inv_spd3(h, inv);
            bool ok = true;
#pragma unroll
            for (int j = 0; j < 6; ++j) ok = ok && isfinite(inv[j]);
            if (!ok) { for (int j = 0; j < 6; ++j) inv[j] = 0.0; }   // rank-deficient landmark: not eliminated
            ig[0] = inv[0] * g[0] + inv[1] * g[1] + inv[2] * g[2];
            ig[1] = inv[1] * g[0] + inv[3] * g[1] + inv[4] * g[2];
            ig[2] = inv[2] * g[0] + inv[4] * g[1] + inv[5] * g[2];
        }
        // ---- pass 2
        for (int baseA = lo; baseA < hi; baseA += 32) {
            const int i = baseA + lane;
            const bool act = i < hi;
            if (!single && act) {
                const int pk = d.pose[i];
                double pose[kPoseStride];
                load_pose(d, pk, pose);
                obs_evaluate<true>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
                pf = d.free_map[pk];
            }
            const int pA = act ? pf : -1;
            double W[18], Y[18];
#pragma unroll
            for (int j = 0; j < 18; ++j) { W[j] = 0.0; Y[j] = 0.0; }
            if (pA >= 0) {
                // S_pp += F'F ; rhs_p += F'(r - E ig)
                const double s0 = r[0] - (E[0] * ig[0] + E[1] * ig[1] + E[2] * ig[2]);
                const double s1 = r[1] - (E[3] * ig[0] + E[4] * ig[1] + E[5] * ig[2]);
#pragma unroll
                for (int a = 0; a < 6; ++a) {
#pragma unroll
                    for (int b = 0; b < 6; ++b)
                        atomicAdd(S + (size_t)(6 * pA + a) * n + 6 * pA + b, F[a] * F[b] + F[6 + a] * F[6 + b]);
                    atomicAdd(rhs + 6 * pA + a, F[a] * s0 + F[6 + a] * s1);
                }
                if (!lfix) {
#pragma unroll
                    for (int a = 0; a < 6; ++a) {
                        W[a * 3 + 0] = F[a] * E[0] + F[6 + a] * E[3];
                        W[a * 3 + 1] = F[a] * E[1] + F[6 + a] * E[4];
                        W[a * 3 + 2] = F[a] * E[2] + F[6 + a] * E[5];
                        Y[a * 3 + 0] = W[a * 3] * inv[0] + W[a * 3 + 1] * inv[1] + W[a * 3 + 2] * inv[2];
                        Y[a * 3 + 1] = W[a * 3] * inv[1] + W[a * 3 + 1] * inv[3] + W[a * 3 + 2] * inv[4];
                        Y[a * 3 + 2] = W[a * 3] * inv[2] + W[a * 3 + 1] * inv[4] + W[a * 3 + 2] * inv[5];
                    }
                }
            }
            if (lfix) continue;
            for (int baseB = lo; baseB < hi; baseB += 32) {
                double WB[18];
                int pB = -1;
                if (baseB == baseA) {
#pragma unroll
                    for (int j = 0; j < 18; ++j) WB[j] = W[j];
                    pB = pA;
                } else {
                    const int k = baseB + lane;
#pragma unroll
                    for (int j = 0; j < 18; ++j) WB[j] = 0.0;
                    if (k < hi) {
                        const int pk = d.pose[k];
                        double pose[kPoseStride], r2[2], F2[12], E2[6];
                        load_pose(d, pk, pose);
                        obs_evaluate<true>(pose, Pw, d.u[k], d.v[k], d.outlier[k] != 0, d.cam, apply_loss, r2, F2, E2);
                        pB = d.free_map[pk];
                        if (pB >= 0) {
#pragma unroll
                            for (int a = 0; a < 6; ++a) {
                                WB[a * 3 + 0] = F2[a] * E2[0] + F2[6 + a] * E2[3];
                                WB[a * 3 + 1] = F2[a] * E2[1] + F2[6 + a] * E2[4];
                                WB[a * 3 + 2] = F2[a] * E2[2] + F2[6 + a] * E2[5];
                            }
                        }
                    }
                }
                const int nB = min(32, hi - baseB);
                for (int k = 0; k < nB; ++k) {
                    const int pk = __shfl_sync(0xffffffffu, pB, k);
                    double Wk[18];
#pragma unroll
                    for (int j = 0; j < 18; ++j) Wk[j] = __shfl_sync(0xffffffffu, WB[j], k);
                    if (pA >= 0 && pk >= 0 && pA <= pk) {
                        double* Sb = S + (size_t)(6 * pA) * n + 6 * pk;
#pragma unroll
                        for (int a = 0; a < 6; ++a)
#pragma unroll
                            for (int b = 0; b < 6; ++b)
                                atomicAdd(Sb + (size_t)a * n + b,
                                          -(Y[a * 3] * Wk[b * 3] + Y[a * 3 + 1] * Wk[b * 3 + 1] + Y[a * 3 + 2] * Wk[b * 3 + 2]));
                    }
                }
            }
        }
    }
    block_sum_to(cost, partials);
}

// mirror the upper block triangle into the lower one and add the pose part of D^2
__global__ void k_finish_S(double* __restrict__ S, int n, const double* __restrict__ D, const int* __restrict__ free_map,
                           int K, bool add_diag) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * n) return;
    const int i = idx / n, j = idx % n;
    const int bi = i / 6, bj = j / 6;
    if (bi > bj) { S[idx] = S[(size_t)j * n + i]; return; }
    if (bi == bj && i > j) { S[idx] = S[(size_t)j * n + i]; return; }   // exact symmetry inside diagonal blocks
    if (i == j && add_diag && D) {
        // find pose k with free_map[k] == bi (K is small relative to n*n work; linear scan per diagonal entry)
        for (int k = 0; k < K; ++k)
            if (free_map[k] == bi) { const double dd = D[6 * (size_t)k + (i - 6 * bi)]; S[idx] += dd * dd; break; }
    }
}

// --------------------------------------------------------------------- host side
static BaDev make_dev(const vio360_ba* ba) {
    BaDev d;
    d.K = ba->K; d.M = ba->M; d.O = ba->O;
    d.u = ba->d_u.as<double>(); d.v = ba->d_v.as<double>();
    d.pose = ba->d_pose.as<int>(); d.point = ba->d_point.as<int>(); d.orig = ba->d_orig.as<int>();
    d.outlier = ba->d_outlier.as<unsigned char>();
    d.lm_ptr = ba->d_lm_ptr.as<int>(); d.pose_ptr = ba->d_pose_ptr.as<int>();
    d.pose_order = ba->d_pose_order.as<int>();
    d.posebuf = ba->d_posebuf.as<double>(); d.points = ba->d_points.as<double>();
    d.free_map = ba->d_free_map.as<int>(); d.point_fixed = ba->d_point_fixed.as<unsigned char>();
    d.D = ba->use_D ? ba->d_D.as<double>() : nullptr;
    d.cam = ba->cam;
    return d;
}

extern "C" int vio360_ba_create(vio360_ctx* ctx, int K, int M, int O, vio360_ba** out) {
    if (!ctx || !out || K <= 0 || M <= 0 || O < 0) return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_create: bad sizes");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    vio360_ba* ba = new vio360_ba();
    ba->ctx = ctx; ba->K = K; ba->M = M; ba->O = O;
    *out = ba;
    return VIO360_OK;
}

extern "C" void vio360_ba_destroy(vio360_ba* ba) {
    if (!ba) return;
    cudaSetDevice(ba->ctx->device);
    cudaStreamSynchronize(ba->ctx->stream);
    DevBuf* bufs[] = {&ba->d_u, &ba->d_v, &ba->d_pose, &ba->d_point, &ba->d_orig, &ba->d_outlier, &ba->d_lm_ptr,
                      &ba->d_pose_ptr, &ba->d_pose_order, &ba->d_Twb, &ba->d_Tcb, &ba->d_posebuf, &ba->d_pose_delta,
                      &ba->d_points, &ba->d_free_map, &ba->d_point_fixed, &ba->d_D, &ba->d_r, &ba->d_Jp, &ba->d_Jl,
                      &ba->d_partials, &ba->d_packed, &ba->d_blocks, &ba->d_chi2, &ba->d_flag,
                      &ba->d_chunks, &ba->d_cov_ptr, &ba->d_cov_idx, &ba->d_free_pose, &ba->d_irregular, &ba->d_partS, &ba->d_partRhs, &ba->d_reach, &ba->d_slots, &ba->d_batch_rng,
                      &ba->d_batch_lm, &ba->d_wide, &ba->d_wide_inv,
                      &ba->d_chol, &ba->d_linv, &ba->d_x, &ba->d_dl, &ba->d_reachF, &ba->d_band,
                      &ba->d_cand_pose_delta, &ba->d_cand_points, &ba->d_cand_posebuf, &ba->d_step, &ba->d_colsq, &ba->d_scale,
                      &ba->d_lm_state, &ba->d_lm_part, &ba->d_blk_part};
    for (DevBuf* b : bufs) b->release();
    peer_release(ba);
    ba->pin.release();
    ba->h_solve.release();
    ba->h_lm.release();
    delete ba;
}

// Partition the shard's landmarks [lm_begin, lm_end) into chunks for k_schur_tile.
//  * regular landmarks (observing poses within kSWin consecutive poses): contiguous landmark ranges whose poses
//    fit one window of kSWin poses -- one CTA per chunk, everything about a landmark happens inside it;
//  * WIDE landmarks (span > kSWin, e.g. a loop closure): poses are cut into groups of kSGrp = kSWin / 2; a landmark
//    joins one job per pair (a <= b) of groups it is seen from.  A job's CTA holds group a in pose slots 0..11 and
//    group b in slots 12..23, gets (E'E + D^2)^-1 of the whole landmark from k_wide_prepare (pass 1) and contributes
//    the S blocks of pose pairs (a, b) only (a == b: also F'F, F'r and the cost of the observations of group a);
//  * landmarks with two observations from one pose (never produced by the reference) stay with k_schur_warp.
// Also builds, per pose, the list of chunk slabs covering its block row (k_schur_reduce).
static void build_chunks(vio360_ba* ba) {
    const int K = ba->K;
    ba->chunks.clear(); ba->irregular.clear(); ba->wide.clear();
    std::vector<int> chunk_lms;                       // landmark ids of all chunks, batch-padded (-1)
    const int nlm = ba->lm_end - ba->lm_begin;
    // a chunk is a whole number of kSMaxLm-landmark batches (a partly filled batch costs a full one), sized so
    // that the shard fills `waves` rounds of one CTA per SM (measured on the 1M-observation window: one round and
    // two rounds tie on one GPU, one round wins on shards -- fewer slabs for k_schur_reduce, no short second round)
    static const int waves = getenv("VIO360_BA_CHUNK_WAVES") ? std::max(1, atoi(getenv("VIO360_BA_CHUNK_WAVES"))) : 1;
    const int nbatch_all = (nlm + kSMaxLm - 1) / kSMaxLm, slots = std::max(1, waves * ba->ctx->num_sms);
    const int target = kSMaxLm * std::max(1, (nbatch_all + slots - 1) / slots);
    std::vector<int> seen(K, -1);
    int nbatch_total = 0;
    auto push_chunk = [&](const std::vector<int>& lms, int pbase, int span, int pbaseB, int spanB, int flags) {
        SchurChunk c{0, (int)lms.size(), pbase, span, nbatch_total, pbaseB, spanB, flags};
        const int nb = ((int)lms.size() + kSMaxLm - 1) / kSMaxLm;
        for (int i = 0; i < nb * kSMaxLm; ++i) chunk_lms.push_back(i < (int)lms.size() ? lms[i] : -1);
        nbatch_total += nb;
        ba->chunks.push_back(c);
    };
    std::vector<int> cur;
    int cmin = K, cmax = -1;
    auto close = [&]() {
        if (!cur.empty() && cmax >= cmin) push_chunk(cur, cmin, cmax - cmin + 1, 0, 0, 0);
        cur.clear(); cmin = K; cmax = -1;
    };
    for (int l = ba->lm_begin; l < ba->lm_end; ++l) {
        const int lo = ba->lm_ptr[l], hi = ba->lm_ptr[l + 1];
        if (hi == lo) continue;                       // unobserved landmark: contributes nothing
        int pmin = K, pmax = -1;
        bool dup = false;
        for (int q = lo; q < hi; ++q) {
            const int p = ba->h_pose[q];
            if (seen[p] == l) dup = true;
            seen[p] = l;
            pmin = std::min(pmin, p); pmax = std::max(pmax, p);
        }
        if (dup) { ba->irregular.push_back(l); continue; }
        if (pmax - pmin + 1 > kSWin) { ba->wide.push_back(l); continue; }
        const int nmin = std::min(cmin, pmin), nmax = std::max(cmax, pmax);
        if (cmax >= cmin && (nmax - nmin + 1 > kSWin || (int)cur.size() >= target)) {
            close();
            cmin = pmin; cmax = pmax;
        } else {
            cmin = nmin; cmax = nmax;
        }
        cur.push_back(l);
    }
    close();
    ba->n_regular_chunks = (int)ba->chunks.size();
    // ---- jobs of the wide landmarks: (group a, group b) -> landmarks seen from both
    if (!ba->wide.empty()) {
        const int ngrp = (K + kSGrp - 1) / kSGrp;
        std::vector<std::vector<int>> job((size_t)ngrp * ngrp);
        std::vector<int> grp;
        for (int l : ba->wide) {
            grp.clear();
            for (int q = ba->lm_ptr[l]; q < ba->lm_ptr[l + 1]; ++q) grp.push_back(ba->h_pose[q] / kSGrp);
            std::sort(grp.begin(), grp.end());
            grp.erase(std::unique(grp.begin(), grp.end()), grp.end());
            for (size_t i = 0; i < grp.size(); ++i)
                for (size_t j = i; j < grp.size(); ++j) job[(size_t)grp[i] * ngrp + grp[j]].push_back(l);
        }
        size_t wide_batches = 0;
        for (const auto& v : job) wide_batches += (v.size() + kSMaxLm - 1) / kSMaxLm;
        const int wtarget = kSMaxLm * (int)std::max<size_t>(1, (wide_batches + slots - 1) / slots);
        std::vector<int> part;
        for (int a = 0; a < ngrp; ++a)
            for (int b2 = a; b2 < ngrp; ++b2) {
                const std::vector<int>& v = job[(size_t)a * ngrp + b2];
                for (size_t i0 = 0; i0 < v.size(); i0 += (size_t)wtarget) {
                    part.assign(v.begin() + i0, v.begin() + std::min(v.size(), i0 + (size_t)wtarget));
                    const int pa = a * kSGrp, sa = std::min(kSGrp, K - pa), pb = b2 * kSGrp, sb = std::min(kSGrp, K - pb);
                    if (a == b2) push_chunk(part, pa, sa, 0, 0, kChunkWide);
                    else push_chunk(part, pa, sa, pb, sb, kChunkWide | kChunkCross);
                }
            }
    }
    // ---- per pose: the slabs covering its block row.  entry = {chunk, slot of the pose, first column pose |
    // columns << 16, slot of the first column}: the slab holds block (pose, q) for first <= q < first + columns
    ba->cov_ptr.assign(K + 1, 0);
    auto for_rows = [&](auto&& fn) {
        for (int c = 0; c < (int)ba->chunks.size(); ++c) {
            const SchurChunk& ch = ba->chunks[c];
            if (ch.flags & kChunkCross) { for (int p = ch.pbase; p < ch.pbase + ch.span; ++p) fn(c, p, p - ch.pbase, ch.pbaseB, ch.spanB, kSGrp); }
            else { for (int p = ch.pbase; p < ch.pbase + ch.span; ++p) fn(c, p, p - ch.pbase, ch.pbase, ch.span, 0); }
        }
    };
    for_rows([&](int, int p, int, int, int, int) { ba->cov_ptr[p + 1]++; });
    for (int p = 0; p < K; ++p) ba->cov_ptr[p + 1] += ba->cov_ptr[p];
    ba->cov_idx.assign(4 * (size_t)std::max(1, ba->cov_ptr[K]), 0);
    std::vector<int> fill(ba->cov_ptr.begin(), ba->cov_ptr.end() - 1);
    ba->reach.assign(K, -1);
    for_rows([&](int c, int p, int slot, int first, int cols, int slot0) {
        int* e = ba->cov_idx.data() + 4 * (size_t)fill[p]++;
        e[0] = c; e[1] = slot; e[2] = first | (cols << 16); e[3] = slot0;
        ba->reach[p] = std::max(ba->reach[p], first + cols - 1);
    });
    // ---- slot table: per batch of kSMaxLm landmarks, the observation held by thread (pose slot, landmark slot)
    ba->batch_lm = chunk_lms;
    if (ba->batch_lm.empty()) ba->batch_lm.assign(kSMaxLm, -1);
    ba->slot_table.assign((size_t)std::max(1, nbatch_total) * kSProd, -1);
    ba->batch_rng.assign((size_t)std::max(1, nbatch_total), 0xff);   // empty: lo = 255 > hi = 0
    for (const SchurChunk& c : ba->chunks)
        for (int i = 0; i < c.lm_end; ++i) {
            const int b = i / kSMaxLm, ls = i % kSMaxLm;
            const int l = ba->batch_lm[(size_t)(c.batch0 + b) * kSMaxLm + ls];
            int* row = ba->slot_table.data() + (size_t)(c.batch0 + b) * kSProd;
            for (int q = ba->lm_ptr[l]; q < ba->lm_ptr[l + 1]; ++q) {
                const int p = ba->h_pose[q];
                int pw = -1;
                if (p >= c.pbase && p < c.pbase + c.span) pw = p - c.pbase;
                else if ((c.flags & kChunkCross) && p >= c.pbaseB && p < c.pbaseB + c.spanB) pw = kSGrp + p - c.pbaseB;
                if (pw < 0) { if (!(c.flags & kChunkWide)) ba->chunk_error = true; continue; }   // a wide landmark's other groups
                if (row[pw * kSMaxLm + ls] != -1) { ba->chunk_error = true; continue; }
                row[pw * kSMaxLm + ls] = q;
                int& rg = ba->batch_rng[c.batch0 + b];   // used pose slots of the batch: lo | hi << 8
                const int lo = std::min(rg & 0xff, pw), hi = std::max((rg >> 8) & 0xff, pw);
                rg = lo | (hi << 8);
            }
        }
    ba->free_pose.clear();
    for (int k = 0; k < K; ++k) if (ba->free_map[k] >= 0) ba->free_pose.push_back(k);
    ba->red_band = 1;
    for (int bi = 0, bj = 0; bi < (int)ba->free_pose.size(); ++bi) {   // reach is not monotone: scan per row
        bj = bi;
        while (bj + 1 < (int)ba->free_pose.size() && ba->free_pose[bj + 1] <= ba->reach[ba->free_pose[bi]]) ++bj;
        ba->red_band = std::max(ba->red_band, bj - bi + 1);
    }
    ba->band_zero_n = -1;
    ba->chunks_dirty = true;
}

static void compute_shard(vio360_ba* ba) {
    // contiguous landmark ranges balanced by sum of (degree^2 + degree)
    const int M = ba->M;
    std::vector<double> w(M + 1, 0.0);
    for (int l = 0; l < M; ++l) {
        const double dg = ba->lm_ptr[l + 1] - ba->lm_ptr[l];
        w[l + 1] = w[l] + dg * dg + 4.0 * dg;
    }
    auto cut = [&](int r) {
        if (r <= 0) return 0;
        if (r >= ba->shard_world) return M;
        const double target = w[M] * r / ba->shard_world;
        return (int)(std::lower_bound(w.begin(), w.end(), target) - w.begin());
    };
    ba->lm_begin = std::min(cut(ba->shard_rank), M);
    ba->lm_end = std::min(cut(ba->shard_rank + 1), M);
    if (ba->lm_end < ba->lm_begin) ba->lm_end = ba->lm_begin;
    build_chunks(ba);
}

extern "C" int vio360_ba_set_shard(vio360_ba* ba, int rank, int world) {
    if (!ba) return VIO360_ERR_INVALID;
    if (world < 1 || rank < 0 || rank >= world) return vio360_fail(ba->ctx, VIO360_ERR_INVALID, "ba_set_shard: bad rank/world");
    ba->shard_rank = rank; ba->shard_world = world;
    if (ba->peer.exported) ba->peer.stale = true;   // the peers hold the old shard's row ranges: detach + export again
    if (ba->has_structure) compute_shard(ba);
    return VIO360_OK;
}

extern "C" int vio360_ba_set_structure(vio360_ba* ba, const int32_t* pose_idx, const int32_t* point_idx,
                                       const double* obs_u, const double* obs_v, const double* T_wb_init,
                                       const double* T_cb, double cols, double rows, const double* info,
                                       double huber_delta, const uint8_t* pose_fixed, const uint8_t* point_fixed) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!pose_idx || !point_idx || !obs_u || !obs_v || !T_wb_init || !T_cb || !info)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_set_structure: null argument");
    const int K = ba->K, M = ba->M, O = ba->O;
    for (int i = 0; i < O; ++i)
        if (pose_idx[i] < 0 || pose_idx[i] >= K || point_idx[i] < 0 || point_idx[i] >= M)
            return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_set_structure: index out of range at observation %d", i);
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    // landmark-major stable ordering (counting sort), pose-major ordering of those
    ba->lm_ptr.assign(M + 1, 0);
    for (int i = 0; i < O; ++i) ba->lm_ptr[point_idx[i] + 1]++;
    for (int l = 0; l < M; ++l) ba->lm_ptr[l + 1] += ba->lm_ptr[l];
    ba->perm.resize(O);
    {
        std::vector<int> fill(ba->lm_ptr.begin(), ba->lm_ptr.end() - 1);
        for (int i = 0; i < O; ++i) ba->perm[fill[point_idx[i]]++] = i;
    }
    std::vector<int> s_pose(O), s_point(O), pose_ptr(K + 1, 0), pose_order(O);
    std::vector<double> s_u(O), s_v(O);
    for (int q = 0; q < O; ++q) {
        const int i = ba->perm[q];
        s_pose[q] = pose_idx[i]; s_point[q] = point_idx[i]; s_u[q] = obs_u[i]; s_v[q] = obs_v[i];
        pose_ptr[pose_idx[i] + 1]++;
    }
    for (int k = 0; k < K; ++k) pose_ptr[k + 1] += pose_ptr[k];
    {
        std::vector<int> fill(pose_ptr.begin(), pose_ptr.end() - 1);
        for (int q = 0; q < O; ++q) pose_order[fill[s_pose[q]]++] = q;
    }
    ba->pose_fixed.assign(K, 0); ba->point_fixed.assign(M, 0);
    if (pose_fixed) ba->pose_fixed.assign(pose_fixed, pose_fixed + K);
    if (point_fixed) ba->point_fixed.assign(point_fixed, point_fixed + M);
    ba->free_map.assign(K, -1);
    ba->n_free = 0;
    for (int k = 0; k < K; ++k) if (!ba->pose_fixed[k]) ba->free_map[k] = ba->n_free++;
    memcpy(ba->Tcb, T_cb, sizeof(ba->Tcb));
    cam_const_init(ba->cam, cols, rows, info, T_cb, huber_delta);
    ba->h_pose = s_pose;
    if (ba->peer.exported) ba->peer.stale = true;   // the exported row ranges / band width describe the old structure
    {   // block columns of the band of S over ALL landmarks (every rank computes the same number): at least kSWin, the
        // chunk window, and as wide as the widest landmark's span of free poses
        int wmax = 1;
        for (int l = 0; l < M; ++l) {
            int fmin = K, fmax = -1;
            for (int q = ba->lm_ptr[l]; q < ba->lm_ptr[l + 1]; ++q) {
                const int f = ba->free_map[s_pose[q]];
                if (f >= 0) { fmin = std::min(fmin, f); fmax = std::max(fmax, f); }
            }
            if (fmax >= fmin) wmax = std::max(wmax, fmax - fmin + 1);
        }
        ba->band_blocks = std::max(kSWin, wmax);
    }
    compute_shard(ba);

    cudaStream_t st = ctx->stream;
    const size_t Os = (size_t)std::max(O, 1);
    VIO_CUDA(ctx, ba->d_u.reserve(Os * 8)); VIO_CUDA(ctx, ba->d_v.reserve(Os * 8));
    VIO_CUDA(ctx, ba->d_pose.reserve(Os * 4)); VIO_CUDA(ctx, ba->d_point.reserve(Os * 4));
    VIO_CUDA(ctx, ba->d_orig.reserve(Os * 4)); VIO_CUDA(ctx, ba->d_outlier.reserve(Os));
    VIO_CUDA(ctx, ba->d_lm_ptr.reserve((size_t)(M + 1) * 4)); VIO_CUDA(ctx, ba->d_pose_ptr.reserve((size_t)(K + 1) * 4));
    VIO_CUDA(ctx, ba->d_pose_order.reserve(Os * 4));
    VIO_CUDA(ctx, ba->d_Twb.reserve((size_t)K * 16 * 8)); VIO_CUDA(ctx, ba->d_Tcb.reserve(16 * 8));
    VIO_CUDA(ctx, ba->d_posebuf.reserve((size_t)K * kPoseStride * 8));
    VIO_CUDA(ctx, ba->d_pose_delta.reserve((size_t)K * 6 * 8)); VIO_CUDA(ctx, ba->d_points.reserve((size_t)M * 3 * 8));
    VIO_CUDA(ctx, ba->d_free_map.reserve((size_t)K * 4)); VIO_CUDA(ctx, ba->d_point_fixed.reserve((size_t)M));
    VIO_CUDA(ctx, ba->d_D.reserve(((size_t)6 * K + (size_t)3 * M) * 8));
    if (O > 0) {
        VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_u.p, s_u.data(), Os * 8, cudaMemcpyHostToDevice, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_v.p, s_v.data(), Os * 8, cudaMemcpyHostToDevice, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_pose.p, s_pose.data(), Os * 4, cudaMemcpyHostToDevice, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_point.p, s_point.data(), Os * 4, cudaMemcpyHostToDevice, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_orig.p, ba->perm.data(), Os * 4, cudaMemcpyHostToDevice, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_pose_order.p, pose_order.data(), Os * 4, cudaMemcpyHostToDevice, st));
        VIO_CUDA(ctx, cudaMemsetAsync(ba->d_outlier.p, 0, Os, st));
    }
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_lm_ptr.p, ba->lm_ptr.data(), (size_t)(M + 1) * 4, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_pose_ptr.p, pose_ptr.data(), (size_t)(K + 1) * 4, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_Twb.p, T_wb_init, (size_t)K * 16 * 8, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_Tcb.p, T_cb, 16 * 8, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_free_map.p, ba->free_map.data(), (size_t)K * 4, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_point_fixed.p, ba->point_fixed.data(), (size_t)M, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));   // host vectors above go out of scope
    ba->has_structure = true; ba->has_params = false; ba->has_eval = false; ba->has_reduced = false; ba->solve_dirty = true;
    return VIO360_OK;
}

extern "C" int vio360_ba_set_outliers(vio360_ba* ba, const uint8_t* outlier) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_structure) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_set_outliers before ba_set_structure");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    if (ba->O == 0) return VIO360_OK;
    if (!outlier) { VIO_CUDA(ctx, cudaMemsetAsync(ba->d_outlier.p, 0, (size_t)ba->O, ctx->stream)); return VIO360_OK; }
    VIO_CUDA(ctx, ba->pin.reserve((size_t)ba->O));
    unsigned char* h = ba->pin.as<unsigned char>();
    for (int q = 0; q < ba->O; ++q) h[q] = outlier[ba->perm[q]];
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_outlier.p, h, (size_t)ba->O, cudaMemcpyHostToDevice, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VIO360_OK;
}

extern "C" int vio360_ba_set_parameters(vio360_ba* ba, const double* pose_delta, const double* points) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_structure) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_set_parameters before ba_set_structure");
    if (!pose_delta || !points) return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_set_parameters: null argument");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t nb_p = (size_t)ba->K * 6 * 8, nb_l = (size_t)ba->M * 3 * 8;
    VIO_CUDA(ctx, ba->pin.reserve(nb_p + nb_l));
    memcpy(ba->pin.p, pose_delta, nb_p);
    memcpy((char*)ba->pin.p + nb_p, points, nb_l);
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_pose_delta.p, ba->pin.p, nb_p, cudaMemcpyHostToDevice, ctx->stream));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_points.p, (char*)ba->pin.p + nb_p, nb_l, cudaMemcpyHostToDevice, ctx->stream));
    k_pose_prepare<<<(ba->K + 63) / 64, 64, 0, ctx->stream>>>(ba->K, ba->d_Twb.as<double>(), ba->d_Tcb.as<double>(),
                                                             ba->d_pose_delta.as<double>(), ba->d_posebuf.as<double>());
    VIO_LAUNCH_CHECK(ctx);
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // pinned staging is reused by the next call
    ba->has_params = true; ba->has_eval = false; ba->has_reduced = false;
    return VIO360_OK;
}

static int sum_partials(vio360_ba* ba, int nblocks, double* dev_out) {
    vio360_ctx* ctx = ba->ctx;
    { KernelTimer kt_(ctx, "k_sum_partials");
    k_sum_partials<<<1, 256, 0, ctx->stream>>>(ba->d_partials.as<double>(), nblocks, dev_out);
    }
    VIO_LAUNCH_CHECK(ctx);
    return VIO360_OK;
}

static int ba_evaluate_impl(vio360_ba* ba, int want_jacobians, int apply_loss, double* cost, bool f32);
extern "C" int vio360_ba_evaluate(vio360_ba* ba, int want_jacobians, int apply_loss, double* cost) {
    return ba_evaluate_impl(ba, want_jacobians, apply_loss, cost, false);
}
extern "C" int vio360_ba_evaluate_f32(vio360_ba* ba, int want_jacobians, int apply_loss, double* cost) {
    return ba_evaluate_impl(ba, want_jacobians, apply_loss, cost, true);
}
static int ba_evaluate_impl(vio360_ba* ba, int want_jacobians, int apply_loss, double* cost, bool f32) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_params) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_evaluate before ba_set_parameters");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const int O = ba->O;
    const size_t Os = (size_t)std::max(O, 1);
    VIO_CUDA(ctx, ba->d_r.reserve(Os * 2 * 8));
    if (want_jacobians) { VIO_CUDA(ctx, ba->d_Jp.reserve(Os * 12 * 8)); VIO_CUDA(ctx, ba->d_Jl.reserve(Os * 6 * 8)); }
    const int nb = std::max(1, (O + 127) / 128);
    VIO_CUDA(ctx, ba->d_partials.reserve((size_t)(nb + 1) * 8));
    BaDev d = make_dev(ba);
    if (f32) {
        KernelTimer kt_(ctx, "k_evaluate_f32");
        if (want_jacobians)
            k_evaluate_f32<true><<<nb, 128, 0, ctx->stream>>>(d, apply_loss != 0, ba->d_r.as<double>(), ba->d_Jp.as<double>(),
                                                             ba->d_Jl.as<double>(), ba->d_partials.as<double>());
        else
            k_evaluate_f32<false><<<nb, 128, 0, ctx->stream>>>(d, apply_loss != 0, ba->d_r.as<double>(), nullptr, nullptr,
                                                              ba->d_partials.as<double>());
    } else if (want_jacobians)
        { KernelTimer kt_(ctx, "k_evaluate");
        k_evaluate<true><<<nb, 128, 0, ctx->stream>>>(d, apply_loss != 0, ba->d_r.as<double>(), ba->d_Jp.as<double>(),
                                                     ba->d_Jl.as<double>(), ba->d_partials.as<double>());
        }
    else
        { KernelTimer kt_(ctx, "k_evaluate");
        k_evaluate<false><<<nb, 128, 0, ctx->stream>>>(d, apply_loss != 0, ba->d_r.as<double>(), nullptr, nullptr,
                                                      ba->d_partials.as<double>());
        }
    VIO_LAUNCH_CHECK(ctx);
    int rc = sum_partials(ba, nb, ba->d_partials.as<double>() + nb);
    if (rc) return rc;
    double h = 0.0;
    VIO_CUDA(ctx, cudaMemcpyAsync(&h, ba->d_partials.as<double>() + nb, 8, cudaMemcpyDeviceToHost, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (cost) *cost = h;
    ba->has_eval = true; ba->has_eval_jac = want_jacobians != 0;
    return VIO360_OK;
}

extern "C" int vio360_ba_get_residuals_jacobians(vio360_ba* ba, double* r, double* Jp, double* Jl) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_eval) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_get_residuals_jacobians before ba_evaluate");
    if ((Jp || Jl) && !ba->has_eval_jac) return vio360_fail(ctx, VIO360_ERR_STATE, "last ba_evaluate had want_jacobians=0");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t O = (size_t)ba->O;
    if (O == 0) return VIO360_OK;
    if (r) VIO_CUDA(ctx, cudaMemcpyAsync(r, ba->d_r.p, O * 16, cudaMemcpyDeviceToHost, ctx->stream));
    if (Jp) VIO_CUDA(ctx, cudaMemcpyAsync(Jp, ba->d_Jp.p, O * 96, cudaMemcpyDeviceToHost, ctx->stream));
    if (Jl) VIO_CUDA(ctx, cudaMemcpyAsync(Jl, ba->d_Jl.p, O * 48, cudaMemcpyDeviceToHost, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VIO360_OK;
}

// want_hpl = false: the LM loop only needs the diagonal blocks and the gradient; the O x 18 cross blocks (138 MB of
// stores for the 1M-observation window) are skipped
static int run_blocks(vio360_ba* ba, int apply_loss, bool want_hpl = true) {
    // layout of d_blocks: Hpp K*36 | bp K*6 | Hll M*9 | bl M*3 | Hpl O*18
    vio360_ctx* ctx = ba->ctx;
    const size_t K = ba->K, M = ba->M, O = ba->O;
    const size_t tot = K * 42 + M * 12 + (want_hpl ? O * 18 : 0);
    VIO_CUDA(ctx, ba->d_blocks.reserve(std::max<size_t>(tot, 1) * 8));
    double* Hpp = ba->d_blocks.as<double>(); double* bp = Hpp + K * 36; double* Hll = bp + K * 6;
    double* bl = Hll + M * 9; double* Hpl = want_hpl ? bl + M * 3 : nullptr;
    BaDev d = make_dev(ba);
    { KernelTimer kt_(ctx, "k_blocks_landmark");
    if (ba->eval_f32) k_blocks_landmark<float><<<(int)((M * 32 + 127) / 128), 128, 0, ctx->stream>>>(d, apply_loss != 0, Hll, bl, Hpl);
    else k_blocks_landmark<double><<<(int)((M * 32 + 127) / 128), 128, 0, ctx->stream>>>(d, apply_loss != 0, Hll, bl, Hpl);
    }
    VIO_LAUNCH_CHECK(ctx);
    { KernelTimer kt_(ctx, "k_blocks_pose");
    const int slices = (int)std::min<size_t>(16, std::max<size_t>(1, (2 * (size_t)ctx->num_sms + K - 1) / K));
    const size_t part_bytes = K * (size_t)slices * 27 * 8, need = part_bytes + K * 4;
    if (slices > 1 && need != ba->blk_part_bytes) {   // layout [partial sums | counters]; the counters start at zero and reset themselves
        VIO_CUDA(ctx, ba->d_blk_part.reserve(need));
        VIO_CUDA(ctx, cudaMemsetAsync(ba->d_blk_part.p, 0, need, ctx->stream));
        ba->blk_part_bytes = need;                    // (a window of another size moves the counters: cleared again)
    }
    double* part = slices > 1 ? ba->d_blk_part.as<double>() : nullptr;
    unsigned int* arrived = slices > 1 ? reinterpret_cast<unsigned int*>(ba->d_blk_part.as<unsigned char>() + part_bytes) : nullptr;
    if (ba->eval_f32) k_blocks_pose<float><<<dim3((unsigned)K, slices), 256, 0, ctx->stream>>>(d, apply_loss != 0, false, Hpp, bp, part, arrived);
    else k_blocks_pose<double><<<dim3((unsigned)K, slices), 256, 0, ctx->stream>>>(d, apply_loss != 0, false, Hpp, bp, part, arrived);
    }
    VIO_LAUNCH_CHECK(ctx);
    return VIO360_OK;
}

extern "C" int vio360_ba_normal_blocks(vio360_ba* ba, int apply_loss, double* Hpp, double* bp, double* Hll,
                                       double* bl, double* Hpl) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_params) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_normal_blocks before ba_set_parameters");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = run_blocks(ba, apply_loss);
    if (rc) return rc;
    const size_t K = ba->K, M = ba->M, O = ba->O;
    double* dHpp = ba->d_blocks.as<double>(); double* dbp = dHpp + K * 36; double* dHll = dbp + K * 6;
    double* dbl = dHll + M * 9; double* dHpl = dbl + M * 3;
    cudaStream_t st = ctx->stream;
    if (Hpp) VIO_CUDA(ctx, cudaMemcpyAsync(Hpp, dHpp, K * 36 * 8, cudaMemcpyDeviceToHost, st));
    if (bp) VIO_CUDA(ctx, cudaMemcpyAsync(bp, dbp, K * 6 * 8, cudaMemcpyDeviceToHost, st));
    if (Hll) VIO_CUDA(ctx, cudaMemcpyAsync(Hll, dHll, M * 9 * 8, cudaMemcpyDeviceToHost, st));
    if (bl) VIO_CUDA(ctx, cudaMemcpyAsync(bl, dbl, M * 3 * 8, cudaMemcpyDeviceToHost, st));
    if (Hpl && O) VIO_CUDA(ctx, cudaMemcpyAsync(Hpl, dHpl, O * 18 * 8, cudaMemcpyDeviceToHost, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    return VIO360_OK;
}

extern "C" int vio360_ba_column_sqnorm(vio360_ba* ba, int apply_loss, double* out) {
    if (!ba || !out) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_params) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_column_sqnorm before ba_set_parameters");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = run_blocks(ba, apply_loss);
    if (rc) return rc;
    const size_t K = ba->K, M = ba->M;
    double* dHpp = ba->d_blocks.as<double>(); double* dHll = dHpp + K * 42;
    std::vector<double> hpp(K * 36), hll(M * 9);
    VIO_CUDA(ctx, cudaMemcpyAsync(hpp.data(), dHpp, K * 36 * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VIO_CUDA(ctx, cudaMemcpyAsync(hll.data(), dHll, M * 9 * 8, cudaMemcpyDeviceToHost, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t k = 0; k < K; ++k) for (int a = 0; a < 6; ++a) out[6 * k + a] = hpp[36 * k + 7 * a];
    for (size_t l = 0; l < M; ++l) for (int a = 0; a < 3; ++a) out[6 * K + 3 * l + a] = hll[9 * l + 4 * a];
    return VIO360_OK;
}

extern "C" int vio360_ba_set_eval_precision(vio360_ba* ba, int fp32) {
    if (!ba) return VIO360_ERR_INVALID;
    ba->eval_f32 = fp32 != 0;
    ba->has_reduced = false;
    return VIO360_OK;
}

extern "C" int vio360_ba_set_lm_diagonal(vio360_ba* ba, const double* lm_diag) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_structure) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_set_lm_diagonal before ba_set_structure");
    ba->use_D = lm_diag != nullptr;
    if (!lm_diag) return VIO360_OK;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t nb = (6 * (size_t)ba->K + 3 * (size_t)ba->M) * 8;
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // pinned staging may still be in flight
    VIO_CUDA(ctx, ba->pin.reserve(nb));
    memcpy(ba->pin.p, lm_diag, nb);
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_D.p, ba->pin.p, nb, cudaMemcpyHostToDevice, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VIO360_OK;
}

extern "C" int vio360_ba_reduced_dim(const vio360_ba* ba) { return ba ? 6 * ba->n_free : VIO360_ERR_INVALID; }

extern "C" int vio360_ba_build_reduced_system(vio360_ba* ba, const double* lm_diag, int apply_loss, double* S,
                                              double* rhs, double* cost) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_params) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_build_reduced_system before ba_set_parameters");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const int n = 6 * ba->n_free;
    const size_t K = ba->K, M = ba->M;
    const size_t packed = (size_t)n * n + n + 1;
    {
        const void* before = ba->d_packed.p;
        VIO_CUDA(ctx, ba->d_packed.reserve(packed * 8));
        if (ba->d_packed.p != before) ba->band_zero_n = -1;
    }
    cudaStream_t st = ctx->stream;
    if (lm_diag) {
        int rcD = vio360_ba_set_lm_diagonal(ba, lm_diag);
        if (rcD) return rcD;
    }
    double* dS = ba->d_packed.as<double>(); double* drhs = dS + (size_t)n * n; double* dcost = drhs + n;
    BaDev d = make_dev(ba);
    if (ba->chunk_error) return vio360_fail(ctx, VIO360_ERR_STATE, "internal: inconsistent chunk table");
    static const bool force_atomic = getenv("VIO360_BA_ATOMIC") != nullptr;   // tuning / A-B switch, read once per process
    const bool use_tiles = !force_atomic;
    const int nch = use_tiles ? (int)ba->chunks.size() : 0;
    const int nirr = use_tiles ? (int)ba->irregular.size() : ba->lm_end - ba->lm_begin;
    const int nb_irr = (nirr * 32 + 127) / 128;
    VIO_CUDA(ctx, ba->d_partials.reserve((size_t)(nch + nb_irr + 2) * 8));
    if (ba->peer.attached && ba->peer.stale)
        return vio360_fail(ctx, VIO360_ERR_STATE, "ba_build_reduced_system: vio360_ba_set_structure / vio360_ba_set_shard ran after vio360_ba_peer_export; the peers hold the old row ranges: vio360_ba_peer_detach, then export and attach again on every rank");
    if (ba->peer.attached && (nirr > 0 || !use_tiles || ba->shard_world != ba->peer_world()))
        return vio360_fail(ctx, VIO360_ERR_STATE, "ba_build_reduced_system: a peer exchange is attached but this build cannot use the block band (landmarks outside the chunk windows, or the shard changed): vio360_ba_peer_detach first");
    if (use_tiles) {
        if (ba->chunks_dirty) {
            VIO_CUDA(ctx, cudaStreamSynchronize(st));
            const size_t nc = std::max<size_t>(1, ba->chunks.size());
            VIO_CUDA(ctx, ba->d_chunks.reserve(nc * sizeof(SchurChunk)));
            VIO_CUDA(ctx, ba->d_cov_ptr.reserve(ba->cov_ptr.size() * 4));
            VIO_CUDA(ctx, ba->d_cov_idx.reserve(ba->cov_idx.size() * 4));
            VIO_CUDA(ctx, ba->d_free_pose.reserve(std::max<size_t>(1, ba->free_pose.size()) * 4));
            VIO_CUDA(ctx, ba->d_irregular.reserve(std::max<size_t>(1, ba->irregular.size()) * 4));
            VIO_CUDA(ctx, ba->d_partS.reserve(nc * kST * kST * 8));
            VIO_CUDA(ctx, ba->d_partRhs.reserve(nc * kST * 8));
            VIO_CUDA(ctx, ba->d_reach.reserve(ba->reach.size() * 4));
            VIO_CUDA(ctx, ba->d_slots.reserve(ba->slot_table.size() * 4));
            VIO_CUDA(ctx, cudaMemcpy(ba->d_slots.p, ba->slot_table.data(), ba->slot_table.size() * 4, cudaMemcpyHostToDevice));
            VIO_CUDA(ctx, ba->d_batch_rng.reserve(ba->batch_rng.size() * 4));
            VIO_CUDA(ctx, cudaMemcpy(ba->d_batch_rng.p, ba->batch_rng.data(), ba->batch_rng.size() * 4, cudaMemcpyHostToDevice));
            VIO_CUDA(ctx, ba->d_batch_lm.reserve(ba->batch_lm.size() * 4));
            VIO_CUDA(ctx, cudaMemcpy(ba->d_batch_lm.p, ba->batch_lm.data(), ba->batch_lm.size() * 4, cudaMemcpyHostToDevice));
            VIO_CUDA(ctx, ba->d_wide.reserve(std::max<size_t>(1, ba->wide.size()) * 4));
            if (!ba->wide.empty()) {
                VIO_CUDA(ctx, cudaMemcpy(ba->d_wide.p, ba->wide.data(), ba->wide.size() * 4, cudaMemcpyHostToDevice));
                VIO_CUDA(ctx, ba->d_wide_inv.reserve((size_t)ba->M * 9 * 8));
            }
            VIO_CUDA(ctx, cudaMemcpy(ba->d_reach.p, ba->reach.data(), ba->reach.size() * 4, cudaMemcpyHostToDevice));
            if (!ba->chunks.empty()) VIO_CUDA(ctx, cudaMemcpy(ba->d_chunks.p, ba->chunks.data(), ba->chunks.size() * sizeof(SchurChunk), cudaMemcpyHostToDevice));
            VIO_CUDA(ctx, cudaMemcpy(ba->d_cov_ptr.p, ba->cov_ptr.data(), ba->cov_ptr.size() * 4, cudaMemcpyHostToDevice));
            VIO_CUDA(ctx, cudaMemcpy(ba->d_cov_idx.p, ba->cov_idx.data(), ba->cov_idx.size() * 4, cudaMemcpyHostToDevice));
            if (!ba->free_pose.empty()) VIO_CUDA(ctx, cudaMemcpy(ba->d_free_pose.p, ba->free_pose.data(), ba->free_pose.size() * 4, cudaMemcpyHostToDevice));
            if (!ba->irregular.empty()) VIO_CUDA(ctx, cudaMemcpy(ba->d_irregular.p, ba->irregular.data(), ba->irregular.size() * 4, cudaMemcpyHostToDevice));
            ba->chunks_dirty = false;
        }
        if (!ba->smem_attr_set) {
            VIO_CUDA(ctx, cudaFuncSetAttribute(k_schur_tile<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSchurSmemBytes));
            VIO_CUDA(ctx, cudaFuncSetAttribute(k_schur_tile<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSchurSmemBytes));
            ba->smem_attr_set = true;
        }
        if (nch > 0 && !ba->wide.empty()) {   // pass 1 of the wide landmarks: (E'E + D^2)^-1 over all their observations
            const int nw = (int)ba->wide.size();
            KernelTimer kt_(ctx, "k_wide_prepare");
            if (ba->eval_f32) k_wide_prepare<float><<<(nw * 32 + 127) / 128, 128, 0, st>>>(d, ba->d_wide.as<int>(), nw, apply_loss != 0, ba->d_wide_inv.as<double>());
            else k_wide_prepare<double><<<(nw * 32 + 127) / 128, 128, 0, st>>>(d, ba->d_wide.as<int>(), nw, apply_loss != 0, ba->d_wide_inv.as<double>());
        }
        if (nch > 0 && !ba->wide.empty()) VIO_LAUNCH_CHECK(ctx);
        if (nch > 0) {
            { KernelTimer kt_(ctx, ba->eval_f32 ? "k_schur_tile_f32" : "k_schur_tile");
            if (ba->eval_f32)
                k_schur_tile<float><<<nch, kSThreads, kSchurSmemBytes, st>>>(d, ba->d_chunks.as<SchurChunk>(), ba->d_slots.as<int>(), ba->d_batch_rng.as<int>(),
                                                                           ba->d_batch_lm.as<int>(), ba->d_wide_inv.as<double>(), apply_loss != 0,
                                                                           ba->d_partS.as<double>(), ba->d_partRhs.as<double>(), ba->d_partials.as<double>());
            else
                k_schur_tile<double><<<nch, kSThreads, kSchurSmemBytes, st>>>(d, ba->d_chunks.as<SchurChunk>(), ba->d_slots.as<int>(), ba->d_batch_rng.as<int>(),
                                                                            ba->d_batch_lm.as<int>(), ba->d_wide_inv.as<double>(), apply_loss != 0,
                                                                            ba->d_partS.as<double>(), ba->d_partRhs.as<double>(), ba->d_partials.as<double>());
            }
            VIO_LAUNCH_CHECK(ctx);
        }
        if (ctx->debug_sync) { fprintf(stderr, "schur: n=%d nch=%d (regular %d) wide=%zu nirr=%d K=%d cov=%zu\n", n, nch, ba->n_regular_chunks, ba->wide.size(), nirr, ba->K, ba->cov_idx.size() / 4); for (auto& c : ba->chunks) fprintf(stderr, "  chunk %d landmarks, poses %d+%d | %d+%d, flags %d, batch0 %d\n", c.lm_end, c.pbase, c.span, c.pbaseB, c.spanB, c.flags, c.batch0); }
        const bool fused = nirr == 0 && n > 0;   // one pass: sum slabs, mirror, LM diagonal, cost
        if (n > 0) {
            const double* Dp = ba->use_D ? ba->d_D.as<double>() : nullptr;
            if (fused) {
                // only the block rows' envelope is summed; the rest of S is zeroed once per structure
                // (nothing else writes there: the solver factors a copy, an allreduce adds zeros)
                // (a sharded window is allreduced in place by the caller, which fills the other ranks' envelopes)
                const bool sharded = ba->shard_world > 1;
                const bool peer = sharded && ba->peer.attached;
                const bool band = sharded || ba->red_band < ba->n_free;
                double* outS = dS; double* outRhs = drhs; double* outCost = dcost;
                int band_w = 0, row0 = 0, nrows = ba->n_free;
                if (peer) {
                    // the band this rank announces in the next vio360_ba_peer_exchange; only the block rows its
                    // landmarks touch are summed (the other ranks never read the rest), the LM diagonal is
                    // added by the exchange kernel
                    band_w = ba->peer.band_w;
                    if (band_w != std::min(ba->n_free, ba->band_blocks) || ba->peer.n_free != ba->n_free)
                        return vio360_fail(ctx, VIO360_ERR_STATE, "ba_build_reduced_system: the structure changed after vio360_ba_peer_export");
                    outS = peer_band(ba, ba->shard_rank, ba->peer.epoch + 1);
                    outRhs = outS + (size_t)ba->n_free * band_w * 36; outCost = outRhs + n;
                    row0 = ba->peer.row_lo[ba->shard_rank]; nrows = ba->peer.row_hi[ba->shard_rank] - row0;
                    ba->band_zero_n = -1;
                } else if (sharded) {
                    // compact exchange buffer: chunk windows are kSWin poses wide on every rank, so block
                    // row bi never reaches past bi + kSWin - 1 (every slot inside the grid is rewritten)
                    band_w = std::min(ba->n_free, ba->band_blocks);
                    const size_t nd = (size_t)ba->n_free * band_w * 36 + n + 1;
                    if (ba->d_band.bytes < nd * 8) {
                        VIO_CUDA(ctx, ba->d_band.reserve(nd * 8));
                        VIO_CUDA(ctx, cudaMemsetAsync(ba->d_band.p, 0, nd * 8, st));   // slots past the last block column
                    }
                    outS = ba->d_band.as<double>(); outRhs = outS + (size_t)ba->n_free * band_w * 36; outCost = outRhs + n;
                    ba->band_zero_n = -1;
                } else if (band && ba->band_zero_n != n) {
                    VIO_CUDA(ctx, cudaMemsetAsync(dS, 0, (size_t)n * n * 8, st));
                    ba->band_zero_n = n;
                }
                ba->band_w = band_w; ba->dense_stale = sharded;
                double* dS = outS; double* drhs = outRhs; double* dcost = outCost;   // shadow: the launches below write here
                const dim3 nblk(sharded ? band_w : (band ? ba->red_band : ba->n_free), nrows);
                const bool add_diag = ba->shard_rank == 0 && !peer;
                if (peer) ba->peer.band_epoch = ba->peer.epoch + 1;
                PeerPublish pub;
                memset(&pub, 0, sizeof(pub));
                if (peer) {   // the reduce releases finished block rows to every rank (ba_peer.cuh)
                    pub.world = ba->peer.world; pub.rank = ba->shard_rank; pub.n_free = ba->n_free; pub.band_w = band_w;
                    pub.epoch = ba->peer.epoch + 1; pub.builds = nrows > 0 ? ++ba->peer.builds : ba->peer.builds;
                    for (int r = 0; r < ba->peer.world; ++r) pub.row_flags[r] = peer_row_flags(ba, r);
                    pub.row_count = peer_row_flags(ba, ba->shard_rank) + (size_t)kMaxPeers * ba->n_free;
                    pub.mode = peer_publish_mode(); pub.row_lo = row0; pub.row_hi = row0 + nrows;
                    if (pub.mode == 3) pub.epoch = 0;   // default: the reduce does not publish, the pull kernel's first CTA announces
                }
                if (nrows > 0) {
                KernelTimer kt_(ctx, "k_schur_reduce");
                // walkers per block pair: enough for one round of four chunks each over the longest per-pose chunk list
                // (few block pairs, i.e. a short window with many chunks per pose: more walkers per pair)
                int max_cov = 0;
                for (int p = 0; p < ba->K; ++p) max_cov = std::max(max_cov, ba->cov_ptr[p + 1] - ba->cov_ptr[p]);
                static const int force_lanes = getenv("VIO360_BA_RED_LANES") ? atoi(getenv("VIO360_BA_RED_LANES")) : 0;
                static const bool no_pdl = getenv("VIO360_BA_NO_PDL") != nullptr;   // A/B switch
                const bool use_pdl = !no_pdl && !ctx->timing && nch > 0;           // event pairs between the kernels would serialise them anyway
                int lanes = (size_t)nblk.x * nblk.y >= 2000 ? 7 : 14;
                if ((size_t)nblk.x * nblk.y >= 2000 && max_cov <= 20) lanes = 5;
                if (force_lanes) lanes = force_lanes;
#define VIO_REDUCE_LAUNCH(L)                                                                                                   \
                VIO_CUDA(ctx, vio360_launch_pdl(use_pdl, k_schur_reduce<true, L>, nblk, dim3(36 * L), 0, st, dS, n, drhs, dcost, (const double*)ba->d_partS.as<double>(), \
                                                (const double*)ba->d_partRhs.as<double>(), (const double*)ba->d_partials.as<double>(), nch,              \
                                                (const SchurChunk*)ba->d_chunks.as<SchurChunk>(), (const int*)ba->d_cov_ptr.as<int>(),                   \
                                                (const int4*)ba->d_cov_idx.as<int4>(), (const int*)ba->d_reach.as<int>(), (const int*)ba->d_free_pose.as<int>(), \
                                                Dp, add_diag, band, band_w, row0, ba->n_free, pub))
                if (lanes <= 3) VIO_REDUCE_LAUNCH(3);
                else if (lanes <= 5) VIO_REDUCE_LAUNCH(5);
                else if (lanes <= 7) VIO_REDUCE_LAUNCH(7);
                else VIO_REDUCE_LAUNCH(14);
#undef VIO_REDUCE_LAUNCH
                }
            } else {
                ba->band_zero_n = -1; ba->band_w = 0; ba->dense_stale = false;
                const dim3 nblk(ba->n_free, ba->n_free);
                KernelTimer kt_(ctx, "k_schur_reduce");
                k_schur_reduce<false, 7><<<nblk, 36 * 7, 0, st>>>(dS, n, drhs, dcost, ba->d_partS.as<double>(), ba->d_partRhs.as<double>(),
                                                            ba->d_partials.as<double>(), nch, ba->d_chunks.as<SchurChunk>(),
                                                            ba->d_cov_ptr.as<int>(), ba->d_cov_idx.as<int4>(), ba->d_reach.as<int>(),
                                                            ba->d_free_pose.as<int>(), Dp, false, false, 0, 0, ba->n_free, PeerPublish{});
            }
            VIO_LAUNCH_CHECK(ctx);
        }
        if (fused) {
            ba->has_reduced = true; ba->reduced_apply_loss = apply_loss != 0;
            if (S || rhs || cost) return vio360_ba_fetch_reduced_system(ba, S, rhs, cost);
            return VIO360_OK;   // asynchronous on the context's stream
        }
        if (nirr > 0) {
            { KernelTimer kt_(ctx, "k_schur_warp");
            k_schur_warp<<<nb_irr, 128, 0, st>>>(d, 0, nirr, ba->d_irregular.as<int>(), apply_loss != 0, dS, n, drhs,
                                                 ba->d_partials.as<double>() + nch);
            }
            VIO_LAUNCH_CHECK(ctx);
        }
    } else {
        VIO_CUDA(ctx, cudaMemsetAsync(ba->d_packed.p, 0, packed * 8, st));
        ba->band_zero_n = -1; ba->band_w = 0; ba->dense_stale = false;
        if (nirr > 0) {
            { KernelTimer kt_(ctx, "k_schur_warp");
            k_schur_warp<<<nb_irr, 128, 0, st>>>(d, ba->lm_begin, ba->lm_end, nullptr, apply_loss != 0, dS, n, drhs, ba->d_partials.as<double>());
            }
            VIO_LAUNCH_CHECK(ctx);
        }
    }
    int rc = sum_partials(ba, nch + nb_irr, dcost);
    if (rc) return rc;
    if (n > 0) {
        { KernelTimer kt_(ctx, "k_finish_S");
        k_finish_S<<<(n * n + 255) / 256, 256, 0, st>>>(dS, n, ba->use_D ? ba->d_D.as<double>() : nullptr,
                                                       ba->d_free_map.as<int>(), ba->K, ba->shard_rank == 0);
        }
        VIO_LAUNCH_CHECK(ctx);
    }
    ba->has_reduced = true; ba->reduced_apply_loss = apply_loss != 0;
    if (S || rhs || cost) return vio360_ba_fetch_reduced_system(ba, S, rhs, cost);
    return VIO360_OK;   // asynchronous on the context's stream
}

// Sharded builds leave their result in the compact band buffer; the dense [S | rhs | cost] view is
// produced on first use (and again by vio360_ba_expand_band after the caller's allreduce).
static int ensure_dense(vio360_ba* ba) {
    if (!ba->dense_stale) return VIO360_OK;
    vio360_ctx* ctx = ba->ctx;
    if (ba->peer.attached)
        return vio360_fail(ctx, VIO360_ERR_STATE, "the sharded build has not been exchanged yet: call vio360_ba_peer_exchange on every rank first");
    const size_t n = 6 * (size_t)ba->n_free, total = n * n + n + 1;
    { KernelTimer kt_(ctx, "k_expand_band");
    k_expand_band<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(ba->d_band.as<double>(), ba->n_free, ba->band_w, ba->d_packed.as<double>());
    }
    VIO_LAUNCH_CHECK(ctx);
    ba->dense_stale = false;
    return VIO360_OK;
}

extern "C" int vio360_ba_reduced_band_dev(vio360_ba* ba, double** band_dev, size_t* n_doubles) {
    if (!ba || !band_dev || !n_doubles) return VIO360_ERR_INVALID;
    if (!ba->has_reduced) return vio360_fail(ba->ctx, VIO360_ERR_STATE, "no reduced system built yet");
    if (ba->peer.attached)
        return vio360_fail(ba->ctx, VIO360_ERR_STATE, "a peer exchange is attached: call vio360_ba_peer_exchange instead of all-reducing the band");
    if (ba->band_w <= 0)
        return vio360_fail(ba->ctx, VIO360_ERR_STATE, "the last build has no compact exchange buffer (no shard set, or landmarks outside the chunk windows): use vio360_ba_reduced_system_dev");
    *band_dev = ba->d_band.as<double>();
    *n_doubles = (size_t)ba->n_free * ba->band_w * 36 + 6 * (size_t)ba->n_free + 1;
    return VIO360_OK;
}

extern "C" int vio360_ba_expand_band(vio360_ba* ba) {
    if (!ba) return VIO360_ERR_INVALID;
    if (!ba->has_reduced || ba->band_w <= 0) return vio360_fail(ba->ctx, VIO360_ERR_STATE, "ba_expand_band: no compact exchange buffer");
    ba->dense_stale = true;
    return ensure_dense(ba);
}


// ---- peer-memory exchange (ba_peer.cuh) --------------------------------------------------------
struct PeerBlob {                 // what vio360_ba_peer_export hands to the other ranks (VIO360_PEER_BLOB_BYTES)
    cudaIpcMemHandle_t ipc;       // 64 bytes
    uint64_t base, nd;
    int32_t pid, device, rank, world, n_free, band_w, row_lo, row_hi;
    uint8_t pad[VIO360_PEER_BLOB_BYTES - 64 - 16 - 32];
};
static_assert(sizeof(PeerBlob) == VIO360_PEER_BLOB_BYTES, "peer blob layout");

extern "C" int vio360_ba_peer_export(vio360_ba* ba, void* blob_out) {
    if (!ba || !blob_out) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_structure) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_export before ba_set_structure");
    if (ba->shard_world < 2 || ba->shard_world > kMaxPeers)
        return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_export: needs a shard with 2..%d ranks (vio360_ba_set_shard)", kMaxPeers);
    if (!ba->irregular.empty() || ba->chunk_error || ba->n_free <= 0)
        return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_export: %zu landmarks of this shard are observed twice from one pose (the reference never builds such a window), so the build has no block band; allreduce the dense buffer instead", ba->irregular.size());
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    peer_release(ba);
    auto& P = ba->peer;
    P.world = ba->shard_world;
    P.n_free = ba->n_free; P.band_w = std::min(ba->n_free, ba->band_blocks);
    P.nd = (size_t)P.n_free * P.band_w * 36 + 6 * (size_t)P.n_free + 1;
    P.nd_pad = (P.nd + 15) & ~(size_t)15;
    P.header_bytes = (kPeerHeaderBytes + ((size_t)(kMaxPeers + 1) * P.n_free + 4) * 4 + 127) & ~(size_t)127;   // row flags, row counters, rows-done counter
    const size_t bytes = P.header_bytes + 2 * P.nd_pad * 8;
    VIO_CUDA(ctx, P.local.reserve(bytes));
    VIO_CUDA(ctx, cudaMemset(P.local.p, 0, bytes));     // flags = epoch 0, bands zero (slots nobody writes stay zero)
    VIO_CUDA(ctx, cudaHostAlloc((void**)&P.h_err, sizeof(int), cudaHostAllocMapped));
    *P.h_err = 0;
    // block rows (free-pose index) covered by this rank's chunks
    int lo = P.n_free, hi = 0;
    for (const SchurChunk& c : ba->chunks)
        for (int p = c.pbase; p < c.pbase + c.span; ++p)
            if (ba->free_map[p] >= 0) { lo = std::min(lo, ba->free_map[p]); hi = std::max(hi, ba->free_map[p] + 1); }
    if (hi <= lo) lo = hi = 0;
    PeerBlob b;
    memset(&b, 0, sizeof(b));
    VIO_CUDA(ctx, cudaIpcGetMemHandle(&b.ipc, P.local.p));
    b.base = (uint64_t)reinterpret_cast<uintptr_t>(P.local.p); b.nd = P.nd;
    b.pid = (int32_t)getpid(); b.device = ctx->device; b.rank = ba->shard_rank; b.world = ba->shard_world;
    b.n_free = P.n_free; b.band_w = P.band_w; b.row_lo = lo; b.row_hi = hi;
    memcpy(blob_out, &b, sizeof(b));
    P.exported = true;
    return VIO360_OK;
}

extern "C" int vio360_ba_peer_attach(vio360_ba* ba, const void* blobs) {
    if (!ba || !blobs) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    auto& P = ba->peer;
    if (!P.exported) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_attach before ba_peer_export");
    if (P.stale) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_attach: the structure or shard changed after vio360_ba_peer_export: export again");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    P.same_device = 0;
    const PeerBlob* B = static_cast<const PeerBlob*>(blobs);
    for (int r = 0; r < P.world; ++r) {
        PeerBlob b;
        memcpy(&b, B + r, sizeof(b));
        if (b.rank != r || b.world != P.world || b.n_free != P.n_free || b.band_w != P.band_w || b.nd != P.nd)
            return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_peer_attach: blob %d does not describe rank %d of this window (rank %d world %d n_free %d band %d)",
                               r, r, b.rank, b.world, b.n_free, b.band_w);
        P.row_lo[r] = b.row_lo; P.row_hi[r] = b.row_hi;
        if (b.device == ctx->device) P.same_device++;   // (device ordinals; CUDA_VISIBLE_DEVICES-remapped processes count as distinct)
        if (r == ba->shard_rank) { P.base[r] = P.local.p; continue; }
        if (b.pid == (int32_t)getpid()) {
            // same process (one host thread per GPU): plain peer access
            if (b.device != ctx->device) {
                cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) return vio360_fail(ctx, VIO360_ERR_CUDA, "ba_peer_attach: no peer access from device %d to %d: %s", ctx->device, b.device, cudaGetErrorString(e));
            }
            P.base[r] = reinterpret_cast<void*>(static_cast<uintptr_t>(b.base));
        } else {
            cudaError_t e = cudaIpcOpenMemHandle(&P.base[r], b.ipc, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) return vio360_fail(ctx, VIO360_ERR_CUDA, "ba_peer_attach: cudaIpcOpenMemHandle of rank %d: %s", r, cudaGetErrorString(e));
            P.ipc_open[r] = true;
        }
    }
    P.attached = true; P.epoch = 0; P.zero_n = -1;
    P.trace = getenv("VIO360_PEER_TRACE") != nullptr;
    return VIO360_OK;
}

extern "C" int vio360_ba_peer_exchange(vio360_ba* ba) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    auto& P = ba->peer;
    if (!P.attached) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_exchange before ba_peer_attach");
    if (P.stale) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_exchange: the structure or shard changed after vio360_ba_peer_export (detach, export and attach again)");
    if (!ba->has_reduced || ba->band_w <= 0) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_exchange: no sharded build to exchange");
    if (*P.h_err) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_exchange: an earlier exchange timed out waiting for a peer rank");
    if (P.band_epoch != P.epoch + 1) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_peer_exchange: no vio360_ba_build_reduced_system since the last exchange");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int n = 6 * ba->n_free;
    double* dS = ba->d_packed.as<double>();
    if (P.zero_ptr != ba->d_packed.p || P.zero_n != n) {    // outside the band S is zero and stays zero
        VIO_CUDA(ctx, cudaMemsetAsync(dS, 0, (size_t)n * n * 8, st));
        P.zero_ptr = ba->d_packed.p; P.zero_n = n;
    }
    const unsigned epoch = ++P.epoch;
    PeerView pv;
    memset(&pv, 0, sizeof(pv));
    pv.world = P.world; pv.rank = ba->shard_rank;
    for (int r = 0; r < P.world; ++r) {
        pv.band[r] = peer_band(ba, r, epoch);
        pv.row_lo[r] = P.row_lo[r]; pv.row_hi[r] = P.row_hi[r];
    }
    pv.row_flags = peer_row_flags(ba, ba->shard_rank);
    if (peer_publish_mode() == 3)
        for (int r = 0; r < P.world; ++r) pv.announce[r] = peer_row_flags(ba, r) + (size_t)ba->shard_rank * ba->n_free + P.row_lo[ba->shard_rank];
    int* d_err = nullptr;
    VIO_CUDA(ctx, cudaHostGetDevicePointer((void**)&d_err, P.h_err, 0));
    // trace words live behind the flags in the exchange header: [CTA 0 wait ns, CTA 0 total ns]
    unsigned long long* d_trace = P.trace ? reinterpret_cast<unsigned long long*>(static_cast<char*>(P.local.p) + 512) : nullptr;
    const size_t total = P.nd;
    // one element per thread where possible: a thread's peer loads are one NVLink round trip, so the kernel
    // is a single latency deep when everything is in flight at once
    // Every CTA polls the arrival flags, so all CTAs of all ranks that share a device must be resident together
    // (a rank whose CTA 0 cannot get an SM slot never announces): cap the grid at this rank's share of the
    // 8 x 256-thread CTAs per SM.
    const unsigned grid = (unsigned)std::min<size_t>((total + 255) / 256, std::max<size_t>(1, (size_t)8 * ctx->num_sms / std::max(1, P.same_device)));
    { KernelTimer kt_(ctx, "k_band_pull");
    static const bool no_pdl = getenv("VIO360_BA_NO_PDL") != nullptr;
    VIO_CUDA(ctx, vio360_launch_pdl(!no_pdl && !ctx->timing, k_band_pull, dim3(grid), dim3(256), 0, st, pv, epoch, ba->n_free, P.band_w,
                                    (const int*)ba->d_free_pose.as<int>(), (const double*)(ba->use_D ? ba->d_D.as<double>() : nullptr), dS, d_err, d_trace));
    }
    VIO_LAUNCH_CHECK(ctx);
    ba->dense_stale = false; ba->band_zero_n = -1;
    return VIO360_OK;
}

extern "C" int vio360_ba_peer_detach(vio360_ba* ba) {
    if (!ba) return VIO360_ERR_INVALID;
    cudaSetDevice(ba->ctx->device);
    cudaStreamSynchronize(ba->ctx->stream);
    if (ba->peer.trace && ba->peer.attached && ba->peer.epoch > 1) {
        unsigned long long h[4] = {0, 0, 0, 0};
        cudaMemcpy(h, static_cast<char*>(ba->peer.local.p) + 512, sizeof(h), cudaMemcpyDeviceToHost);
        const double ne = (double)ba->peer.epoch;
        fprintf(stderr, "[vio360 peer trace] rank %d: %u exchanges, CTA 0 waited %.2f us and ran %.2f us per exchange\n",
                ba->shard_rank, ba->peer.epoch, h[0] / ne * 1e-3, h[1] / ne * 1e-3);
    }
    peer_release(ba);
    return VIO360_OK;
}

extern "C" int vio360_ba_reduced_system_dev(vio360_ba* ba, double** packed_dev, size_t* n_doubles) {
    if (!ba || !packed_dev || !n_doubles) return VIO360_ERR_INVALID;
    if (!ba->has_reduced) return vio360_fail(ba->ctx, VIO360_ERR_STATE, "no reduced system built yet");
    { int rcE = ensure_dense(ba); if (rcE) return rcE; }
    const size_t n = 6 * (size_t)ba->n_free;
    *packed_dev = ba->d_packed.as<double>();
    *n_doubles = n * n + n + 1;
    return VIO360_OK;
}

extern "C" int vio360_ba_fetch_reduced_system(vio360_ba* ba, double* S, double* rhs, double* cost) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_reduced) return vio360_fail(ctx, VIO360_ERR_STATE, "no reduced system built yet");
    { int rcE = ensure_dense(ba); if (rcE) return rcE; }
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n = 6 * (size_t)ba->n_free;
    double* dS = ba->d_packed.as<double>();
    if (S && n) VIO_CUDA(ctx, cudaMemcpyAsync(S, dS, n * n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (rhs && n) VIO_CUDA(ctx, cudaMemcpyAsync(rhs, dS + n * n, n * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (cost) VIO_CUDA(ctx, cudaMemcpyAsync(cost, dS + n * n + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ba->peer.h_err && *ba->peer.h_err)
        return vio360_fail(ctx, VIO360_ERR_STATE, "the peer exchange timed out waiting for another rank: the reduced system is incomplete");
    return VIO360_OK;
}

extern "C" int vio360_ba_chi2(vio360_ba* ba, double thr, double* chi2, uint8_t* flag) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_params) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_chi2 before ba_set_parameters");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t O = (size_t)ba->O;
    if (O == 0) return VIO360_OK;
    VIO_CUDA(ctx, ba->d_chi2.reserve(O * 8)); VIO_CUDA(ctx, ba->d_flag.reserve(O));
    BaDev d = make_dev(ba);
    k_chi2<<<(int)((O + 127) / 128), 128, 0, ctx->stream>>>(d, thr, ba->d_chi2.as<double>(), ba->d_flag.as<unsigned char>());
    VIO_LAUNCH_CHECK(ctx);
    if (chi2) VIO_CUDA(ctx, cudaMemcpyAsync(chi2, ba->d_chi2.p, O * 8, cudaMemcpyDeviceToHost, ctx->stream));
    if (flag) VIO_CUDA(ctx, cudaMemcpyAsync(flag, ba->d_flag.p, O, cudaMemcpyDeviceToHost, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VIO360_OK;
}

// ------------------------------------------------------------------ PnP (pose only)
// PnPFactor::Evaluate / compute_chi_square (Factors.cpp:33-210, 212-265): same projection
// and Jacobian as the BA factor with the point held constant and one shared pose.
__global__ void __launch_bounds__(128) k_pnp(int N, const double* __restrict__ obs_uv, const double* __restrict__ pts,
                                             const unsigned char* __restrict__ outlier, const double* __restrict__ posebuf,
                                             CamConst cam, double* __restrict__ r_out, double* __restrict__ Jp_out,
                                             unsigned char* __restrict__ valid, double* __restrict__ chi2,
                                             double* __restrict__ partials /* [gridDim.x][28] */) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double acc[28];
#pragma unroll
    for (int j = 0; j < 28; ++j) acc[j] = 0.0;
    if (i < N) {
        double pose[kPoseStride];
#pragma unroll
        for (int j = 0; j < kPoseStride; ++j) pose[j] = __ldg(posebuf + j);
        const double Pw[3] = {pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2]};
        const double u = obs_uv[2 * (size_t)i], v = obs_uv[2 * (size_t)i + 1];
        const bool outl = outlier && outlier[i];
        ObsProj p;
        obs_project(pose, Pw, u, v, cam, p);
        const bool bad = !outl && p.branch == 1;        // Evaluate returns false, residual untouched (:74-80)
        double r[2] = {0.0, 0.0}, F[12], E[6];
        double cost = 0.0;
        if (!bad) cost = obs_evaluate<true>(pose, Pw, u, v, outl, cam, true, r, F, E);
        else { for (int j = 0; j < 12; ++j) F[j] = 0.0; }
        r_out[2 * (size_t)i] = r[0]; r_out[2 * (size_t)i + 1] = r[1];
        if (Jp_out) for (int j = 0; j < 12; ++j) Jp_out[12 * (size_t)i + j] = F[j];
        valid[i] = bad ? 0 : 1;
        if (chi2) {   // ignores the outlier flag (:212-265); DBL_MAX for a point at the centre
            double c;
            if (p.branch == 1) c = 1.7976931348623157e308;
            else {
                const double a0 = p.du * cam.info[0] + p.dv * cam.info[2], a1 = p.du * cam.info[1] + p.dv * cam.info[3];
                c = a0 * p.du + a1 * p.dv;
            }
            chi2[i] = c;
        }
        if (!bad) {
            int t = 0;
#pragma unroll
            for (int a = 0; a < 6; ++a)
#pragma unroll
                for (int b = a; b < 6; ++b) acc[t++] = F[a] * F[b] + F[6 + a] * F[6 + b];
#pragma unroll
            for (int a = 0; a < 6; ++a) acc[21 + a] = F[a] * r[0] + F[6 + a] * r[1];
            acc[27] = cost;
        }
    }
    __shared__ double sm[4][28];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int j = 0; j < 28; ++j) {
        const double s = warp_sum(acc[j]);
        if (lane == 0) sm[w][j] = s;
    }
    __syncthreads();
    if (threadIdx.x < 28)
        partials[28 * (size_t)blockIdx.x + threadIdx.x] = sm[0][threadIdx.x] + sm[1][threadIdx.x] + sm[2][threadIdx.x] + sm[3][threadIdx.x];
}

__global__ void k_pnp_reduce(const double* __restrict__ partials, int nb, double* __restrict__ out28) {
    const int j = threadIdx.x;
    if (j >= 28) return;
    double s = 0.0;
    for (int b = 0; b < nb; ++b) s += partials[28 * (size_t)b + j];
    out28[j] = s;
}

extern "C" int vio360_pnp_evaluate(vio360_ctx* ctx, int N, const double* obs_uv, const double* world_pts,
                                   const uint8_t* outlier, const double* T_wb_init, const double* T_cb, double cols,
                                   double rows, const double* info, double huber_delta, const double* pose_delta,
                                   double* r, double* Jp, uint8_t* valid, double* H, double* b, double* cost,
                                   double* chi2) {
    if (!ctx) return VIO360_ERR_INVALID;
    if (N <= 0 || !obs_uv || !world_pts || !T_wb_init || !T_cb || !info || !pose_delta || !r || !valid)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "pnp_evaluate: bad arguments");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t n = (size_t)N;
    const int nb = (N + 127) / 128;
    // one scratch allocation: inputs | outputs
    const size_t in_d = n * 2 + n * 3 + 16 + 16 + 6;          // doubles
    const size_t out_d = n * 2 + n * 12 + n + (size_t)nb * 28 + 28 + kPoseStride;
    void* scratch_p = nullptr;
    VIO_CUDA(ctx, vio360_scratch(ctx, (in_d + out_d) * 8 + n * 2, &scratch_p));   // the context's scratch: no allocation per call
    double* d_obs = static_cast<double*>(scratch_p); double* d_pts = d_obs + n * 2; double* d_Twb = d_pts + n * 3;
    double* d_Tcb = d_Twb + 16; double* d_xi = d_Tcb + 16;
    double* d_r = d_xi + 6; double* d_J = d_r + n * 2; double* d_chi = d_J + n * 12; double* d_part = d_chi + n;
    double* d_out28 = d_part + (size_t)nb * 28; double* d_pose = d_out28 + 28;
    unsigned char* d_outl = reinterpret_cast<unsigned char*>(d_obs + in_d + out_d); unsigned char* d_valid = d_outl + n;
    auto cleanup = [&]() {};
#define PNP_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return vio360_fail(ctx, VIO360_ERR_CUDA, "%s -> %s", #call, cudaGetErrorString(e_)); } } while (0)
    // inputs and outputs travel through the context's pinned staging buffer: one copy in (plus the outlier flags), one
    // out (plus the valid flags) instead of eleven pageable copies of a few hundred bytes each
    const size_t out_copy_d = n * 15 + (size_t)nb * 28 + 28;          // r | J | chi2 | partials | H, b, cost
    void* pin_p = nullptr;
    PNP_CUDA(vio360_pinned(ctx, (in_d + out_copy_d) * 8 + 2 * n, &pin_p));
    double* h_in = static_cast<double*>(pin_p); double* h_out = h_in + in_d;
    unsigned char* h_flags = reinterpret_cast<unsigned char*>(h_out + out_copy_d);
    memcpy(h_in, obs_uv, n * 16); memcpy(h_in + n * 2, world_pts, n * 24);
    memcpy(h_in + n * 5, T_wb_init, 128); memcpy(h_in + n * 5 + 16, T_cb, 128); memcpy(h_in + n * 5 + 32, pose_delta, 48);
    PNP_CUDA(cudaMemcpyAsync(d_obs, h_in, in_d * 8, cudaMemcpyHostToDevice, st));
    if (outlier) { memcpy(h_flags, outlier, n); PNP_CUDA(cudaMemcpyAsync(d_outl, h_flags, n, cudaMemcpyHostToDevice, st)); }
    CamConst cam;
    cam_const_init(cam, cols, rows, info, T_cb, huber_delta);
    k_pose_prepare<<<1, 32, 0, st>>>(1, d_Twb, d_Tcb, d_xi, d_pose);
    ctx->launches++;
    k_pnp<<<nb, 128, 0, st>>>(N, d_obs, d_pts, outlier ? d_outl : nullptr, d_pose, cam, d_r, d_J, d_valid, d_chi, d_part);
    ctx->launches++;
    k_pnp_reduce<<<1, 32, 0, st>>>(d_part, nb, d_out28);
    ctx->launches++;
    PNP_CUDA(cudaGetLastError());
    PNP_CUDA(cudaMemcpyAsync(h_out, d_r, out_copy_d * 8, cudaMemcpyDeviceToHost, st));
    PNP_CUDA(cudaMemcpyAsync(h_flags + n, d_valid, n, cudaMemcpyDeviceToHost, st));
    PNP_CUDA(cudaStreamSynchronize(st));
    memcpy(r, h_out, n * 16);
    if (Jp) memcpy(Jp, h_out + n * 2, n * 96);
    if (chi2) memcpy(chi2, h_out + n * 14, n * 8);
    memcpy(valid, h_flags + n, n);
    const double* h28 = h_out + n * 15 + (size_t)nb * 28;
#undef PNP_CUDA
    cleanup();
    if (H) {
        int t = 0;
        for (int a = 0; a < 6; ++a) for (int c = a; c < 6; ++c) { H[a * 6 + c] = h28[t]; H[c * 6 + a] = h28[t]; ++t; }
    }
    if (b) for (int a = 0; a < 6; ++a) b[a] = h28[21 + a];
    if (cost) *cost = h28[27];
    return VIO360_OK;
}

// ---- solve the reduced system and back-substitute on the device ----
// Device part of the solve: x (6 n_free, free-pose order) in d_x, the factorisation status behind it, and, when
// want_points, the back-substituted landmark step in d_dl.  Nothing is copied to the host.
static int solve_reduced_dev(vio360_ba* ba, bool want_points) {
    vio360_ctx* ctx = ba->ctx;
    { int rcE = ensure_dense(ba); if (rcE) return rcE; }
    cudaStream_t st = ctx->stream;
    const int nb = ba->n_free, K = ba->K, M = ba->M;
    const size_t n = 6 * (size_t)nb;
    const size_t panel_bytes = n * 6 * sizeof(double);   // rows below the diagonal block + the rhs row
    if (panel_bytes > (size_t)200 * 1024) return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_solve_reduced: %d free poses exceed the single-CTA solver (max 710)", nb);
    VIO_CUDA(ctx, ba->d_x.reserve(n * 8 + 16));
    if (nb == 0) VIO_CUDA(ctx, cudaMemsetAsync(ba->d_x.p, 0, 16, st));
    if (nb > 0) {
        if (ba->solve_dirty) {
            // envelope of S in free-pose index space: pose p shares a landmark with poses up to reach[p]
            // (union over ALL landmarks, so it also covers the all-reduced system of a sharded window)
            std::vector<int> rp(K);
            for (int p = 0; p < K; ++p) rp[p] = p;
            for (int l = 0; l < M; ++l) {
                int pmax = -1;
                for (int q = ba->lm_ptr[l]; q < ba->lm_ptr[l + 1]; ++q) pmax = std::max(pmax, ba->h_pose[q]);
                for (int q = ba->lm_ptr[l]; q < ba->lm_ptr[l + 1]; ++q) rp[ba->h_pose[q]] = std::max(rp[ba->h_pose[q]], pmax);
            }
            std::vector<int> upto(K);   // number of free poses with index <= p
            int cnt = 0;
            for (int p = 0; p < K; ++p) { if (ba->free_map[p] >= 0) ++cnt; upto[p] = cnt; }
            std::vector<int> reachF(nb);
            int run = 0;
            for (int p = 0; p < K; ++p) {
                const int a = ba->free_map[p];
                if (a < 0) continue;
                run = std::max(run, std::max(a, upto[rp[p]] - 1));   // monotone: the factor's fill stays inside
                reachF[a] = run;
            }
            VIO_CUDA(ctx, ba->d_reachF.reserve((size_t)nb * 4));
            VIO_CUDA(ctx, cudaMemcpy(ba->d_reachF.p, reachF.data(), (size_t)nb * 4, cudaMemcpyHostToDevice));
            ba->max_reach = 0;
            for (int a = 0; a < nb; ++a) ba->max_reach = std::max(ba->max_reach, reachF[a] - a);
            ba->solve_dirty = false;
        }
        VIO_CUDA(ctx, ba->d_linv.reserve((size_t)nb * 36 * 8));
        VIO_CUDA(ctx, ba->d_x.reserve(n * 8 + 16));
        const double* dS = ba->d_packed.as<double>();
        VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_x.p, dS + n * n, n * 8, cudaMemcpyDeviceToDevice, st));
        if (!ba->chol_attr_set) {
            VIO_CUDA(ctx, cudaFuncSetAttribute(k_chol_skyline, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
            VIO_CUDA(ctx, cudaFuncSetAttribute(k_chol_band, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
            VIO_CUDA(ctx, cudaFuncSetAttribute(k_chol_band_reg, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
            ba->chol_attr_set = true;
        }
        int* d_status = reinterpret_cast<int*>(ba->d_x.as<double>() + n);
        const size_t band_smem = ((size_t)kCholWb * kCholWb * 36 + ((kCholWb - 1) * 6 + 1) * 6 + n) * sizeof(double);
        if (ba->max_reach < kCholWb && band_smem <= (size_t)220 * 1024) {
            // banded window: trailing matrix in shared memory, S itself is only read
            VIO_CUDA(ctx, ba->d_chol.reserve(((size_t)nb * (kCholWb - 1) * 36 + (size_t)nb * 42 + (size_t)nb * 36) * 8));
            double* Lcol = ba->d_chol.as<double>();
            double* Ldiag = Lcol + (size_t)nb * (kCholWb - 1) * 36;
            double* Linv = Ldiag + (size_t)nb * 42;
            // Windows of up to 24 free poses (the reference's local BA: 10 keyframes) fit the shared-memory window whole
            // and finish in a few columns: 25 us against 34 us for the register-window kernel, whose set-up they do not
            // amortise.  Larger systems: registers (0.45 ms against 0.81 ms at 199 free poses).  VIO360_CHOL_SMEM=1: A/B switch.
            static const bool force_smem_window = getenv("VIO360_CHOL_SMEM") != nullptr;
            const bool smem_window = force_smem_window || nb <= kCholWb;
            { KernelTimer kt_(ctx, "k_chol_band");
            if (smem_window)
                k_chol_band<<<1, kCholThreads, band_smem, st>>>(dS, nb, ba->d_reachF.as<int>(), Lcol, Lcol + (size_t)nb * (kCholWb - 1) * 36,
                                                               ba->d_x.as<double>(), d_status);
            else
                k_chol_band_reg<<<1, kChol2Threads, (size_t)(2 * 6 * kCholWb * 6 + kChol2Threads * 36 + n) * sizeof(double), st>>>(
                    dS, nb, Lcol, Ldiag, ba->d_x.as<double>(), d_status);
            }
            VIO_LAUNCH_CHECK(ctx);
            KernelTimer kt_(ctx, "k_chol_band_back");
            if (smem_window)
                k_chol_band_back<<<1, 32, n * sizeof(double), st>>>(nb, ba->d_reachF.as<int>(), Lcol, Ldiag, ba->d_x.as<double>());
            else
                k_chol_band_back_col<<<1, kCholBackThreads, 2 * n * sizeof(double), st>>>(nb, Lcol, Ldiag, Linv, ba->d_x.as<double>());
        } else {
            VIO_CUDA(ctx, ba->d_chol.reserve(n * n * 8));
            VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_chol.p, dS, n * n * 8, cudaMemcpyDeviceToDevice, st));
            KernelTimer kt_(ctx, "k_chol_skyline");
            k_chol_skyline<<<1, kCholThreads, panel_bytes, st>>>(ba->d_chol.as<double>(), nb, ba->d_reachF.as<int>(), ba->d_linv.as<double>(),
                                                                ba->d_x.as<double>(), d_status);
        }
        VIO_LAUNCH_CHECK(ctx);
    }
    if (want_points && M > 0) {
        VIO_CUDA(ctx, ba->d_dl.reserve(3 * (size_t)M * 8));
        VIO_CUDA(ctx, cudaMemsetAsync(ba->d_dl.p, 0, 3 * (size_t)M * 8, st));
        const int nl = ba->lm_end - ba->lm_begin;
        if (nl > 0) {
            BaDev d = make_dev(ba);
            KernelTimer kt_(ctx, "k_back_substitute");
            k_back_substitute<<<(nl + 3) / 4, 128, 0, st>>>(d, ba->lm_begin, ba->lm_end, ba->reduced_apply_loss, ba->d_x.as<double>(), ba->d_dl.as<double>());
        }
        VIO_LAUNCH_CHECK(ctx);
    }
    return VIO360_OK;
}

extern "C" int vio360_ba_solve_reduced(vio360_ba* ba, double* delta_pose, double* delta_point) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_reduced) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_solve_reduced before ba_build_reduced_system");
    if (!delta_pose) return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_solve_reduced: delta_pose is required");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int nb = ba->n_free, K = ba->K, M = ba->M;
    const size_t n = 6 * (size_t)nb;
    VIO_CUDA(ctx, ba->h_solve.reserve((n + 3 * (size_t)M + 8) * 8));
    double* hx = ba->h_solve.as<double>();
    { int rcS = solve_reduced_dev(ba, delta_point != nullptr); if (rcS) return rcS; }
    if (nb > 0) VIO_CUDA(ctx, cudaMemcpyAsync(hx, ba->d_x.p, n * 8 + 8, cudaMemcpyDeviceToHost, st));
    if (delta_point && M > 0) VIO_CUDA(ctx, cudaMemcpyAsync(hx + n + 1, ba->d_dl.p, 3 * (size_t)M * 8, cudaMemcpyDeviceToHost, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    if (nb > 0 && *reinterpret_cast<const int*>(hx + n) != 0)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_solve_reduced: the reduced system is not positive definite");
    for (int p = 0; p < K; ++p) {
        const int a = ba->free_map[p];
        for (int k = 0; k < 6; ++k) delta_pose[6 * (size_t)p + k] = a >= 0 ? hx[6 * (size_t)a + k] : 0.0;
    }
    if (delta_point && M > 0) memcpy(delta_point, hx + n + 1, 3 * (size_t)M * 8);
    return VIO360_OK;
}

// ---- the LM iteration of ceres::Solve on the device-built reduced system ----
extern "C" int vio360_lm_default_options(vio360_lm_options* o) {
    if (!o) return VIO360_ERR_INVALID;
    o->max_num_iterations = 50;                 // Optimizer.cpp:76 (options.max_num_iterations)
    o->initial_trust_region_radius = 1e4;       // ceres::Solver::Options defaults (include/ceres/solver.h)
    o->max_trust_region_radius = 1e16;
    o->min_trust_region_radius = 1e-32;
    o->min_relative_decrease = 1e-3;
    o->min_lm_diagonal = 1e-6;
    o->max_lm_diagonal = 1e32;
    o->function_tolerance = 1e-6;
    o->gradient_tolerance = 1e-10;
    o->parameter_tolerance = 1e-8;
    o->jacobi_scaling = 1;
    o->max_num_consecutive_invalid_steps = 5;
    o->apply_loss = 1;
    return VIO360_OK;
}

extern "C" int vio360_ba_lm_solve(vio360_ba* ba, const vio360_lm_options* opt, double* pose_delta, double* points,
                                  vio360_lm_summary* summary, double* trace, int trace_capacity) {
    if (!ba) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_structure) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_lm_solve before ba_set_structure");
    if (!opt || !pose_delta || !points || !summary) return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_lm_solve: null argument");
    if (ba->shard_world > 1) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_lm_solve runs on an unsharded window (vio360_ba_set_shard(0, 1))");
    if (!ba->irregular.empty()) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_lm_solve: %zu landmarks are observed twice from one pose", ba->irregular.size());
    int rc = vio360_ba_set_parameters(ba, pose_delta, points);     // x0 to the device, pose chain at x0
    if (rc) return rc;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const int K = ba->K, M = ba->M, O = ba->O, nvar = 6 * K + 3 * M;
    const size_t Ks = (size_t)K, Ms = (size_t)M;
    VIO_CUDA(ctx, ba->d_cand_pose_delta.reserve(Ks * 6 * 8)); VIO_CUDA(ctx, ba->d_cand_points.reserve(Ms * 3 * 8));
    VIO_CUDA(ctx, ba->d_cand_posebuf.reserve(Ks * kPoseStride * 8));
    VIO_CUDA(ctx, ba->d_step.reserve((size_t)nvar * 8)); VIO_CUDA(ctx, ba->d_colsq.reserve((size_t)nvar * 8));
    VIO_CUDA(ctx, ba->d_scale.reserve((size_t)nvar * 8)); VIO_CUDA(ctx, ba->d_lm_state.reserve(2 * sizeof(LmState)));
    const int nbo = std::max(1, (O + 127) / 128), nbc = std::min(64, std::max(1, (nvar + kLmRedThreads - 1) / kLmRedThreads));
    const int nbl = std::min(148, std::max(1, (nvar + kLmRedThreads - 1) / kLmRedThreads));   // CTAs of k_lm_linearized
    VIO_CUDA(ctx, ba->d_lm_part.reserve((size_t)(2 * nbo + nbc + 2 * nbl + 8) * 8));
    VIO_CUDA(ctx, ba->h_lm.reserve(sizeof(LmState)));
    VIO_CUDA(ctx, ba->d_x.reserve(6 * (size_t)std::max(ba->n_free, 1) * 8 + 16));
    double* part_step = ba->d_lm_part.as<double>(); double* part_model = part_step + nbc; double* part_cost = part_model + nbo;
    double* part_lin = part_cost + nbo;
    LmState* d_state = ba->d_lm_state.as<LmState>();
    LmState* h_state = ba->h_lm.as<LmState>();
    LmOptionsDev od;
    od.min_diag = opt->min_lm_diagonal; od.max_diag = opt->max_lm_diagonal; od.max_radius = opt->max_trust_region_radius;
    od.min_radius = opt->min_trust_region_radius; od.min_relative_decrease = opt->min_relative_decrease;
    od.function_tolerance = opt->function_tolerance; od.gradient_tolerance = opt->gradient_tolerance;
    od.parameter_tolerance = opt->parameter_tolerance; od.jacobi_scaling = opt->jacobi_scaling;
    od.max_invalid = opt->max_num_consecutive_invalid_steps;
    const bool loss = opt->apply_loss != 0;
    auto read_state = [&]() -> int {
        VIO_CUDA(ctx, cudaMemcpyAsync(h_state, d_state, sizeof(LmState), cudaMemcpyDeviceToHost, st));
        VIO_CUDA(ctx, cudaStreamSynchronize(st));
        return VIO360_OK;
    };
    // iteration 0 (IterationZero, trust_region_minimizer.cc:181-224): cost at x0
    LmState init;
    memset(&init, 0, sizeof(init));
    init.radius = opt->initial_trust_region_radius; init.decrease_factor = 2.0; init.step_successful = 1; init.step_valid = 1; init.relinearize = 1;
    {
        BaDev d = make_dev(ba);
        KernelTimer kt_(ctx, "k_lm_cost");
        k_lm_cost<<<nbo, 128, 0, st>>>(d, loss, part_cost);
    }
    VIO_LAUNCH_CHECK(ctx);
    std::vector<double> h_part((size_t)nbo);
    VIO_CUDA(ctx, cudaMemcpyAsync(h_part.data(), part_cost, (size_t)nbo * 8, cudaMemcpyDeviceToHost, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    {   // the initial cost only enters comparisons with tolerances 1e-3 / 1e-6: summed on the host in index order
        double c0 = 0.0;
        for (int i = 0; i < nbo; ++i) c0 += h_part[i];
        init.x_cost = c0;
    }
    VIO_CUDA(ctx, cudaMemcpyAsync(d_state, &init, sizeof(LmState), cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));      // `init` is a stack object
    memset(summary, 0, sizeof(*summary));
    summary->initial_cost = init.x_cost;
    int n_trace = 0;
    auto push_trace = [&](const LmState& s, double cost) {
        if (!trace || n_trace >= trace_capacity) return;
        double* t = trace + 8 * (size_t)n_trace++;
        t[0] = cost; t[1] = s.cost_change; t[2] = s.gradient_max_norm; t[3] = s.step_norm; t[4] = s.relative_decrease;
        t[5] = s.radius; t[6] = (double)(s.step_valid ? (s.step_successful ? 1 : 0) : -1); t[7] = s.model_cost_change;
    };
    bool first = true, relinearize = true;
    int termination = 0;
    LmState last = init;
    for (;;) {
        if (relinearize) {
            // Evaluator::Evaluate with Jacobian at x: the Gauss-Newton blocks give the squared column norms and J'r
            rc = run_blocks(ba, loss ? 1 : 0, false);
            if (rc) return rc;
            const double* Hpp = ba->d_blocks.as<double>(); const double* bp = Hpp + Ks * 36; const double* Hll = bp + Ks * 6; const double* bl = Hll + Ms * 9;
            { KernelTimer kt_(ctx, "k_lm_linearized");
            k_lm_linearized<<<nbl, kLmRedThreads, 0, st>>>(K, M, Hpp, bp, Hll, bl, ba->d_free_map.as<int>(), ba->d_point_fixed.as<unsigned char>(),
                                                         ba->d_pose_delta.as<double>(), ba->d_points.as<double>(), ba->d_colsq.as<double>(),
                                                         ba->d_scale.as<double>(), first ? 1 : 0, od.jacobi_scaling, part_lin);
            }
            VIO_LAUNCH_CHECK(ctx);
            k_lm_linearized_finish<<<1, kLmRedThreads, 0, st>>>(part_lin, nbl, d_state, od.gradient_tolerance);
            VIO_LAUNCH_CHECK(ctx);
            relinearize = false;
            if (first) {   // iteration 0: its trace row needs the gradient norm; later re-linearisations are not read back --
                           // k_lm_linearized_finish makes the gradient test (:674-692) and a terminated state makes the
                           // step below a no-op that the read after k_lm_decide reports
                if ((rc = read_state())) return rc;
                last = *h_state;
                push_trace(last, last.x_cost);
                first = false;
                if (last.termination != 0) { termination = last.termination; break; }
            }
        }
        // ---- ComputeTrustRegionStep: LM diagonal, fused build of the reduced system, Cholesky, back substitution
        { KernelTimer kt_(ctx, "k_lm_diagonal");
        k_lm_diagonal<<<(nvar + 255) / 256, 256, 0, st>>>(nvar, ba->d_colsq.as<double>(), ba->d_scale.as<double>(), d_state, od.min_diag, od.max_diag, ba->d_D.as<double>());
        }
        VIO_LAUNCH_CHECK(ctx);
        ba->use_D = true;
        if ((rc = vio360_ba_build_reduced_system(ba, nullptr, loss ? 1 : 0, nullptr, nullptr, nullptr))) return rc;
        if ((rc = solve_reduced_dev(ba, true))) return rc;
        const int* d_status = reinterpret_cast<const int*>(ba->d_x.as<double>() + 6 * (size_t)ba->n_free);
        { KernelTimer kt_(ctx, "k_lm_candidate");
        k_lm_candidate<<<nbc, kLmRedThreads, 0, st>>>(K, M, ba->d_free_map.as<int>(), ba->d_point_fixed.as<unsigned char>(), ba->d_x.as<double>(),
                                                     ba->d_dl.as<double>(), ba->d_pose_delta.as<double>(), ba->d_points.as<double>(), ba->d_step.as<double>(),
                                                     ba->d_cand_pose_delta.as<double>(), ba->d_cand_points.as<double>(), part_step);
        }
        VIO_LAUNCH_CHECK(ctx);
        k_pose_prepare<<<(K + 63) / 64, 64, 0, st>>>(K, ba->d_Twb.as<double>(), ba->d_Tcb.as<double>(), ba->d_cand_pose_delta.as<double>(), ba->d_cand_posebuf.as<double>());
        VIO_LAUNCH_CHECK(ctx);
        BaDev d = make_dev(ba);
        { KernelTimer kt_(ctx, "k_lm_model_change");
        k_lm_model_change<<<nbo, 128, 0, st>>>(d, loss, ba->d_step.as<double>(), part_model);
        }
        VIO_LAUNCH_CHECK(ctx);
        BaDev dc = d;
        dc.posebuf = ba->d_cand_posebuf.as<double>(); dc.points = ba->d_cand_points.as<double>();
        { KernelTimer kt_(ctx, "k_lm_cost");
        k_lm_cost<<<nbo, 128, 0, st>>>(dc, loss, part_cost);
        }
        VIO_LAUNCH_CHECK(ctx);
        { KernelTimer kt_(ctx, "k_lm_decide");
        k_lm_decide<<<1, kLmRedThreads, 0, st>>>(part_step, nbc, part_model, nbo, part_cost, nbo, d_status, od, opt->max_num_iterations, d_state, d_state + 1);
        }
        VIO_LAUNCH_CHECK(ctx);
        if ((rc = read_state())) return rc;
        last = *h_state;
        if (last.termination == 1) { termination = 1; break; }   // set by the re-linearisation of the previous step: no step was taken
        push_trace(last, last.step_successful ? last.x_cost : last.candidate_cost);
        if (last.step_successful) {   // x = candidate_x (HandleSuccessfulStep): swap the current and candidate buffers
            std::swap(ba->d_pose_delta, ba->d_cand_pose_delta); std::swap(ba->d_points, ba->d_cand_points); std::swap(ba->d_posebuf, ba->d_cand_posebuf);
            relinearize = true;
        }
        if (last.termination != 0) { termination = last.termination; break; }
    }
    // the reduced system / evaluation caches belong to an earlier point
    ba->has_eval = false; ba->has_reduced = false; ba->has_params = true;
    VIO_CUDA(ctx, cudaMemcpyAsync(pose_delta, ba->d_pose_delta.p, Ks * 6 * 8, cudaMemcpyDeviceToHost, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(points, ba->d_points.p, Ms * 3 * 8, cudaMemcpyDeviceToHost, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    summary->iterations = last.iteration; summary->num_successful_steps = last.num_successful + 1;   // iteration 0 counts (Ceres)
    summary->num_unsuccessful_steps = last.num_unsuccessful; summary->termination = termination;
    summary->final_cost = last.x_cost; summary->final_radius = last.radius; summary->final_gradient_max_norm = last.gradient_max_norm;
    summary->n_trace = n_trace;
    if (termination == -1) return vio360_fail(ctx, VIO360_ERR_INVALID, "ba_lm_solve: %d consecutive invalid steps (the reduced system is not positive definite or the model does not decrease)", last.num_invalid);
    return VIO360_OK;
}

// y += J x at the current parameters (see k_jacobian_right_multiply): what a ceres::internal::SparseMatrix standing for
// the device-resident Jacobian answers to RightMultiply (host/ceres_adapters).
extern "C" int vio360_ba_jacobian_right_multiply(vio360_ba* ba, int apply_loss, const double* x, double* y) {
    if (!ba || !x || !y) return VIO360_ERR_INVALID;
    vio360_ctx* ctx = ba->ctx;
    if (!ba->has_params) return vio360_fail(ctx, VIO360_ERR_STATE, "ba_jacobian_right_multiply before ba_set_parameters");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t nv = 6 * (size_t)ba->K + 3 * (size_t)ba->M, O = (size_t)ba->O;
    if (O == 0) return VIO360_OK;
    cudaStream_t st = ctx->stream;
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    VIO_CUDA(ctx, ba->d_step.reserve(nv * 8)); VIO_CUDA(ctx, ba->d_r.reserve(O * 16));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_step.p, x, nv * 8, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(ba->d_r.p, y, O * 16, cudaMemcpyHostToDevice, st));
    ba->has_eval = false;   // d_r is borrowed
    { KernelTimer kt_(ctx, "k_jacobian_right_multiply");
    k_jacobian_right_multiply<<<(int)((O + 127) / 128), 128, 0, st>>>(make_dev(ba), apply_loss != 0, ba->d_step.as<double>(), ba->d_r.as<double>());
    }
    VIO_LAUNCH_CHECK(ctx);
    VIO_CUDA(ctx, cudaMemcpyAsync(y, ba->d_r.p, O * 16, cudaMemcpyDeviceToHost, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    return VIO360_OK;
}

// ---- batched DLT triangulation (Estimator.cpp:1082-1139) ----
extern "C" int vio360_triangulate(vio360_ctx* ctx, const float* bearing1, const float* bearing2, const float* T1w, const float* T2w, int n,
                                  float* points, uint8_t* valid) {
    if (!ctx) return VIO360_ERR_INVALID;
    if (n < 0 || !T1w || !T2w || (n > 0 && (!bearing1 || !bearing2 || !points || !valid)))
        return vio360_fail(ctx, VIO360_ERR_INVALID, "triangulate: bad arguments");
    if (n == 0) return VIO360_OK;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t nf = 3 * (size_t)n;
    void* scratch_p = nullptr;
    VIO_CUDA(ctx, vio360_scratch(ctx, nf * 4 * 3 + (size_t)n, &scratch_p));
    float* d_b1 = static_cast<float*>(scratch_p); float* d_b2 = d_b1 + nf; float* d_pts = d_b2 + nf;
    unsigned char* d_valid = reinterpret_cast<unsigned char*>(d_pts + nf);
    VIO_CUDA(ctx, cudaMemcpyAsync(d_b1, bearing1, nf * 4, cudaMemcpyHostToDevice, st));
    VIO_CUDA(ctx, cudaMemcpyAsync(d_b2, bearing2, nf * 4, cudaMemcpyHostToDevice, st));
    Mat4f A, B;
    memcpy(A.m, T1w, sizeof(A.m)); memcpy(B.m, T2w, sizeof(B.m));
    { KernelTimer kt_(ctx, "k_triangulate");
    k_triangulate<<<(n + 127) / 128, 128, 0, st>>>(n, d_b1, d_b2, A, B, d_pts, d_valid);
    }
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return vio360_fail(ctx, VIO360_ERR_CUDA, "k_triangulate: %s", cudaGetErrorString(e));
    cudaMemcpyAsync(points, d_pts, nf * 4, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(valid, d_valid, (size_t)n, cudaMemcpyDeviceToHost, st);
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return vio360_fail(ctx, VIO360_ERR_CUDA, "triangulate: %s", cudaGetErrorString(e));
    return VIO360_OK;
}
