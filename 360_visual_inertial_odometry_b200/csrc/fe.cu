// fe.cu -- path A host side: device pyramid cache, batched tracking and detection.
// Compiled with --fmad=false (see fe_lk.cuh / fe_detect.cuh); explicit fmaf() calls mark
// the places where OpenCV's dispatched code fuses.
#include <algorithm>
#include <mutex>

#include <time.h>

#include "common.cuh"
#include "fe_detect.cuh"
#include "fe_grid.cuh"
#include "fe_ransac.cuh"
#include "fe_lk.cuh"
#include "fe_pyr.cuh"
#include "fe_seq.cuh"

using namespace vio360;

static uint64_t g_fe_generation = 0;   // bumped by every vio360_fe_configure: sequences of an older state are invalid

struct FeSlot {
    uint64_t key = 0;
    uint64_t last_use = 0;
    uint64_t batch_tag = 0;      // slot is pinned by the batch with this tag
    bool has_pyramid = false;
    bool reserved = false;       // owned by a vio360_seq: never evicted by the keyed cache
    bool loaded = false;         // level 0 has been aliased / its upload has been enqueued
    uint64_t pending_tag = 0;    // batch whose upload list holds this slot's frame (not enqueued yet)
    DevBuf mem;                  // levels 0..L-1 (level 0 unused when aliased)
    uint8_t* own[kMaxLevels] = {};
    const uint8_t* p[kMaxLevels] = {};
    int pitch[kMaxLevels] = {};
};

struct vio360_fe_state {
    bool configured = false;
    uint64_t generation = 0;
    DevBuf d_seq;                // per-call tables of the per-frame chain (fe_seq_host.inc)
    PinBuf h_seq;
    int W = 0, H = 0, nlevels = 0, win = 21, max_iters = 30;
    double eps2 = 0, min_eig = 0;
    float polar_ratio = 0.15f;
    int margin = 20, max_points = 0, max_batch = 0;
    int lw[kMaxLevels] = {}, lh[kMaxLevels] = {}, lpitch[kMaxLevels] = {};
    size_t loff[kMaxLevels] = {}, slot_bytes = 0;
    std::vector<FeSlot> slots;
    std::unordered_map<uint64_t, int> by_key;
    uint64_t clock = 0, batch_tag = 0;
    // tracking pools
    DevBuf d_jobs, d_pyrjobs, d_pts_in, d_pts_out, d_status, d_err, d_keep;
    DevBuf d_lk_queue;           // k_lk_track's work counter {next item, finished warps}; zero between launches
    cudaEvent_t det_ev[2] = {nullptr, nullptr};   // fork / join of the detector's second stream
    PinBuf h_pts_in, h_out;
    // ring of pinned descriptor staging so that device-space calls never wait for the stream
    struct Stage { PinBuf jobs, pyrjobs; cudaEvent_t ev = nullptr; bool used = false; };
    Stage stage[4];
    DevBuf d_jobs_ring[4], d_pyrjobs_ring[4];
    unsigned stage_idx = 0;
    Stage* cur_stage = nullptr;
    int cur_stage_i = 0;
    // detection pools
    DevBuf d_djobs, d_eig, d_mask, d_small, d_cand, d_grid, d_corners, d_edge;
    PinBuf h_djobs, h_dout;
    DevBuf d_occ;
    PinBuf h_occ;
    // rotation RANSAC pools
    DevBuf d_ransac;
    PinBuf h_ransac;
    int det_batch_cap = 0, max_cand = 1 << 17;   // initial candidate capacity per image (grown on demand)
    int eig_jobs = 0;                             // images the response-map pool holds (none while only the fused detector runs)
    size_t cand_cap = 0;
    int lk_mb = 2;               // occupancy variant of k_lk_track (tuning knob: VIO360_LK_MB)
    // host-space batches: frames are uploaded on a second stream in groups of h2d_group jobs, and a group's
    // pyramid + LK launches wait only for its own frames, so the kernels run behind the PCIe copies
    cudaStream_t copy_stream = nullptr;
    cudaStream_t pyr_stream = nullptr;   // device-space batches: pyramids of the next group beside the LK kernel of this one
    std::vector<cudaEvent_t> grp_ev;
    int h2d_group = 4;           // VIO360_FE_H2D_GROUP (3840x1920 x 64 pairs: 4 -> 6959, 8 -> 6801, 16 -> 6823, off -> 6348 frames/s)
};

void vio360_fe_release(vio360_ctx* ctx) {
    vio360_fe_state* s = ctx->fe;
    if (!s) return;
    for (auto& sl : s->slots) sl.mem.release();
    s->d_seq.release(); s->h_seq.release();
    DevBuf* db[] = {&s->d_jobs, &s->d_pyrjobs, &s->d_pts_in, &s->d_pts_out, &s->d_status, &s->d_err, &s->d_keep,
                    &s->d_djobs, &s->d_eig, &s->d_mask, &s->d_small, &s->d_cand, &s->d_grid,
                    &s->d_corners, &s->d_ransac, &s->d_occ, &s->d_lk_queue, &s->d_edge};
    for (DevBuf* b : db) b->release();
    PinBuf* pb[] = {&s->h_pts_in, &s->h_out, &s->h_djobs, &s->h_dout, &s->h_ransac, &s->h_occ};
    for (PinBuf* b : pb) b->release();
    for (auto& sg : s->stage) { sg.jobs.release(); sg.pyrjobs.release(); if (sg.ev) cudaEventDestroy(sg.ev); }
    for (auto& d : s->d_jobs_ring) d.release();
    for (auto& d : s->d_pyrjobs_ring) d.release();
    for (cudaEvent_t e : s->grp_ev) cudaEventDestroy(e);
    for (cudaEvent_t e : s->det_ev) if (e) cudaEventDestroy(e);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    if (s->pyr_stream) cudaStreamDestroy(s->pyr_stream);
    delete s;
    ctx->fe = nullptr;
}

static inline int align_up(int v, int a) { return (v + a - 1) / a * a; }

extern "C" int vio360_fe_configure(vio360_ctx* ctx, int width, int height, int max_level, int win, int max_iters,
                                   double eps, double min_eig_thr, float polar_ratio, int boundary_margin,
                                   int max_points, int max_batch) {
    if (!ctx) return VIO360_ERR_INVALID;
    if (width < 32 || height < 32 || max_level < 0 || max_level >= kMaxLevels || max_points <= 0 || max_batch <= 0)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_configure: bad size/level arguments");
    if (win != kWin)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_configure: only the reference's 21x21 window is built (FeatureTracker.cpp:33)");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    vio360_fe_release(ctx);
    vio360_fe_state* s = new vio360_fe_state();
    ctx->fe = s;
    s->generation = ++g_fe_generation;
    s->W = width; s->H = height; s->win = win;
    // calcOpticalFlowPyrLK: maxCount clamped to [0,100], epsilon to [0,10] then squared
    s->max_iters = std::min(std::max(max_iters, 0), 100);
    const double e = std::min(std::max(eps, 0.0), 10.0);
    s->eps2 = e * e;
    s->min_eig = min_eig_thr;
    s->polar_ratio = polar_ratio; s->margin = boundary_margin;
    s->max_points = max_points; s->max_batch = max_batch;
    // buildOpticalFlowPyramid: stop before a level that does not exceed the window
    int w = width, h = height, L = 0;
    size_t off = 0;
    for (int l = 0; l <= max_level; ++l) {
        s->lw[l] = w; s->lh[l] = h; s->lpitch[l] = align_up(w, 128);
        s->loff[l] = off; off += (size_t)s->lpitch[l] * h;
        L = l + 1;
        const int nw = (w + 1) / 2, nh = (h + 1) / 2;
        if (nw <= win || nh <= win) break;
        w = nw; h = nh;
    }
    s->nlevels = L; s->slot_bytes = off;
    s->slots.resize((size_t)2 * max_batch + 2);
    for (auto& sl : s->slots) {
        VIO_CUDA(ctx, sl.mem.reserve(s->slot_bytes));
        for (int l = 0; l < L; ++l) { sl.own[l] = sl.mem.as<uint8_t>() + s->loff[l]; sl.p[l] = sl.own[l]; sl.pitch[l] = s->lpitch[l]; }
    }
    const size_t np = (size_t)max_points * max_batch;
    for (int i = 0; i < 4; ++i) {
        VIO_CUDA(ctx, s->stage[i].jobs.reserve(sizeof(TrackJobDev) * (size_t)max_batch));
        VIO_CUDA(ctx, s->stage[i].pyrjobs.reserve(sizeof(PyrJob) * (size_t)2 * max_batch * kMaxLevels));
        VIO_CUDA(ctx, cudaEventCreateWithFlags(&s->stage[i].ev, cudaEventDisableTiming));
        VIO_CUDA(ctx, s->d_jobs_ring[i].reserve(sizeof(TrackJobDev) * (size_t)max_batch));
        VIO_CUDA(ctx, s->d_pyrjobs_ring[i].reserve(sizeof(PyrJob) * (size_t)2 * max_batch * kMaxLevels));
    }
    VIO_CUDA(ctx, s->d_pts_in.reserve(np * 8)); VIO_CUDA(ctx, s->d_pts_out.reserve(np * 8));
    VIO_CUDA(ctx, s->d_status.reserve(np)); VIO_CUDA(ctx, s->d_err.reserve(np * 4)); VIO_CUDA(ctx, s->d_keep.reserve(np));
    VIO_CUDA(ctx, s->h_pts_in.reserve(np * 8));
    VIO_CUDA(ctx, s->h_out.reserve(np * (8 + 1 + 4 + 1)));
    for (cudaEvent_t& e : s->det_ev) VIO_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    VIO_CUDA(ctx, s->d_lk_queue.reserve(256));
    VIO_CUDA(ctx, cudaMemsetAsync(s->d_lk_queue.p, 0, 256, ctx->stream));
    if (const char* e = getenv("VIO360_LK_MB")) s->lk_mb = atoi(e);
    if (const char* e = getenv("VIO360_FE_H2D_GROUP")) s->h2d_group = std::max(1, atoi(e));
    VIO_CUDA(ctx, cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
    {
        int lo = 0, hi = 0;
        VIO_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&lo, &hi));
        VIO_CUDA(ctx, cudaStreamCreateWithPriority(&s->pyr_stream, cudaStreamNonBlocking, hi));   // short kernels: let them in first
    }
    if (const char* e = getenv("VIO360_DET_CAND")) s->max_cand = std::max(1, atoi(e));   // initial candidate capacity per image
    s->configured = true;
    return VIO360_OK;
}

// Picks the next pinned descriptor stage; waits only for the copies that last used it.
static int begin_stage(vio360_ctx* ctx) {
    vio360_fe_state* s = ctx->fe;
    s->cur_stage_i = (int)(s->stage_idx++ & 3u);
    s->cur_stage = &s->stage[s->cur_stage_i];
    if (s->cur_stage->used) VIO_CUDA(ctx, cudaEventSynchronize(s->cur_stage->ev));
    s->cur_stage->used = true;
    return VIO360_OK;
}
static int end_stage(vio360_ctx* ctx) {
    VIO_CUDA(ctx, cudaEventRecord(ctx->fe->cur_stage->ev, ctx->stream));
    return VIO360_OK;
}

// ---- pyramid cache -------------------------------------------------------------------
static int acquire_slot(vio360_ctx* ctx, uint64_t key, bool* hit, int* out_idx) {
    vio360_fe_state* s = ctx->fe;
    *hit = false;
    if (key != 0) {
        auto it = s->by_key.find(key);
        if (it != s->by_key.end()) {
            FeSlot& sl = s->slots[it->second];
            sl.last_use = ++s->clock; sl.batch_tag = s->batch_tag;
            // a slot whose frame never got as far as the copy queue (an earlier call failed between acquiring
            // the slot and enqueuing the upload) holds stale pixels: treat the hit as a miss
            *hit = sl.loaded || sl.pending_tag == s->batch_tag;
            if (!*hit) { sl.has_pyramid = false; sl.p[0] = sl.own[0]; sl.pitch[0] = s->lpitch[0]; }
            *out_idx = it->second;
            return VIO360_OK;
        }
    }
    int best = -1;
    for (int i = 0; i < (int)s->slots.size(); ++i) {
        const FeSlot& sl = s->slots[i];
        if (sl.batch_tag == s->batch_tag || sl.reserved) continue;   // in use by this batch / owned by a sequence
        if (best < 0 || sl.last_use < s->slots[best].last_use) best = i;
    }
    if (best < 0) return vio360_fail(ctx, VIO360_ERR_STATE, "pyramid cache exhausted (batch larger than max_batch?)");
    FeSlot& sl = s->slots[best];
    if (sl.key != 0) s->by_key.erase(sl.key);
    sl.key = key; sl.last_use = ++s->clock; sl.batch_tag = s->batch_tag; sl.has_pyramid = false;
    sl.loaded = false; sl.pending_tag = 0;
    sl.p[0] = sl.own[0]; sl.pitch[0] = s->lpitch[0];
    if (key != 0) s->by_key[key] = best;
    *out_idx = best;
    return VIO360_OK;
}

static int load_frame(vio360_ctx* ctx, int slot, const uint8_t* img, size_t pitch, int mem_space, cudaStream_t h2d_stream = nullptr) {
    vio360_fe_state* s = ctx->fe;
    FeSlot& sl = s->slots[slot];
    sl.loaded = true;
    if (mem_space == VIO360_MEM_DEVICE) {
        if ((reinterpret_cast<uintptr_t>(img) & 3) == 0 && (pitch & 3) == 0) {
            sl.p[0] = img; sl.pitch[0] = (int)pitch;   // alias: no copy of level 0
            return VIO360_OK;
        }
        VIO_CUDA(ctx, cudaMemcpy2DAsync(sl.own[0], s->lpitch[0], img, pitch, s->W, s->H, cudaMemcpyDeviceToDevice, ctx->stream));
    } else {
        cudaStream_t cs = h2d_stream ? h2d_stream : ctx->stream;
        if (pitch == (size_t)s->W && s->lpitch[0] == s->W) {   // contiguous on both sides: one linear DMA (cut at registration boundaries)
            VIO_CUDA(ctx, vio360_h2d_linear(ctx, sl.own[0], img, (size_t)s->W * s->H, cs));
        } else if (ctx->registered.empty()) {
            VIO_CUDA(ctx, cudaMemcpy2DAsync(sl.own[0], s->lpitch[0], img, pitch, s->W, s->H, cudaMemcpyHostToDevice, cs));
        } else {
            // a pitched view: one 2-D copy when its byte span lies inside one registration or outside all of them, else row by row
            const uintptr_t p0 = reinterpret_cast<uintptr_t>(img), p1 = p0 + (size_t)(s->H - 1) * pitch + (size_t)s->W;
            auto it = ctx->registered.upper_bound(p0);
            bool one_piece;
            if (it != ctx->registered.begin() && std::prev(it)->second > p0) one_piece = std::prev(it)->second >= p1;   // starts inside a registration
            else one_piece = it == ctx->registered.end() || it->first >= p1;                                        // starts in pageable memory
            if (one_piece) {
                VIO_CUDA(ctx, cudaMemcpy2DAsync(sl.own[0], s->lpitch[0], img, pitch, s->W, s->H, cudaMemcpyHostToDevice, cs));
            } else {
                for (int y = 0; y < s->H; ++y)
                    VIO_CUDA(ctx, vio360_h2d_linear(ctx, sl.own[0] + (size_t)y * s->lpitch[0], img + (size_t)y * pitch, (size_t)s->W, cs));
            }
        }
    }
    sl.p[0] = sl.own[0]; sl.pitch[0] = s->lpitch[0];
    return VIO360_OK;
}

// build levels 1..L-1 for the listed slots, one launch per level for the whole batch.
// Three steps so that a caller can ship the descriptors of several groups in one copy BEFORE it queues bulk
// uploads (host-to-device copies of all streams share the copy engine's queue in issue order):
// stage_pyr_jobs fills the pinned table, upload_pyr_jobs copies a range, launch_pyramids launches.
static void stage_pyr_jobs(vio360_ctx* ctx, const std::vector<int>& todo, size_t job_off) {
    vio360_fe_state* s = ctx->fe;
    const int nb = (int)todo.size();
    PyrJob* hj = s->cur_stage->pyrjobs.as<PyrJob>() + job_off;
    for (int l = 0; l + 1 < s->nlevels; ++l)
        for (int b = 0; b < nb; ++b) {
            FeSlot& sl = s->slots[todo[b]];
            PyrJob& j = hj[(size_t)l * nb + b];
            j.src = sl.p[l]; j.spitch = sl.pitch[l]; j.dst = sl.own[l + 1]; j.dpitch = s->lpitch[l + 1];
        }
}
static int upload_pyr_jobs(vio360_ctx* ctx, size_t job_off, size_t count) {
    vio360_fe_state* s = ctx->fe;
    if (count == 0) return VIO360_OK;
    VIO_CUDA(ctx, cudaMemcpyAsync(s->d_pyrjobs_ring[s->cur_stage_i].as<PyrJob>() + job_off, s->cur_stage->pyrjobs.as<PyrJob>() + job_off,
                                  sizeof(PyrJob) * count, cudaMemcpyHostToDevice, ctx->stream));
    return VIO360_OK;
}
static int launch_pyramids(vio360_ctx* ctx, const std::vector<int>& todo, size_t job_off, cudaStream_t on = nullptr) {
    vio360_fe_state* s = ctx->fe;
    const cudaStream_t pst = on ? on : ctx->stream;
    if (todo.empty() || s->nlevels <= 1) { for (int i : todo) s->slots[i].has_pyramid = true; return VIO360_OK; }
    const int nb = (int)todo.size();
    const PyrJob* hj = s->cur_stage->pyrjobs.as<PyrJob>() + job_off;    // job_off: several groups share one descriptor stage
    PyrJob* dj = s->d_pyrjobs_ring[s->cur_stage_i].as<PyrJob>() + job_off;
    for (int l = 0; l + 1 < s->nlevels; ++l) {
        const int dw = s->lw[l + 1], dh = s->lh[l + 1];
        // fast path: 8-column lanes need w % 8 == 0 and 8-byte aligned source rows
        bool fast = (s->lw[l] % 8) == 0 && s->lw[l] >= 16 && (s->lpitch[l + 1] % 4) == 0;
        for (int b = 0; b < nb && fast; ++b) {
            const PyrJob& j = hj[(size_t)l * nb + b];
            fast = (reinterpret_cast<uintptr_t>(j.src) % 8) == 0 && (j.spitch % 8) == 0 &&
                   (reinterpret_cast<uintptr_t>(j.dst) % 4) == 0;
        }
        if (fast) {
            const int nwx = (s->lw[l] + kPyrWarpCols - 1) / kPyrWarpCols, wpb = kPyrFastThreads / 32;
            // shorter strips while the launch has fewer than ~24 warps per SM
            int strip = kPyrStripRows;
            while (strip > 4 && (long long)nwx * ((dh + strip - 1) / strip) * nb < 24ll * ctx->num_sms) strip >>= 1;
            const int nsy = (dh + strip - 1) / strip, total = nwx * nsy * nb;
            { KernelTimer kt_(ctx, "k_pyr_down");
            k_pyr_down8<<<(total + wpb - 1) / wpb, kPyrFastThreads, 0, pst>>>(dj + (size_t)l * nb, s->lw[l], s->lh[l], dw, dh, nwx, nsy, total, strip);
            }
        } else {
            dim3 grid((dw + kPyrTW - 1) / kPyrTW, (dh + kPyrTH - 1) / kPyrTH, nb);
            { KernelTimer kt_(ctx, "k_pyr_down");
            k_pyr_down<<<grid, 256, 0, pst>>>(dj + (size_t)l * nb, s->lw[l], s->lh[l], dw, dh);
            }
        }
        VIO_LAUNCH_CHECK(ctx);
    }
    for (int i : todo) s->slots[i].has_pyramid = true;
    return VIO360_OK;
}
static int build_pyramids(vio360_ctx* ctx, const std::vector<int>& todo) {
    vio360_fe_state* s = ctx->fe;
    if (!todo.empty() && s->nlevels > 1) {
        stage_pyr_jobs(ctx, todo, 0);
        int rc = upload_pyr_jobs(ctx, 0, todo.size() * (size_t)(s->nlevels - 1));
        if (rc) return rc;
    }
    return launch_pyramids(ctx, todo, 0);
}

static void fill_levels(const vio360_fe_state* s, const FeSlot& sl, FrameLevels& f) {
    for (int l = 0; l < kMaxLevels; ++l) { f.p[l] = l < s->nlevels ? sl.p[l] : nullptr; f.pitch[l] = l < s->nlevels ? sl.pitch[l] : 0; }
}

static LkParams lk_params(const vio360_fe_state* s) {
    LkParams P;
    P.nlevels = s->nlevels;
    for (int l = 0; l < kMaxLevels; ++l) { P.w[l] = s->lw[l]; P.h[l] = s->lh[l]; }
    P.max_iters = s->max_iters; P.eps2 = s->eps2; P.min_eig = s->min_eig;
    P.polar_thr = s->polar_ratio; P.margin = (float)s->margin;
    return P;
}

// one LK launch for nj jobs of up to max_n points each
static int launch_lk(vio360_ctx* ctx, const TrackJobDev* dj, int nj, int max_n, const LkParams& lp) {
    vio360_fe_state* s = ctx->fe;
    cudaStream_t st = ctx->stream;
    { KernelTimer kt_(ctx, "k_lk_track");
    const int n_items = max_n * nj;
    unsigned* queue = s->d_lk_queue.as<unsigned>();
    auto launch = [&](auto kern, int warps) -> cudaError_t {
        const size_t smem = (size_t)warps * kLkWarpWords * sizeof(uint32_t);
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        static const int carve = getenv("VIO360_LK_CARVE") ? atoi(getenv("VIO360_LK_CARVE")) : -1;   // percent of the 228 KB; -1: the driver's choice
        if (carve >= 0) {
            e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            if (e != cudaSuccess) return e;
        }
        kern<<<dim3((max_n + warps - 1) / warps, nj), warps * 32, smem, st>>>(dj, lp, max_n, n_items, queue);
        return cudaSuccess;
    };
    // persistent form: as many CTAs as the device holds at once (warps_per_sm warps per SM), never more than the work
    auto launch_persist = [&](auto kern, int warps, int warps_per_sm) -> cudaError_t {
        const size_t smem = (size_t)warps * kLkWarpWords * sizeof(uint32_t);
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        const int ctas = std::min((n_items + warps - 1) / warps, ctx->num_sms * (warps_per_sm / warps));
        kern<<<dim3(ctas, 1), warps * 32, smem, st>>>(dj, lp, max_n, n_items, queue);
        return cudaSuccess;
    };
    cudaError_t e;
    switch (s->lk_mb) {   // tuning knob VIO360_LK_MB; the measured variants are listed in DESIGN.md section 8
        case 3: e = launch(k_lk_track<7, 96, false>, 7); break;    // 3 CTAs / SM: 1.16 ms per step against 0.93 (spills)
        case 4: e = launch(k_lk_track<6, 80, false>, 6); break;    // 4 CTAs / SM: 1.02 ms per step
        case 7: e = launch(k_lk_track<4, 128, false>, 4); break;
        case 8: e = launch(k_lk_track<2, 128, false>, 2); break;
        case 10: e = launch(k_lk_track<8, 128, false>, 8); break;   // round 1's shape: 0.726-0.744 ms
        case 11: e = launch_persist(k_lk_track<4, 128, true>, 4, 16); break;
        case 12: e = launch_persist(k_lk_track<1, 128, true>, 1, 16); break;
        case 18: e = launch_persist(k_lk_track<8, 128, true>, 8, 16); break;
        // One-warp CTAs: a slot is refilled as soon as its feature is done (8-warp CTAs idle a quarter of the warp
        // slots waiting for their slowest feature).  64 x 1000 features at 3840x1920: 0.686 ms; persistent work-queue
        // forms 0.73-0.83 ms (the queue state costs registers, which spills), 4-warp CTAs 0.712, 2-warp 0.783.
        default: e = launch(k_lk_track<1, 128, false>, 1); break;
    }
    if (e != cudaSuccess) return vio360_fail(ctx, VIO360_ERR_CUDA, "k_lk_track: %s", cudaGetErrorString(e));
    }
    VIO_LAUNCH_CHECK(ctx);
    return VIO360_OK;
}

extern "C" int vio360_fe_track_batch(vio360_ctx* ctx, const vio360_track_job* jobs, int n_jobs, int mem_space) {
    if (!ctx) return VIO360_ERR_INVALID;
    vio360_fe_state* s = ctx->fe;
    if (!s || !s->configured) return vio360_fail(ctx, VIO360_ERR_STATE, "fe_track before fe_configure");
    if (!jobs || n_jobs < 0 || n_jobs > s->max_batch) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_track_batch: bad job count");
    if (n_jobs == 0) return VIO360_OK;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    // host-space calls share one pinned point staging block: wait for the previous call
    if (mem_space == VIO360_MEM_HOST) VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    int rc0 = begin_stage(ctx);
    if (rc0) return rc0;
    timespec ts_begin; clock_gettime(CLOCK_MONOTONIC, &ts_begin);
    s->batch_tag++;
    // host-space batches are cut into groups of h2d_group jobs: a group's frames are uploaded on the copy
    // stream, its pyramid and LK launches wait only for those copies (device-space batches are one group)
    static const bool force_piped = getenv("VIO360_FE_H2D_FORCE") != nullptr;   // A/B probe: copy stream even for one group
    const bool piped = mem_space == VIO360_MEM_HOST && (n_jobs > s->h2d_group || force_piped);
    // device-space batches, OPT-IN (VIO360_FE_DEV_GROUPS=4): the pyramids of all groups run back to back on a second
    // (high priority) stream and a group's LK launch waits only for its own pyramids, so the HBM-bound pyramid
    // kernels of group g + 1 share the machine with the issue-bound LK kernel of group g.  Measured on B200 (64 x
    // 3840x1920, profiles/r02_bench_run6*.json): 1.027 ms per step against 0.950 ms with one group -- the LK
    // kernel holds every register of the SMs (2 CTAs x 256 threads x 128), so the pyramid CTAs only displace LK
    // CTAs, and four LK launches pay four tails.  Kept for other shapes; off by default.
    static const bool no_overlap = getenv("VIO360_FE_NO_OVERLAP") != nullptr;
    static const int dev_groups = getenv("VIO360_FE_DEV_GROUPS") ? std::max(1, atoi(getenv("VIO360_FE_DEV_GROUPS"))) : 1;
    const bool dev_piped = mem_space == VIO360_MEM_DEVICE && n_jobs >= 2 * dev_groups && dev_groups > 1 && !ctx->timing && !no_overlap;
    const int gsize = piped ? s->h2d_group : (dev_piped ? (n_jobs + dev_groups - 1) / dev_groups : n_jobs), ngroups = (n_jobs + gsize - 1) / gsize;
    std::vector<std::vector<int>> todo(ngroups);
    std::vector<int> queued;
    struct Upload { int slot, group; const uint8_t* img; size_t pitch; };
    std::vector<Upload> uploads;
    if (piped || dev_piped)
        while ((int)s->grp_ev.size() < ngroups + 1) {
            cudaEvent_t e;
            VIO_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            s->grp_ev.push_back(e);
        }
    TrackJobDev* hj = s->cur_stage->jobs.as<TrackJobDev>();
    TrackJobDev* dj = s->d_jobs_ring[s->cur_stage_i].as<TrackJobDev>();
    int max_n = 0;
    size_t off = 0;
    // every job is validated before the first cache slot is taken: a rejected batch leaves the cache untouched
    for (int b = 0; b < n_jobs; ++b) {
        const vio360_track_job& j = jobs[b];
        if (j.n < 0 || j.n > s->max_points) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_track_batch: job %d has %d points (max %d)", b, j.n, s->max_points);
        if (!j.prev || !j.cur || (j.n > 0 && (!j.prev_xy || !j.cur_xy || !j.status)))
            return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_track_batch: job %d has a null pointer", b);
        if (j.prev_pitch < (size_t)s->W || j.cur_pitch < (size_t)s->W) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_track_batch: pitch < width");
    }
    for (int b = 0; b < n_jobs; ++b) {
        const vio360_track_job& j = jobs[b];
        int sp, sc; bool hit;
        const uint8_t* imgs[2] = {j.prev, j.cur};
        const size_t pitches[2] = {j.prev_pitch, j.cur_pitch};
        const uint64_t keys[2] = {j.prev_key, j.cur_key};
        int idx[2];
        const int g = b / gsize;
        for (int f = 0; f < 2; ++f) {
            int rc = acquire_slot(ctx, keys[f], &hit, &idx[f]);
            if (rc) return rc;
            if (!hit) {
                if (mem_space == VIO360_MEM_DEVICE) {       // alias (or device copy) now: the level table below needs it
                    rc = load_frame(ctx, idx[f], imgs[f], pitches[f], mem_space);
                    if (rc) return rc;
                } else {
                    uploads.push_back({idx[f], g, imgs[f], pitches[f]});   // queued after the descriptor copies
                    s->slots[idx[f]].pending_tag = s->batch_tag;
                }
            } else if (mem_space == VIO360_MEM_DEVICE && s->slots[idx[f]].p[0] != s->slots[idx[f]].own[0]) {
                // level 0 of a device-space frame is ALIASED, never copied: a keyed hit re-uses the cached levels
                // 1..L-1 but reads level 0 through the pointer of THIS call (the caller may have moved the frame;
                // see the lifetime rule in vio360.h)
                if ((reinterpret_cast<uintptr_t>(imgs[f]) & 3) == 0 && (pitches[f] & 3) == 0) {
                    s->slots[idx[f]].p[0] = imgs[f]; s->slots[idx[f]].pitch[0] = (int)pitches[f];
                } else if ((rc = load_frame(ctx, idx[f], imgs[f], pitches[f], mem_space))) return rc;
            }
            if (!s->slots[idx[f]].has_pyramid && std::find(queued.begin(), queued.end(), idx[f]) == queued.end()) {
                queued.push_back(idx[f]); todo[g].push_back(idx[f]);
            }
        }
        sp = idx[0]; sc = idx[1];
        TrackJobDev& d = hj[b];
        fill_levels(s, s->slots[sp], d.prev); fill_levels(s, s->slots[sc], d.cur);
        d.n = j.n; d.n_ptr = nullptr;
        if (mem_space == VIO360_MEM_DEVICE) {
            d.prev_xy = j.prev_xy; d.cur_xy = j.cur_xy; d.status = j.status; d.err = j.err; d.keep = j.keep;
        } else {
            d.prev_xy = s->d_pts_in.as<float>() + 2 * off; d.cur_xy = s->d_pts_out.as<float>() + 2 * off;
            d.status = s->d_status.as<uint8_t>() + off; d.err = j.err ? s->d_err.as<float>() + off : nullptr;
            d.keep = j.keep ? s->d_keep.as<uint8_t>() + off : nullptr;
            if (j.n) memcpy(s->h_pts_in.as<float>() + 2 * off, j.prev_xy, (size_t)j.n * 8);
        }
        off += (size_t)j.n;
        max_n = std::max(max_n, j.n);
    }
    cudaStream_t st = ctx->stream;
    int rc = VIO360_OK;
    static const bool fe_trace = getenv("VIO360_FE_TRACE") != nullptr;
    auto now_us = []() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; };
    const double t_copies = fe_trace ? now_us() : 0.0;
    if (max_n > 0) {
        VIO_CUDA(ctx, cudaMemcpyAsync(dj, hj, sizeof(TrackJobDev) * (size_t)n_jobs, cudaMemcpyHostToDevice, st));
        if (mem_space == VIO360_MEM_HOST)
            VIO_CUDA(ctx, cudaMemcpyAsync(s->d_pts_in.p, s->h_pts_in.p, off * 8, cudaMemcpyHostToDevice, st));
    }
    // an error return below must not leave uploads of the caller's frames in flight on the copy stream
    struct CopyGuard { cudaStream_t st; bool armed; ~CopyGuard() { if (armed) cudaStreamSynchronize(st); } } copy_guard{s->copy_stream, piped};
    // descriptors first, bulk frame uploads second: host-to-device copies of both streams are served in issue
    // order, so a descriptor copy queued behind the frames would hold every kernel back until the last frame
    const size_t per = (size_t)std::max(0, s->nlevels - 1);
    size_t pyr_off = 0;
    for (int g = 0; g < ngroups; ++g) { stage_pyr_jobs(ctx, todo[g], pyr_off); pyr_off += todo[g].size() * per; }
    if ((rc = upload_pyr_jobs(ctx, 0, pyr_off))) return rc;
    {
        size_t u = 0;
        for (int g = 0; g < ngroups; ++g) {
            for (; u < uploads.size() && uploads[u].group == g; ++u)
                if ((rc = load_frame(ctx, uploads[u].slot, uploads[u].img, uploads[u].pitch, mem_space, piped ? s->copy_stream : nullptr))) return rc;
            if (piped) VIO_CUDA(ctx, cudaEventRecord(s->grp_ev[g], s->copy_stream));
        }
    }
    const LkParams lp = lk_params(s);
    pyr_off = 0;
    if (dev_piped) {
        // the side stream starts after everything already queued on the main stream (descriptor copies of this call,
        // the previous call's LK kernels, which may still read the slots this call rebuilds)
        VIO_CUDA(ctx, cudaEventRecord(s->grp_ev[ngroups], st));
        VIO_CUDA(ctx, cudaStreamWaitEvent(s->pyr_stream, s->grp_ev[ngroups], 0));
        size_t po = 0;
        for (int g = 0; g < ngroups; ++g) {
            if ((rc = launch_pyramids(ctx, todo[g], po, s->pyr_stream))) return rc;
            po += todo[g].size() * per;
            VIO_CUDA(ctx, cudaEventRecord(s->grp_ev[g], s->pyr_stream));
        }
    }
    for (int g = 0; g < ngroups; ++g) {
        const int j0 = g * gsize, nj = std::min(gsize, n_jobs - j0);
        if (piped || dev_piped) VIO_CUDA(ctx, cudaStreamWaitEvent(st, s->grp_ev[g], 0));
        if (!dev_piped && (rc = launch_pyramids(ctx, todo[g], pyr_off))) return rc;
        pyr_off += todo[g].size() * per;
        if (max_n == 0) continue;
        if ((rc = launch_lk(ctx, dj + j0, nj, max_n, lp))) return rc;
    }
    if (max_n == 0) {
        if (mem_space == VIO360_MEM_HOST) VIO_CUDA(ctx, cudaStreamSynchronize(st));   // the caller's frames are free again
        return end_stage(ctx);
    }
    if ((rc = end_stage(ctx))) return rc;
    const double t_kernels = fe_trace ? now_us() : 0.0;
    if (mem_space == VIO360_MEM_HOST) {
        // one pinned block: [xy off*8 | err off*4 | status off | keep off]
        uint8_t* ho = s->h_out.as<uint8_t>();
        VIO_CUDA(ctx, cudaMemcpyAsync(ho, s->d_pts_out.p, off * 8, cudaMemcpyDeviceToHost, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ho + off * 8, s->d_err.p, off * 4, cudaMemcpyDeviceToHost, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ho + off * 12, s->d_status.p, off, cudaMemcpyDeviceToHost, st));
        VIO_CUDA(ctx, cudaMemcpyAsync(ho + off * 13, s->d_keep.p, off, cudaMemcpyDeviceToHost, st));
        VIO_CUDA(ctx, cudaStreamSynchronize(st));
        copy_guard.armed = false;      // the main stream waited for every upload group
        if (fe_trace) fprintf(stderr, "[vio360 fe trace] %d jobs: copies enqueued in %.0f us, kernels in %.0f us, stream idle %.0f us later\n",
                              n_jobs, t_copies - (ts_begin.tv_sec * 1e6 + ts_begin.tv_nsec * 1e-3), t_kernels - t_copies, now_us() - t_kernels);
        size_t o = 0;
        for (int b = 0; b < n_jobs; ++b) {
            const vio360_track_job& j = jobs[b];
            if (j.n) {
                memcpy(j.cur_xy, ho + o * 8, (size_t)j.n * 8);
                memcpy(j.status, ho + off * 12 + o, (size_t)j.n);
                if (j.err) memcpy(j.err, ho + off * 8 + o * 4, (size_t)j.n * 4);
                if (j.keep) memcpy(j.keep, ho + off * 13 + o, (size_t)j.n);
            }
            o += (size_t)j.n;
        }
    }
    return VIO360_OK;
}

extern "C" int vio360_fe_track(vio360_ctx* ctx, const uint8_t* prev, size_t prev_pitch, uint64_t prev_key,
                               const uint8_t* cur, size_t cur_pitch, uint64_t cur_key, const float* prev_xy, int n,
                               float* cur_xy, uint8_t* status, float* err, int* n_tracked) {
    if (!ctx) return VIO360_ERR_INVALID;
    if (n == 0) { if (n_tracked) *n_tracked = 0; return VIO360_OK; }   // FeatureTracker.cpp:233-235
    vio360_track_job j;
    j.prev = prev; j.prev_pitch = prev_pitch; j.prev_key = prev_key;
    j.cur = cur; j.cur_pitch = cur_pitch; j.cur_key = cur_key;
    j.prev_xy = prev_xy; j.n = n; j.cur_xy = cur_xy; j.status = status; j.err = err; j.keep = nullptr;
    const int rc = vio360_fe_track_batch(ctx, &j, 1, VIO360_MEM_HOST);
    if (rc) return rc;
    if (n_tracked) { int c = 0; for (int i = 0; i < n; ++i) c += status[i] != 0; *n_tracked = c; }
    return VIO360_OK;
}

extern "C" int vio360_fe_pyramid(vio360_ctx* ctx, const uint8_t* img, size_t pitch, int n_levels, uint8_t* const* out,
                                 const size_t* out_pitch) {
    if (!ctx) return VIO360_ERR_INVALID;
    vio360_fe_state* s = ctx->fe;
    if (!s || !s->configured) return vio360_fail(ctx, VIO360_ERR_STATE, "fe_pyramid before fe_configure");
    if (!img || !out || !out_pitch || n_levels < 1 || n_levels > s->nlevels || pitch < (size_t)s->W)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_pyramid: bad arguments");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    int rc = begin_stage(ctx);
    if (rc) return rc;
    s->batch_tag++;
    bool hit; int idx;
    if ((rc = acquire_slot(ctx, 0, &hit, &idx))) return rc;
    if ((rc = load_frame(ctx, idx, img, pitch, VIO360_MEM_HOST))) return rc;
    if ((rc = build_pyramids(ctx, std::vector<int>{idx}))) return rc;
    if ((rc = end_stage(ctx))) return rc;
    const FeSlot& sl = s->slots[idx];
    for (int l = 0; l < n_levels; ++l)
        VIO_CUDA(ctx, cudaMemcpy2DAsync(out[l], out_pitch[l], sl.p[l], sl.pitch[l], s->lw[l], s->lh[l], cudaMemcpyDeviceToHost, ctx->stream));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return VIO360_OK;
}

// ---- detection -----------------------------------------------------------------------
// candidate segments: max_cand keys per image, det_cand_jobs images
static int reserve_candidates(vio360_ctx* ctx, size_t per_image, int n_jobs) {
    vio360_fe_state* s = ctx->fe;
    s->max_cand = (int)std::max<size_t>(per_image, (size_t)s->max_cand);
    const size_t cap = (size_t)s->max_cand * n_jobs;
    if (cap > s->cand_cap) { VIO_CUDA(ctx, s->d_cand.reserve(cap * 8)); s->cand_cap = cap; }
    return VIO360_OK;
}

static bool detect_fused_requested() {
    const char* e = getenv("VIO360_DET_FUSED");   // read per call: the tests switch it
    return e && atoi(e) != 0;
}

static int ensure_detect_pools(vio360_ctx* ctx, int n_jobs, int max_corners, bool need_map = true) {
    vio360_fe_state* s = ctx->fe;
    const size_t WH = (size_t)s->W * s->H;
    if (need_map && n_jobs > s->eig_jobs) {   // 4*W*H per image: not taken by the fused detector
        VIO_CUDA(ctx, s->d_eig.reserve(WH * 4 * n_jobs));
        s->eig_jobs = n_jobs;
    }
    if (n_jobs > s->det_batch_cap) {
        VIO_CUDA(ctx, s->d_small.reserve(16 * (size_t)(n_jobs + 1)));
        VIO_CUDA(ctx, s->d_djobs.reserve(sizeof(DetectJobDev) * (size_t)n_jobs));
        VIO_CUDA(ctx, s->h_djobs.reserve(sizeof(DetectJobDev) * (size_t)n_jobs));
        s->det_batch_cap = n_jobs;
    }
    int rc = reserve_candidates(ctx, (size_t)s->max_cand, n_jobs);
    if (rc) return rc;
    const size_t cap = (size_t)(max_corners > 0 ? max_corners : 0);
    VIO_CUDA(ctx, s->d_corners.reserve(cap * 8 * n_jobs));
    VIO_CUDA(ctx, s->h_dout.reserve(cap * 8 * n_jobs + 16 * (size_t)(n_jobs + 1)));
    return VIO360_OK;
}

static DetectParams detect_params(const vio360_fe_state* s) {
    DetectParams P;
    P.W = s->W; P.H = s->H;
    P.top_rows = (int)(s->H * s->polar_ratio);                    // Camera.cpp:104
    P.bottom_start = (int)(s->H * (1.0f - s->polar_ratio));       // Camera.cpp:105
    P.margin = s->margin;
    P.idx_bits = 1;
    while (((size_t)1 << P.idx_bits) < (size_t)s->W * s->H) ++P.idx_bits;
    P.max_cand = (unsigned int)s->max_cand;
    return P;
}

// half-width per row offset of cv::circle(img, c, radius, color, -1): OpenCV's midpoint loop
// (imgproc/src/drawing.cpp, Circle() with fill) draws rows c.y +- dy over [c.x - dx, c.x + dx] and
// rows c.y +- dx over [c.x - dy, c.x + dy]
static DiscTable disc_table(int radius) {
    DiscTable T;
    T.radius = radius;
    for (int i = 0; i <= kMaxDiscRadius; ++i) T.hw[i] = -1;
    int err = 0, dx = radius, dy = 0, plus = 1, minus = (radius << 1) - 1;
    while (dx >= dy) {
        T.hw[dy] = (signed char)std::max<int>(T.hw[dy], dx);
        T.hw[dx] = (signed char)std::max<int>(T.hw[dx], dy);
        dy++;
        err += plus;
        plus += 2;
        const int mask = (err <= 0) - 1;
        err -= minus & mask;
        dx += mask;
        minus -= mask & 2;
    }
    return T;
}

static void launch_min_eig(vio360_ctx* ctx, const DetectJobDev* dj, const DetectParams& P, int n_jobs, bool any_mask, cudaStream_t on = nullptr) {
    const int nwx = (P.W + kEigWarpCols - 1) / kEigWarpCols;
    const cudaStream_t st = on ? on : ctx->stream;
    KernelTimer kt_(ctx, "k_min_eig");
    // few frames: the march is latency bound (rows x one row's latency) and one column per lane halves that latency;
    // a batch that fills the 14 warps per SM of the 2-column form twice over issues fewer instructions with it
    const char* env_narrow = getenv("VIO360_EIG_NARROW");   // read per call: the tests switch it
    const int force_narrow = env_narrow ? atoi(env_narrow) : -1;
    const bool narrow = force_narrow >= 0 ? force_narrow != 0 : (long long)nwx * n_jobs < (long long)ctx->num_sms * 14;
    if (narrow) {
        const int nw1 = (P.W + kEigWarpCols1 - 1) / kEigWarpCols1;
        const dim3 grid1((nw1 + kEigWarps - 1) / kEigWarps, n_jobs);
        if (any_mask) k_min_eig_narrow<true><<<grid1, kEigWarps * 32, 0, st>>>(dj, P, nw1);
        else k_min_eig_narrow<false><<<grid1, kEigWarps * 32, 0, st>>>(dj, P, nw1);
        return;
    }
    const dim3 grid((nwx + kEigWarps - 1) / kEigWarps, n_jobs);
    if (any_mask) k_min_eig<true><<<grid, kEigWarps * 32, 0, st>>>(dj, P, nwx);
    else k_min_eig<false><<<grid, kEigWarps * 32, 0, st>>>(dj, P, nwx);
}

// K4 + K5 for a batch whose job table is already on the device: response map, candidates, lazy sort + greedy filter.
// one group of images: response map, candidates, lazy sort + greedy filter, on stream `st`
static int launch_detect_group(vio360_ctx* ctx, const DetectJobDev* dj, const DetectParams& P, int n_jobs, bool any_mask, int max_corners,
                               double quality, double min_dist, bool smem_grid, size_t grid_bytes, float* pool_xy, cudaStream_t st) {
    vio360_fe_state* s = ctx->fe;
    // Fused form (VIO360_DET_FUSED=1): no response map.  The march emits the candidates of its strip's interior columns
    // itself, k_seam_candidates those of the strip seams, k_sort_select applies the final threshold.  Bit exact like the
    // default, a seventh of its DRAM traffic and no 4*W*H buffer per image -- and 10 % slower on 64 frames of 3840x1920
    // (3.34 against 3.02 ms): the march is bound by instruction issue and the test costs it 58 instructions per strip
    // row (1.94 -> 2.68 ms), more than the k_candidates4 it replaces (0.54 -> 0.16 ms).  Kept as an option for callers
    // that are short of memory; default off.
    const int nwx60 = (P.W + kEigWarpCols - 1) / kEigWarpCols;
    const bool fused = detect_fused_requested();
    if (fused) {
        const size_t edge_elems = (size_t)n_jobs * nwx60 * P.H;
        if (s->d_edge.reserve(edge_elems * sizeof(float4)) != cudaSuccess) return vio360_fail(ctx, VIO360_ERR_CUDA, "fe_detect: seam buffer allocation failed");
        float4* edge = s->d_edge.as<float4>() + (size_t)(dj - s->d_djobs.as<DetectJobDev>()) * nwx60 * P.H;   // a group's own part
        const dim3 grid((nwx60 + kEigWarps - 1) / kEigWarps, n_jobs);
        { KernelTimer kt_(ctx, "k_min_eig");
        if (any_mask) k_min_eig_cand<true><<<grid, kEigWarps * 32, 0, st>>>(dj, P, nwx60, quality, edge);
        else k_min_eig_cand<false><<<grid, kEigWarps * 32, 0, st>>>(dj, P, nwx60, quality, edge);
        }
        VIO_LAUNCH_CHECK(ctx);
        if (nwx60 > 1) {
            KernelTimer kt_(ctx, "k_candidates");
            k_seam_candidates<<<dim3((P.H + 127) / 128, nwx60 - 1, n_jobs), 128, 0, st>>>(dj, P, nwx60, quality, edge);
        }
        VIO_LAUNCH_CHECK(ctx);
    } else {
    launch_min_eig(ctx, dj, P, n_jobs, any_mask, st);
    VIO_LAUNCH_CHECK(ctx);
    { KernelTimer kt_(ctx, "k_candidates");
    if ((s->W & 3) == 0)
        k_candidates4<<<dim3((s->W / 4 + kCand4Threads - 1) / kCand4Threads, (s->H - 2 + kCand4TY - 1) / kCand4TY, n_jobs), kCand4Threads, 0, st>>>(dj, P, quality);
    else
        k_candidates<<<dim3((s->W - 2 + kCandTX - 1) / kCandTX, (s->H - 2 + kCandTY - 1) / kCandTY, n_jobs), kCandTX, 0, st>>>(dj, P, quality);
    }
    VIO_LAUNCH_CHECK(ctx);
    }
    // shared memory: the cell grid (when it fits) + one tranche of keys, a power of two
    const size_t budget = (size_t)200 * 1024 - (smem_grid ? grid_bytes : 0);
    int tranche = kSelMaxTranche;
    while (tranche > 1024 && (size_t)tranche * 8 > budget) tranche >>= 1;
    const size_t smem = (smem_grid ? grid_bytes : 0) + (size_t)tranche * 8;
    { KernelTimer kt_(ctx, "k_sort_select");
    if (smem_grid) {
        VIO_CUDA(ctx, cudaFuncSetAttribute(k_sort_select<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        k_sort_select<true><<<n_jobs, kSelThreads, smem, st>>>(dj, P, max_corners, min_dist, tranche, pool_xy, fused ? quality : 0.0);
    } else {
        VIO_CUDA(ctx, cudaFuncSetAttribute(k_sort_select<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        k_sort_select<false><<<n_jobs, kSelThreads, smem, st>>>(dj, P, max_corners, min_dist, tranche, pool_xy, fused ? quality : 0.0);
    }
    }
    VIO_LAUNCH_CHECK(ctx);
    return VIO360_OK;
}

// K4 + K5 for a batch whose job table is already on the device.  The three kernels want different things -- k_min_eig
// is bound by instruction issue, k_candidates by HBM, k_sort_select is one latency-bound CTA per image -- so cutting the
// batch into groups that alternate between two streams (a group's candidates / sort beside the next group's response
// maps) was tried: 64 frames of 3840x1920 take 3.55 ms in one group, 3.60 in 2, 4.06 in 4, 5.56 in 8 -- the smaller
// k_min_eig grids lose more than the overlap gains.  VIO360_DET_GROUPS=n keeps the experiment reachable; default 1.
static int launch_detect_pipeline(vio360_ctx* ctx, const DetectJobDev* dj, const DetectParams& P, int n_jobs, bool any_mask, int max_corners,
                                  double quality, double min_dist, bool smem_grid, size_t grid_bytes) {
    vio360_fe_state* s = ctx->fe;
    cudaStream_t st = ctx->stream;
    float* pool_xy = s->d_corners.as<float>();
    static const int want_groups = getenv("VIO360_DET_GROUPS") ? std::max(1, atoi(getenv("VIO360_DET_GROUPS"))) : 1;
    const int groups = (ctx->timing || n_jobs < 8) ? 1 : std::min(want_groups, n_jobs);
    if (groups == 1) return launch_detect_group(ctx, dj, P, n_jobs, any_mask, max_corners, quality, min_dist, smem_grid, grid_bytes, pool_xy, st);
    VIO_CUDA(ctx, cudaEventRecord(s->det_ev[0], st));
    VIO_CUDA(ctx, cudaStreamWaitEvent(s->pyr_stream, s->det_ev[0], 0));
    for (int g = 0; g < groups; ++g) {
        const int j0 = (int)((long long)n_jobs * g / groups), j1 = (int)((long long)n_jobs * (g + 1) / groups);
        if (j1 <= j0) continue;
        int rc = launch_detect_group(ctx, dj + j0, P, j1 - j0, any_mask, max_corners, quality, min_dist, smem_grid, grid_bytes,
                                     pool_xy + 2 * (size_t)P.max_cand * j0, (g & 1) ? s->pyr_stream : st);
        if (rc) return rc;
    }
    VIO_CUDA(ctx, cudaEventRecord(s->det_ev[1], s->pyr_stream));
    VIO_CUDA(ctx, cudaStreamWaitEvent(st, s->det_ev[1], 0));
    return VIO360_OK;
}

extern "C" int vio360_fe_detect_batch(vio360_ctx* ctx, const vio360_detect_job* jobs, int n_jobs, int max_corners,
                                      double quality, double min_dist, int mem_space) {
    if (!ctx) return VIO360_ERR_INVALID;
    vio360_fe_state* s = ctx->fe;
    if (!s || !s->configured) return vio360_fail(ctx, VIO360_ERR_STATE, "fe_detect before fe_configure");
    if (!jobs || n_jobs < 0 || n_jobs > s->max_batch) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_detect_batch: bad job count");
    if (n_jobs == 0) return VIO360_OK;
    const size_t WH = (size_t)s->W * s->H;
    if (s->W < 3 || s->H < 3 || s->W > 65535 || s->H > 65535) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_detect_batch: image size out of range");
    if (max_corners > 0 && (size_t)max_corners > WH) max_corners = (int)WH;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    int rc = ensure_detect_pools(ctx, n_jobs, max_corners, !detect_fused_requested());
    if (rc) return rc;
    DetectParams P = detect_params(s);
    cudaStream_t st = ctx->stream;
    const size_t cap = (size_t)(max_corners > 0 ? max_corners : 0);
    const int cell = min_dist >= 1.0 ? (int)lrint(min_dist) : 1;
    const int gw = (s->W + cell - 1) / cell, gh = (s->H + cell - 1) / cell;
    const size_t grid_bytes = min_dist >= 1.0 ? (size_t)gw * gh * sizeof(uint4) : 0;
    const bool smem_grid = grid_bytes + 8 * 1024 <= (size_t)200 * 1024;   // the cell grid and at least a 1024-key tranche
    if (!smem_grid) {
        VIO_CUDA(ctx, s->d_grid.reserve(grid_bytes * n_jobs));
        VIO_CUDA(ctx, cudaMemsetAsync(s->d_grid.p, 0, grid_bytes * n_jobs, st));
    }
    s->batch_tag++;
    bool any_mask = false, any_occ = false;
    int occ_radius = -1, max_occ = 0;
    size_t occ_floats = 0;
    for (int b = 0; b < n_jobs; ++b) {
        const bool occ = jobs[b].occupied_xy && jobs[b].n_occupied > 0;
        any_mask = any_mask || jobs[b].mask || occ;
        if (!occ) continue;
        if (jobs[b].occupied_radius < 0 || jobs[b].occupied_radius > kMaxDiscRadius || (occ_radius >= 0 && occ_radius != jobs[b].occupied_radius))
            return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_detect_batch: occupied_radius must be one value in [0, %d] per batch", kMaxDiscRadius);
        occ_radius = jobs[b].occupied_radius;
        any_occ = true;
        max_occ = std::max(max_occ, jobs[b].n_occupied);
        occ_floats += 2 * (size_t)jobs[b].n_occupied;
    }
    // a device-side mask per image is needed when the mask comes from the host or discs are drawn into it
    if (any_mask && (mem_space == VIO360_MEM_HOST || any_occ)) VIO_CUDA(ctx, s->d_mask.reserve(WH * n_jobs));
    if (any_occ) {
        VIO_CUDA(ctx, s->d_occ.reserve(occ_floats * 4 + (size_t)n_jobs * 16));
        VIO_CUDA(ctx, s->h_occ.reserve(occ_floats * 4 + (size_t)n_jobs * 16));
    }
    // [pointers n_jobs x 8 | counts n_jobs x 4 | pad | points]
    const float** h_occ_ptr = s->h_occ.as<const float*>();
    int* h_occ_n = any_occ ? reinterpret_cast<int*>(h_occ_ptr + n_jobs) : nullptr;
    float* h_occ_xy = any_occ ? reinterpret_cast<float*>(s->h_occ.as<uint8_t>() + (size_t)n_jobs * 16) : nullptr;
    float* d_occ_xy = any_occ ? reinterpret_cast<float*>(s->d_occ.as<uint8_t>() + (size_t)n_jobs * 16) : nullptr;
    DetectJobDev* hj = s->h_djobs.as<DetectJobDev>();
    for (int b = 0; b < n_jobs; ++b) {
        const vio360_detect_job& j = jobs[b];
        if (!j.img || !j.out_xy || !j.out_n || j.pitch < (size_t)s->W) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_detect_batch: job %d invalid", b);
        bool hit; int idx;
        if ((rc = acquire_slot(ctx, j.key, &hit, &idx))) return rc;
        if (!hit && (rc = load_frame(ctx, idx, j.img, j.pitch, mem_space))) return rc;
        DetectJobDev& d = hj[b];
        d.img = s->slots[idx].p[0]; d.pitch = s->slots[idx].pitch[0];
        d.mask = nullptr; d.mask_pitch = 0;
        const bool occ = j.occupied_xy && j.n_occupied > 0;
        if (j.mask || occ) {
            if (mem_space == VIO360_MEM_DEVICE && !occ) { d.mask = j.mask; d.mask_pitch = (int)j.mask_pitch; }
            else {   // private copy: uploaded, or the caller's device mask / all-valid with discs drawn into it
                uint8_t* dm = s->d_mask.as<uint8_t>() + WH * b;
                if (j.mask) VIO_CUDA(ctx, cudaMemcpy2DAsync(dm, s->W, j.mask, j.mask_pitch, s->W, s->H,
                                                           mem_space == VIO360_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
                else VIO_CUDA(ctx, cudaMemsetAsync(dm, 255, WH, st));
                d.mask = dm; d.mask_pitch = s->W;
            }
        }
        if (any_occ) {
            h_occ_n[b] = occ ? j.n_occupied : 0;
            h_occ_ptr[b] = nullptr;
            if (occ && mem_space == VIO360_MEM_DEVICE) h_occ_ptr[b] = j.occupied_xy;
            else if (occ) {
                memcpy(h_occ_xy, j.occupied_xy, (size_t)j.n_occupied * 8);
                h_occ_ptr[b] = d_occ_xy;
                h_occ_xy += 2 * (size_t)j.n_occupied; d_occ_xy += 2 * (size_t)j.n_occupied;
            }
        }
        d.eig = s->d_eig.as<float>() + WH * b;
        unsigned int* small = s->d_small.as<unsigned int>() + 4 * b;
        d.max_bits = small; d.n_cand = small + 1; d.out_n = (int*)(small + 2);
        d.max_corners_dev = nullptr;
        d.grid = smem_grid ? nullptr : (uint4*)(s->d_grid.as<uint8_t>() + grid_bytes * b);
        d.out_xy = mem_space == VIO360_MEM_DEVICE ? j.out_xy : (max_corners > 0 ? s->d_corners.as<float>() + cap * 2 * b : nullptr);
    }
    const DetectJobDev* dj = s->d_djobs.as<DetectJobDev>();
    if (max_corners <= 0 && mem_space == VIO360_MEM_HOST) VIO_CUDA(ctx, s->d_corners.reserve((size_t)s->max_cand * 8 * n_jobs));
    if (any_occ) {
        VIO_CUDA(ctx, cudaMemcpyAsync(s->d_occ.p, s->h_occ.p, (size_t)n_jobs * 16 + (mem_space == VIO360_MEM_DEVICE ? 0 : occ_floats * 4), cudaMemcpyHostToDevice, st));
    }
    bool discs_drawn = false;
    // One pass = response map, candidates into per-image segments, per-image lazy sort + greedy filter: no
    // library kernel and no read-back in between.  The pass is repeated only when an image had more candidates
    // than its segment holds (seen in the counts that come back with the results; nothing is ever dropped).
    for (;;) {
        for (int b = 0; b < n_jobs; ++b) hj[b].cand = s->d_cand.as<unsigned long long>() + (size_t)s->max_cand * b;
        VIO_CUDA(ctx, cudaMemcpyAsync(s->d_djobs.p, hj, sizeof(DetectJobDev) * (size_t)n_jobs, cudaMemcpyHostToDevice, st));
        VIO_CUDA(ctx, cudaMemsetAsync(s->d_small.p, 0, 16 * (size_t)(n_jobs + 1), st));
        if (!smem_grid) VIO_CUDA(ctx, cudaMemsetAsync(s->d_grid.p, 0, grid_bytes * n_jobs, st));
        if (any_occ && !discs_drawn) {
            const DiscTable T = disc_table(occ_radius);
            { KernelTimer kt_(ctx, "k_occupancy_discs");
            k_occupancy_discs<<<dim3((max_occ + kDiscWarps - 1) / kDiscWarps, n_jobs), kDiscWarps * 32, 0, st>>>(dj, s->d_occ.as<const float*>(), reinterpret_cast<const int*>(s->d_occ.as<const float*>() + n_jobs), s->W, s->H, T);
            }
            VIO_LAUNCH_CHECK(ctx);
            discs_drawn = true;
        }
        P.max_cand = (unsigned int)s->max_cand;
        if ((rc = launch_detect_pipeline(ctx, dj, P, n_jobs, any_mask, max_corners, quality, min_dist, smem_grid, grid_bytes))) return rc;
        uint8_t* ho = s->h_dout.as<uint8_t>();
        const size_t small_bytes = 16 * (size_t)(n_jobs + 1);
        VIO_CUDA(ctx, cudaMemcpyAsync(ho, s->d_small.p, small_bytes, cudaMemcpyDeviceToHost, st));
        if (mem_space == VIO360_MEM_HOST && max_corners > 0)
            VIO_CUDA(ctx, cudaMemcpyAsync(ho + small_bytes, s->d_corners.p, cap * 8 * n_jobs, cudaMemcpyDeviceToHost, st));
        VIO_CUDA(ctx, cudaStreamSynchronize(st));
        unsigned int worst = 0;
        for (int b = 0; b < n_jobs; ++b) worst = std::max(worst, ((const unsigned int*)ho)[4 * b + 1]);
        if (worst <= (unsigned int)s->max_cand) break;
        if ((rc = reserve_candidates(ctx, (size_t)worst + worst / 4, n_jobs))) return rc;
        if (max_corners <= 0 && mem_space == VIO360_MEM_HOST) VIO_CUDA(ctx, s->d_corners.reserve((size_t)s->max_cand * 8 * n_jobs));
    }
    {
        const uint8_t* ho = s->h_dout.as<uint8_t>();
        const size_t small_bytes = 16 * (size_t)(n_jobs + 1);
        for (int b = 0; b < n_jobs; ++b) {
            const int n = ((const int*)ho)[4 * b + 2];
            *jobs[b].out_n = n;
            if (mem_space != VIO360_MEM_HOST || n <= 0) continue;
            if (max_corners > 0) memcpy(jobs[b].out_xy, ho + small_bytes + cap * 8 * b, (size_t)n * 8);
            else VIO_CUDA(ctx, cudaMemcpyAsync(jobs[b].out_xy, s->d_corners.as<float>() + 2 * (size_t)s->max_cand * b, (size_t)n * 8, cudaMemcpyDeviceToHost, st));
        }
        if (mem_space == VIO360_MEM_HOST && max_corners <= 0) VIO_CUDA(ctx, cudaStreamSynchronize(st));
    }
    return VIO360_OK;
}

extern "C" int vio360_fe_detect(vio360_ctx* ctx, const uint8_t* img, size_t pitch, uint64_t key, const uint8_t* mask,
                                size_t mask_pitch, int max_corners, double quality, double min_dist, float* out_xy,
                                int* out_n) {
    vio360_detect_job j;
    j.img = img; j.pitch = pitch; j.key = key; j.mask = mask; j.mask_pitch = mask_pitch; j.out_xy = out_xy; j.out_n = out_n;
    j.occupied_xy = nullptr; j.n_occupied = 0; j.occupied_radius = 0;
    return vio360_fe_detect_batch(ctx, &j, 1, max_corners, quality, min_dist, VIO360_MEM_HOST);
}

extern "C" int vio360_fe_min_eigen_val(vio360_ctx* ctx, const uint8_t* img, size_t pitch, float* out) {
    if (!ctx) return VIO360_ERR_INVALID;
    vio360_fe_state* s = ctx->fe;
    if (!s || !s->configured) return vio360_fail(ctx, VIO360_ERR_STATE, "fe_min_eigen_val before fe_configure");
    if (!img || !out || pitch < (size_t)s->W) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_min_eigen_val: bad arguments");
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    VIO_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    int rc = ensure_detect_pools(ctx, 1, 1);
    if (rc) return rc;
    s->batch_tag++;
    bool hit; int idx;
    if ((rc = acquire_slot(ctx, 0, &hit, &idx))) return rc;
    if ((rc = load_frame(ctx, idx, img, pitch, VIO360_MEM_HOST))) return rc;
    cudaStream_t st = ctx->stream;
    VIO_CUDA(ctx, cudaMemsetAsync(s->d_small.p, 0, 16, st));
    DetectJobDev* hj = s->h_djobs.as<DetectJobDev>();
    memset(hj, 0, sizeof(DetectJobDev));
    hj->img = s->slots[idx].p[0]; hj->pitch = s->slots[idx].pitch[0];
    hj->eig = s->d_eig.as<float>();
    hj->max_bits = s->d_small.as<unsigned int>(); hj->n_cand = hj->max_bits + 1; hj->out_n = (int*)(hj->max_bits + 2);
    VIO_CUDA(ctx, cudaMemcpyAsync(s->d_djobs.p, hj, sizeof(DetectJobDev), cudaMemcpyHostToDevice, st));
    launch_min_eig(ctx, s->d_djobs.as<DetectJobDev>(), detect_params(s), 1, false);
    VIO_LAUNCH_CHECK(ctx);
    VIO_CUDA(ctx, cudaMemcpyAsync(out, s->d_eig.p, (size_t)s->W * s->H * 4, cudaMemcpyDeviceToHost, st));
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    return VIO360_OK;
}

// ---- rotation-only RANSAC -------------------------------------------
extern "C" int vio360_fe_ransac_rotation_batch(vio360_ctx* ctx, const vio360_ransac_job* jobs, int n_jobs, int iters,
                                               float threshold_deg, int mem_space) {
    if (!ctx) return VIO360_ERR_INVALID;
    vio360_fe_state* s = ctx->fe;
    if (!s || !s->configured) return vio360_fail(ctx, VIO360_ERR_STATE, "fe_ransac before fe_configure");
    if (!jobs || n_jobs < 0 || iters < 0) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_ransac: bad arguments");
    if (n_jobs == 0) return VIO360_OK;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    // pool layout per job (4-byte units): [prev 2n | cur 2n | samples 3I | b1 3n | b2 3n | R 9I | counts I | best 2 | mask n bytes]
    auto a4 = [](size_t v) { return (v + 3) / 4 * 4; };
    size_t words = 0, hwords = 0;
    int max_n = 0;
    for (int b = 0; b < n_jobs; ++b) {
        const vio360_ransac_job& j = jobs[b];
        if (j.n < 0 || (j.n > 0 && (!j.prev_xy || !j.cur_xy || !j.inlier_mask)) || !j.best_inliers || (j.n >= 3 && iters > 0 && !j.samples))
            return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_ransac: job %d invalid", b);
        const size_t n = (size_t)j.n, I = (size_t)iters;
        words += 4 * n + 3 * I + 6 * n + 9 * I + I + 4 + a4(n) / 4 + 8;
        hwords += 4 * n + 3 * I + I + 4 + a4(n) / 4 + 8;
        max_n = std::max(max_n, j.n);
    }
    VIO_CUDA(ctx, s->d_ransac.reserve(words * 4 + sizeof(RansacJobDev) * (size_t)n_jobs));
    VIO_CUDA(ctx, s->h_ransac.reserve(hwords * 4 + sizeof(RansacJobDev) * (size_t)n_jobs));
    RansacJobDev* hj = s->h_ransac.as<RansacJobDev>();
    RansacJobDev* dj = s->d_ransac.as<RansacJobDev>();
    uint32_t* dw = reinterpret_cast<uint32_t*>(dj + n_jobs);
    uint32_t* hw = reinterpret_cast<uint32_t*>(hj + n_jobs);
    struct Back { uint32_t* h_counts; uint32_t* h_best; uint8_t* h_mask; };
    std::vector<Back> back(n_jobs);
    for (int b = 0; b < n_jobs; ++b) {
        const vio360_ransac_job& j = jobs[b];
        const size_t n = (size_t)j.n, I = (size_t)iters;
        RansacJobDev& d = hj[b];
        d.n = j.n;
        if (mem_space == VIO360_MEM_DEVICE) {
            d.prev_xy = j.prev_xy; d.cur_xy = j.cur_xy; d.samples = j.samples; d.mask = j.inlier_mask;
        } else {
            memcpy(hw, j.prev_xy, n * 8); VIO_CUDA(ctx, cudaMemcpyAsync(dw, hw, n * 8, cudaMemcpyHostToDevice, st));
            d.prev_xy = (const float*)dw; dw += 2 * n; hw += 2 * n;
            memcpy(hw, j.cur_xy, n * 8); VIO_CUDA(ctx, cudaMemcpyAsync(dw, hw, n * 8, cudaMemcpyHostToDevice, st));
            d.cur_xy = (const float*)dw; dw += 2 * n; hw += 2 * n;
            if (j.n >= 3 && I) { memcpy(hw, j.samples, I * 12); VIO_CUDA(ctx, cudaMemcpyAsync(dw, hw, I * 12, cudaMemcpyHostToDevice, st)); }
            d.samples = (const int*)dw; dw += 3 * I; hw += 3 * I;
        }
        d.b1 = (float*)dw; dw += 3 * n;
        d.b2 = (float*)dw; dw += 3 * n;
        d.R = (float*)dw; dw += 9 * I;
        d.counts = (int*)dw; dw += I;
        d.best = (int*)dw; dw += 4;
        back[b].h_counts = hw; hw += I;
        back[b].h_best = hw; hw += 4;
        if (mem_space != VIO360_MEM_DEVICE) { d.mask = (uint8_t*)dw; dw += a4(n) / 4; }
        back[b].h_mask = (uint8_t*)hw; hw += a4(n) / 4;
        // validate the sample list here: an index outside [0, n) would read outside the bearings
        if (mem_space != VIO360_MEM_DEVICE && j.n >= 3)
            for (size_t q = 0; q < 3 * I; ++q)
                if (j.samples[q] < 0 || j.samples[q] >= j.n) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_ransac: job %d sample %zu out of range", b, q);
    }
    VIO_CUDA(ctx, cudaMemcpyAsync(dj, hj, sizeof(RansacJobDev) * (size_t)n_jobs, cudaMemcpyHostToDevice, st));
    const float thr = (float)((double)threshold_deg * 3.14159265358979323846 / 180.0);   // FeatureTracker.cpp:305
    if (max_n > 0) {
        { KernelTimer kt_(ctx, "k_ransac_bearings");
        k_ransac_bearings<<<dim3((max_n + 255) / 256, n_jobs), 256, 0, st>>>(dj, s->W, s->H);
        }
        VIO_LAUNCH_CHECK(ctx);
        if (iters > 0) {
            KernelTimer kt_(ctx, "k_ransac_hypotheses");
            k_ransac_hypotheses<<<dim3((iters + kRansacWarps - 1) / kRansacWarps, n_jobs), kRansacWarps * 32, 0, st>>>(dj, iters, thr);
        }
        VIO_LAUNCH_CHECK(ctx);
    }
    { KernelTimer kt_(ctx, "k_ransac_pick");
    k_ransac_pick<<<n_jobs, 256, 0, st>>>(dj, iters, thr);
    }
    VIO_LAUNCH_CHECK(ctx);
    for (int b = 0; b < n_jobs; ++b) {
        const vio360_ransac_job& j = jobs[b];
        VIO_CUDA(ctx, cudaMemcpyAsync(back[b].h_best, hj[b].best, 8, cudaMemcpyDeviceToHost, st));
        if (j.counts && iters > 0 && mem_space != VIO360_MEM_DEVICE) VIO_CUDA(ctx, cudaMemcpyAsync(back[b].h_counts, hj[b].counts, (size_t)iters * 4, cudaMemcpyDeviceToHost, st));
        if (j.counts && iters > 0 && mem_space == VIO360_MEM_DEVICE) VIO_CUDA(ctx, cudaMemcpyAsync(j.counts, hj[b].counts, (size_t)iters * 4, cudaMemcpyDeviceToDevice, st));
        if (mem_space != VIO360_MEM_DEVICE && j.n > 0) VIO_CUDA(ctx, cudaMemcpyAsync(back[b].h_mask, hj[b].mask, (size_t)j.n, cudaMemcpyDeviceToHost, st));
    }
    VIO_CUDA(ctx, cudaStreamSynchronize(st));
    for (int b = 0; b < n_jobs; ++b) {
        const vio360_ransac_job& j = jobs[b];
        *j.best_inliers = (int)back[b].h_best[0];
        if (mem_space != VIO360_MEM_DEVICE) {
            if (j.n > 0) memcpy(j.inlier_mask, back[b].h_mask, (size_t)j.n);
            if (j.counts && iters > 0) memcpy(j.counts, back[b].h_counts, (size_t)iters * 4);
        }
    }
    return VIO360_OK;
}

extern "C" int vio360_fe_ransac_rotation(vio360_ctx* ctx, const float* prev_xy, const float* cur_xy, int n, const int32_t* samples,
                                         int iters, float threshold_deg, uint8_t* inlier_mask, int* best_inliers) {
    vio360_ransac_job j;
    j.prev_xy = prev_xy; j.cur_xy = cur_xy; j.n = n; j.samples = samples; j.inlier_mask = inlier_mask; j.best_inliers = best_inliers;
    j.counts = nullptr;
    return vio360_fe_ransac_rotation_batch(ctx, &j, 1, iters, threshold_deg, VIO360_MEM_HOST);
}

// ---- rotation RANSAC without any allocation (no counterpart in the reference) -----------------------
// vio360_fe_ransac_rotation_noalloc: the same three kernels as vio360_fe_ransac_rotation_batch on ONE job whose
// inputs and outputs are the caller's device arrays, with every intermediate (job descriptor, bearings,
// hypothesis rotations, inlier counts) in the file-scope array below instead of the context's growable pools.
// Added for callers that may not call cudaMalloc (code under CUDA graph
// capture, serving loops, harnesses that forbid it).  The array costs kRansacStaticWords * 4 bytes (about 137 KB)
// of device memory on every GPU that loads the library, until the process exits.
//
// g_ransac_static names the array for DEVICE code.  In host code the same name is only a handle: the device
// address is asked for once per device (cudaGetSymbolAddress with that device current) and that address is
// what the kernels receive.
//
// Nothing in the scratch has to start at zero: the set-up kernel rewrites the descriptor, k_ransac_bearings
// rewrites bearings [0, n), k_ransac_hypotheses rewrites R and counts [0, iters), and k_ransac_pick reads only
// those ranges; whatever an earlier call left beyond them is never read.  Every call on a device shares the
// array: two calls that overlap on the GPU (different streams / contexts of one device) race.
constexpr int kRansacStaticMaxPoints = VIO360_RANSAC_NOALLOC_MAX_POINTS;
constexpr int kRansacStaticMaxIters = VIO360_RANSAC_NOALLOC_MAX_ITERS;
constexpr int kRansacStaticJobWords = 64;    // descriptor, padded to 256 bytes
constexpr size_t kRansacStaticWords = kRansacStaticJobWords + 6 * (size_t)kRansacStaticMaxPoints + 10 * (size_t)kRansacStaticMaxIters;
static_assert(sizeof(RansacJobDev) <= kRansacStaticJobWords * 4, "descriptor slot");
static __device__ __align__(256) uint32_t g_ransac_static[kRansacStaticWords];

constexpr int kRansacStaticMaxDevices = 64;
static std::mutex g_ransac_static_mutex;
static uint32_t* g_ransac_static_dev[kRansacStaticMaxDevices] = {};   // device address of g_ransac_static, per device ordinal

// Fills the descriptor inside the scratch (layout in words: job 64 | b1 3N | b2 3N | R 9I | counts I, N / I = the maxima).
__global__ void k_ransac_static_setup(uint32_t* __restrict__ scratch, const float* prev_xy, const float* cur_xy, int n, const int* samples,
                                      uint8_t* mask, int* best) {
    RansacJobDev jb;
    jb.prev_xy = prev_xy; jb.cur_xy = cur_xy; jb.n = n; jb.samples = samples;
    uint32_t* w = scratch + kRansacStaticJobWords;
    jb.b1 = reinterpret_cast<float*>(w); w += 3 * kRansacStaticMaxPoints;
    jb.b2 = reinterpret_cast<float*>(w); w += 3 * kRansacStaticMaxPoints;
    jb.R = reinterpret_cast<float*>(w); w += 9 * kRansacStaticMaxIters;
    jb.counts = reinterpret_cast<int*>(w);
    jb.mask = mask; jb.best = best;
    *reinterpret_cast<RansacJobDev*>(scratch) = jb;
}

extern "C" int vio360_fe_ransac_rotation_noalloc(vio360_ctx* ctx, const float* prev_xy, const float* cur_xy, int n, const int32_t* samples,
                                                 int iters, float threshold_deg, uint8_t* inlier_mask, int32_t* best) {
    if (!ctx) return VIO360_ERR_INVALID;
    vio360_fe_state* s = ctx->fe;
    if (!s || !s->configured) return vio360_fail(ctx, VIO360_ERR_STATE, "fe_ransac_rotation_noalloc before fe_configure");
    if (n < 0 || iters < 0 || !best || (n > 0 && (!prev_xy || !cur_xy || !inlier_mask)) || (n >= 3 && iters > 0 && !samples))
        return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_ransac_rotation_noalloc: bad arguments");
    if (n > kRansacStaticMaxPoints || iters > kRansacStaticMaxIters)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_ransac_rotation_noalloc: n = %d / iters = %d exceed the static scratch (at most %d points, %d hypotheses); "
                           "use vio360_fe_ransac_rotation_batch for larger problems", n, iters, kRansacStaticMaxPoints, kRansacStaticMaxIters);
    if (ctx->device < 0 || ctx->device >= kRansacStaticMaxDevices)
        return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_ransac_rotation_noalloc: device ordinal %d out of range", ctx->device);
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t* scratch;
    {   // once per device, with that device current
        std::lock_guard<std::mutex> lock(g_ransac_static_mutex);
        if (!g_ransac_static_dev[ctx->device]) {
            void* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_ransac_static);
            if (e != cudaSuccess || !p)
                return vio360_fail(ctx, VIO360_ERR_CUDA, "fe_ransac_rotation_noalloc: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
            g_ransac_static_dev[ctx->device] = static_cast<uint32_t*>(p);
        }
        scratch = g_ransac_static_dev[ctx->device];
    }
    cudaStream_t st = ctx->stream;
    const RansacJobDev* dj = reinterpret_cast<const RansacJobDev*>(scratch);
    const float thr = (float)((double)threshold_deg * 3.14159265358979323846 / 180.0);   // FeatureTracker.cpp:305
    k_ransac_static_setup<<<1, 1, 0, st>>>(scratch, prev_xy, cur_xy, n, samples, inlier_mask, best);
    VIO_LAUNCH_CHECK(ctx);
    if (n > 0) {
        k_ransac_bearings<<<dim3((n + 255) / 256, 1), 256, 0, st>>>(dj, s->W, s->H);
        VIO_LAUNCH_CHECK(ctx);
        if (iters > 0) {
            k_ransac_hypotheses<<<dim3((iters + kRansacWarps - 1) / kRansacWarps, 1), kRansacWarps * 32, 0, st>>>(dj, iters, thr);
            VIO_LAUNCH_CHECK(ctx);
        }
    }
    k_ransac_pick<<<1, 256, 0, st>>>(dj, iters, thr);
    VIO_LAUNCH_CHECK(ctx);
    return VIO360_OK;
}

// ---- feature-list filters between tracking and detection -------------------------
// mode 0: RemoveClusteredFeatures, mode 1: LimitFeaturesPerGrid.  Frames are rows of `stride` features.
static int grid_filter(vio360_ctx* ctx, int mode, const float* xy, const int32_t* track, const int* n, int n_frames, int stride,
                       int width, int height, int cols, int rows, float std_ratio, int max_per_cell, uint8_t* keep, int mem_space) {
    if (!ctx) return VIO360_ERR_INVALID;
    vio360_fe_state* s = ctx->fe;
    if (!s || !s->configured) return vio360_fail(ctx, VIO360_ERR_STATE, "fe grid filter before fe_configure");
    if (!xy || !n || !keep || n_frames < 0 || stride < 0 || width <= 0 || height <= 0 || cols <= 0 || rows <= 0 || (mode == 1 && (!track || max_per_cell < 0)))
        return vio360_fail(ctx, VIO360_ERR_INVALID, "fe grid filter: bad arguments");
    if ((size_t)cols * rows >= 0xffff) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe grid filter: more than 65534 cells");
    if (n_frames == 0) return VIO360_OK;
    int max_n = 0;
    for (int f = 0; f < n_frames; ++f) {
        if (n[f] < 0 || n[f] > stride || n[f] > kGridMaxPoints) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe grid filter: frame %d has %d features (stride %d, max %d)", f, n[f], stride, kGridMaxPoints);
        max_n = std::max(max_n, n[f]);
    }
    if (max_n == 0) return VIO360_OK;
    VIO_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    VIO_CUDA(ctx, cudaStreamSynchronize(st));          // the pinned pool below may still be in flight
    const size_t tot = (size_t)n_frames * stride;
    const bool host = mem_space == VIO360_MEM_HOST;
    // pool: [jobs | xy 2*tot floats | track tot ints | keep tot bytes] (the last three only for host-space calls)
    const size_t job_bytes = (sizeof(GridJobDev) * (size_t)n_frames + 15) & ~(size_t)15;
    const size_t bytes = job_bytes + (host ? tot * 8 + tot * 4 + ((tot + 15) & ~(size_t)15) : 0);
    VIO_CUDA(ctx, s->d_ransac.reserve(bytes));
    VIO_CUDA(ctx, s->h_ransac.reserve(bytes));
    uint8_t* hp = s->h_ransac.as<uint8_t>();
    uint8_t* dp = s->d_ransac.as<uint8_t>();
    GridJobDev* hj = reinterpret_cast<GridJobDev*>(hp);
    const float* dxy = xy; const int32_t* dtr = track; uint8_t* dkeep = keep;
    if (host) {
        memcpy(hp + job_bytes, xy, tot * 8);
        if (track) memcpy(hp + job_bytes + tot * 8, track, tot * 4);
        dxy = reinterpret_cast<const float*>(dp + job_bytes);
        dtr = reinterpret_cast<const int32_t*>(dp + job_bytes + tot * 8);
        dkeep = dp + job_bytes + tot * 12;
    }
    for (int f = 0; f < n_frames; ++f) {
        hj[f].xy = dxy + 2 * (size_t)f * stride; hj[f].track = dtr ? dtr + (size_t)f * stride : nullptr;
        hj[f].keep = dkeep + (size_t)f * stride; hj[f].n = n[f];
    }
    VIO_CUDA(ctx, cudaMemcpyAsync(dp, hp, host ? job_bytes + tot * 12 : job_bytes, cudaMemcpyHostToDevice, st));
    const int np = (max_n + 1) & ~1;
    if (mode == 0) {
        const size_t smem = (size_t)np * 2 + (size_t)cols * rows;
        if (smem > 200 * 1024) return vio360_fail(ctx, VIO360_ERR_INVALID, "fe_cluster_filter: grid too large");
        VIO_CUDA(ctx, cudaFuncSetAttribute(k_cluster_filter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem, 48 * 1024)));
        KernelTimer kt_(ctx, "k_cluster_filter");
        k_cluster_filter<<<n_frames, kGridThreads, smem, st>>>(reinterpret_cast<const GridJobDev*>(dp), width, height, cols, rows, std_ratio);
    } else {
        const size_t smem = (size_t)np * 2 + (size_t)max_n * 4;
        VIO_CUDA(ctx, cudaFuncSetAttribute(k_grid_cap, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(smem, 48 * 1024)));
        KernelTimer kt_(ctx, "k_grid_cap");
        k_grid_cap<<<n_frames, kGridThreads, smem, st>>>(reinterpret_cast<const GridJobDev*>(dp), width, height, cols, rows, max_per_cell);
    }
    VIO_LAUNCH_CHECK(ctx);
    if (host) {
        VIO_CUDA(ctx, cudaMemcpyAsync(hp + job_bytes + tot * 12, dp + job_bytes + tot * 12, tot, cudaMemcpyDeviceToHost, st));
        VIO_CUDA(ctx, cudaStreamSynchronize(st));
        for (int f = 0; f < n_frames; ++f) memcpy(keep + (size_t)f * stride, hp + job_bytes + tot * 12 + (size_t)f * stride, (size_t)n[f]);
    }
    return VIO360_OK;
}

extern "C" int vio360_fe_cluster_filter(vio360_ctx* ctx, const float* xy, const int* n, int n_frames, int stride, int width, int height,
                                        int grid_cols, int grid_rows, float std_ratio, uint8_t* keep, int mem_space) {
    return grid_filter(ctx, 0, xy, nullptr, n, n_frames, stride, width, height, grid_cols, grid_rows, std_ratio, 0, keep, mem_space);
}

extern "C" int vio360_fe_grid_cap(vio360_ctx* ctx, const float* xy, const int32_t* track_count, const int* n, int n_frames, int stride,
                                  int width, int height, int grid_cols, int grid_rows, int max_per_cell, uint8_t* keep, int mem_space) {
    return grid_filter(ctx, 1, xy, track_count, n, n_frames, stride, width, height, grid_cols, grid_rows, 0.f, max_per_cell, keep, mem_space);
}

#include "fe_seq_host.inc"
