/*
 * vio360.h -- C ABI of libvio360_cuda.so: the B200 (sm_100a) implementation of the two
 * data-parallel hot paths of 93won/360_visual_inertial_odometry.
 *
 *   path A  ERP feature front-end  (reference: src/processing/FeatureTracker.{h,cpp})
 *   path B  BA factor evaluation   (reference: src/optimization/Factors.{h,cpp} + the
 *           Schur block accumulation of vendored Ceres 2.0.0)
 *
 * Conventions: plain pointers and sizes only; every function returns VIO360_OK (0) or
 * a negative error code and never throws; vio360_last_error() gives the message.  One
 * context = one device + one stream; a context is not re-entrant.  Pointers are HOST
 * memory unless the parameter is documented as device memory or a VIO360_MEM_DEVICE
 * flag is passed.  There is NO CPU fallback: every entry point fails with
 * VIO360_ERR_CUDA when no sm_100-class device is usable.
 *
 * "file:line" citations are relative to the reference repository root.
 */
#ifndef VIO360_H
#define VIO360_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VIO360_OK 0
#define VIO360_ERR_INVALID (-1) /* bad argument                                  */
#define VIO360_ERR_CUDA (-2)    /* CUDA runtime error / no usable device          */
#define VIO360_ERR_STATE (-3)   /* call order violated (e.g. evaluate before set) */
#define VIO360_ERR_NOMEM (-4)

#define VIO360_MEM_HOST 0
#define VIO360_MEM_DEVICE 1

typedef struct vio360_ctx vio360_ctx;
typedef struct vio360_ba vio360_ba;

/* ------------------------------------------------------------------ context ---- */
int vio360_version(void);
int vio360_create(int device, vio360_ctx** out);
void vio360_destroy(vio360_ctx* ctx);
const char* vio360_last_error(const vio360_ctx* ctx); /* ctx may be NULL: create() errors */
int vio360_synchronize(vio360_ctx* ctx);
void* vio360_stream(vio360_ctx* ctx);                 /* the context's cudaStream_t */
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
long long vio360_launch_count(const vio360_ctx* ctx);
/* Per-kernel device timing: when enabled, every kernel launch of this context is bracketed
 * by CUDA events on the context's stream.  vio360_kernel_time() synchronises and returns
 * the accumulated duration and launch count of the kernel called `name` (e.g.
 * "k_lk_track", "k_pyr_down", "k_schur_tile").                                        */
int vio360_enable_kernel_timing(vio360_ctx* ctx, int on);
int vio360_kernel_time(vio360_ctx* ctx, const char* name, double* total_ms, long long* launches);
int vio360_reset_kernel_times(vio360_ctx* ctx);
/* measured FP64 / FP32 FMA pipe peak of the device in TFLOP/s (dependent-FMA loops on
 * all SMs), the denominator SURVEY.md section 8-D asks for.                        */
int vio360_probe_fma_peak(vio360_ctx* ctx, double* fp64_tflops, double* fp32_tflops);
/* Host frame buffers.  The reference hands cv::Mat images in pageable host memory to
 * FeatureTracker::TrackFeatures (src/processing/FeatureTracker.cpp:61; the caller's cv::imread / ring buffer,
 * app/main.cpp).  A copy from pageable memory is staged by the driver at a fraction of the PCIe rate;
 * page-locking the caller's buffer ONCE (cudaHostRegister) lets every later VIO360_MEM_HOST call copy straight
 * from it at the full rate, without the caller having to allocate pinned memory itself.  `ptr` / `bytes` need no
 * alignment: the whole pages INSIDE the range are page-locked (its first and last page may be shared with other
 * allocations and stay pageable; the library cuts its frame copies at those boundaries, the two fringes of less
 * than a page each travel as pageable copies).  Registering the same pointer again returns VIO360_OK; a range
 * that overlaps memory page-locked by someone else is an error.  Unregister (same pointer) before freeing the
 * buffer; vio360_destroy unregisters what is left.                                                       */
int vio360_host_register(vio360_ctx* ctx, const void* ptr, size_t bytes);
int vio360_host_unregister(vio360_ctx* ctx, const void* ptr);

/* ------------------------------------------------------- path A: front-end ---- */
/* Replaces the constructor state of vio_360::FeatureTracker
 * (src/processing/FeatureTracker.cpp:24-59): LK window / levels / criteria
 * (:33-35, hard-coded 21x21, maxLevel 3, 30 iterations, eps 0.01f), the polar mask
 * (src/database/Camera.cpp:100-118) and the boundary mask (:47-58), which are evaluated
 * analytically on the device instead of being materialised.  eps is the TermCriteria
 * epsilon before OpenCV squares it; min_eig_thr is calcOpticalFlowPyrLK's
 * minEigThreshold.  max_points / max_batch size the device pools.                    */
int vio360_fe_configure(vio360_ctx* ctx, int width, int height, int max_level, int win,
                        int max_iters, double eps, double min_eig_thr, float polar_ratio,
                        int boundary_margin, int max_points, int max_batch);

/* One frame pair.  `*_key` != 0 lets the context keep the frame's device pyramid and
 * re-use it when the same key shows up again (the reference rebuilds both pyramids on
 * every call, FeatureTracker.cpp:238); key 0 = never cache.                          */
typedef struct {
    const uint8_t* prev; size_t prev_pitch; uint64_t prev_key;
    const uint8_t* cur;  size_t cur_pitch;  uint64_t cur_key;
    const float* prev_xy;  /* n x 2, AoS like std::vector<cv::Point2f>              */
    int n;
    float* cur_xy;         /* n x 2 out                                             */
    uint8_t* status;       /* n out, 0/1 (cv::calcOpticalFlowPyrLK status)          */
    float* err;            /* n out or NULL                                         */
    uint8_t* keep;         /* n out or NULL: status && !polar && !near-boundary,
                              the post filter of FeatureTracker.cpp:118-126          */
} vio360_track_job;

/* Replaces FeatureTracker::TrackOpticalFlow (src/processing/FeatureTracker.cpp:228-251,
 * i.e. cv::calcOpticalFlowPyrLK(prev, cur, prevPts, nextPts, status, err, Size(21,21),
 * 3, TermCriteria(EPS|COUNT,30,0.01), 0, 0.01)).  n == 0 returns OK without touching
 * the outputs (:233-235).  *n_tracked (may be NULL) = count of status != 0 (:242-249). */
int vio360_fe_track(vio360_ctx* ctx, const uint8_t* prev, size_t prev_pitch,
                    uint64_t prev_key, const uint8_t* cur, size_t cur_pitch,
                    uint64_t cur_key, const float* prev_xy, int n, float* cur_xy,
                    uint8_t* status, float* err, int* n_tracked);
/* Batched form for independent frame pairs (BASELINE config 4: one job per sequence).
 * mem_space says where ALL pointers inside the jobs live (the job array itself is
 * host memory).  Lifetime of device-space frames: a frame whose base and pitch are
 * 4-byte aligned is NOT copied -- level 0 is read through the pointer passed to THIS
 * call (also on a keyed cache hit, which only re-uses levels 1..L-1), so the buffer must
 * stay valid and unchanged until the call's kernels have run (vio360_synchronize), and a
 * key must always name the same pixels.  Unaligned device frames and host frames are
 * copied.  All jobs are validated before the cache is touched.                        */
int vio360_fe_track_batch(vio360_ctx* ctx, const vio360_track_job* jobs, int n_jobs,
                          int mem_space);

typedef struct {
    const uint8_t* img; size_t pitch; uint64_t key;
    const uint8_t* mask; size_t mask_pitch; /* optional extra mask (the occupancy mask of
                                               FeatureTracker::CreateFeatureMask, :386-402);
                                               NULL = base mask only                  */
    float* out_xy;   /* max_corners x 2 out, integer coordinates as float           */
    int* out_n;      /* out: corners found                                          */
    /* Instead of rasterising and uploading the occupancy mask, a job may pass the positions
     * of the existing valid features (n_occupied x 2 floats, same memory space as the image)
     * and the disc radius static_cast<int>(m_min_distance), in [0, 64]: the device zeroes the
     * pixels cv::circle(mask, pt, radius, 0, -1) would (FeatureTracker.cpp:397-398; centre
     * rounded as cv::Point(pt), OpenCV's midpoint-circle fill) on top of `mask`, or of an
     * all-valid mask when mask is NULL.  NULL = no discs.                              */
    const float* occupied_xy;
    int n_occupied;
    int occupied_radius;
} vio360_detect_job;

/* Replaces FeatureTracker::DetectNewFeatures (src/processing/FeatureTracker.cpp:208-226,
 * i.e. cv::goodFeaturesToTrack(image, corners, max_features, quality, min_distance,
 * polar & boundary & mask, 3, false, 0.04)).                                         */
int vio360_fe_detect(vio360_ctx* ctx, const uint8_t* img, size_t pitch, uint64_t key,
                     const uint8_t* mask, size_t mask_pitch, int max_corners,
                     double quality, double min_dist, float* out_xy, int* out_n);
int vio360_fe_detect_batch(vio360_ctx* ctx, const vio360_detect_job* jobs, int n_jobs,
                           int max_corners, double quality, double min_dist, int mem_space);

/* Integer pyramid of one frame (cv::buildOpticalFlowPyramid levels without aprons):
 * level l is written to out[l] with pitch out_pitch[l]; sizes are ((w+1)>>1,(h+1)>>1)
 * per level.  Exposed for the bit-exact pyramid parity tests.                         */
int vio360_fe_pyramid(vio360_ctx* ctx, const uint8_t* img, size_t pitch, int n_levels,
                      uint8_t* const* out, const size_t* out_pitch);
/* cv::cornerMinEigenVal(img, dst, 3, 3) response map (float, W x H, dense pitch). */
int vio360_fe_min_eigen_val(vio360_ctx* ctx, const uint8_t* img, size_t pitch, float* out);

/* Rotation-only RANSAC outlier rejection on tracked features.
 * Replaces FeatureTracker::RejectOutliersRotationRANSAC / EstimateRotation /
 * ComputeRotationInliers (src/processing/FeatureTracker.cpp:253-379) with
 * Camera::PixelToBearing / AngularDistance (src/database/Camera.cpp:22-47, 89-98).
 * The reference draws its 3-point samples from std::random_device (:273-275); the draw stays
 * on the host and is passed in: samples is iters x 3 distinct point indices per row.
 * threshold_deg is m_ransac_threshold (degrees).  Outputs: inlier_mask[n] (0/1) and
 * *best_inliers (the function's return value): n < 3 -> all ones, n; no hypothesis with an
 * inlier -> all ones, 0.  counts (iters ints, optional) receives every hypothesis' inlier
 * count, -1 where |det R - 1| > 0.1 rejected it (:300-302).                              */
typedef struct vio360_ransac_job {
    const float* prev_xy;
    const float* cur_xy;
    int n;
    const int32_t* samples;
    uint8_t* inlier_mask;
    int* best_inliers;       /* host memory in both memory spaces */
    int32_t* counts;         /* may be NULL */
} vio360_ransac_job;
int vio360_fe_ransac_rotation(vio360_ctx* ctx, const float* prev_xy, const float* cur_xy, int n,
                              const int32_t* samples, int iters, float threshold_deg,
                              uint8_t* inlier_mask, int* best_inliers);
int vio360_fe_ransac_rotation_batch(vio360_ctx* ctx, const vio360_ransac_job* jobs, int n_jobs,
                                    int iters, float threshold_deg, int mem_space);

/* NO COUNTERPART IN THE REFERENCE: added for callers that
 * may not call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  One rotation-RANSAC job, same kernels and results as vio360_fe_ransac_rotation_batch in
 * VIO360_MEM_DEVICE, but the call allocates nothing (no cudaMalloc / cudaFree / pool growth, no host
 * staging): its intermediates live in a static device array of the library, sized for at most
 * VIO360_RANSAC_NOALLOC_MAX_POINTS points and VIO360_RANSAC_NOALLOC_MAX_ITERS hypotheses (about
 * 137 KB, held on every GPU that loads the library until the process exits).  Larger n or iters
 * -> VIO360_ERR_INVALID with a message; use the batch call for those.
 * ALL arrays are DEVICE memory of the context's device: prev_xy / cur_xy n x 2 floats, samples
 * iters x 3 indices in [0, n) (not validated, as in the device space of the batch call),
 * inlier_mask n bytes, best 2 ints: [0] the winner's inlier count (the reference function's return
 * value; n when n < 3, 0 when no hypothesis has an inlier), [1] its hypothesis index or -1.
 * The kernels are enqueued on the context's stream and the call returns without synchronising:
 * read the outputs after vio360_synchronize (or in stream order).  Requires vio360_fe_configure
 * (image size).  All calls on one device share the static array, which keeps the previous call's
 * contents (no part of it needs clearing, every range read is rewritten first): calls that overlap
 * on the GPU -- two contexts / streams of one device -- race and must be serialised by the caller. */
#define VIO360_RANSAC_NOALLOC_MAX_POINTS 4096
#define VIO360_RANSAC_NOALLOC_MAX_ITERS 1024
int vio360_fe_ransac_rotation_noalloc(vio360_ctx* ctx, const float* prev_xy, const float* cur_xy, int n,
                                      const int32_t* samples, int iters, float threshold_deg,
                                      uint8_t* inlier_mask, int32_t* best);

/* the feature-list filters FeatureTracker::TrackFeatures runs between tracking and
 * detection, on the device so that the point list can stay there.  Frames are rows of `stride`
 * features (xy: n_frames x stride x 2 floats, keep: n_frames x stride bytes, n[f] <= stride, host
 * array in both memory spaces); keep[i] = 1 for features the reference keeps, in its order.
 *  - vio360_fe_cluster_filter replaces FeatureTracker::RemoveClusteredFeatures
 *    (src/processing/FeatureTracker.cpp:404-497): grid cells (grid_cols x grid_rows over width x height,
 *    :420-436) holding >= 4 features whose positional standard deviation (float, summed in feature
 *    order, :459-476) is below std_ratio * cell diagonal lose all their features; frames with fewer
 *    than 4 features are left alone (:411).
 *  - vio360_fe_grid_cap replaces Frame::AssignFeaturesToGrid + Frame::LimitFeaturesPerGrid
 *    (src/database/Frame.cpp:108-202): features outside the image are dropped (:125-128); a cell keeps
 *    its max_per_cell features with the highest track_count; ties keep feature order (std::sort of a
 *    cell of <= 16 indices is libstdc++'s stable insertion sort; for larger cells with ties the
 *    reference's pick is implementation-defined and this call stays with feature order).          */
int vio360_fe_cluster_filter(vio360_ctx* ctx, const float* xy, const int* n, int n_frames, int stride,
                             int width, int height, int grid_cols, int grid_rows, float std_ratio,
                             uint8_t* keep, int mem_space);
int vio360_fe_grid_cap(vio360_ctx* ctx, const float* xy, const int32_t* track_count, const int* n,
                       int n_frames, int stride, int width, int height, int grid_cols, int grid_rows,
                       int max_per_cell, uint8_t* keep, int mem_space);

/* SURVEY 8-F rank 2, complete: one call per frame that replaces FeatureTracker::TrackFeatures
 * (src/processing/FeatureTracker.cpp:61-206) for a whole sequence, with the feature list RESIDENT ON THE DEVICE
 * between the stages and between the frames:
 *   TrackOpticalFlow (:110) -> post filter (:118-126) -> RejectOutliersRotationRANSAC (:131) -> a Feature per
 *   inlier with its id carried and track_count + 1 (:139-163) -> RemoveClusteredFeatures (:168) ->
 *   Frame::AssignFeaturesToGrid + LimitFeaturesPerGrid (:171-172; Frame.cpp:108-202) -> if count < max_features:
 *   DetectNewFeatures(image, CreateFeatureMask(features)) (:176-199), new ids from m_next_feature_id, grid cap again.
 * The first frame of a sequence (no previous frame, :68-92) is the same chain on an empty list.
 * Only the new frame goes in and only the final feature list comes out; no intermediate count or list
 * visits the host.  The shared_ptr<Feature> observation bookkeeping stays with the caller (ids are returned).
 *
 * The reference seeds its RANSAC from std::random_device (:273-274).  Here the caller passes
 * ransac_iterations x 3 random 32-bit words per frame; a word r selects index (r * n) >> 32 among the n
 * filtered points, and an index equal to an earlier one of its triple moves to (index + 1) % n.           */
typedef struct vio360_seq vio360_seq;
typedef struct {
    int max_features; float min_distance; float quality_level;   /* FeatureTracker ctor (Estimator.cpp:47-53)            */
    int grid_cols, grid_rows, max_features_per_grid;             /* Frame::SetGridParameters (Estimator.cpp:572-576)      */
    int remove_clustered; float clustered_std_ratio;             /* ConfigUtils visualization_highlight_clustered_grid,
                                                                    visualization_clustered_std_ratio (FeatureTracker.cpp:407-423) */
    int ransac_iterations; float ransac_threshold_deg;           /* FeatureTracker.cpp:36-37 (1000, 2.0)                  */
} vio360_seq_config;
/* vio360_fe_configure must have been called (max_points >= 2 * max_features); a later vio360_fe_configure invalidates
 * the sequence.  Limit: the list kernels keep one frame's list (positions, cells, member lists) in shared memory,
 * 21 bytes per entry of the 2 * max_features capacity + 4 bytes per grid cell <= 200 KB, i.e. about 4700 features
 * at 20 x 10 cells (the reference uses 500 .. 1000); larger values are rejected here.                          */
int vio360_seq_create(vio360_ctx* ctx, const vio360_seq_config* cfg, vio360_seq** out);
void vio360_seq_destroy(vio360_seq* seq);
int vio360_seq_reset(vio360_seq* seq);          /* forget the features and the previous frame; ids restart at 0 */
int vio360_seq_capacity(const vio360_seq* seq);  /* entries the output arrays of a frame job must hold (2 * max_features) */
typedef struct {
    vio360_seq* seq;
    const uint8_t* image; size_t pitch;   /* the new frame (CV_8UC1).  Device-space frames with 4-byte aligned base and
                                             pitch are aliased, not copied: they must stay valid and unchanged until the
                                             NEXT frame of the sequence has been processed (it is that call's previous frame) */
    const uint32_t* ransac_words;         /* ransac_iterations x 3                                                       */
    int32_t* ids; float* xy; int32_t* track_count;   /* out: the frame's features in the reference's order (capacity entries) */
    int* n_features; int* n_tracked; int* n_detected; /* out, HOST memory in both spaces: list length, m_num_tracked, m_num_detected */
} vio360_frame_job;
/* One frame of each of n_jobs independent sequences (BASELINE config 4: one job per sequence). */
int vio360_fe_track_frame_batch(vio360_ctx* ctx, const vio360_frame_job* jobs, int n_jobs, int mem_space);

/* ------------------------------------------- path B: BA factor evaluation ---- */
/* A window = what Optimizer::RunLocalBA / RunBA assemble (src/optimization/
 * Optimizer.cpp:741-840): K poses, M landmarks, O observations.                     */
int vio360_ba_create(vio360_ctx* ctx, int K, int M, int O, vio360_ba** out);
void vio360_ba_destroy(vio360_ba* ba);

/* Replaces the O heap-allocated BAFactor objects (src/optimization/Factors.h:198-203,
 * built at Optimizer.cpp:795-840) by one SoA upload.  T_wb_init is K x 16 row-major
 * (m_T_wb_init), T_cb 16 row-major (m_Tcb), info the 2x2 information matrix shared by
 * all factors (Optimizer.cpp:824), huber_delta the ceres::HuberLoss parameter
 * (Optimizer.cpp:835; <= 0 = no loss function).  pose_fixed / point_fixed mirror
 * Problem::SetParameterBlockConstant (Optimizer.cpp:854-877); NULL = none.           */
int vio360_ba_set_structure(vio360_ba* ba, const int32_t* pose_idx, const int32_t* point_idx,
                            const double* obs_u, const double* obs_v,
                            const double* T_wb_init, const double* T_cb, double cols,
                            double rows, const double* info, double huber_delta,
                            const uint8_t* pose_fixed, const uint8_t* point_fixed);
/* BAFactor::set_outlier for every factor at once (Factors.h:137; Optimizer.cpp:903-906).
 * NULL clears all flags.                                                            */
int vio360_ba_set_outliers(vio360_ba* ba, const uint8_t* outlier);
/* Multi-GPU: this instance only processes the landmarks of shard `rank` of `world`
 * (contiguous landmark ranges balanced by sum of squared degrees).  Partial reduced
 * systems of all ranks sum to the full one.  Default (0, 1).                        */
int vio360_ba_set_shard(vio360_ba* ba, int rank, int world);

/* The evaluation point: pose_delta K x 6 ([rho, phi], LieUtils.cpp:305-308) and
 * points M x 3 -- what Ceres copies into the user's parameter blocks before
 * EvaluationCallback::PrepareForEvaluation (program_evaluator.h:155-160).  Runs the
 * per-pose chain of Factors.cpp:357-364 once per pose instead of once per factor.   */
int vio360_ba_set_parameters(vio360_ba* ba, const double* pose_delta, const double* points);

/* Replaces the ParallelFor over ResidualBlock::Evaluate -> BAFactor::Evaluate
 * (program_evaluator.h:187-258; Factors.cpp:327-542) incl. HuberLoss + Corrector
 * (residual_block.cc:161-195) when apply_loss != 0.  *cost = sum of 0.5*rho(s).      */
int vio360_ba_evaluate(vio360_ba* ba, int want_jacobians, int apply_loss, double* cost);
/* The optional fp32 mode of the north star: the same per-observation math (Factors.cpp:327-542 with the
 * Huber correction) evaluated in float on the FP32 pipe from the fp64 prepared poses; results are stored
 * as double and read with vio360_ba_get_residuals_jacobians.  Parity bar 1e-4 relative (max-norm). */
int vio360_ba_evaluate_f32(vio360_ba* ba, int want_jacobians, int apply_loss, double* cost);
/* The fp32 mode for the FUSED paths: with fp32 != 0, vio360_ba_build_reduced_system, vio360_ba_normal_blocks,
 * vio360_ba_column_sqnorm (and the builds inside vio360_ba_lm_solve) evaluate the per-observation math in float on the
 * FP32 pipe and widen the rows; the sums, the per-landmark 3x3 Cholesky and the Schur products stay in double.
 * Bars (written in tests/test_gpu_ba.py): Hessian blocks 1e-4 relative; the reduced system, which amplifies the
 * Jacobian rounding by the conditioning of E'E, 1e-3.  Default 0 (double, 1e-9).                                 */
int vio360_ba_set_eval_precision(vio360_ba* ba, int fp32);
/* Copy out the last evaluate(): r O x 2, Jp O x 12 (2x6 row-major), Jl O x 6 (2x3
 * row-major), in the caller's observation order.  Any pointer may be NULL.          */
int vio360_ba_get_residuals_jacobians(vio360_ba* ba, double* r, double* Jp, double* Jl);

/* Gauss-Newton blocks of J^T J and J^T r at the current parameters (fused with the
 * factor evaluation, nothing materialised): Hpp K x 36, bp K x 6, Hll M x 9, bl M x 3,
 * Hpl O x 18 (F^T E per observation, 6x3).  Blocks of fixed parameters are zero.
 * Any pointer may be NULL.                                                          */
int vio360_ba_normal_blocks(vio360_ba* ba, int apply_loss, double* Hpp, double* bp,
                            double* Hll, double* bl, double* Hpl);

/* Number of rows of the reduced camera system: 6 * (number of free poses). */
int vio360_ba_reduced_dim(const vio360_ba* ba);
/* Replaces SchurEliminator<2,3,6>::Eliminate (thirdparty/ceres-solver/internal/ceres/
 * schur_eliminator_impl.h:179-308) fused with the factor evaluation: builds
 * S = F'F + Dp^2 - F'E (E'E + Dl^2)^-1 E'F and rhs = F'r - F'E (E'E + Dl^2)^-1 E'r at
 * the current parameters.  lm_diag (6K + 3M, pose part first) is the Levenberg-Marquardt
 * diagonal of levenberg_marquardt_strategy.cc:78-89; NULL = keep the diagonal installed by
 * vio360_ba_set_lm_diagonal (none by default).  S (n x n
 * dense row-major, full symmetric) / rhs (n) / cost may be NULL: the result then stays
 * on the device (see vio360_ba_reduced_system_dev).  With a shard set, the result is
 * the shard's partial sum.                                                           */
int vio360_ba_build_reduced_system(vio360_ba* ba, const double* lm_diag, int apply_loss,
                                   double* S, double* rhs, double* cost);
/* Installs (or, with NULL, clears) the LM diagonal used by later build calls. */
int vio360_ba_set_lm_diagonal(vio360_ba* ba, const double* lm_diag);
/* Device view of the last reduced system as ONE packed buffer [S n*n | rhs n | cost 1]
 * so that a single ncclAllReduce(sum, double) combines the shards.
 * Writable (in-place allreduce) only when a shard is set; without one treat it as
 * read-only: the library keeps the zeros outside S's block envelope across builds. */
int vio360_ba_reduced_system_dev(vio360_ba* ba, double** packed_dev, size_t* n_doubles);
/* Block-sparse exchange payload of a SHARDED build (SURVEY.md 8-E "block-sparse upper-triangular S"):
 * [n_free block rows][w block columns from the diagonal][6 x 6 row-major] | rhs n | cost 1, with
 * w = min(n_free, 24) identical on every rank (1.4 MB instead of 11.4 MB for 200 keyframes).
 * Allreduce it in place (sum, double), then call vio360_ba_expand_band: the dense view, the fetch
 * and the solver then see the summed system.  VIO360_ERR_STATE when the last build was not
 * sharded or had landmarks outside the chunk windows (then allreduce the dense buffer instead). */
int vio360_ba_reduced_band_dev(vio360_ba* ba, double** band_dev, size_t* n_doubles);
int vio360_ba_expand_band(vio360_ba* ba);
/* Peer-memory exchange of a sharded build over NVLink / NVSwitch: replaces the
 * allreduce + vio360_ba_expand_band pair by ONE kernel per rank that announces its partial band,
 * waits for the other ranks, pulls the bands covering each block straight out of their memory,
 * sums them in rank order (bit-identical on all ranks), adds the pose part of the LM diagonal and
 * writes the dense [S | rhs | cost].  No NCCL call, no staging copy.  There is no counterpart in
 * the reference (single process, Ceres' mutex-guarded cells, schur_eliminator_impl.h:211).
 *   1. after vio360_ba_set_structure + vio360_ba_set_shard(rank, world <= 8), every rank calls
 *      vio360_ba_peer_export and hands its VIO360_PEER_BLOB_BYTES blob to all ranks (any transport:
 *      MPI, torch.distributed all_gather, a socket; inside one process just an array);
 *   2. every rank calls vio360_ba_peer_attach with the world's blobs in rank order (other processes'
 *      buffers are mapped with cudaIpcOpenMemHandle, same-process ones by peer access), then the
 *      caller synchronises the ranks once (all attached before the first exchange);
 *   3. per LM iteration: vio360_ba_build_reduced_system(..., NULL, NULL, NULL) then
 *      vio360_ba_peer_exchange, on every rank, the same number of times (collective semantics);
 *      the dense view, the fetch and the solver then see the summed system.
 * VIO360_ERR_STATE from export when landmarks of the shard fit no chunk window (allreduce the dense
 * buffer instead).  A rank that never arrives makes the others fail after ~2 s (reported by the
 * next fetch / exchange), it cannot hang the device.  Ranks must stay alive until all have detached. */
#define VIO360_PEER_BLOB_BYTES 128
int vio360_ba_peer_export(vio360_ba* ba, void* blob_out);
int vio360_ba_peer_attach(vio360_ba* ba, const void* blobs);
int vio360_ba_peer_exchange(vio360_ba* ba);
int vio360_ba_peer_detach(vio360_ba* ba);
/* Copy the (possibly all-reduced) packed device buffer back to the host. */
int vio360_ba_fetch_reduced_system(vio360_ba* ba, double* S, double* rhs, double* cost);

/* diag(J^T J) per parameter (6K + 3M), the input of the LM diagonal. */
int vio360_ba_column_sqnorm(vio360_ba* ba, int apply_loss, double* out);

/* consume the reduced system on the device.  Solves S x = rhs of the last
 * vio360_ba_build_reduced_system by a block-envelope Cholesky (replaces the dense LLT of Ceres'
 * DenseSchurComplementSolver::SolveReducedLinearSystem, schur_complement_solver.cc) and
 * back-substitutes the landmarks, y_l = (E'E + D^2)^-1 (E'r - E'F x)
 * (SchurEliminator::BackSubstitute, schur_eliminator_impl.h:310-380).  delta_pose is K x 6
 * (zeros for constant poses), delta_point M x 3 or NULL (zeros for constant points and for
 * landmarks outside this rank's shard).  Sign convention: exactly the x, y of S x = rhs as
 * Ceres forms them (the caller negates as LevenbergMarquardtStrategy::ComputeStep does).
 * With shards, all-reduce the packed system first; every rank then holds the same x.     */
int vio360_ba_solve_reduced(vio360_ba* ba, double* delta_pose, double* delta_point);

/* the whole Levenberg-Marquardt solve of Optimizer::RunLocalBA / RunBA
 * (src/optimization/Optimizer.cpp:883-886: ceres::Solve with LEVENBERG_MARQUARDT, SPARSE_SCHUR, max 50
 * iterations, HuberLoss) as ONE call on the device-resident window: per iteration the Jacobi-scaled LM diagonal
 * (levenberg_marquardt_strategy.cc:78-89; trust_region_minimizer.cc:262-276), the fused build of the reduced
 * camera system, its Cholesky solve and the landmark back substitution (schur_eliminator_impl.h:310-380), the
 * step -y (:103-123), the model cost change (trust_region_minimizer.cc:430-441), the candidate cost, the
 * parameter / function tolerance tests (:717-761), step acceptance and the radius update
 * (:767-829; levenberg_marquardt_strategy.cc:148-161), gradient / radius / iteration termination (:302-356).
 * Parameters, Jacobian blocks and systems never leave the device; one 128-byte state record per iteration does.
 * Options mirror ceres::Solver::Options (include/ceres/solver.h defaults; vio360_lm_default_options fills them with
 * what Optimizer::SetupSolverOptions leaves in place, Optimizer.cpp:73-81).  pose_delta (K x 6) and points (M x 3)
 * are the start on entry and the solution on return.  trace (optional) receives 8 doubles per iteration, iteration
 * 0 first: cost, cost_change, |gradient|_max, |step|, relative_decrease, radius, accepted (1 / 0 / -1 invalid),
 * model_cost_change.  Follow with vio360_ba_chi2 for the outlier pass of Optimizer.cpp:898-915.                   */
typedef struct {
    int max_num_iterations;
    double initial_trust_region_radius, max_trust_region_radius, min_trust_region_radius;
    double min_relative_decrease, min_lm_diagonal, max_lm_diagonal;
    double function_tolerance, gradient_tolerance, parameter_tolerance;
    int jacobi_scaling, max_num_consecutive_invalid_steps;
    int apply_loss;              /* HuberLoss as installed by vio360_ba_set_structure (Optimizer.cpp:835) */
} vio360_lm_options;
typedef struct {
    int iterations, num_successful_steps, num_unsuccessful_steps;
    int termination;             /* 1 gradient tolerance, 2 function tolerance, 3 parameter tolerance, 4 minimum trust
                                    region radius (all four: ceres CONVERGENCE), 5 max iterations (NO_CONVERGENCE), -1 failure */
    double initial_cost, final_cost, final_radius, final_gradient_max_norm;
    int n_trace;
} vio360_lm_summary;
int vio360_lm_default_options(vio360_lm_options* opt);
int vio360_ba_lm_solve(vio360_ba* ba, const vio360_lm_options* opt, double* pose_delta, double* points,
                       vio360_lm_summary* summary, double* trace, int trace_capacity);

/* y += J x at the current parameters, J = the (loss-corrected when apply_loss) Jacobian of all residual blocks with the
 * columns of constant blocks removed: x is 6K + 3M (pose part first), y is O x 2 in the caller's observation order.
 * LinearOperator::RightMultiply (internal/ceres/linear_operator.h:46) for the SparseMatrix stand-in of
 * host/ceres_adapters -- the Jacobian itself is never materialised.                                            */
int vio360_ba_jacobian_right_multiply(vio360_ba* ba, int apply_loss, const double* x, double* y);

/* Replaces the per-factor BAFactor::compute_chi_square loop (Factors.cpp:544-612;
 * Optimizer.cpp:898-915): chi2 O (may be NULL), flag O = chi2 > thr (may be NULL).   */
int vio360_ba_chi2(vio360_ba* ba, double thr, double* chi2, uint8_t* flag);

/* ------------------------------------------------ path B (pose only): PnP ---- */
/* Replaces PnPFactor::Evaluate / compute_chi_square (Factors.cpp:33-210, 212-265) for
 * the N factors of one Optimizer::SolvePnP round (Optimizer.cpp:162-177): one shared
 * pose delta, constant world points.  r N x 2, Jp N x 12 (may be NULL), valid N (0 where
 * Evaluate returns false, :74-80), H 36 / b 6 = sum of J^T J / J^T r with Huber
 * (may be NULL), chi2 N (may be NULL).                                              */
int vio360_pnp_evaluate(vio360_ctx* ctx, int N, const double* obs_uv, const double* world_pts,
                        const uint8_t* outlier, const double* T_wb_init, const double* T_cb,
                        double cols, double rows, const double* info, double huber_delta,
                        const double* pose_delta, double* r, double* Jp, uint8_t* valid,
                        double* H, double* b, double* cost, double* chi2);

/* batched DLT triangulation of the matches of one keyframe pair.  Replaces the
 * per-match Estimator::TriangulateSinglePoint (src/processing/Estimator.cpp:1082-1139, called at :1218).
 * bearing1/bearing2: n x 3 unit bearings in the two camera frames; T1w, T2w: 4 x 4 row-major
 * world-to-camera transforms (Eigen::Matrix4f values); points: n x 3; valid[i] = the function's bool
 * (|w| >= 1e-10 and finite coordinates), points of invalid matches are zero.  fp32 as the reference. */
int vio360_triangulate(vio360_ctx* ctx, const float* bearing1, const float* bearing2, const float* T1w,
                       const float* T2w, int n, float* points, uint8_t* valid);

#ifdef __cplusplus
}
#endif
#endif /* VIO360_H */
