/*
 * oracle/ba_ref.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, dependency free) of the reference's bundle-adjustment
 * factor evaluation path ("hot path B", SURVEY.md section 8-A').  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * link or call this.  The shipped library (libvio360_cuda.so) never does.
 *
 * PARITY UNPINNED: the reference ships no tests / golden vectors for this path and
 * cannot be compiled here (needs Eigen + glog + OpenCV C++; none are installed).
 * The restatement is cross-checked instead by (1) an independent numpy mirror using
 * LAPACK SVD, (2) finite differences over T*exp(eps), (3) mpmath 50-digit closed
 * forms -- see tests/test_oracle_ba.py.
 *
 * Every function cites the reference lines it follows (paths relative to the
 * reference root).  Matrices are row-major.
 */
#ifndef VIO360_ORACLE_BA_REF_H
#define VIO360_ORACLE_BA_REF_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { double R[9]; double t[3]; } vo_se3;

/* ---- src/util/LieUtils.{h,cpp} (double precision classes) ------------------- */
/* 3x3 SVD, singular values descending (Eigen::JacobiSVD contract). A = U S V^T   */
void vo_svd3(const double A[9], double U[9], double S[3], double V[9]);
/* SO3d::SO3d(const Matrix3d&): closest rotation U V^T, det fix. LieUtils.cpp:275-288 */
void vo_so3_project(const double M[9], double R[9]);
/* SO3d::Exp. LieUtils.cpp:203-219 */
void vo_so3_exp(const double w[3], double R[9]);
/* SE3d::SE3d(const Matrix4d&). LieUtils.cpp:292-295 */
void vo_se3_from_matrix(const double T[16], vo_se3* out);
/* SE3d::exp. LieUtils.cpp:305-333 */
void vo_se3_exp(const double xi[6], vo_se3* out);
/* SE3d::operator*. LieUtils.h:268-271 (+ SO3d::operator* :207-209) */
void vo_se3_mul(const vo_se3* a, const vo_se3* b, vo_se3* out);
/* SE3d::inverse. LieUtils.h:279-282 (+ SO3d::Inverse :217-219) */
void vo_se3_inverse(const vo_se3* a, vo_se3* out);

/* ---- src/optimization/Factors.{h,cpp} -------------------------------------- */
typedef struct {
    double obs[2];        /* m_observation                          */
    double cols, rows;    /* m_camera_params                        */
    double Tcb[16];       /* m_Tcb  (body -> camera), row-major 4x4 */
    double Twb_init[16];  /* m_T_wb_init                            */
    double info[4];       /* m_information 2x2                      */
    int    is_outlier;    /* m_is_outlier                           */
    double world_point[3];/* PnPFactor::m_world_point (PnP only)    */
} vo_factor;

/* BAFactor::Evaluate, Factors.cpp:327-542. jac_pose (2x6) / jac_point (2x3) may be
 * NULL independently; want_jac==0 models `jacobians == nullptr`. returns 1 (true). */
int vo_ba_evaluate(const vo_factor* f, const double pose6[6], const double pt[3],
                   double r[2], int want_jac, double* jac_pose, double* jac_point);
/* PnPFactor::Evaluate, Factors.cpp:33-210. returns 0 for the L<1e-10 `return false`. */
int vo_pnp_evaluate(const vo_factor* f, const double pose6[6],
                    double r[2], double* jac_pose);
/* BAFactor::compute_chi_square, Factors.cpp:544-612 */
double vo_ba_chi2(const vo_factor* f, const double pose6[6], const double pt[3]);
/* PnPFactor::compute_chi_square, Factors.cpp:212-265 */
double vo_pnp_chi2(const vo_factor* f, const double pose6[6]);

/* ---- vendored Ceres 2.0.0 pieces on the path -------------------------------- */
/* HuberLoss::Evaluate, thirdparty/ceres-solver/internal/ceres/loss_function.cc:48-62 */
void vo_huber(double a, double s, double rho[3]);
/* ResidualBlock::Evaluate tail (cost + Corrector), residual_block.cc:133-197 and
 * corrector.cc:42-160.  r (nres) and the Jacobian blocks are corrected in place.
 * returns the block's cost 0.5*rho[0].                                           */
double vo_apply_loss(double huber_a, double* r, int nres,
                     double* jac_a, int na, double* jac_b, int nb);

/* ---- window level (ProgramEvaluator + SchurEliminator + LM diagonal) -------- */
typedef struct {
    int K, M, O;
    const int* pose_idx;         /* O */
    const int* point_idx;        /* O */
    const double* obs_u;         /* O */
    const double* obs_v;         /* O */
    const unsigned char* outlier;/* O or NULL */
    const double* Twb_init;      /* K x 16 */
    double Tcb[16];
    double cols, rows;
    double info[4];
    double huber_a;              /* <=0 : no loss function */
    const unsigned char* pose_fixed;  /* K or NULL */
    const unsigned char* point_fixed; /* M or NULL */
} vo_window;

/* One ProgramEvaluator::Evaluate pass (program_evaluator.h:138-260) executed as the
 * reference does it: per factor, full pose chain (6 SVD projections) each time.
 * Outputs (any may be NULL): r[O*2], Jp[O*12], Jl[O*6] after loss correction, cost.
 * Jacobian blocks of constant parameter blocks are still written here (Ceres would
 * pass NULL for them); the Schur routine below ignores them. nthreads>=1 (pthreads). */
int vo_window_evaluate(const vo_window* w, const double* pose_delta, const double* points,
                       int apply_loss, double* r, double* Jp, double* Jl, double* cost,
                       int nthreads);
/* Same arithmetic with the pose chain hoisted (once per pose) -- the "hoisted" CPU
 * variant of SURVEY.md section 8-D, so that a GPU speed-up is not credited with a
 * purely algorithmic saving.                                                        */
int vo_window_evaluate_hoisted(const vo_window* w, const double* pose_delta,
                               const double* points, int apply_loss, double* r,
                               double* Jp, double* Jl, double* cost, int nthreads);

/* Gauss-Newton blocks of J^T J and J^T r from corrected r/Jp/Jl:
 * Hpp[K*36], bp[K*6], Hll[M*9], bl[M*3], Hpl[O*18] (per observation F^T E, 6x3).
 * Fixed blocks get zeros.                                                           */
void vo_normal_blocks(const vo_window* w, const double* r, const double* Jp,
                      const double* Jl, double* Hpp, double* bp, double* Hll,
                      double* bl, double* Hpl);

/* SchurEliminator<2,3,6>::Eliminate (schur_eliminator_impl.h:179-308 and helpers)
 * on the corrected r/Jp/Jl.  D (6K + 3M, pose part first; NULL = zeros) is the LM
 * diagonal.  Free poses are numbered in increasing pose order; n = 6*Kfree.
 * S is n x n dense row-major, FULL symmetric; rhs is n.  free_map[K] (may be NULL)
 * receives the reduced index of each pose (-1 if fixed).  returns n.               */
int vo_schur_eliminate(const vo_window* w, const double* r, const double* Jp,
                       const double* Jl, const double* D, double* S, double* rhs,
                       int* free_map);
/* LevenbergMarquardtStrategy::ComputeStep diagonal
 * (levenberg_marquardt_strategy.cc:78-89): D = sqrt(clamp(colnorm2,1e-6,1e32)/radius) */
void vo_lm_diagonal(const vo_window* w, const double* Jp, const double* Jl,
                    double radius, double* D);
/* chi-square per observation as the Optimizer loops do (Optimizer.cpp:898-915):
 * chi2[O], flag[O] = chi2 > thr.                                                   */
void vo_window_chi2(const vo_window* w, const double* pose_delta, const double* points,
                    double thr, double* chi2, unsigned char* flag);

/* Estimator::TriangulateSinglePoint (src/processing/Estimator.cpp:1082-1139) for n matches of one
 * keyframe pair, fp32.  T1w/T2w 4x4 row-major.  Eigen::JacobiSVD<Matrix4f> is restated by a one-sided
 * Jacobi SVD (last right singular vector; unique up to sign).  valid[i] = the function's return.   */
void vo_triangulate(const float* b1, const float* b2, const float* T1w, const float* T2w, int n,
                    float* points, unsigned char* valid);

#ifdef __cplusplus
}
#endif
#endif
