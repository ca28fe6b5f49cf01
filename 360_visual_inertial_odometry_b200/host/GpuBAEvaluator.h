// GpuBAEvaluator.h -- the Ceres seam for path B.
//
//   GpuBAEvaluator      : ceres::EvaluationCallback   (thirdparty/ceres-solver/include/ceres/
//                         evaluation_callback.h:36-76; registered through
//                         Problem::Options::evaluation_callback, problem.h:170-185)
//   GpuBackedBAFactor   : ceres::SizedCostFunction<2, 6, 3>, replaces factor::BAFactor
//                         (src/optimization/Factors.h:148-204) one to one: same Evaluate /
//                         compute_chi_square / set_outlier / is_outlier.
//
// Ceres calls PrepareForEvaluation() once, single threaded, after it has copied the state
// into the user's parameter blocks (program_evaluator.h:155-160); the evaluator gathers the
// blocks, runs the CUDA factor evaluation for the whole window and downloads residuals and
// Jacobians.  Ceres then calls Evaluate() per residual block from its worker threads
// (Optimizer.cpp:79 sets 4); GpuBackedBAFactor::Evaluate only copies its rows out, so it is
// thread safe and touches no CUDA state.  The un-corrected residuals/Jacobians are served
// (Ceres applies HuberLoss itself, residual_block.cc:161-195).
//
// Built against the real Ceres headers when they are on the include path
// (-DVIO360_WITH_CERES); otherwise against look-alike base classes with the identical
// virtual signatures so the seam is still compiled and exercised.
#pragma once

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "../../include/vio360.h"

#ifdef VIO360_WITH_CERES
#include "ceres/evaluation_callback.h"
#include "ceres/sized_cost_function.h"
#else
namespace ceres {
class EvaluationCallback {
public:
    virtual ~EvaluationCallback() {}
    virtual void PrepareForEvaluation(bool evaluate_jacobians, bool new_evaluation_point) = 0;
};
class CostFunction {
public:
    virtual ~CostFunction() {}
    virtual bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const = 0;
};
template <int kNumResiduals, int... Ns>
class SizedCostFunction : public CostFunction {};
}  // namespace ceres
#endif

namespace vio_360 {

class GpuBAEvaluator;

// One residual block.  Constructed by GpuBAEvaluator::AddFactor; owned by the ceres::Problem
// exactly like the reference's `new factor::BAFactor(...)` (Optimizer.cpp:829).
class GpuBackedBAFactor : public ceres::SizedCostFunction<2, 6, 3> {
public:
    bool Evaluate(double const* const* parameters, double* residuals, double** jacobians) const override;
    double compute_chi_square(double const* const* parameters) const;   // Factors.cpp:544-612
    void set_outlier(bool is_outlier);                                   // Factors.h:137
    bool is_outlier() const;
    int index() const { return m_index; }

private:
    friend class GpuBAEvaluator;
    GpuBackedBAFactor(GpuBAEvaluator* owner, int index) : m_owner(owner), m_index(index) {}
    GpuBAEvaluator* m_owner;
    int m_index;
};

class GpuBAEvaluator : public ceres::EvaluationCallback {
public:
    // cols/rows: factor::CameraParameters (Factors.h:41-46); T_cb: row-major 4x4 (m_Tcb);
    // info: 2x2 information matrix (Optimizer.cpp:824); huber_delta: Optimizer.cpp:835.
    GpuBAEvaluator(double cols, double rows, const double T_cb[16], const double info[4], double huber_delta,
                   int device = 0);
    ~GpuBAEvaluator() override;

    // Problem assembly, mirroring Optimizer.cpp:771-877.  Pointers are the caller's live
    // parameter storage (pose_params[i].data(), point_params[j].data()).
    int AddPose(double* pose_delta6, const double T_wb_init[16]);
    int AddPoint(double* point3);
    GpuBackedBAFactor* AddFactor(int pose, int point, double obs_u, double obs_v);
    void SetPoseConstant(int pose);     // problem.SetParameterBlockConstant, Optimizer.cpp:854-858
    void SetPointConstant(int point);   // Optimizer.cpp:871-873
    void Finalize();                    // uploads the structure; call once after assembly

    // ceres::EvaluationCallback
    void PrepareForEvaluation(bool evaluate_jacobians, bool new_evaluation_point) override;

    // The blocks Ceres' SchurEliminator would compute (schur_eliminator_impl.h:179-308), built on
    // the device at the current parameter values.  lm_diag: 6K+3M or nullptr.  S is n x n, n = ReducedDim().
    int ReducedDim() const;
    void BuildReducedSystem(const double* lm_diag, double* S, double* rhs, double* cost);
    // per-factor chi-square at the current parameter values (Optimizer.cpp:898-915)
    void ComputeChiSquare(double threshold, std::vector<double>& chi2, std::vector<uint8_t>& outlier_flags);
    // Drop-in for `ceres::Solve(options, &problem, &summary)` at Optimizer.cpp:883-886 when the window's factors are
    // all BAFactors: the Levenberg-Marquardt iteration Ceres would run (trust_region_minimizer.cc,
    // levenberg_marquardt_strategy.cc, SPARSE_SCHUR) executed on the device by vio360_ba_lm_solve, starting at the
    // values in the user's parameter blocks and writing the solution back into them.  opt == nullptr: the options
    // Optimizer::SetupSolverOptions leaves in place (vio360_lm_default_options).
    vio360_lm_summary SolveLM(const vio360_lm_options* opt = nullptr);

    int NumFactors() const { return (int)m_pose_idx.size(); }

private:
    friend class GpuBackedBAFactor;
    void check(int rc, const char* what) const;
    void upload_parameters();
    vio360_ctx* m_ctx = nullptr;
    vio360_ba* m_ba = nullptr;
    double m_cols, m_rows, m_Tcb[16], m_info[4], m_huber;
    std::vector<double*> m_pose_ptrs, m_point_ptrs;
    std::vector<double> m_Twb_init, m_obs_u, m_obs_v;
    std::vector<int32_t> m_pose_idx, m_point_idx;
    std::vector<uint8_t> m_pose_fixed, m_point_fixed, m_outlier;
    bool m_outliers_dirty = false, m_have_jacobians = false, m_finalized = false, m_eval_valid = false;
    std::vector<double> m_params_pose, m_params_point;   // gathered evaluation point
    std::vector<double> m_r, m_Jp, m_Jl, m_chi2;          // results served to the factors
    std::vector<double> m_chi2_pose, m_chi2_point;        // the point m_chi2 was computed at
    std::vector<double> m_eval_pose, m_eval_point;        // the point m_r / m_Jp / m_Jl were computed at
    std::vector<uint8_t> m_flags;
};

}  // namespace vio_360
