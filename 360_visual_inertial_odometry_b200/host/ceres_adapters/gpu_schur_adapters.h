// gpu_schur_adapters.h -- SURVEY 8-F rank 3, the "Ceres receives the blocks" seam: a custom
// ceres::internal::Evaluator and ceres::internal::LinearSolver that hand the trust-region loop of the VENDORED
// Ceres 2.0.0 the device-built reduced camera system instead of a materialised Jacobian.
//
//   GpuJacobian          : ceres::internal::SparseMatrix   (internal/ceres/sparse_matrix.h:64-104) -- a token that stands
//                          for the device-resident Jacobian.  It owns no values; SquaredColumnNorm comes from the
//                          device's Gauss-Newton blocks, RightMultiply from vio360_ba_jacobian_right_multiply,
//                          ScaleColumns records the Jacobi scaling (trust_region_minimizer.cc:262-276).
//   GpuSchurEvaluator    : ceres::internal::Evaluator      (internal/ceres/evaluator.h:59-160) -- replaces
//                          ProgramEvaluator<BlockEvaluatePreparer, BlockJacobianWriter> (program_evaluator.h:138-260):
//                          cost / residuals / gradient from the device; `jacobian` only marks the linearisation point.
//   GpuSchurLinearSolver : ceres::internal::LinearSolver   (internal/ceres/linear_solver.h:99-330) -- replaces
//                          SparseSchurComplementSolver (schur_complement_solver.cc): per_solve_options.D is the LM
//                          diagonal of LevenbergMarquardtStrategy::ComputeStep (levenberg_marquardt_strategy.cc:78-101);
//                          the reduced system is built, factorised and back-substituted on the device.
//
// Installation (thirdparty/ceres-solver is patched in two places, shown in INTEGRATION.md):
//   trust_region_preprocessor.cc  SetupEvaluator():     pp->evaluator.reset(new GpuSchurEvaluator(window, program));
//                                 SetupLinearSolver():  pp->linear_solver.reset(new GpuSchurLinearSolver(window));
//
// BUILD STATUS: these classes derive from Ceres INTERNAL headers, which include Eigen.  This container has no Eigen
// (SURVEY F5), so the file is NOT compiled here and not part of libvio360_host.so; it is written against the
// vendored headers at the cited lines and compiles only inside the reference's own build (-DVIO360_CERES_INTERNAL).
// The same iteration, without Ceres in the loop, is vio360_ba_lm_solve / GpuBAEvaluator::SolveLM, which IS built
// and tested (tests/test_gpu_ba.py::test_lm_solve_vs_oracle).
#pragma once
#ifdef VIO360_CERES_INTERNAL

#include <map>
#include <string>
#include <vector>

#include "../../../include/vio360.h"
#include "ceres/evaluator.h"
#include "ceres/linear_solver.h"
#include "ceres/sparse_matrix.h"

namespace vio_360 {

// The window as the adapters see it: the C-ABI handle plus the maps between Ceres' reduced program (constant blocks
// removed, parameter blocks in elimination order: points first, then poses -- Optimizer.cpp:883 SPARSE_SCHUR with
// automatic ordering) and the window's (pose-major, 6K + 3M) layout.
struct GpuWindow {
    vio360_ba* ba = nullptr;
    int K = 0, M = 0, O = 0;
    std::vector<int> state_offset_pose;    // K: offset of pose k in Ceres' state vector, -1 when constant
    std::vector<int> state_offset_point;   // M
    std::vector<double> const_pose, const_point;   // values of the constant blocks (K x 6, M x 3)
    int num_parameters = 0, num_residuals = 0;
};

class GpuJacobian : public ceres::internal::SparseMatrix {
public:
    explicit GpuJacobian(GpuWindow* w) : w_(w), scale_((size_t)w->num_parameters, 1.0) {}
    void RightMultiply(const double* x, double* y) const override;   // y += J diag(scale) x
    void LeftMultiply(const double* x, double* y) const override;    // not used on the SPARSE_SCHUR path
    void SquaredColumnNorm(double* x) const override;                // diag(J'J) scaled
    void ScaleColumns(const double* scale) override;
    void SetZero() override {}
    void ToDenseMatrix(ceres::internal::Matrix* dense_matrix) const override;
    void ToTextFile(FILE* file) const override;
    double* mutable_values() override { return nullptr; }
    const double* values() const override { return nullptr; }
    int num_rows() const override { return w_->num_residuals; }
    int num_cols() const override { return w_->num_parameters; }
    int num_nonzeros() const override { return 18 * w_->O; }
    const std::vector<double>& scale() const { return scale_; }
    void ResetScale() { scale_.assign(scale_.size(), 1.0); }

private:
    GpuWindow* w_;
    std::vector<double> scale_;   // accumulated column scaling, Ceres state order
};

class GpuSchurEvaluator : public ceres::internal::Evaluator {
public:
    explicit GpuSchurEvaluator(GpuWindow* w) : w_(w) {}
    ceres::internal::SparseMatrix* CreateJacobian() const override { return new GpuJacobian(w_); }
    bool Evaluate(const EvaluateOptions& evaluate_options, const double* state, double* cost, double* residuals,
                  double* gradient, ceres::internal::SparseMatrix* jacobian) override;
    bool Plus(const double* state, const double* delta, double* state_plus_delta) const override;   // plain addition: the
                                                    // reference attaches no LocalParameterization (Optimizer.cpp:771-773)
    int NumParameters() const override { return w_->num_parameters; }
    int NumEffectiveParameters() const override { return w_->num_parameters; }
    int NumResiduals() const override { return w_->num_residuals; }

private:
    void upload(const double* state);
    GpuWindow* w_;
    std::vector<double> pose_, point_;
};

class GpuSchurLinearSolver : public ceres::internal::LinearSolver {
public:
    explicit GpuSchurLinearSolver(GpuWindow* w) : w_(w) {}
    Summary Solve(ceres::internal::LinearOperator* A, const double* b, const PerSolveOptions& per_solve_options, double* x) override;

private:
    GpuWindow* w_;
    std::vector<double> D_, dp_, dl_;
};

}  // namespace vio_360
#endif  // VIO360_CERES_INTERNAL
