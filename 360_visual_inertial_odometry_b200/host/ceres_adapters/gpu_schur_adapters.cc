// gpu_schur_adapters.cc -- see gpu_schur_adapters.h (not compiled in this container: needs Ceres internals + Eigen).
#ifdef VIO360_CERES_INTERNAL
#include "gpu_schur_adapters.h"

#include <cmath>
#include <cstring>
#include <stdexcept>

namespace vio_360 {
namespace {

// Ceres state vector (reduced program) -> window layout; constant blocks keep their stored values
void gather(const GpuWindow& w, const double* state, std::vector<double>& pose, std::vector<double>& point) {
    pose = w.const_pose; point = w.const_point;
    for (int k = 0; k < w.K; ++k)
        if (w.state_offset_pose[k] >= 0) std::memcpy(&pose[6 * (size_t)k], state + w.state_offset_pose[k], 6 * sizeof(double));
    for (int l = 0; l < w.M; ++l)
        if (w.state_offset_point[l] >= 0) std::memcpy(&point[3 * (size_t)l], state + w.state_offset_point[l], 3 * sizeof(double));
}
// window-layout vector (6K + 3M) <-> Ceres-order vector (num_parameters)
void to_window(const GpuWindow& w, const double* v, std::vector<double>& out) {
    out.assign(6 * (size_t)w.K + 3 * (size_t)w.M, 0.0);
    for (int k = 0; k < w.K; ++k)
        if (w.state_offset_pose[k] >= 0) std::memcpy(&out[6 * (size_t)k], v + w.state_offset_pose[k], 6 * sizeof(double));
    for (int l = 0; l < w.M; ++l)
        if (w.state_offset_point[l] >= 0) std::memcpy(&out[6 * (size_t)w.K + 3 * (size_t)l], v + w.state_offset_point[l], 3 * sizeof(double));
}
void from_window(const GpuWindow& w, const double* win, double* v) {
    for (int k = 0; k < w.K; ++k)
        if (w.state_offset_pose[k] >= 0) std::memcpy(v + w.state_offset_pose[k], win + 6 * (size_t)k, 6 * sizeof(double));
    for (int l = 0; l < w.M; ++l)
        if (w.state_offset_point[l] >= 0) std::memcpy(v + w.state_offset_point[l], win + 6 * (size_t)w.K + 3 * (size_t)l, 3 * sizeof(double));
}
void check(int rc, const char* what) {
    if (rc != VIO360_OK) throw std::runtime_error(std::string(what) + " failed");
}

}  // namespace

// ---- GpuJacobian
void GpuJacobian::RightMultiply(const double* x, double* y) const {
    std::vector<double> xs((size_t)w_->num_parameters), xw;
    for (size_t i = 0; i < xs.size(); ++i) xs[i] = x[i] * scale_[i];
    to_window(*w_, xs.data(), xw);
    check(vio360_ba_jacobian_right_multiply(w_->ba, /*apply_loss=*/1, xw.data(), y), "vio360_ba_jacobian_right_multiply");
}
void GpuJacobian::LeftMultiply(const double*, double*) const {
    throw std::logic_error("GpuJacobian::LeftMultiply: not needed by TrustRegionMinimizer + LevenbergMarquardtStrategy + Schur");
}
void GpuJacobian::SquaredColumnNorm(double* x) const {
    std::vector<double> c(6 * (size_t)w_->K + 3 * (size_t)w_->M);
    check(vio360_ba_column_sqnorm(w_->ba, 1, c.data()), "vio360_ba_column_sqnorm");
    from_window(*w_, c.data(), x);
    for (int i = 0; i < w_->num_parameters; ++i) x[i] *= scale_[(size_t)i] * scale_[(size_t)i];
}
void GpuJacobian::ScaleColumns(const double* scale) {
    for (size_t i = 0; i < scale_.size(); ++i) scale_[i] *= scale[i];
}
void GpuJacobian::ToDenseMatrix(ceres::internal::Matrix*) const { throw std::logic_error("GpuJacobian is not materialised"); }
void GpuJacobian::ToTextFile(FILE*) const { throw std::logic_error("GpuJacobian is not materialised"); }

// ---- GpuSchurEvaluator (program_evaluator.h:138-260)
void GpuSchurEvaluator::upload(const double* state) {
    gather(*w_, state, pose_, point_);
    check(vio360_ba_set_parameters(w_->ba, pose_.data(), point_.data()), "vio360_ba_set_parameters");
}
bool GpuSchurEvaluator::Evaluate(const EvaluateOptions& evaluate_options, const double* state, double* cost, double* residuals,
                                 double* gradient, ceres::internal::SparseMatrix* jacobian) {
    upload(state);
    const int loss = evaluate_options.apply_loss_function ? 1 : 0;
    double c = 0.0;
    check(vio360_ba_evaluate(w_->ba, /*want_jacobians=*/0, loss, &c), "vio360_ba_evaluate");
    if (cost) *cost = c;
    if (residuals) check(vio360_ba_get_residuals_jacobians(w_->ba, residuals, nullptr, nullptr), "vio360_ba_get_residuals_jacobians");
    if (gradient) {   // g = J' r from the Gauss-Newton blocks (bp, bl)
        std::vector<double> bp(6 * (size_t)w_->K), bl(3 * (size_t)w_->M), g(bp.size() + bl.size());
        check(vio360_ba_normal_blocks(w_->ba, loss, nullptr, bp.data(), nullptr, bl.data(), nullptr), "vio360_ba_normal_blocks");
        std::memcpy(g.data(), bp.data(), bp.size() * sizeof(double));
        std::memcpy(g.data() + bp.size(), bl.data(), bl.size() * sizeof(double));
        from_window(*w_, g.data(), gradient);
    }
    if (jacobian) static_cast<GpuJacobian*>(jacobian)->ResetScale();   // a fresh, unscaled Jacobian at `state`
    return true;
}
bool GpuSchurEvaluator::Plus(const double* state, const double* delta, double* state_plus_delta) const {
    for (int i = 0; i < w_->num_parameters; ++i) state_plus_delta[i] = state[i] + delta[i];
    return true;
}

// ---- GpuSchurLinearSolver: solve (J s)' (J s) y + D'D y = (J s)' b for y (schur_complement_solver.cc; b is the residual
// vector the evaluator produced at the same point, which the device recomputes inside the fused build)
ceres::internal::LinearSolver::Summary GpuSchurLinearSolver::Solve(ceres::internal::LinearOperator* A, const double* /*b*/,
                                                                  const PerSolveOptions& per_solve_options, double* x) {
    Summary summary;
    const GpuJacobian* J = static_cast<const GpuJacobian*>(A);
    const std::vector<double>& s = J->scale();
    // the device solves the UNSCALED system: regulariser |D y|^2 with y = x_unscaled / s  ->  D_unscaled = D / s
    std::vector<double> Ds((size_t)w_->num_parameters);
    for (size_t i = 0; i < Ds.size(); ++i) Ds[i] = per_solve_options.D ? per_solve_options.D[i] / s[i] : 0.0;
    to_window(*w_, Ds.data(), D_);
    if (vio360_ba_build_reduced_system(w_->ba, D_.data(), 1, nullptr, nullptr, nullptr) != VIO360_OK) {
        summary.termination_type = ceres::LINEAR_SOLVER_FATAL_ERROR; summary.message = "vio360_ba_build_reduced_system failed";
        return summary;
    }
    dp_.assign(6 * (size_t)w_->K, 0.0); dl_.assign(3 * (size_t)w_->M, 0.0);
    if (vio360_ba_solve_reduced(w_->ba, dp_.data(), dl_.data()) != VIO360_OK) {   // not positive definite: recoverable (a larger D)
        summary.termination_type = ceres::LINEAR_SOLVER_FAILURE; summary.message = "reduced camera system is not positive definite";
        return summary;
    }
    std::vector<double> win(dp_);
    win.insert(win.end(), dl_.begin(), dl_.end());
    from_window(*w_, win.data(), x);
    for (int i = 0; i < w_->num_parameters; ++i) x[i] /= s[(size_t)i];           // back to the scaled unknown Ceres expects
    summary.termination_type = ceres::LINEAR_SOLVER_SUCCESS; summary.num_iterations = 1; summary.message = "Success";
    return summary;
}

}  // namespace vio_360
#endif  // VIO360_CERES_INTERNAL
