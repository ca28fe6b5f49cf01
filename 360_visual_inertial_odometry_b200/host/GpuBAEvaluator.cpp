// GpuBAEvaluator.cpp -- see GpuBAEvaluator.h.
#include "GpuBAEvaluator.h"

#include <cstring>
#include <stdexcept>

namespace vio_360 {

void GpuBAEvaluator::check(int rc, const char* what) const {
    if (rc != VIO360_OK) throw std::runtime_error(std::string(what) + ": " + vio360_last_error(m_ctx));
}

GpuBAEvaluator::GpuBAEvaluator(double cols, double rows, const double T_cb[16], const double info[4],
                               double huber_delta, int device)
    : m_cols(cols), m_rows(rows), m_huber(huber_delta) {
    std::memcpy(m_Tcb, T_cb, sizeof(m_Tcb));
    std::memcpy(m_info, info, sizeof(m_info));
    if (vio360_create(device, &m_ctx) != VIO360_OK)
        throw std::runtime_error(std::string("GpuBAEvaluator: ") + vio360_last_error(nullptr));
}

GpuBAEvaluator::~GpuBAEvaluator() {
    vio360_ba_destroy(m_ba);
    vio360_destroy(m_ctx);
}

int GpuBAEvaluator::AddPose(double* pose_delta6, const double T_wb_init[16]) {
    m_pose_ptrs.push_back(pose_delta6);
    m_Twb_init.insert(m_Twb_init.end(), T_wb_init, T_wb_init + 16);
    m_pose_fixed.push_back(0);
    return (int)m_pose_ptrs.size() - 1;
}

int GpuBAEvaluator::AddPoint(double* point3) {
    m_point_ptrs.push_back(point3);
    m_point_fixed.push_back(0);
    return (int)m_point_ptrs.size() - 1;
}

GpuBackedBAFactor* GpuBAEvaluator::AddFactor(int pose, int point, double obs_u, double obs_v) {
    if (m_finalized) throw std::runtime_error("GpuBAEvaluator::AddFactor after Finalize");
    m_pose_idx.push_back(pose); m_point_idx.push_back(point);
    m_obs_u.push_back(obs_u); m_obs_v.push_back(obs_v);
    m_outlier.push_back(0);
    return new GpuBackedBAFactor(this, (int)m_pose_idx.size() - 1);
}

void GpuBAEvaluator::SetPoseConstant(int pose) { m_pose_fixed.at((size_t)pose) = 1; }
void GpuBAEvaluator::SetPointConstant(int point) { m_point_fixed.at((size_t)point) = 1; }

void GpuBAEvaluator::Finalize() {
    const int K = (int)m_pose_ptrs.size(), M = (int)m_point_ptrs.size(), O = NumFactors();
    check(vio360_ba_create(m_ctx, K, M, O, &m_ba), "vio360_ba_create");
    check(vio360_ba_set_structure(m_ba, m_pose_idx.data(), m_point_idx.data(), m_obs_u.data(), m_obs_v.data(),
                                  m_Twb_init.data(), m_Tcb, m_cols, m_rows, m_info, m_huber, m_pose_fixed.data(),
                                  m_point_fixed.data()),
          "vio360_ba_set_structure");
    m_params_pose.resize((size_t)K * 6); m_params_point.resize((size_t)M * 3);
    m_r.assign((size_t)O * 2, 0.0); m_Jp.assign((size_t)O * 12, 0.0); m_Jl.assign((size_t)O * 6, 0.0);
    m_finalized = true;
}

void GpuBAEvaluator::upload_parameters() {
    for (size_t k = 0; k < m_pose_ptrs.size(); ++k) std::memcpy(&m_params_pose[6 * k], m_pose_ptrs[k], 6 * sizeof(double));
    for (size_t j = 0; j < m_point_ptrs.size(); ++j) std::memcpy(&m_params_point[3 * j], m_point_ptrs[j], 3 * sizeof(double));
    if (m_outliers_dirty) {
        check(vio360_ba_set_outliers(m_ba, m_outlier.data()), "vio360_ba_set_outliers");
        m_outliers_dirty = false;
    }
    check(vio360_ba_set_parameters(m_ba, m_params_pose.data(), m_params_point.data()), "vio360_ba_set_parameters");
}

void GpuBAEvaluator::PrepareForEvaluation(bool evaluate_jacobians, bool new_evaluation_point) {
    if (!m_finalized) throw std::runtime_error("GpuBAEvaluator::PrepareForEvaluation before Finalize");
    if (!new_evaluation_point && (m_have_jacobians || !evaluate_jacobians) && !m_outliers_dirty) return;   // cached
    upload_parameters();
    double cost = 0.0;
    check(vio360_ba_evaluate(m_ba, evaluate_jacobians ? 1 : 0, /*apply_loss=*/0, &cost), "vio360_ba_evaluate");
    check(vio360_ba_get_residuals_jacobians(m_ba, m_r.data(), evaluate_jacobians ? m_Jp.data() : nullptr,
                                            evaluate_jacobians ? m_Jl.data() : nullptr),
          "vio360_ba_get_residuals_jacobians");
    m_have_jacobians = evaluate_jacobians;
    m_eval_pose = m_params_pose; m_eval_point = m_params_point;
    m_eval_valid = true;
}

int GpuBAEvaluator::ReducedDim() const { return vio360_ba_reduced_dim(m_ba); }

void GpuBAEvaluator::BuildReducedSystem(const double* lm_diag, double* S, double* rhs, double* cost) {
    upload_parameters();
    if (!lm_diag) check(vio360_ba_set_lm_diagonal(m_ba, nullptr), "vio360_ba_set_lm_diagonal");
    check(vio360_ba_build_reduced_system(m_ba, lm_diag, /*apply_loss=*/1, S, rhs, cost), "vio360_ba_build_reduced_system");
}

void GpuBAEvaluator::ComputeChiSquare(double threshold, std::vector<double>& chi2, std::vector<uint8_t>& outlier_flags) {
    upload_parameters();
    chi2.resize((size_t)NumFactors()); outlier_flags.resize((size_t)NumFactors());
    check(vio360_ba_chi2(m_ba, threshold, chi2.data(), outlier_flags.data()), "vio360_ba_chi2");
    m_chi2 = chi2;
    m_chi2_pose = m_params_pose; m_chi2_point = m_params_point;
}

vio360_lm_summary GpuBAEvaluator::SolveLM(const vio360_lm_options* opt) {
    if (!m_finalized) throw std::runtime_error("GpuBAEvaluator::SolveLM before Finalize");
    vio360_lm_options def;
    if (!opt) { check(vio360_lm_default_options(&def), "vio360_lm_default_options"); opt = &def; }
    for (size_t k = 0; k < m_pose_ptrs.size(); ++k) std::memcpy(&m_params_pose[6 * k], m_pose_ptrs[k], 6 * sizeof(double));
    for (size_t j = 0; j < m_point_ptrs.size(); ++j) std::memcpy(&m_params_point[3 * j], m_point_ptrs[j], 3 * sizeof(double));
    if (m_outliers_dirty) { check(vio360_ba_set_outliers(m_ba, m_outlier.data()), "vio360_ba_set_outliers"); m_outliers_dirty = false; }
    vio360_lm_summary sm;
    check(vio360_ba_lm_solve(m_ba, opt, m_params_pose.data(), m_params_point.data(), &sm, nullptr, 0), "vio360_ba_lm_solve");
    // ceres::Solve leaves the solution in the user's parameter blocks (constant blocks unchanged: their step is zero)
    for (size_t k = 0; k < m_pose_ptrs.size(); ++k) std::memcpy(m_pose_ptrs[k], &m_params_pose[6 * k], 6 * sizeof(double));
    for (size_t j = 0; j < m_point_ptrs.size(); ++j) std::memcpy(m_point_ptrs[j], &m_params_point[3 * j], 3 * sizeof(double));
    m_have_jacobians = false;   // cached r / J belong to an earlier point
    m_r.assign(m_r.size(), 0.0);
    m_eval_valid = false;
    return sm;
}

// ------------------------------------------------------------------ the per-block facade
// The rows served are those of the point the last PrepareForEvaluation uploaded.  Ceres passes pointers into ITS state
// vector (ResidualBlock::Evaluate, residual_block.cc:81-88), so the check is by value: a caller that evaluates the factor
// at another point (Problem::Evaluate elsewhere, a gradient checker, the reference's post-solve loops with explicit
// pointers, Optimizer.cpp:220,432,662,906) gets an exception instead of stale numbers.
bool GpuBackedBAFactor::Evaluate(double const* const* parameters, double* residuals, double** jacobians) const {
    const GpuBAEvaluator& e = *m_owner;
    const size_t i = (size_t)m_index;
    if (!e.m_eval_valid) throw std::runtime_error("GpuBackedBAFactor::Evaluate: no evaluation point uploaded (register the GpuBAEvaluator as Problem::Options::evaluation_callback)");
    if (parameters && parameters[0] && parameters[1] &&
        (std::memcmp(parameters[0], &e.m_eval_pose[6 * (size_t)e.m_pose_idx[i]], 6 * sizeof(double)) != 0 ||
         std::memcmp(parameters[1], &e.m_eval_point[3 * (size_t)e.m_point_idx[i]], 3 * sizeof(double)) != 0))
        throw std::runtime_error("GpuBackedBAFactor::Evaluate: called at a point other than the one GpuBAEvaluator::PrepareForEvaluation uploaded");
    residuals[0] = e.m_r[2 * i]; residuals[1] = e.m_r[2 * i + 1];
    if (jacobians) {
        if (jacobians[0]) std::memcpy(jacobians[0], &e.m_Jp[12 * i], 12 * sizeof(double));
        if (jacobians[1]) std::memcpy(jacobians[1], &e.m_Jl[6 * i], 6 * sizeof(double));
    }
    return true;   // BAFactor::Evaluate returns true on every branch (Factors.cpp:327-542)
}

double GpuBackedBAFactor::compute_chi_square(double const* const* parameters) const {
    // served from the batch computed by GpuBAEvaluator::ComputeChiSquare (one kernel for all factors)
    const GpuBAEvaluator& e = *m_owner;
    const size_t i = (size_t)m_index;
    if (e.m_chi2.size() != (size_t)e.NumFactors())
        throw std::runtime_error("GpuBackedBAFactor::compute_chi_square: call GpuBAEvaluator::ComputeChiSquare first");
    if (parameters && parameters[0] && parameters[1] &&
        (std::memcmp(parameters[0], &e.m_chi2_pose[6 * (size_t)e.m_pose_idx[i]], 6 * sizeof(double)) != 0 ||
         std::memcmp(parameters[1], &e.m_chi2_point[3 * (size_t)e.m_point_idx[i]], 3 * sizeof(double)) != 0))
        throw std::runtime_error("GpuBackedBAFactor::compute_chi_square: the parameters differ from the point GpuBAEvaluator::ComputeChiSquare used");
    return m_owner->m_chi2[(size_t)m_index];
}

void GpuBackedBAFactor::set_outlier(bool is_outlier) {
    m_owner->m_outlier[(size_t)m_index] = is_outlier ? 1 : 0;
    m_owner->m_outliers_dirty = true;
}

bool GpuBackedBAFactor::is_outlier() const { return m_owner->m_outlier[(size_t)m_index] != 0; }

}  // namespace vio_360
