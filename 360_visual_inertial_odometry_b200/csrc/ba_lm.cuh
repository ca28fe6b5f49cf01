// ba_lm.cuh -- included by ba.cu.  SURVEY 8-F rank 3 / 7.3-8: the Levenberg-Marquardt iteration Ceres runs for
// Optimizer::RunLocalBA (src/optimization/Optimizer.cpp:883-886: LM, SPARSE_SCHUR, <= 50 iterations), on the
// device-built reduced camera system, with the parameters, Jacobian blocks and systems resident on the device:
//   TrustRegionMinimizer::Minimize           thirdparty/ceres-solver/internal/ceres/trust_region_minimizer.cc:67-139
//   LevenbergMarquardtStrategy::ComputeStep  levenberg_marquardt_strategy.cc:66-146 (diagonal, solve J y = r, step = -y)
//   ::StepAccepted / ::StepRejected          :148-161 (radius update)
//   TrustRegionStepEvaluator::StepQuality    trust_region_step_evaluator.cc:52-67 (monotonic steps)
//   Jacobi scaling                           trust_region_minimizer.cc:262-276 (Solver::Options::jacobi_scaling = true)
// These kernels are the glue around the existing ones (k_blocks_*, k_schur_tile / k_schur_reduce, k_chol_band,
// k_back_substitute, k_evaluate): per-parameter vectors, three reductions and the accept / reject decision.
#pragma once

namespace vio360 {

struct LmOptionsDev {
    double min_diag, max_diag, max_radius, min_radius, min_relative_decrease;
    double function_tolerance, gradient_tolerance, parameter_tolerance;
    int jacobi_scaling, max_invalid;
};

struct LmState {           // device-resident scalars of one solve; a copy goes to the host after every iteration
    double radius, decrease_factor;
    double x_cost, candidate_cost, model_cost_change;
    double x_norm, step_norm, gradient_max_norm;
    double relative_decrease, cost_change;
    int iteration, num_successful, num_unsuccessful, num_invalid;
    int step_valid, step_successful;   // of the iteration just finished
    int termination;       // 0 running, 1 gradient tolerance, 2 function tolerance, 3 parameter tolerance, 4 min radius,
                           // 5 max iterations, -1 failure (consecutive invalid steps / not positive definite)
    int relinearize;       // the next iteration needs new Jacobian blocks (the last step was accepted)
};

constexpr int kLmRedThreads = 256;

// deterministic CTA sum of `v` -> out[blockIdx.x * stride + slot]
__device__ __forceinline__ double lm_block_sum(double v) {
    __shared__ double sm[kLmRedThreads / 32];
    __syncthreads();
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += sm[w];
    return t;   // valid in thread 0
}
__device__ __forceinline__ double lm_block_max(double v) {
    __shared__ double sm[kLmRedThreads / 32];
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t = fmax(t, sm[w]);
    return t;
}

// After k_blocks_*: squared column norms of J (diagonals of Hpp / Hll, zero for constant blocks), the max norm of
// the gradient J'r over the free parameters, |x| over the free parameters; on the first linearisation also the
// Jacobi scaling 1 / (1 + sqrt(|column|^2)).   The vectors are 6K + 3M long (151 200 entries for the 200-keyframe
// window): a grid of CTAs leaves per-CTA partials (max, sum), k_lm_linearized_finish combines them in index order.
__global__ void __launch_bounds__(kLmRedThreads) k_lm_linearized(int K, int M, const double* __restrict__ Hpp, const double* __restrict__ bp,
                                                                 const double* __restrict__ Hll, const double* __restrict__ bl,
                                                                 const int* __restrict__ free_map, const unsigned char* __restrict__ point_fixed,
                                                                 const double* __restrict__ pose_delta, const double* __restrict__ points,
                                                                 double* __restrict__ colsq, double* __restrict__ scale, int first, int jacobi_scaling,
                                                                 double* __restrict__ partials) {
    const int np = 6 * K, n = np + 3 * M;
    double gmax = 0.0, x2 = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double c, g, x;
        bool is_free;
        if (i < np) {
            const int k = i / 6, a = i - 6 * k;
            is_free = free_map[k] >= 0;
            c = Hpp[36 * (size_t)k + 7 * a]; g = bp[i]; x = pose_delta[i];
        } else {
            const int j = i - np, l = j / 3, a = j - 3 * l;
            is_free = point_fixed[l] == 0;
            c = Hll[9 * (size_t)l + 4 * a]; g = bl[j]; x = points[j];
        }
        colsq[i] = is_free ? c : 0.0;
        if (first) scale[i] = jacobi_scaling ? 1.0 / (1.0 + sqrt(is_free ? c : 0.0)) : 1.0;
        if (is_free) { gmax = fmax(gmax, fabs(g)); x2 += x * x; }
    }
    const double gm = lm_block_max(gmax);
    const double xs = lm_block_sum(x2);
    if (threadIdx.x == 0) { partials[blockIdx.x] = gm; partials[gridDim.x + blockIdx.x] = xs; }
}
// The gradient test of trust_region_minimizer.cc:674-692 is made here, so that the host does not have to read the state
// back between the re-linearisation and the next step: a state that has terminated makes k_lm_decide a no-op.
__global__ void __launch_bounds__(kLmRedThreads) k_lm_linearized_finish(const double* __restrict__ partials, int nparts, LmState* __restrict__ st,
                                                                        double gradient_tolerance) {
    double gmax = 0.0, x2 = 0.0;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) { gmax = fmax(gmax, partials[i]); x2 += partials[nparts + i]; }
    const double gm = lm_block_max(gmax);
    const double xs = lm_block_sum(x2);
    if (threadIdx.x == 0) {
        st->gradient_max_norm = gm; st->x_norm = sqrt(xs);
        if (st->termination == 0 && gm <= gradient_tolerance) st->termination = 1;   // GradientToleranceReached
    }
}

// LevenbergMarquardtStrategy::ComputeStep (:78-89) on the Jacobi-scaled Jacobian, expressed for the unscaled
// system the fused build solves: J s has column norms^2 c s^2; D_scaled = sqrt(clamp(c s^2) / radius); the
// regulariser |D_scaled y|^2 of the scaled unknown y = x / s is |(D_scaled / s) x|^2.
__global__ void k_lm_diagonal(int n, const double* __restrict__ colsq, const double* __restrict__ scale, const LmState* __restrict__ st,
                              double min_diag, double max_diag, double* __restrict__ D) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double s = scale[i];
    const double d = fmin(fmax(colsq[i] * s * s, min_diag), max_diag);
    D[i] = sqrt(d / st->radius) / s;
}

// step = -y (:123), candidate = x + step (Evaluator::Plus without local parameterisations: the reference's pose
// blocks are plain 6-vectors, Optimizer.cpp:771-773), partial sums of |step|^2 over the free parameters.
__global__ void __launch_bounds__(kLmRedThreads) k_lm_candidate(int K, int M, const int* __restrict__ free_map, const unsigned char* __restrict__ point_fixed,
                                                                const double* __restrict__ xsol, const double* __restrict__ dl,
                                                                const double* __restrict__ pose_delta, const double* __restrict__ points,
                                                                double* __restrict__ step, double* __restrict__ cand_pose_delta, double* __restrict__ cand_points,
                                                                double* __restrict__ partials) {
    const int np = 6 * K, n = np + 3 * M;
    double s2 = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double d = 0.0;
        if (i < np) {
            const int k = i / 6, fm = free_map[k];
            if (fm >= 0) d = -xsol[6 * fm + (i - 6 * k)];
            cand_pose_delta[i] = pose_delta[i] + d;
        } else {
            const int j = i - np;
            if (point_fixed[j / 3] == 0) d = -dl[j];
            cand_points[j] = points[j] + d;
        }
        step[i] = d;
        s2 += d * d;
    }
    const double t = lm_block_sum(s2);
    if (threadIdx.x == 0) partials[blockIdx.x] = t;
}

// model_cost_change = -(J step)'(f + J step / 2) (trust_region_minimizer.cc:430-441), J and f after the loss
// correction, one thread per observation; partial sums per CTA.
__global__ void __launch_bounds__(128) k_lm_model_change(BaDev d, bool apply_loss, const double* __restrict__ step, double* __restrict__ partials) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double acc = 0.0;
    if (i < d.O) {
        const int p = d.pose[i], l = d.point[i];
        double pose[kPoseStride];
#pragma unroll
        for (int j = 0; j < kPoseStride; ++j) pose[j] = __ldg(d.posebuf + kPoseStride * (size_t)p + j);
        const double Pw[3] = {__ldg(d.points + 3 * (size_t)l), __ldg(d.points + 3 * (size_t)l + 1), __ldg(d.points + 3 * (size_t)l + 2)};
        double r[2], F[12], E[6];
        obs_evaluate<true>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
        const double* sp = step + 6 * (size_t)p;
        const double* sl = step + 6 * (size_t)d.K + 3 * (size_t)l;
        double m0 = 0.0, m1 = 0.0;
        if (d.free_map[p] >= 0) {
#pragma unroll
            for (int k = 0; k < 6; ++k) { m0 += F[k] * sp[k]; m1 += F[6 + k] * sp[k]; }
        }
        if (d.point_fixed[l] == 0) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { m0 += E[k] * sl[k]; m1 += E[3 + k] * sl[k]; }
        }
        acc = -(m0 * (r[0] + 0.5 * m0) + m1 * (r[1] + 0.5 * m1));
    }
    block_sum_to(acc, partials);
}

// y += J x for the Ceres SparseMatrix adapter (LinearOperator::RightMultiply; used by the trust-region loop for the
// model cost change): x is 6K + 3M (constant blocks ignored), y is O x 2 in the caller's observation order.
__global__ void __launch_bounds__(128) k_jacobian_right_multiply(BaDev d, bool apply_loss, const double* __restrict__ x, double* __restrict__ y) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d.O) return;
    const int p = d.pose[i], l = d.point[i];
    double pose[kPoseStride];
#pragma unroll
    for (int j = 0; j < kPoseStride; ++j) pose[j] = __ldg(d.posebuf + kPoseStride * (size_t)p + j);
    const double Pw[3] = {__ldg(d.points + 3 * (size_t)l), __ldg(d.points + 3 * (size_t)l + 1), __ldg(d.points + 3 * (size_t)l + 2)};
    double r[2], F[12], E[6];
    obs_evaluate<true>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
    double m0 = 0.0, m1 = 0.0;
    if (d.free_map[p] >= 0) {
#pragma unroll
        for (int k = 0; k < 6; ++k) { m0 += F[k] * x[6 * (size_t)p + k]; m1 += F[6 + k] * x[6 * (size_t)p + k]; }
    }
    if (d.point_fixed[l] == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { m0 += E[k] * x[6 * (size_t)d.K + 3 * (size_t)l + k]; m1 += E[3 + k] * x[6 * (size_t)d.K + 3 * (size_t)l + k]; }
    }
    const size_t o = (size_t)d.orig[i];
    y[2 * o] += m0; y[2 * o + 1] += m1;
}

// cost of the candidate point (Evaluator::Evaluate with residuals / jacobian null): partial sums per CTA
__global__ void __launch_bounds__(128) k_lm_cost(BaDev d, bool apply_loss, double* __restrict__ partials) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double cost = 0.0;
    if (i < d.O) {
        const int p = d.pose[i], l = d.point[i];
        double pose[kPoseStride];
#pragma unroll
        for (int j = 0; j < kPoseStride; ++j) pose[j] = __ldg(d.posebuf + kPoseStride * (size_t)p + j);
        const double Pw[3] = {__ldg(d.points + 3 * (size_t)l), __ldg(d.points + 3 * (size_t)l + 1), __ldg(d.points + 3 * (size_t)l + 2)};
        double r[2], F[1], E[1];
        cost = obs_evaluate<false>(pose, Pw, d.u[i], d.v[i], d.outlier[i] != 0, d.cam, apply_loss, r, F, E);
    }
    block_sum_to(cost, partials);
}

// The decision of one trust-region iteration (trust_region_minimizer.cc:95-138 and the Finalize... checks :302-356).
// sums[0] = sum |step|^2 partials (n0), sums[1] = model-change partials (n1), sums[2] = candidate-cost partials (n2).
__global__ void __launch_bounds__(kLmRedThreads) k_lm_decide(const double* __restrict__ p0, int n0, const double* __restrict__ p1, int n1,
                                                             const double* __restrict__ p2, int n2, const int* __restrict__ chol_status,
                                                             LmOptionsDev opt, int max_iterations, LmState* __restrict__ st, LmState* __restrict__ host_copy) {
    double a = 0.0, b = 0.0, c = 0.0;
    for (int i = threadIdx.x; i < n0; i += blockDim.x) a += p0[i];
    for (int i = threadIdx.x; i < n1; i += blockDim.x) b += p1[i];
    for (int i = threadIdx.x; i < n2; i += blockDim.x) c += p2[i];
    const double step2 = lm_block_sum(a), model = lm_block_sum(b), cand = lm_block_sum(c);
    if (threadIdx.x != 0) return;
    LmState s = *st;
    if (s.termination != 0) { *host_copy = s; return; }   // the gradient test after the last re-linearisation ended the solve
    s.iteration += 1;
    s.step_norm = sqrt(step2); s.model_cost_change = model; s.candidate_cost = cand;
    s.step_valid = (chol_status[0] == 0) && (model > 0.0);        // LINEAR_SOLVER_FAILURE or model_cost_change <= 0
    s.step_successful = 0; s.relinearize = 0;
    s.relative_decrease = 0.0; s.cost_change = 0.0;
    if (!s.step_valid) {                                           // HandleInvalidStep (:452-487)
        s.num_invalid += 1;
        if (s.num_invalid >= opt.max_invalid) s.termination = -1;
        else { s.radius = s.radius / s.decrease_factor; s.decrease_factor *= 2.0; }   // StepIsInvalid = StepRejected(0)
        s.num_unsuccessful += 1;
    } else {
        s.num_invalid = 0;
        const double step_tol = opt.parameter_tolerance * (s.x_norm + opt.parameter_tolerance);
        s.cost_change = s.x_cost - cand;
        if (s.step_norm <= step_tol) s.termination = 3;                                          // ParameterToleranceReached (:717-738)
        else if (fabs(s.cost_change) <= opt.function_tolerance * s.x_cost) s.termination = 2;   // FunctionToleranceReached (:741-761)
        else {
            s.relative_decrease = s.cost_change / model;                                         // StepQuality, monotonic
            if (s.relative_decrease > opt.min_relative_decrease) {                                // HandleSuccessfulStep
                s.step_successful = 1; s.relinearize = 1; s.num_successful += 1;
                s.x_cost = cand;
                const double q = 2.0 * s.relative_decrease - 1.0;
                s.radius = s.radius / fmax(1.0 / 3.0, 1.0 - q * q * q);
                s.radius = fmin(opt.max_radius, s.radius);
                s.decrease_factor = 2.0;
            } else {
                s.radius = s.radius / s.decrease_factor; s.decrease_factor *= 2.0;
                s.num_unsuccessful += 1;
            }
        }
    }
    // FinalizeIterationAndCheckIfMinimizerCanContinue; the gradient test of a successful step needs the NEW gradient and
    // is made by the host after the re-linearisation
    if (s.termination == 0) {
        if (s.iteration >= max_iterations) s.termination = 5;
        else if (s.radius <= opt.min_radius) s.termination = 4;
    }
    *st = s;
    *host_copy = s;
}

}  // namespace vio360
