// tests/emu/ba_math_emu.cpp -- TEST HARNESS ONLY (never linked into libvio360_cuda.so).
// Compiles csrc/ba_math.cuh, the arithmetic the CUDA kernels execute, with the host
// compiler so that its results can be compared with the oracle on a machine without a
// GPU.  It mirrors what k_pose_prepare / k_evaluate / k_chi2 do per thread.
#include "../../360_visual_inertial_odometry_b200/csrc/ba_math.cuh"

using namespace vio360;

extern "C" {

void emu_pose_prepare(int K, const double* Twb, const double* Tcb, const double* delta, double* out) {
    for (int k = 0; k < K; ++k) pose_prepare(Twb + 16 * k, Tcb, delta + 6 * k, out + kPoseStride * k);
}

double emu_evaluate(int O, const int* pose_idx, const int* point_idx, const double* u, const double* v,
                    const unsigned char* outlier, const double* posebuf, const double* points, double cols,
                    double rows, const double* info, const double* Tcb, double huber_a, int apply_loss,
                    double* r, double* Jp, double* Jl, double* chi2) {
    CamConst cam;
    cam_const_init(cam, cols, rows, info, Tcb, huber_a);
    double cost = 0.0;
    for (int i = 0; i < O; ++i) {
        const double* pose = posebuf + kPoseStride * pose_idx[i];
        const double* Pw = points + 3 * point_idx[i];
        cost += obs_evaluate<true>(pose, Pw, u[i], v[i], outlier && outlier[i], cam, apply_loss != 0,
                                   r + 2 * i, Jp + 12 * i, Jl + 6 * i);
        if (chi2) chi2[i] = obs_chi2(pose, Pw, u[i], v[i], outlier && outlier[i], cam);
    }
    return cost;
}

// the fp32 mode (k_evaluate_f32): poses, points and observations rounded to float, float math, widened results
double emu_evaluate_f32(int O, const int* pose_idx, const int* point_idx, const double* u, const double* v,
                        const unsigned char* outlier, const double* posebuf, const double* points, double cols,
                        double rows, const double* info, const double* Tcb, double huber_a, int apply_loss,
                        double* r, double* Jp, double* Jl) {
    CamConst cam;
    cam_const_init(cam, cols, rows, info, Tcb, huber_a);
    double cost = 0.0;
    for (int i = 0; i < O; ++i) {
        float pose[kPoseStride], Pw[3], rf[2], F[12], E[6];
        for (int j = 0; j < kPoseStride; ++j) pose[j] = (float)posebuf[kPoseStride * pose_idx[i] + j];
        for (int j = 0; j < 3; ++j) Pw[j] = (float)points[3 * point_idx[i] + j];
        cost += (double)obs_evaluate<true>(pose, Pw, (float)u[i], (float)v[i], outlier && outlier[i], cam, apply_loss != 0, rf, F, E);
        r[2 * i] = rf[0]; r[2 * i + 1] = rf[1];
        for (int j = 0; j < 12; ++j) Jp[12 * i + j] = F[j];
        for (int j = 0; j < 6; ++j) Jl[6 * i + j] = E[j];
    }
    return cost;
}

void emu_so3_project(const double* M, double* R) { so3_project(M, R); }
void emu_inv_spd3(const double* m6, double* inv6) { inv_spd3(m6, inv6); }
}
