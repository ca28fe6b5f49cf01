// ba_math.cuh -- per-pose and per-observation arithmetic of the BA factor path, written
// once as host/device inline code.  The CUDA kernels in ba.cu are the only product
// callers; tests/emu compiles the same header with g++ to check the arithmetic against
// the oracle on machines without a GPU (that harness is test code, not a fallback).
//
// Reference semantics restated here (reference-root relative):
//   pose chain          src/optimization/Factors.cpp:357-364, src/util/LieUtils.cpp:203-333
//   residual/Jacobians  src/optimization/Factors.cpp:369-534 (BA), :65-199 (PnP)
//   Huber + Corrector   thirdparty/ceres-solver/internal/ceres/loss_function.cc:48-62,
//                       corrector.cc:84-88 (rho'' <= 0 branch), residual_block.cc:161-195
#pragma once

#include <math.h>

#if defined(__CUDACC__)
#define VIO_HD __host__ __device__ __forceinline__
#else
#define VIO_HD inline
#endif

namespace vio360 {

constexpr double kPi = 3.14159265358979323846;
constexpr double kInvPi = 0.31830988618379067154;
constexpr double kInv2Pi = 0.15915494309189533577;

// 1/sqrt(x): the device intrinsic is ~1 ulp; every use below is followed by a 1e-9 tolerance.
VIO_HD double vio_rsqrt(double x) {
#if defined(__CUDA_ARCH__)
    return rsqrt(x);
#else
    return 1.0 / sqrt(x);
#endif
}
constexpr int kPoseStride = 33;  // doubles per prepared pose: Rcw 9, tcw 3, Rbw 9, tbw 3, Mcp 9

// ---------------------------------------------------------------- 3x3 helpers
VIO_HD void m3_mul(const double* A, const double* B, double* C) {
    double T[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            T[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
#pragma unroll
    for (int i = 0; i < 9; ++i) C[i] = T[i];
}
VIO_HD void m3_mul_bt(const double* A, const double* B, double* C) {
    double T[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            T[i * 3 + j] = A[i * 3] * B[j * 3] + A[i * 3 + 1] * B[j * 3 + 1] + A[i * 3 + 2] * B[j * 3 + 2];
#pragma unroll
    for (int i = 0; i < 9; ++i) C[i] = T[i];
}
VIO_HD void m3_vec(const double* A, const double* v, double* o) {
    const double a = A[0] * v[0] + A[1] * v[1] + A[2] * v[2];
    const double b = A[3] * v[0] + A[4] * v[1] + A[5] * v[2];
    const double c = A[6] * v[0] + A[7] * v[1] + A[8] * v[2];
    o[0] = a; o[1] = b; o[2] = c;
}
VIO_HD double m3_det(const double* A) {
    return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) +
           A[2] * (A[3] * A[7] - A[4] * A[6]);
}

// Closest rotation in the sense of SO3d::SO3d(const Matrix3d&) (LieUtils.cpp:275-288):
// U V^T of the SVD, third column of U negated when the determinant is negative.
// One-sided Jacobi on the columns; singular values sorted descending so that "column
// 2" is the smallest singular direction, as with Eigen::JacobiSVD.
VIO_HD void so3_project(const double* M, double* R) {
    double B[9], V[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) { B[i] = M[i]; V[i] = (i % 4 == 0) ? 1.0 : 0.0; }
    for (int sweep = 0; sweep < 60; ++sweep) {
        double off = 0.0;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                double alpha = 0, beta = 0, gamma = 0;
                for (int k = 0; k < 3; ++k) {
                    alpha += B[k * 3 + p] * B[k * 3 + p];
                    beta += B[k * 3 + q] * B[k * 3 + q];
                    gamma += B[k * 3 + p] * B[k * 3 + q];
                }
                if (gamma == 0.0) continue;
                const double lim = sqrt(alpha * beta);
                if (fabs(gamma) <= 1e-17 * lim) continue;
                off = fmax(off, fabs(gamma) / (lim > 0 ? lim : 1.0));
                const double zeta = (beta - alpha) / (2.0 * gamma);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int k = 0; k < 3; ++k) {
                    const double bp = B[k * 3 + p], bq = B[k * 3 + q];
                    B[k * 3 + p] = c * bp - s * bq;
                    B[k * 3 + q] = s * bp + c * bq;
                    const double vp = V[k * 3 + p], vq = V[k * 3 + q];
                    V[k * 3 + p] = c * vp - s * vq;
                    V[k * 3 + q] = s * vp + c * vq;
                }
            }
        if (off < 1e-16) break;
    }
    double sv[3];
    for (int j = 0; j < 3; ++j)
        sv[j] = sqrt(B[j] * B[j] + B[3 + j] * B[3 + j] + B[6 + j] * B[6 + j]);
    int o0 = 0, o1 = 1, o2 = 2, tmp;
    if (sv[o1] > sv[o0]) { tmp = o0; o0 = o1; o1 = tmp; }
    if (sv[o2] > sv[o0]) { tmp = o0; o0 = o2; o2 = tmp; }
    if (sv[o2] > sv[o1]) { tmp = o1; o1 = o2; o2 = tmp; }
    const int ord[3] = {o0, o1, o2};
    double U[9], Vs[9];
    for (int j = 0; j < 3; ++j) {
        const int c = ord[j];
        for (int k = 0; k < 3; ++k) {
            U[k * 3 + j] = (sv[c] > 0.0) ? B[k * 3 + c] / sv[c] : 0.0;
            Vs[k * 3 + j] = V[k * 3 + c];
        }
    }
    if (!(sv[ord[2]] > 0.0) && sv[ord[1]] > 0.0) {
        U[2] = U[3] * U[7] - U[6] * U[4];
        U[5] = U[6] * U[1] - U[0] * U[7];
        U[8] = U[0] * U[4] - U[3] * U[1];
    }
    m3_mul_bt(U, Vs, R);
    if (m3_det(R) < 0.0) {
        U[2] = -U[2]; U[5] = -U[5]; U[8] = -U[8];
        m3_mul_bt(U, Vs, R);
    }
}

VIO_HD void hat3(const double* v, double* S) {
    S[0] = 0.0; S[1] = -v[2]; S[2] = v[1];
    S[3] = v[2]; S[4] = 0.0; S[5] = -v[0];
    S[6] = -v[1]; S[7] = v[0]; S[8] = 0.0;
}

// SE3d::exp (LieUtils.cpp:305-333) incl. SO3d::Exp (:203-219)
VIO_HD void se3_exp(const double* xi, double* R, double* t) {
    const double* rho = xi;
    const double* phi = xi + 3;
    const double theta = sqrt(phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2]);
    double M[9], K[9], KK[9];
    if (theta < 1e-10) {
        hat3(phi, K);
        for (int i = 0; i < 9; ++i) M[i] = K[i] + ((i % 4 == 0) ? 1.0 : 0.0);
        so3_project(M, R);
        t[0] = rho[0]; t[1] = rho[1]; t[2] = rho[2];
        return;
    }
    const double theta_inv = 1.0 / theta;
    const double k[3] = {phi[0] * theta_inv, phi[1] * theta_inv, phi[2] * theta_inv};
    hat3(k, K);
    m3_mul(K, K, KK);
    const double st = sin(theta), ct = cos(theta);
    for (int i = 0; i < 9; ++i) M[i] = ((i % 4 == 0) ? 1.0 : 0.0) + st * K[i] + (1.0 - ct) * KK[i];
    so3_project(M, R);
    double P[9], PP[9], Vm[9];
    hat3(phi, P);
    m3_mul(P, P, PP);
    const double theta2 = theta * theta;
    const double a = (1.0 - ct) / theta2, b = (theta - st) / (theta2 * theta);
    for (int i = 0; i < 9; ++i) Vm[i] = ((i % 4 == 0) ? 1.0 : 0.0) + a * P[i] + b * PP[i];
    m3_vec(Vm, rho, t);
}

// The per-pose chain of Factors.cpp:357-364, once per pose.  out: kPoseStride doubles.
VIO_HD void pose_prepare(const double* Twb_init /*16*/, const double* Tcb /*16*/,
                         const double* xi /*6*/, double* out) {
    double R0[9], M[9], Rd[9], td[3], Rwb[9], twb[3], v[3];
    const double R0raw[9] = {Twb_init[0], Twb_init[1], Twb_init[2], Twb_init[4], Twb_init[5],
                             Twb_init[6], Twb_init[8], Twb_init[9], Twb_init[10]};
    so3_project(R0raw, R0);                         // SE3d T_wb_init(m_T_wb_init)
    se3_exp(xi, Rd, td);                            // SE3d::exp(delta_xi)
    m3_mul(R0, Rd, M);
    so3_project(M, Rwb);                            // T_wb_init * delta_T
    m3_vec(R0, td, v);
    twb[0] = Twb_init[3] + v[0]; twb[1] = Twb_init[7] + v[1]; twb[2] = Twb_init[11] + v[2];
    double* Rcw = out; double* tcw = out + 9; double* Rbw = out + 12; double* tbw = out + 21;
    double* Mcp = out + 24;
    const double Rt[9] = {Rwb[0], Rwb[3], Rwb[6], Rwb[1], Rwb[4], Rwb[7], Rwb[2], Rwb[5], Rwb[8]};
    so3_project(Rt, Rbw);                           // Twb.inverse()
    const double nt[3] = {-twb[0], -twb[1], -twb[2]};
    m3_vec(Rbw, nt, tbw);
    const double Rcb_raw[9] = {Tcb[0], Tcb[1], Tcb[2], Tcb[4], Tcb[5], Tcb[6], Tcb[8], Tcb[9], Tcb[10]};
    double Rcb[9];
    so3_project(Rcb_raw, Rcb);                      // SE3d Tcb(m_Tcb)
    m3_mul(Rcb, Rbw, M);
    so3_project(M, Rcw);                            // Tcb * Tbw
    m3_vec(Rcb, tbw, v);
    tcw[0] = Tcb[3] + v[0]; tcw[1] = Tcb[7] + v[1]; tcw[2] = Tcb[11] + v[2];
    m3_mul(Rcb_raw, Rbw, Mcp);                      // J_camera_point, Factors.cpp:527 (raw m_Tcb block)
}

// Constants shared by every factor of a window.
struct CamConst {
    double cols, rows;
    double c2pi;      // cols / (2 pi)
    double rpi;       // rows / pi
    double Lc[4];     // Eigen::LLT(info).matrixL(), or identity when the LLT fails
    double info[4];
    double Rcb[9];    // raw rotation block of m_Tcb (Jacobians use the un-projected one)
    double huber_a;   // <= 0 : no loss
};

VIO_HD void cam_const_init(CamConst& c, double cols, double rows, const double* info,
                           const double* Tcb, double huber_a) {
    c.cols = cols; c.rows = rows;
    c.c2pi = cols / (2.0 * kPi);
    c.rpi = rows / kPi;
    for (int i = 0; i < 4; ++i) c.info[i] = info[i];
    c.Lc[0] = 1.0; c.Lc[1] = 0.0; c.Lc[2] = 0.0; c.Lc[3] = 1.0;
    if (info[0] > 0.0) {   // Factors.cpp:444-455 : unweighted when the LLT fails
        const double l00 = sqrt(info[0]), l10 = info[2] / l00, d = info[3] - l10 * l10;
        if (d > 0.0) { c.Lc[0] = l00; c.Lc[1] = 0.0; c.Lc[2] = l10; c.Lc[3] = sqrt(d); }
    }
    const double R[9] = {Tcb[0], Tcb[1], Tcb[2], Tcb[4], Tcb[5], Tcb[6], Tcb[8], Tcb[9], Tcb[10]};
    for (int i = 0; i < 9; ++i) c.Rcb[i] = R[i];
    c.huber_a = huber_a;
}

// Scalar-type helpers: the per-observation math below is written once for T = double (the parity path) and
// T = float (the optional fp32 mode of the north star, tolerance 1e-4).
VIO_HD float vio_rsqrt(float x) {
#ifdef __CUDA_ARCH__
    return rsqrtf(x);
#else
    return 1.0f / sqrtf(x);
#endif
}
VIO_HD double m_atan2(double y, double x) { return atan2(y, x); }
VIO_HD float m_atan2(float y, float x) { return atan2f(y, x); }
VIO_HD double m_asin(double x) { return asin(x); }
VIO_HD float m_asin(float x) { return asinf(x); }
VIO_HD double m_sqrt(double x) { return sqrt(x); }
VIO_HD float m_sqrt(float x) { return sqrtf(x); }
VIO_HD double m_abs(double x) { return fabs(x); }
VIO_HD float m_abs(float x) { return fabsf(x); }
VIO_HD double m_tiny(double) { return 2.2250738585072014e-308; }   // std::numeric_limits<T>::min()
VIO_HD float m_tiny(float) { return 1.17549435e-38f; }

// Pixel error of one observation: returns branch 0 ok / 1 tiny depth / 2 large error.
template <typename T>
struct ObsProjT { T x, y, z, L_sq, inv_L, du, dv; int branch; };
using ObsProj = ObsProjT<double>;

// Divisions and square roots of Factors.cpp:369-419 are restated with one reciprocal square
// root and multiplications (the FP64 pipe has no divider; a division costs ~17 DFMA).
template <typename T>
VIO_HD void obs_project(const T* pose, const T* Pw, T u_obs, T v_obs, const CamConst& c, ObsProjT<T>& p) {
    const T* Rcw = pose; const T* tcw = pose + 9;
    const T cols = (T)c.cols, rows = (T)c.rows;
    p.x = Rcw[0] * Pw[0] + Rcw[1] * Pw[1] + Rcw[2] * Pw[2] + tcw[0];
    p.y = Rcw[3] * Pw[0] + Rcw[4] * Pw[1] + Rcw[5] * Pw[2] + tcw[1];
    p.z = Rcw[6] * Pw[0] + Rcw[7] * Pw[1] + Rcw[8] * Pw[2] + tcw[2];
    p.L_sq = p.x * p.x + p.y * p.y + p.z * p.z;
    p.branch = 0; p.du = T(0); p.dv = T(0); p.inv_L = T(0);
    if (p.L_sq < T(1e-20)) { p.branch = 1; return; }          // L < 1e-10
    p.inv_L = vio_rsqrt(p.L_sq);
    const T theta = m_atan2(p.x, p.z);
    const T phi = -m_asin(p.y * p.inv_L);
    const T u = cols * (T(0.5) + theta * (T)kInv2Pi);
    const T v = rows * (T(0.5) - phi * (T)kInvPi);
    T du = u_obs - u;
    if (du > cols * T(0.5)) du -= cols;
    else if (du < -cols * T(0.5)) du += cols;
    p.du = du; p.dv = v_obs - v;
    if (m_abs(p.du) > T(100) || m_abs(p.dv) > T(100)) p.branch = 2;
}

// BAFactor::Evaluate for one observation with a prepared pose, followed by the Huber
// correction.  F = d r / d pose (2x6 row-major), E = d r / d point (2x3 row-major).
// returns the block's cost 0.5 * rho(s).
template <bool WANT_JAC, typename T>
VIO_HD T obs_evaluate(const T* pose, const T* Pw, T u_obs, T v_obs, bool outlier, const CamConst& c, bool apply_loss,
                      T* r, T* F, T* E) {
    bool zero_j = false;
    const T Lc0 = (T)c.Lc[0], Lc1 = (T)c.Lc[1], Lc2 = (T)c.Lc[2], Lc3 = (T)c.Lc[3];
    if (outlier) {                       // Factors.cpp:330-344
        r[0] = T(640); r[1] = T(480); zero_j = true;
    } else {
        ObsProjT<T> p;
        obs_project(pose, Pw, u_obs, v_obs, c, p);
        if (p.branch == 1) { r[0] = T(640); r[1] = T(360); zero_j = true; }        // :378-394
        else if (p.branch == 2) { r[0] = T(100); r[1] = T(100); zero_j = true; }   // :427-441
        else {
            r[0] = Lc0 * p.du + Lc1 * p.dv;
            r[1] = Lc2 * p.du + Lc3 * p.dv;
            if (WANT_JAC) {
                const T xz_sq = p.x * p.x + p.z * p.z;
                const T L_sq = p.L_sq;
                if (xz_sq < T(1e-10) || L_sq < T(1e-10)) zero_j = true;                // :466-476
                else {
                    const T c2pi = (T)c.c2pi, rpi = (T)c.rpi;
                    const T inv_xz = vio_rsqrt(xz_sq);
                    const T xz_norm = xz_sq * inv_xz;
                    const T inv_xz_sq = inv_xz * inv_xz, inv_L_sq = p.inv_L * p.inv_L;
                    const T inv_den = inv_L_sq * inv_xz;                  // 1 / (L^2 * |xz|)
                    T Jp0[3], Jp1[3];
                    Jp0[0] = -c2pi * p.z * inv_xz_sq; Jp0[1] = T(0); Jp0[2] = c2pi * p.x * inv_xz_sq;
                    Jp1[0] = rpi * (p.x * p.y) * inv_den;
                    Jp1[1] = -rpi * xz_norm * inv_L_sq;
                    Jp1[2] = rpi * (p.y * p.z) * inv_den;
                    T Jw0[3], Jw1[3];
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        Jw0[j] = Lc0 * Jp0[j] + Lc1 * Jp1[j];
                        Jw1[j] = Lc2 * Jp0[j] + Lc3 * Jp1[j];
                    }
                    const T* Rbw = pose + 12; const T* tbw = pose + 21;
                    const T* Mcp = pose + 24;
                    const T Pb0 = Rbw[0] * Pw[0] + Rbw[1] * Pw[1] + Rbw[2] * Pw[2] + tbw[0];
                    const T Pb1 = Rbw[3] * Pw[0] + Rbw[4] * Pw[1] + Rbw[5] * Pw[2] + tbw[1];
                    const T Pb2 = Rbw[6] * Pw[0] + Rbw[7] * Pw[1] + Rbw[8] * Pw[2] + tbw[2];
                    // A = Jw * Rcb ; translation block = -A ; rotation block rows = a x Pb
                    T A0[3], A1[3];
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        A0[j] = Jw0[0] * (T)c.Rcb[j] + Jw0[1] * (T)c.Rcb[3 + j] + Jw0[2] * (T)c.Rcb[6 + j];
                        A1[j] = Jw1[0] * (T)c.Rcb[j] + Jw1[1] * (T)c.Rcb[3 + j] + Jw1[2] * (T)c.Rcb[6 + j];
                    }
                    F[0] = -A0[0]; F[1] = -A0[1]; F[2] = -A0[2];
                    F[3] = A0[1] * Pb2 - A0[2] * Pb1;
                    F[4] = A0[2] * Pb0 - A0[0] * Pb2;
                    F[5] = A0[0] * Pb1 - A0[1] * Pb0;
                    F[6] = -A1[0]; F[7] = -A1[1]; F[8] = -A1[2];
                    F[9] = A1[1] * Pb2 - A1[2] * Pb1;
                    F[10] = A1[2] * Pb0 - A1[0] * Pb2;
                    F[11] = A1[0] * Pb1 - A1[1] * Pb0;
#pragma unroll
                    for (int j = 0; j < 3; ++j) {
                        E[j] = Jw0[0] * Mcp[j] + Jw0[1] * Mcp[3 + j] + Jw0[2] * Mcp[6 + j];
                        E[3 + j] = Jw1[0] * Mcp[j] + Jw1[1] * Mcp[3 + j] + Jw1[2] * Mcp[6 + j];
                    }
                }
            }
        }
    }
    if (WANT_JAC && zero_j) {
#pragma unroll
        for (int j = 0; j < 12; ++j) F[j] = T(0);
#pragma unroll
        for (int j = 0; j < 6; ++j) E[j] = T(0);
    }
    // ResidualBlock::Evaluate tail: cost + Corrector (rho'' <= 0 for Huber)
    const T s = r[0] * r[0] + r[1] * r[1];
    const T ha = (T)c.huber_a;
    if (!apply_loss || !(c.huber_a > 0.0)) return T(0.5) * s;
    const T b = ha * ha;
    if (s > b) {
        const T rs = vio_rsqrt(s), rt = s * rs;
        const T rho0 = T(2) * ha * rt - b;
        const T rho1c = ha * rs;
        const T rho1 = rho1c > m_tiny(T(0)) ? rho1c : m_tiny(T(0));
        const T sc = m_sqrt(rho1);
        if (WANT_JAC && !zero_j) {
#pragma unroll
            for (int j = 0; j < 12; ++j) F[j] *= sc;
#pragma unroll
            for (int j = 0; j < 6; ++j) E[j] *= sc;
        }
        r[0] *= sc; r[1] *= sc;
        return T(0.5) * rho0;
    }
    return T(0.5) * s;
}

// BAFactor::compute_chi_square (Factors.cpp:544-612) with a prepared pose.
VIO_HD double obs_chi2(const double* pose, const double* Pw, double u_obs, double v_obs,
                       bool outlier, const CamConst& c) {
    if (outlier) return 0.0;
    ObsProj p;
    obs_project(pose, Pw, u_obs, v_obs, c, p);
    if (p.branch == 1) return 1000.0;
    const double a0 = p.du * c.info[0] + p.dv * c.info[2];
    const double a1 = p.du * c.info[1] + p.dv * c.info[3];
    return a0 * p.du + a1 * p.dv;
}

// Inverse of a symmetric positive definite 3x3 given by its upper triangle
// m = {a00,a01,a02,a11,a12,a22} via Cholesky (InvertPSDMatrix, invert_psd_matrix.h).
VIO_HD void inv_spd3(const double* m, double* inv /*6, same packing*/) {
    const double l00 = sqrt(m[0]);
    const double l10 = m[1] / l00, l20 = m[2] / l00;
    const double l11 = sqrt(m[3] - l10 * l10);
    const double l21 = (m[4] - l20 * l10) / l11;
    const double l22 = sqrt(m[5] - l20 * l20 - l21 * l21);
    const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
    const double i10 = -l10 * i00 * i11;
    const double i21 = -l21 * i11 * i22;
    const double i20 = -(l20 * i00 + l21 * i10) * i22;
    inv[0] = i00 * i00 + i10 * i10 + i20 * i20;
    inv[1] = i10 * i11 + i20 * i21;
    inv[2] = i20 * i22;
    inv[3] = i11 * i11 + i21 * i21;
    inv[4] = i21 * i22;
    inv[5] = i22 * i22;
}

}  // namespace vio360
