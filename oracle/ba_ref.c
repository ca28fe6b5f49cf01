/*
 * oracle/ba_ref.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  See ba_ref.h.
 * PARITY UNPINNED (no reference tests / golden vectors; reference not buildable here).
 *
 * Plain C restatement of:
 *   src/util/LieUtils.cpp:203-219, 275-288, 292-295, 305-333   (SO3d / SE3d, double)
 *   src/util/LieUtils.h:207-219, 268-283                      (compose / inverse)
 *   src/optimization/Factors.cpp:33-210, 212-265, 327-542, 544-612
 *   thirdparty/ceres-solver/internal/ceres/loss_function.cc:48-62      (HuberLoss)
 *   thirdparty/ceres-solver/internal/ceres/corrector.cc:42-160         (Corrector)
 *   thirdparty/ceres-solver/internal/ceres/residual_block.cc:133-197
 *   thirdparty/ceres-solver/internal/ceres/schur_eliminator_impl.h:179-560
 *   thirdparty/ceres-solver/internal/ceres/levenberg_marquardt_strategy.cc:78-89
 */
#include "ba_ref.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

static const double kEpsilonD = 1e-10; /* LieUtils.h:24 */

/* ------------------------------------------------------------------ 3x3 helpers */
static void mat3_mul(const double A[9], const double B[9], double C[9]) {
    double T[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            T[i * 3 + j] = A[i * 3 + 0] * B[0 * 3 + j] + A[i * 3 + 1] * B[1 * 3 + j] +
                           A[i * 3 + 2] * B[2 * 3 + j];
    memcpy(C, T, sizeof(T));
}
static void mat3_mul_bt(const double A[9], const double B[9], double C[9]) { /* A*B^T */
    double T[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j)
            T[i * 3 + j] = A[i * 3 + 0] * B[j * 3 + 0] + A[i * 3 + 1] * B[j * 3 + 1] +
                           A[i * 3 + 2] * B[j * 3 + 2];
    memcpy(C, T, sizeof(T));
}
static void mat3_vec(const double A[9], const double v[3], double o[3]) {
    double t0 = A[0] * v[0] + A[1] * v[1] + A[2] * v[2];
    double t1 = A[3] * v[0] + A[4] * v[1] + A[5] * v[2];
    double t2 = A[6] * v[0] + A[7] * v[1] + A[8] * v[2];
    o[0] = t0; o[1] = t1; o[2] = t2;
}
static double mat3_det(const double A[9]) {
    return A[0] * (A[4] * A[8] - A[5] * A[7]) - A[1] * (A[3] * A[8] - A[5] * A[6]) +
           A[2] * (A[3] * A[7] - A[4] * A[6]);
}
static void hat3(const double v[3], double S[9]) { /* LieUtils.cpp:365-371 (Hatd) */
    S[0] = 0.0;   S[1] = -v[2]; S[2] = v[1];
    S[3] = v[2];  S[4] = 0.0;   S[5] = -v[0];
    S[6] = -v[1]; S[7] = v[0];  S[8] = 0.0;
}

/* One-sided (Hestenes) Jacobi SVD of a 3x3, singular values sorted descending.
 * Eigen::JacobiSVD (two-sided, QR-preconditioned) returns the same factorisation up
 * to simultaneous column signs of U and V and ~1e-16 rounding; U*V^T, the only
 * quantity the reference consumes (LieUtils.cpp:281,286), is invariant to both.   */
void vo_svd3(const double A[9], double U[9], double S[3], double V[9]) {
    double B[9]; /* working copy, columns get orthogonalised: B = A * V */
    memcpy(B, A, sizeof(B));
    for (int i = 0; i < 9; ++i) V[i] = (i % 4 == 0) ? 1.0 : 0.0;
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
                double lim = sqrt(alpha * beta);
                if (fabs(gamma) <= 1e-17 * lim) continue;
                off = fmax(off, fabs(gamma) / (lim > 0 ? lim : 1.0));
                double zeta = (beta - alpha) / (2.0 * gamma);
                double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
                for (int k = 0; k < 3; ++k) {
                    double bp = B[k * 3 + p], bq = B[k * 3 + q];
                    B[k * 3 + p] = c * bp - s * bq;
                    B[k * 3 + q] = s * bp + c * bq;
                    double vp = V[k * 3 + p], vq = V[k * 3 + q];
                    V[k * 3 + p] = c * vp - s * vq;
                    V[k * 3 + q] = s * vp + c * vq;
                }
            }
        if (off < 1e-16) break;
    }
    double sv[3];
    for (int j = 0; j < 3; ++j)
        sv[j] = sqrt(B[0 * 3 + j] * B[0 * 3 + j] + B[1 * 3 + j] * B[1 * 3 + j] +
                     B[2 * 3 + j] * B[2 * 3 + j]);
    int ord[3] = {0, 1, 2};
    for (int i = 0; i < 2; ++i)
        for (int j = i + 1; j < 3; ++j)
            if (sv[ord[j]] > sv[ord[i]]) { int t = ord[i]; ord[i] = ord[j]; ord[j] = t; }
    double Vs[9];
    for (int j = 0; j < 3; ++j) {
        int c = ord[j];
        S[j] = sv[c];
        for (int k = 0; k < 3; ++k) {
            U[k * 3 + j] = (sv[c] > 0.0) ? B[k * 3 + c] / sv[c] : 0.0;
            Vs[k * 3 + j] = V[k * 3 + c];
        }
    }
    memcpy(V, Vs, sizeof(Vs));
    /* complete U for rank-deficient input so that it stays orthogonal */
    if (S[2] <= 0.0 && S[1] > 0.0) {
        U[0 * 3 + 2] = U[1 * 3 + 0] * U[2 * 3 + 1] - U[2 * 3 + 0] * U[1 * 3 + 1];
        U[1 * 3 + 2] = U[2 * 3 + 0] * U[0 * 3 + 1] - U[0 * 3 + 0] * U[2 * 3 + 1];
        U[2 * 3 + 2] = U[0 * 3 + 0] * U[1 * 3 + 1] - U[1 * 3 + 0] * U[0 * 3 + 1];
    }
}

/* LieUtils.cpp:275-288 */
void vo_so3_project(const double M[9], double R[9]) {
    double U[9], S[3], V[9];
    vo_svd3(M, U, S, V);
    mat3_mul_bt(U, V, R);              /* m_matrix = U * V.transpose()      :281 */
    if (mat3_det(R) < 0.0) {           /* :284 */
        U[2] = -U[2]; U[5] = -U[5]; U[8] = -U[8]; /* U.col(2) *= -1       :285 */
        mat3_mul_bt(U, V, R);          /* :286 */
    }
}

/* LieUtils.cpp:203-219 */
void vo_so3_exp(const double w[3], double R[9]) {
    const double theta = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    double M[9], K[9];
    if (theta < kEpsilonD) {           /* :206-209  R ~ I + hat(omega) */
        hat3(w, K);
        for (int i = 0; i < 9; ++i) M[i] = K[i] + ((i % 4 == 0) ? 1.0 : 0.0);
        vo_so3_project(M, R);
        return;
    }
    const double theta_inv = 1.0 / theta;               /* :211 */
    const double k[3] = {w[0] * theta_inv, w[1] * theta_inv, w[2] * theta_inv};
    double KK[9];
    hat3(k, K);
    mat3_mul(K, K, KK);
    const double s = sin(theta), c1 = 1.0 - cos(theta); /* :216-218 */
    for (int i = 0; i < 9; ++i) M[i] = ((i % 4 == 0) ? 1.0 : 0.0) + s * K[i] + c1 * KK[i];
    vo_so3_project(M, R);
}

/* LieUtils.cpp:292-295 */
void vo_se3_from_matrix(const double T[16], vo_se3* out) {
    double M[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};
    vo_so3_project(M, out->R);
    out->t[0] = T[3]; out->t[1] = T[7]; out->t[2] = T[11];
}

/* LieUtils.cpp:305-333 */
void vo_se3_exp(const double xi[6], vo_se3* out) {
    const double* rho = xi;      /* head<3> : translation part :307 */
    const double* phi = xi + 3;  /* tail<3> : rotation part    :308 */
    vo_so3_exp(phi, out->R);     /* :310 */
    const double theta = sqrt(phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2]);
    if (theta < kEpsilonD) {     /* :315-317 */
        out->t[0] = rho[0]; out->t[1] = rho[1]; out->t[2] = rho[2];
        return;
    }
    double P[9], PP[9], Vm[9];
    hat3(phi, P);
    mat3_mul(P, P, PP);
    const double theta2 = theta * theta, st = sin(theta), ct = cos(theta);
    const double a = (1.0 - ct) / theta2;               /* :325 */
    const double b = (theta - st) / (theta2 * theta);   /* :326 */
    for (int i = 0; i < 9; ++i) Vm[i] = ((i % 4 == 0) ? 1.0 : 0.0) + a * P[i] + b * PP[i];
    mat3_vec(Vm, rho, out->t);                          /* :328 */
}

/* LieUtils.h:268-271: SE3d(m_rotation * other.m_rotation, t + R * other.t);
 * SO3d::operator* re-projects the product (LieUtils.h:207-209).                    */
void vo_se3_mul(const vo_se3* a, const vo_se3* b, vo_se3* out) {
    double M[9], Rt[3];
    mat3_mul(a->R, b->R, M);
    mat3_vec(a->R, b->t, Rt);
    vo_se3 o;
    vo_so3_project(M, o.R);
    for (int i = 0; i < 3; ++i) o.t[i] = a->t[i] + Rt[i];
    *out = o;
}

/* LieUtils.h:279-282: R_inv = SO3d(R^T) (re-projected, :217-219); t = R_inv * (-t) */
void vo_se3_inverse(const vo_se3* a, vo_se3* out) {
    double Mt[9] = {a->R[0], a->R[3], a->R[6], a->R[1], a->R[4], a->R[7],
                    a->R[2], a->R[5], a->R[8]};
    vo_se3 o;
    vo_so3_project(Mt, o.R);
    const double nt[3] = {-a->t[0], -a->t[1], -a->t[2]};
    mat3_vec(o.R, nt, o.t);
    *out = o;
}

/* ------------------------------------------------------------ factor evaluation */
typedef struct { vo_se3 Tbw, Tcw; } vo_chain;

/* The pose chain of Factors.cpp:357-364 (BA) == :48-62 (PnP): six SO(3) projections. */
static void pose_chain(const double Twb_init[16], const double Tcb[16],
                       const double pose6[6], vo_chain* c) {
    vo_se3 T_wb_init, delta_T, Twb, Tcb_se3;
    vo_se3_from_matrix(Twb_init, &T_wb_init);  /* SE3d T_wb_init(m_T_wb_init)     :357 */
    vo_se3_exp(pose6, &delta_T);               /* SE3d::exp(delta_xi)             :358 */
    vo_se3_mul(&T_wb_init, &delta_T, &Twb);    /* right perturbation              :359 */
    vo_se3_inverse(&Twb, &c->Tbw);             /* Tbw = Twb.inverse()             :362 */
    vo_se3_from_matrix(Tcb, &Tcb_se3);         /* SE3d Tcb(m_Tcb)                 :363 */
    vo_se3_mul(&Tcb_se3, &c->Tbw, &c->Tcw);    /* Tcw = Tcb * Tbw                 :364 */
}

/* Eigen::LLT<Matrix2d>: lower Cholesky factor; returns 0 when not positive definite
 * (Eigen::NumericalIssue).                                                          */
static int llt2(const double A[4], double L[4]) {
    if (!(A[0] > 0.0)) return 0;
    const double l00 = sqrt(A[0]);
    const double l10 = A[2] / l00;
    const double d = A[3] - l10 * l10;
    if (!(d > 0.0)) return 0;
    L[0] = l00; L[1] = 0.0; L[2] = l10; L[3] = sqrt(d);
    return 1;
}

typedef struct {
    double pc[3], L, du, dv;
    int branch; /* 0 ok, 1 tiny L, 2 large error */
} vo_proj;

/* projection + wrapped pixel error, Factors.cpp:369-425 (BA) == :65-113 (PnP) */
static void project_error(const vo_factor* f, const vo_se3* Tcw, const double pw[3],
                          vo_proj* p) {
    mat3_vec(Tcw->R, pw, p->pc);
    for (int i = 0; i < 3; ++i) p->pc[i] += Tcw->t[i];
    const double x = p->pc[0], y = p->pc[1], z = p->pc[2];
    p->L = sqrt(x * x + y * y + z * z);
    p->branch = 0;
    if (p->L < 1e-10) { p->branch = 1; return; }
    const double cols = f->cols, rows = f->rows;
    const double theta = atan2(x, z);
    const double phi = -asin(y / p->L);
    const double u = cols * (0.5 + theta / (2.0 * M_PI));
    const double v = rows * (0.5 - phi / M_PI);
    double du = f->obs[0] - u;
    const double dv = f->obs[1] - v;
    if (du > cols / 2.0) du -= cols;
    else if (du < -cols / 2.0) du += cols;
    p->du = du; p->dv = dv;
    if (fabs(du) > 100.0 || fabs(dv) > 100.0) p->branch = 2;
}

/* J_proj_camera, Factors.cpp:481-488 (BA) == :169-176 (PnP) */
static void proj_jacobian(double cols, double rows, const double pc[3], double xz_sq,
                          double L_sq, double xz_norm, double J[6]) {
    const double x = pc[0], y = pc[1], z = pc[2];
    J[0] = -cols / (2.0 * M_PI) * z / xz_sq;
    J[1] = 0.0;
    J[2] = cols / (2.0 * M_PI) * x / xz_sq;
    J[3] = rows / M_PI * (x * y) / (L_sq * xz_norm);
    J[4] = -rows / M_PI * xz_norm / L_sq;
    J[5] = rows / M_PI * (y * z) / (L_sq * xz_norm);
}

static void mat23_mul33(const double A[6], const double B[9], double C[6]) {
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 3; ++j)
            C[i * 3 + j] = A[i * 3 + 0] * B[0 * 3 + j] + A[i * 3 + 1] * B[1 * 3 + j] +
                           A[i * 3 + 2] * B[2 * 3 + j];
}
static void lmul2(const double L[4], const double A[6], double C[6]) { /* L(2x2)*A(2x3) */
    for (int j = 0; j < 3; ++j) {
        C[j] = L[0] * A[j] + L[1] * A[3 + j];
        C[3 + j] = L[2] * A[j] + L[3] * A[3 + j];
    }
}

/* Factors.cpp:327-542 */
int vo_ba_evaluate(const vo_factor* f, const double pose6[6], const double pt[3],
                   double r[2], int want_jac, double* jac_pose, double* jac_point) {
    if (!want_jac) { jac_pose = NULL; jac_point = NULL; }
    if (f->is_outlier) {                                   /* :330-344 */
        r[0] = 640.0; r[1] = 480.0;
        if (jac_pose) memset(jac_pose, 0, 12 * sizeof(double));
        if (jac_point) memset(jac_point, 0, 6 * sizeof(double));
        return 1;
    }
    vo_chain c;
    pose_chain(f->Twb_init, f->Tcb, pose6, &c);
    vo_proj p;
    project_error(f, &c.Tcw, pt, &p);
    if (p.branch == 1) {                                   /* :378-394 */
        r[0] = 640.0; r[1] = 360.0;
        if (jac_pose) memset(jac_pose, 0, 12 * sizeof(double));
        if (jac_point) memset(jac_point, 0, 6 * sizeof(double));
        return 1;
    }
    if (p.branch == 2) {                                   /* :427-441 */
        r[0] = 100.0; r[1] = 100.0;
        if (jac_pose) memset(jac_pose, 0, 12 * sizeof(double));
        if (jac_point) memset(jac_point, 0, 6 * sizeof(double));
        return 1;
    }
    double Lc[4];
    const int llt_ok = llt2(f->info, Lc);                  /* :444-455 */
    if (llt_ok) {
        r[0] = Lc[0] * p.du + Lc[1] * p.dv;
        r[1] = Lc[2] * p.du + Lc[3] * p.dv;
    } else {
        r[0] = p.du; r[1] = p.dv;
    }
    if (!want_jac) return 1;                               /* :458 */
    const double xz_sq = p.pc[0] * p.pc[0] + p.pc[2] * p.pc[2];
    const double L_sq = p.L * p.L;
    const double xz_norm = sqrt(xz_sq);
    if (xz_sq < 1e-10 || L_sq < 1e-10) {                   /* :466-476 */
        if (jac_pose) memset(jac_pose, 0, 12 * sizeof(double));
        if (jac_point) memset(jac_point, 0, 6 * sizeof(double));
        return 1;
    }
    double Jproj[6], Jw[6];
    proj_jacobian(f->cols, f->rows, p.pc, xz_sq, L_sq, xz_norm, Jproj);
    if (llt_ok) lmul2(Lc, Jproj, Jw);                      /* :491-497 */
    else memcpy(Jw, Jproj, sizeof(Jw));
    /* raw (un-projected) rotation block of m_Tcb, :506,509,527 */
    const double Rcb[9] = {f->Tcb[0], f->Tcb[1], f->Tcb[2], f->Tcb[4], f->Tcb[5],
                           f->Tcb[6], f->Tcb[8], f->Tcb[9], f->Tcb[10]};
    if (jac_pose) {                                        /* :500-521 */
        double Pb[3], hatPb[9], Jrot3[9], Jt[6], Jr[6], nRcb[9];
        mat3_vec(c.Tbw.R, pt, Pb);
        for (int i = 0; i < 3; ++i) Pb[i] += c.Tbw.t[i];
        hat3(Pb, hatPb);
        for (int i = 0; i < 9; ++i) nRcb[i] = -Rcb[i];
        mat3_mul(Rcb, hatPb, Jrot3);
        mat23_mul33(Jw, nRcb, Jt);
        mat23_mul33(Jw, Jrot3, Jr);
        for (int i = 0; i < 2; ++i)
            for (int j = 0; j < 3; ++j) {
                jac_pose[i * 6 + j] = Jt[i * 3 + j];
                jac_pose[i * 6 + 3 + j] = Jr[i * 3 + j];
            }
    }
    if (jac_point) {                                       /* :524-533 */
        double Jcp[9];
        mat3_mul(Rcb, c.Tbw.R, Jcp);
        mat23_mul33(Jw, Jcp, jac_point);
    }
    return 1;
}

/* Factors.cpp:33-210 */
int vo_pnp_evaluate(const vo_factor* f, const double pose6[6], double r[2],
                    double* jac_pose) {
    if (f->is_outlier) {                                   /* :35-44 */
        r[0] = 640.0; r[1] = 480.0;
        if (jac_pose) memset(jac_pose, 0, 12 * sizeof(double));
        return 1;
    }
    vo_chain c;
    pose_chain(f->Twb_init, f->Tcb, pose6, &c);
    vo_proj p;
    project_error(f, &c.Tcw, f->world_point, &p);
    if (p.branch == 1) {                                   /* :74-80: residual untouched */
        if (jac_pose) memset(jac_pose, 0, 12 * sizeof(double));
        return 0;
    }
    if (p.branch == 2) {                                   /* :115-125 */
        r[0] = 100.0; r[1] = 100.0;
        if (jac_pose) memset(jac_pose, 0, 12 * sizeof(double));
        return 1;
    }
    double Lc[4];
    const int llt_ok = llt2(f->info, Lc);                  /* :128-139 */
    if (llt_ok) {
        r[0] = Lc[0] * p.du + Lc[1] * p.dv;
        r[1] = Lc[2] * p.du + Lc[3] * p.dv;
    } else {
        r[0] = p.du; r[1] = p.dv;
    }
    if (!jac_pose) return 1;
    const double xz_sq = p.pc[0] * p.pc[0] + p.pc[2] * p.pc[2];
    const double L_sq = p.L * p.L;
    const double xz_norm = sqrt(xz_sq);
    if (xz_sq < 1e-10 || L_sq < 1e-10) {                   /* :161-164 */
        memset(jac_pose, 0, 12 * sizeof(double));
        return 1;
    }
    double Jproj[6];
    proj_jacobian(f->cols, f->rows, p.pc, xz_sq, L_sq, xz_norm, Jproj);
    const double Rcb[9] = {f->Tcb[0], f->Tcb[1], f->Tcb[2], f->Tcb[4], f->Tcb[5],
                           f->Tcb[6], f->Tcb[8], f->Tcb[9], f->Tcb[10]};
    double Pb[3], hatPb[9], nRcb[9], Jt[6], JR[6], Jr[6], U[12];
    mat3_vec(c.Tbw.R, f->world_point, Pb);                 /* :148 */
    for (int i = 0; i < 3; ++i) Pb[i] += c.Tbw.t[i];
    for (int i = 0; i < 9; ++i) nRcb[i] = -Rcb[i];
    mat23_mul33(Jproj, nRcb, Jt);                          /* :181 */
    hat3(Pb, hatPb);
    mat23_mul33(Jproj, Rcb, JR);                           /* :185: (J * Rcb) * hatPb */
    mat23_mul33(JR, hatPb, Jr);
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 3; ++j) {
            U[i * 6 + j] = Jt[i * 3 + j];
            U[i * 6 + 3 + j] = Jr[i * 3 + j];
        }
    if (llt_ok) {                                          /* :192-197 */
        for (int j = 0; j < 6; ++j) {
            jac_pose[j] = Lc[0] * U[j] + Lc[1] * U[6 + j];
            jac_pose[6 + j] = Lc[2] * U[j] + Lc[3] * U[6 + j];
        }
    } else {
        memcpy(jac_pose, U, sizeof(U));
    }
    return 1;
}

/* shared tail of the two compute_chi_square functions: e^T * info * e */
static double chi2_of(const vo_factor* f, const vo_proj* p) {
    const double e0 = p->du, e1 = p->dv;
    const double a0 = e0 * f->info[0] + e1 * f->info[2];   /* error^T * Information */
    const double a1 = e0 * f->info[1] + e1 * f->info[3];
    return a0 * e0 + a1 * e1;
}

/* Factors.cpp:544-612.  The leading Evaluate() call at :551 has no side effect on
 * the returned value and is not repeated here.                                      */
double vo_ba_chi2(const vo_factor* f, const double pose6[6], const double pt[3]) {
    if (f->is_outlier) return 0.0;                         /* :545-547 */
    vo_chain c;
    pose_chain(f->Twb_init, f->Tcb, pose6, &c);
    vo_proj p;
    project_error(f, &c.Tcw, pt, &p);
    if (p.branch == 1) return 1000.0;                      /* :581-583 */
    return chi2_of(f, &p);
}

/* Factors.cpp:212-265 (ignores the outlier flag) */
double vo_pnp_chi2(const vo_factor* f, const double pose6[6]) {
    vo_chain c;
    pose_chain(f->Twb_init, f->Tcb, pose6, &c);
    vo_proj p;
    project_error(f, &c.Tcw, f->world_point, &p);
    if (p.branch == 1) return DBL_MAX;                     /* :233-235 */
    return chi2_of(f, &p);
}

/* ------------------------------------------------------------------- Ceres loss */
/* loss_function.cc:48-62 (HuberLoss, a_ = a, b_ = a*a) */
void vo_huber(double a, double s, double rho[3]) {
    const double b = a * a;
    if (s > b) {
        const double r = sqrt(s);
        rho[0] = 2.0 * a * r - b;
        rho[1] = fmax(DBL_MIN, a / r);
        rho[2] = -rho[1] / (2.0 * s);
    } else {
        rho[0] = s; rho[1] = 1.0; rho[2] = 0.0;
    }
}

/* residual_block.cc:133-197 + corrector.cc:42-160.  For rho[2] <= 0 (always true for
 * Huber) the Corrector reduces to a scale by sqrt(rho[1]) (corrector.cc:84-88,
 * 112-116, 143-147); the rank-one branch is restated for completeness.              */
double vo_apply_loss(double huber_a, double* r, int nres, double* jac_a, int na,
                     double* jac_b, int nb) {
    double sq = 0.0;
    for (int i = 0; i < nres; ++i) sq += r[i] * r[i];
    if (!(huber_a > 0.0)) return 0.5 * sq;      /* no loss function: residual_block.cc:150 */
    double rho[3];
    vo_huber(huber_a, sq, rho);
    const double cost = 0.5 * rho[0];
    const double sqrt_rho1 = sqrt(rho[1]);
    double residual_scaling, alpha_sq_norm;
    if (sq == 0.0 || rho[2] <= 0.0) {
        residual_scaling = sqrt_rho1; alpha_sq_norm = 0.0;
    } else {
        const double D = 1.0 + 2.0 * sq * rho[2] / rho[1];
        const double alpha = 1.0 - sqrt(D);
        residual_scaling = sqrt_rho1 / (1 - alpha);
        alpha_sq_norm = alpha / sq;
    }
    double* blocks[2] = {jac_a, jac_b};
    const int widths[2] = {na, nb};
    for (int b = 0; b < 2; ++b) {               /* Jacobians first (residual_block.cc:180) */
        double* J = blocks[b];
        if (!J) continue;
        const int nc = widths[b];
        if (alpha_sq_norm == 0.0) {
            for (int i = 0; i < nres * nc; ++i) J[i] *= sqrt_rho1;
        } else {
            for (int c = 0; c < nc; ++c) {
                double rtj = 0.0;
                for (int i = 0; i < nres; ++i) rtj += J[i * nc + c] * r[i];
                for (int i = 0; i < nres; ++i)
                    J[i * nc + c] = sqrt_rho1 * (J[i * nc + c] - alpha_sq_norm * r[i] * rtj);
            }
        }
    }
    for (int i = 0; i < nres; ++i) r[i] *= residual_scaling; /* residual_block.cc:193 */
    return cost;
}

/* ---------------------------------------------------------------- window level */
static void make_factor(const vo_window* w, int i, vo_factor* f) {
    f->obs[0] = w->obs_u[i]; f->obs[1] = w->obs_v[i];
    f->cols = w->cols; f->rows = w->rows;
    memcpy(f->Tcb, w->Tcb, sizeof(f->Tcb));
    memcpy(f->Twb_init, w->Twb_init + 16 * (size_t)w->pose_idx[i], sizeof(f->Twb_init));
    memcpy(f->info, w->info, sizeof(f->info));
    f->is_outlier = w->outlier ? (int)w->outlier[i] : 0;
    f->world_point[0] = f->world_point[1] = f->world_point[2] = 0.0;
}

typedef struct {
    const vo_window* w;
    const double* pose_delta;
    const double* points;
    int apply_loss, hoisted, tid, nthreads;
    double *r, *Jp, *Jl;
    double cost;
    const vo_chain* chains; /* hoisted only */
} eval_job;

/* hoisted variant: identical arithmetic to vo_ba_evaluate after the pose chain */
static void ba_evaluate_with_chain(const vo_factor* f, const vo_chain* c,
                                   const double pt[3], double r[2], double* jac_pose,
                                   double* jac_point) {
    if (f->is_outlier) {
        r[0] = 640.0; r[1] = 480.0;
        memset(jac_pose, 0, 12 * sizeof(double)); memset(jac_point, 0, 6 * sizeof(double));
        return;
    }
    vo_proj p;
    project_error(f, &c->Tcw, pt, &p);
    if (p.branch) {
        r[0] = p.branch == 1 ? 640.0 : 100.0; r[1] = p.branch == 1 ? 360.0 : 100.0;
        memset(jac_pose, 0, 12 * sizeof(double)); memset(jac_point, 0, 6 * sizeof(double));
        return;
    }
    double Lc[4];
    const int llt_ok = llt2(f->info, Lc);
    if (llt_ok) {
        r[0] = Lc[0] * p.du + Lc[1] * p.dv; r[1] = Lc[2] * p.du + Lc[3] * p.dv;
    } else { r[0] = p.du; r[1] = p.dv; }
    const double xz_sq = p.pc[0] * p.pc[0] + p.pc[2] * p.pc[2];
    const double L_sq = p.L * p.L, xz_norm = sqrt(xz_sq);
    if (xz_sq < 1e-10 || L_sq < 1e-10) {
        memset(jac_pose, 0, 12 * sizeof(double)); memset(jac_point, 0, 6 * sizeof(double));
        return;
    }
    double Jproj[6], Jw[6];
    proj_jacobian(f->cols, f->rows, p.pc, xz_sq, L_sq, xz_norm, Jproj);
    if (llt_ok) lmul2(Lc, Jproj, Jw); else memcpy(Jw, Jproj, sizeof(Jw));
    const double Rcb[9] = {f->Tcb[0], f->Tcb[1], f->Tcb[2], f->Tcb[4], f->Tcb[5],
                           f->Tcb[6], f->Tcb[8], f->Tcb[9], f->Tcb[10]};
    double Pb[3], hatPb[9], Jrot3[9], Jt[6], Jr[6], nRcb[9], Jcp[9];
    mat3_vec(c->Tbw.R, pt, Pb);
    for (int i = 0; i < 3; ++i) Pb[i] += c->Tbw.t[i];
    hat3(Pb, hatPb);
    for (int i = 0; i < 9; ++i) nRcb[i] = -Rcb[i];
    mat3_mul(Rcb, hatPb, Jrot3);
    mat23_mul33(Jw, nRcb, Jt);
    mat23_mul33(Jw, Jrot3, Jr);
    for (int i = 0; i < 2; ++i)
        for (int j = 0; j < 3; ++j) {
            jac_pose[i * 6 + j] = Jt[i * 3 + j];
            jac_pose[i * 6 + 3 + j] = Jr[i * 3 + j];
        }
    mat3_mul(Rcb, c->Tbw.R, Jcp);
    mat23_mul33(Jw, Jcp, jac_point);
}

static void* eval_worker(void* arg) {
    eval_job* j = (eval_job*)arg;
    const vo_window* w = j->w;
    const int O = w->O;
    const int lo = (int)((long long)O * j->tid / j->nthreads);
    const int hi = (int)((long long)O * (j->tid + 1) / j->nthreads);
    double cost = 0.0;
    for (int i = lo; i < hi; ++i) {
        vo_factor f;
        make_factor(w, i, &f);
        double r[2], Jp[12], Jl[6];
        const double* pose6 = j->pose_delta + 6 * (size_t)w->pose_idx[i];
        const double* pt = j->points + 3 * (size_t)w->point_idx[i];
        if (j->hoisted)
            ba_evaluate_with_chain(&f, &j->chains[w->pose_idx[i]], pt, r, Jp, Jl);
        else
            vo_ba_evaluate(&f, pose6, pt, r, 1, Jp, Jl);
        cost += vo_apply_loss(j->apply_loss ? w->huber_a : 0.0, r, 2, Jp, 6, Jl, 3);
        if (j->r) memcpy(j->r + 2 * (size_t)i, r, sizeof(r));
        if (j->Jp) memcpy(j->Jp + 12 * (size_t)i, Jp, sizeof(Jp));
        if (j->Jl) memcpy(j->Jl + 6 * (size_t)i, Jl, sizeof(Jl));
    }
    j->cost = cost;
    return NULL;
}

static int window_evaluate_impl(const vo_window* w, const double* pose_delta,
                                const double* points, int apply_loss, double* r,
                                double* Jp, double* Jl, double* cost, int nthreads,
                                int hoisted) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    vo_chain* chains = NULL;
    if (hoisted) {
        chains = (vo_chain*)malloc(sizeof(vo_chain) * (size_t)w->K);
        for (int k = 0; k < w->K; ++k)
            pose_chain(w->Twb_init + 16 * (size_t)k, w->Tcb, pose_delta + 6 * (size_t)k,
                       &chains[k]);
    }
    eval_job jobs[256];
    pthread_t th[256];
    for (int t = 0; t < nthreads; ++t) {
        eval_job j = {w, pose_delta, points, apply_loss, hoisted, t, nthreads,
                      r, Jp, Jl, 0.0, chains};
        jobs[t] = j;
    }
    for (int t = 1; t < nthreads; ++t) pthread_create(&th[t], NULL, eval_worker, &jobs[t]);
    eval_worker(&jobs[0]);
    for (int t = 1; t < nthreads; ++t) pthread_join(th[t], NULL);
    double c = 0.0;
    for (int t = 0; t < nthreads; ++t) c += jobs[t].cost;
    if (cost) *cost = c;
    free(chains);
    return 1;
}

int vo_window_evaluate(const vo_window* w, const double* pose_delta, const double* points,
                       int apply_loss, double* r, double* Jp, double* Jl, double* cost,
                       int nthreads) {
    return window_evaluate_impl(w, pose_delta, points, apply_loss, r, Jp, Jl, cost,
                                nthreads, 0);
}
int vo_window_evaluate_hoisted(const vo_window* w, const double* pose_delta,
                               const double* points, int apply_loss, double* r,
                               double* Jp, double* Jl, double* cost, int nthreads) {
    return window_evaluate_impl(w, pose_delta, points, apply_loss, r, Jp, Jl, cost,
                                nthreads, 1);
}

void vo_normal_blocks(const vo_window* w, const double* r, const double* Jp,
                      const double* Jl, double* Hpp, double* bp, double* Hll, double* bl,
                      double* Hpl) {
    memset(Hpp, 0, sizeof(double) * 36 * (size_t)w->K);
    memset(bp, 0, sizeof(double) * 6 * (size_t)w->K);
    memset(Hll, 0, sizeof(double) * 9 * (size_t)w->M);
    memset(bl, 0, sizeof(double) * 3 * (size_t)w->M);
    memset(Hpl, 0, sizeof(double) * 18 * (size_t)w->O);
    for (int i = 0; i < w->O; ++i) {
        const int p = w->pose_idx[i], l = w->point_idx[i];
        const int pf = w->pose_fixed && w->pose_fixed[p];
        const int lf = w->point_fixed && w->point_fixed[l];
        const double* F = Jp + 12 * (size_t)i;
        const double* E = Jl + 6 * (size_t)i;
        const double* ri = r + 2 * (size_t)i;
        if (!pf) {
            for (int a = 0; a < 6; ++a) {
                for (int b = 0; b < 6; ++b)
                    Hpp[36 * (size_t)p + a * 6 + b] += F[a] * F[b] + F[6 + a] * F[6 + b];
                bp[6 * (size_t)p + a] += F[a] * ri[0] + F[6 + a] * ri[1];
            }
        }
        if (!lf) {
            for (int a = 0; a < 3; ++a) {
                for (int b = 0; b < 3; ++b)
                    Hll[9 * (size_t)l + a * 3 + b] += E[a] * E[b] + E[3 + a] * E[3 + b];
                bl[3 * (size_t)l + a] += E[a] * ri[0] + E[3 + a] * ri[1];
            }
        }
        if (!pf && !lf)
            for (int a = 0; a < 6; ++a)
                for (int b = 0; b < 3; ++b)
                    Hpl[18 * (size_t)i + a * 3 + b] = F[a] * E[b] + F[6 + a] * E[3 + b];
    }
}

/* InvertPSDMatrix<3> with assume_full_rank (invert_psd_matrix.h): Cholesky solve of I */
static void inv_spd3(const double A[9], double Ainv[9]) {
    const double l00 = sqrt(A[0]);
    const double l10 = A[3] / l00, l20 = A[6] / l00;
    const double l11 = sqrt(A[4] - l10 * l10);
    const double l21 = (A[7] - l20 * l10) / l11;
    const double l22 = sqrt(A[8] - l20 * l20 - l21 * l21);
    /* inverse of L (lower) */
    const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
    const double i10 = -l10 * i00 * i11;
    const double i21 = -l21 * i11 * i22;
    const double i20 = -(l20 * i00 + l21 * i10) * i22;
    /* A^-1 = Linv^T Linv */
    Ainv[0] = i00 * i00 + i10 * i10 + i20 * i20;
    Ainv[1] = Ainv[3] = i10 * i11 + i20 * i21;
    Ainv[2] = Ainv[6] = i20 * i22;
    Ainv[4] = i11 * i11 + i21 * i21;
    Ainv[5] = Ainv[7] = i21 * i22;
    Ainv[8] = i22 * i22;
}

/* schur_eliminator_impl.h:179-308 (+ :383-560).  Chunks = landmarks.               */
int vo_schur_eliminate(const vo_window* w, const double* r, const double* Jp,
                       const double* Jl, const double* D, double* S, double* rhs,
                       int* free_map_out) {
    const int K = w->K, M = w->M, O = w->O;
    int* fm = (int*)malloc(sizeof(int) * (size_t)K);
    int nf = 0;
    for (int k = 0; k < K; ++k) fm[k] = (w->pose_fixed && w->pose_fixed[k]) ? -1 : nf++;
    const int n = 6 * nf;
    memset(S, 0, sizeof(double) * (size_t)n * n);
    memset(rhs, 0, sizeof(double) * (size_t)n);
    if (D) /* :196-218: diagonal of the f-blocks */
        for (int k = 0; k < K; ++k)
            if (fm[k] >= 0)
                for (int a = 0; a < 6; ++a) {
                    const double d = D[6 * (size_t)k + a];
                    S[(size_t)(6 * fm[k] + a) * n + 6 * fm[k] + a] += d * d;
                }
    /* bucket observations per landmark (stable: increasing observation index) */
    int* cnt = (int*)calloc((size_t)M + 1, sizeof(int));
    for (int i = 0; i < O; ++i) cnt[w->point_idx[i] + 1]++;
    for (int l = 0; l < M; ++l) cnt[l + 1] += cnt[l];
    int* order = (int*)malloc(sizeof(int) * (size_t)(O > 0 ? O : 1));
    int* fill = (int*)malloc(sizeof(int) * (size_t)(M > 0 ? M : 1));
    for (int l = 0; l < M; ++l) fill[l] = cnt[l];
    for (int i = 0; i < O; ++i) order[fill[w->point_idx[i]]++] = i;
    double* EtF = (double*)malloc(sizeof(double) * 18 * (size_t)(K > 0 ? K : 1));
    unsigned char* seen = (unsigned char*)malloc((size_t)(K > 0 ? K : 1));
    int* plist = (int*)malloc(sizeof(int) * (size_t)(K > 0 ? K : 1));
    for (int l = 0; l < M; ++l) {
        const int lo = cnt[l], hi = cnt[l + 1];
        if (lo == hi) continue;
        const int lfix = w->point_fixed && w->point_fixed[l];
        if (lfix) { /* NoEBlockRowsUpdate :305-307 : S += F'F, rhs += F'b */
            for (int q = lo; q < hi; ++q) {
                const int i = order[q], p = fm[w->pose_idx[i]];
                if (p < 0) continue;
                const double* F = Jp + 12 * (size_t)i;
                const double* ri = r + 2 * (size_t)i;
                for (int a = 0; a < 6; ++a) {
                    for (int b = 0; b < 6; ++b)
                        S[(size_t)(6 * p + a) * n + 6 * p + b] += F[a] * F[b] + F[6 + a] * F[6 + b];
                    rhs[6 * p + a] += F[a] * ri[0] + F[6 + a] * ri[1];
                }
            }
            continue;
        }
        double ete[9] = {0}, g[3] = {0};
        if (D) for (int a = 0; a < 3; ++a) {
            const double d = D[6 * (size_t)K + 3 * (size_t)l + a];
            ete[a * 4] = d * d;
        }
        int np = 0;
        memset(seen, 0, (size_t)K);
        /* ChunkDiagonalBlockAndGradient :445-505 */
        for (int q = lo; q < hi; ++q) {
            const int i = order[q], pk = w->pose_idx[i], p = fm[pk];
            const double* F = Jp + 12 * (size_t)i;
            const double* E = Jl + 6 * (size_t)i;
            const double* ri = r + 2 * (size_t)i;
            if (p >= 0) /* EBlockRowOuterProduct: S += F'F */
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 6; ++b)
                        S[(size_t)(6 * p + a) * n + 6 * p + b] += F[a] * F[b] + F[6 + a] * F[6 + b];
            for (int a = 0; a < 3; ++a) {
                for (int b = 0; b < 3; ++b) ete[a * 3 + b] += E[a] * E[b] + E[3 + a] * E[3 + b];
                g[a] += E[a] * ri[0] + E[3 + a] * ri[1];
            }
            if (p >= 0) {
                if (!seen[pk]) { seen[pk] = 1; plist[np++] = pk; memset(EtF + 18 * (size_t)pk, 0, 18 * sizeof(double)); }
                double* B = EtF + 18 * (size_t)pk; /* 3x6 */
                for (int a = 0; a < 3; ++a)
                    for (int b = 0; b < 6; ++b) B[a * 6 + b] += E[a] * F[b] + E[3 + a] * F[6 + b];
            }
        }
        double inv[9], ig[3];
        inv_spd3(ete, inv);
        mat3_vec(inv, g, ig);
        /* UpdateRhs :383-426 */
        for (int q = lo; q < hi; ++q) {
            const int i = order[q], p = fm[w->pose_idx[i]];
            if (p < 0) continue;
            const double* F = Jp + 12 * (size_t)i;
            const double* E = Jl + 6 * (size_t)i;
            const double* ri = r + 2 * (size_t)i;
            const double s0 = ri[0] - (E[0] * ig[0] + E[1] * ig[1] + E[2] * ig[2]);
            const double s1 = ri[1] - (E[3] * ig[0] + E[4] * ig[1] + E[5] * ig[2]);
            for (int a = 0; a < 6; ++a) rhs[6 * p + a] += F[a] * s0 + F[6 + a] * s1;
        }
        /* ChunkOuterProduct :512-560 : S(p1,p2) -= B1' inv B2 for every pose pair */
        for (int a1 = 0; a1 < np; ++a1) {
            const double* B1 = EtF + 18 * (size_t)plist[a1];
            double B1tI[18]; /* 6x3 */
            for (int a = 0; a < 6; ++a)
                for (int b = 0; b < 3; ++b)
                    B1tI[a * 3 + b] = B1[0 * 6 + a] * inv[0 * 3 + b] + B1[1 * 6 + a] * inv[1 * 3 + b] +
                                      B1[2 * 6 + a] * inv[2 * 3 + b];
            const int p1 = fm[plist[a1]];
            for (int a2 = 0; a2 < np; ++a2) {
                const double* B2 = EtF + 18 * (size_t)plist[a2];
                const int p2 = fm[plist[a2]];
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 6; ++b)
                        S[(size_t)(6 * p1 + a) * n + 6 * p2 + b] -=
                            B1tI[a * 3 + 0] * B2[0 * 6 + b] + B1tI[a * 3 + 1] * B2[1 * 6 + b] +
                            B1tI[a * 3 + 2] * B2[2 * 6 + b];
            }
        }
    }
    if (free_map_out) memcpy(free_map_out, fm, sizeof(int) * (size_t)K);
    free(fm); free(cnt); free(order); free(fill); free(EtF); free(seen); free(plist);
    return n;
}

/* levenberg_marquardt_strategy.cc:78-89 */
void vo_lm_diagonal(const vo_window* w, const double* Jp, const double* Jl, double radius,
                    double* D) {
    const int K = w->K, M = w->M;
    const size_t np = 6 * (size_t)K + 3 * (size_t)M;
    for (size_t i = 0; i < np; ++i) D[i] = 0.0;
    for (int i = 0; i < w->O; ++i) {
        const int p = w->pose_idx[i], l = w->point_idx[i];
        const double* F = Jp + 12 * (size_t)i;
        const double* E = Jl + 6 * (size_t)i;
        if (!(w->pose_fixed && w->pose_fixed[p]))
            for (int a = 0; a < 6; ++a) D[6 * (size_t)p + a] += F[a] * F[a] + F[6 + a] * F[6 + a];
        if (!(w->point_fixed && w->point_fixed[l]))
            for (int a = 0; a < 3; ++a)
                D[6 * (size_t)K + 3 * (size_t)l + a] += E[a] * E[a] + E[3 + a] * E[3 + a];
    }
    for (size_t i = 0; i < np; ++i) {
        double v = D[i];
        if (v < 1e-6) v = 1e-6;
        if (v > 1e32) v = 1e32;
        D[i] = sqrt(v / radius);
    }
}

/* Optimizer.cpp:898-915 */
void vo_window_chi2(const vo_window* w, const double* pose_delta, const double* points,
                    double thr, double* chi2, unsigned char* flag) {
    for (int i = 0; i < w->O; ++i) {
        vo_factor f;
        make_factor(w, i, &f);
        const double c = vo_ba_chi2(&f, pose_delta + 6 * (size_t)w->pose_idx[i],
                                    points + 3 * (size_t)w->point_idx[i]);
        if (chi2) chi2[i] = c;
        if (flag) flag[i] = (unsigned char)(c > thr);
    }
}

/* ------------------------------------------------------------------ DLT triangulation */
void vo_triangulate(const float* b1, const float* b2, const float* T1, const float* T2, int n,
                    float* points, unsigned char* valid) {
    for (int i = 0; i < n; ++i) {
        const float* u = b1 + 3 * i; const float* v = b2 + 3 * i;
        float B[16], V[16];
        for (int c = 0; c < 4; ++c) {                         /* Estimator.cpp:1098-1101 */
            B[0 * 4 + c] = u[0] * T1[2 * 4 + c] - u[2] * T1[0 * 4 + c];
            B[1 * 4 + c] = u[1] * T1[2 * 4 + c] - u[2] * T1[1 * 4 + c];
            B[2 * 4 + c] = v[0] * T2[2 * 4 + c] - v[2] * T2[0 * 4 + c];
            B[3 * 4 + c] = v[1] * T2[2 * 4 + c] - v[2] * T2[1 * 4 + c];
        }
        for (int k = 0; k < 16; ++k) V[k] = (k % 5 == 0) ? 1.f : 0.f;
        for (int sweep = 0; sweep < 30; ++sweep) {
            float off = 0.f;
            for (int p = 0; p < 3; ++p)
                for (int q = p + 1; q < 4; ++q) {
                    float alpha = 0.f, beta = 0.f, gamma = 0.f;
                    for (int k = 0; k < 4; ++k) {
                        alpha += B[k * 4 + p] * B[k * 4 + p];
                        beta += B[k * 4 + q] * B[k * 4 + q];
                        gamma += B[k * 4 + p] * B[k * 4 + q];
                    }
                    const float lim = sqrtf(alpha * beta);
                    if (gamma == 0.f || fabsf(gamma) <= 1e-9f * lim) continue;
                    const float rel = fabsf(gamma) / (lim > 0.f ? lim : 1.f);
                    if (rel > off) off = rel;
                    const float zeta = (beta - alpha) / (2.f * gamma);
                    const float t = (zeta >= 0.f ? 1.f : -1.f) / (fabsf(zeta) + sqrtf(1.f + zeta * zeta));
                    const float c = 1.f / sqrtf(1.f + t * t), s = c * t;
                    for (int k = 0; k < 4; ++k) {
                        const float bp = B[k * 4 + p], bq = B[k * 4 + q];
                        B[k * 4 + p] = c * bp - s * bq; B[k * 4 + q] = s * bp + c * bq;
                        const float vp = V[k * 4 + p], vq = V[k * 4 + q];
                        V[k * 4 + p] = c * vp - s * vq; V[k * 4 + q] = s * vp + c * vq;
                    }
                }
            if (off < 1e-7f) break;
        }
        int best = 0;
        float smin = FLT_MAX;
        for (int j = 0; j < 4; ++j) {
            const float sv = B[j] * B[j] + B[4 + j] * B[4 + j] + B[8 + j] * B[8 + j] + B[12 + j] * B[12 + j];
            if (sv < smin) { smin = sv; best = j; }
        }
        const float h0 = V[best], h1 = V[4 + best], h2 = V[8 + best], h3 = V[12 + best];
        int ok = !(fabsf(h3) < 1e-10f);                       /* :1108-1110 */
        const float x = h0 / h3, y = h1 / h3, z = h2 / h3;    /* :1113 */
        ok = ok && isfinite(x) && isfinite(y) && isfinite(z); /* :1119-1121 */
        points[3 * i] = ok ? x : 0.f; points[3 * i + 1] = ok ? y : 0.f; points[3 * i + 2] = ok ? z : 0.f;
        valid[i] = (unsigned char)ok;
    }
}
