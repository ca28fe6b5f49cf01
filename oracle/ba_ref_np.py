"""oracle/ba_ref_np.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Independent numpy mirror of oracle/ba_ref.c (LAPACK SVD instead of the hand-written
Jacobi SVD, vectorised over observations) used only to cross-check the C restatement.
PARITY UNPINNED: the reference has no tests/golden vectors for this path.

Follows (reference root relative):
  src/util/LieUtils.cpp:203-219,275-288,305-333 ; src/util/LieUtils.h:207-219,268-283
  src/optimization/Factors.cpp:327-542 (BAFactor::Evaluate), :544-612 (chi-square)
  thirdparty/ceres-solver/internal/ceres/loss_function.cc:48-62, corrector.cc:42-160
"""
import numpy as np


def so3_project(M):
    """SO3d::SO3d(const Matrix3d&)  LieUtils.cpp:275-288 (batched over leading dims)."""
    M = np.asarray(M, dtype=np.float64)
    U, _, Vt = np.linalg.svd(M)
    R = U @ Vt
    neg = np.linalg.det(R) < 0.0
    if np.any(neg):
        U = U.copy()
        U[neg, :, 2] *= -1.0
        R = U @ Vt
    return R


def hat(v):
    v = np.asarray(v, dtype=np.float64)
    S = np.zeros(v.shape[:-1] + (3, 3))
    S[..., 0, 1] = -v[..., 2]; S[..., 0, 2] = v[..., 1]
    S[..., 1, 0] = v[..., 2]; S[..., 1, 2] = -v[..., 0]
    S[..., 2, 0] = -v[..., 1]; S[..., 2, 1] = v[..., 0]
    return S


def so3_exp(w):
    """SO3d::Exp  LieUtils.cpp:203-219 (one vector)."""
    w = np.asarray(w, dtype=np.float64)
    theta = np.sqrt(w @ w)
    if theta < 1e-10:
        return so3_project(np.eye(3) + hat(w))
    k = w * (1.0 / theta)
    K = hat(k)
    return so3_project(np.eye(3) + np.sin(theta) * K + (1.0 - np.cos(theta)) * (K @ K))


def se3_exp(xi):
    """SE3d::exp  LieUtils.cpp:305-333 -> (R, t)."""
    xi = np.asarray(xi, dtype=np.float64)
    rho, phi = xi[:3], xi[3:]
    R = so3_exp(phi)
    theta = np.sqrt(phi @ phi)
    if theta < 1e-10:
        return R, rho.copy()
    P = hat(phi)
    th2 = theta * theta
    V = np.eye(3) + (1.0 - np.cos(theta)) / th2 * P + (theta - np.sin(theta)) / (th2 * theta) * (P @ P)
    return R, V @ rho


def pose_chain(Twb_init, Tcb, xi):
    """Factors.cpp:357-364 -> (R_cw, t_cw, R_bw, t_bw)."""
    R0 = so3_project(Twb_init[:3, :3]); t0 = Twb_init[:3, 3]
    Rd, td = se3_exp(xi)
    Rwb = so3_project(R0 @ Rd); twb = t0 + R0 @ td
    Rbw = so3_project(Rwb.T); tbw = Rbw @ (-twb)
    Rcb = so3_project(Tcb[:3, :3]); tcb = Tcb[:3, 3]
    Rcw = so3_project(Rcb @ Rbw); tcw = tcb + Rcb @ tbw
    return Rcw, tcw, Rbw, tbw


def huber(a, s):
    """HuberLoss::Evaluate, loss_function.cc:48-62 (vectorised) -> rho0, rho1."""
    s = np.asarray(s, dtype=np.float64)
    b = a * a
    big = s > b
    r = np.sqrt(np.where(big, s, 1.0))
    rho0 = np.where(big, 2.0 * a * r - b, s)
    rho1 = np.where(big, np.maximum(np.finfo(np.float64).tiny, a / r), 1.0)
    return rho0, rho1


def window_evaluate(win, pose_delta, points, apply_loss=True):
    """Vectorised BAFactor::Evaluate over a window + Huber/Corrector.
    win: dict from synthetic.make_ba_window. returns r[O,2], Jp[O,2,6], Jl[O,2,3], cost."""
    K = win["K"]; O = win["O"]
    cols, rows = float(win["cols"]), float(win["rows"])
    Tcb = np.asarray(win["Tcb"], dtype=np.float64)
    chains = [pose_chain(win["Twb_init"][k], Tcb, pose_delta[k]) for k in range(K)]
    Rcw = np.stack([c[0] for c in chains]); tcw = np.stack([c[1] for c in chains])
    Rbw = np.stack([c[2] for c in chains]); tbw = np.stack([c[3] for c in chains])
    pi, li = win["pose_idx"], win["point_idx"]
    Pw = points[li]
    Pc = np.einsum("oij,oj->oi", Rcw[pi], Pw) + tcw[pi]
    x, y, z = Pc[:, 0], Pc[:, 1], Pc[:, 2]
    L = np.sqrt(x * x + y * y + z * z)
    tiny = L < 1e-10
    Ls = np.where(tiny, 1.0, L)
    theta = np.arctan2(x, z)
    phi = -np.arcsin(np.clip(y / Ls, -1.0, 1.0))
    u = cols * (0.5 + theta / (2.0 * np.pi))
    v = rows * (0.5 - phi / np.pi)
    du = win["obs_u"] - u
    dv = win["obs_v"] - v
    du = np.where(du > cols / 2.0, du - cols, np.where(du < -cols / 2.0, du + cols, du))
    large = (np.abs(du) > 100.0) | (np.abs(dv) > 100.0)
    info = np.asarray(win["info"], dtype=np.float64).reshape(2, 2)
    try:
        Lc = np.linalg.cholesky(info)
    except np.linalg.LinAlgError:
        Lc = np.eye(2)
    e = np.stack([du, dv], axis=1)
    r = e @ Lc.T
    xz2 = x * x + z * z
    L2 = L * L
    xzn = np.sqrt(xz2)
    sing = (xz2 < 1e-10) | (L2 < 1e-10)
    xz2s = np.where(sing, 1.0, xz2); L2s = np.where(sing, 1.0, L2); xzns = np.where(sing, 1.0, xzn)
    Jproj = np.zeros((O, 2, 3))
    Jproj[:, 0, 0] = -cols / (2.0 * np.pi) * z / xz2s
    Jproj[:, 0, 2] = cols / (2.0 * np.pi) * x / xz2s
    Jproj[:, 1, 0] = rows / np.pi * (x * y) / (L2s * xzns)
    Jproj[:, 1, 1] = -rows / np.pi * xzns / L2s
    Jproj[:, 1, 2] = rows / np.pi * (y * z) / (L2s * xzns)
    Jw = np.einsum("ij,ojk->oik", Lc, Jproj)
    Rcb_raw = Tcb[:3, :3]
    Pb = np.einsum("oij,oj->oi", Rbw[pi], Pw) + tbw[pi]
    Jcam_pose = np.concatenate([np.broadcast_to(-Rcb_raw, (O, 3, 3)), Rcb_raw @ hat(Pb)], axis=2)
    Jp = Jw @ Jcam_pose
    Jl = Jw @ (Rcb_raw @ Rbw[pi])
    outl = np.asarray(win.get("outlier", np.zeros(O, np.uint8))).astype(bool)
    zero_j = outl | tiny | large | sing
    Jp[zero_j] = 0.0; Jl[zero_j] = 0.0
    r[large] = (100.0, 100.0)
    r[tiny] = (640.0, 360.0)
    r[outl] = (640.0, 480.0)
    sq = np.sum(r * r, axis=1)
    if apply_loss and win["huber_a"] > 0:
        rho0, rho1 = huber(win["huber_a"], sq)
        sc = np.sqrt(rho1)
        r = r * sc[:, None]; Jp = Jp * sc[:, None, None]; Jl = Jl * sc[:, None, None]
        cost = 0.5 * np.sum(rho0)
    else:
        cost = 0.5 * np.sum(sq)
    return r, Jp, Jl, cost


def schur_dense(win, r, Jp, Jl, D=None):
    """Dense reference of the reduced camera system: builds the full J, forms
    H = J^T J + D^2 and g = J^T r, then S = Hpp - Hpl Hll^-1 Hlp with numpy.linalg.
    An independent route to what SchurEliminator::Eliminate computes."""
    K, M, O = win["K"], win["M"], win["O"]
    pf = np.asarray(win["pose_fixed"]).astype(bool); lf = np.asarray(win["point_fixed"]).astype(bool)
    fm = np.full(K, -1); fm[~pf] = np.arange(np.count_nonzero(~pf))
    lm = np.full(M, -1); lm[~lf] = np.arange(np.count_nonzero(~lf))
    n_p = 6 * int(np.count_nonzero(~pf)); n_l = 3 * int(np.count_nonzero(~lf))
    J = np.zeros((2 * O, n_p + n_l))
    for i in range(O):
        p = fm[win["pose_idx"][i]]; l = lm[win["point_idx"][i]]
        if p >= 0:
            J[2 * i:2 * i + 2, 6 * p:6 * p + 6] = Jp[i]
        if l >= 0:
            J[2 * i:2 * i + 2, n_p + 3 * l:n_p + 3 * l + 3] = Jl[i]
    H = J.T @ J
    g = J.T @ r.reshape(-1)
    if D is not None:
        Dp = D[:6 * K].reshape(K, 6)[~pf].reshape(-1)
        Dl = D[6 * K:].reshape(M, 3)[~lf].reshape(-1)
        H[np.arange(n_p + n_l), np.arange(n_p + n_l)] += np.concatenate([Dp, Dl]) ** 2
    Hpp, Hpl, Hll = H[:n_p, :n_p], H[:n_p, n_p:], H[n_p:, n_p:]
    # landmarks without observations: singular rows; they do not couple, drop them
    used = np.abs(Hll).sum(axis=1) > 0
    Hll_u = Hll[np.ix_(used, used)]
    X = np.linalg.solve(Hll_u, np.concatenate([Hpl[:, used].T, g[n_p:][used][:, None]], axis=1))
    S = Hpp - Hpl[:, used] @ X[:, :-1]
    rhs = g[:n_p] - Hpl[:, used] @ X[:, -1]
    return S, rhs
