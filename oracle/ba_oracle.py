"""oracle/ba_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes binding of oracle/ba_ref.c (libvio360_oracle_ba.so).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_up = C.POINTER(C.c_ubyte)


class _Window(C.Structure):
    _fields_ = [("K", C.c_int), ("M", C.c_int), ("O", C.c_int),
                ("pose_idx", _ip), ("point_idx", _ip), ("obs_u", _dp), ("obs_v", _dp),
                ("outlier", _up), ("Twb_init", _dp), ("Tcb", C.c_double * 16),
                ("cols", C.c_double), ("rows", C.c_double), ("info", C.c_double * 4),
                ("huber_a", C.c_double), ("pose_fixed", _up), ("point_fixed", _up)]


class _Factor(C.Structure):
    _fields_ = [("obs", C.c_double * 2), ("cols", C.c_double), ("rows", C.c_double),
                ("Tcb", C.c_double * 16), ("Twb_init", C.c_double * 16),
                ("info", C.c_double * 4), ("is_outlier", C.c_int),
                ("world_point", C.c_double * 3)]


def build(force=False):
    so = os.path.join(_HERE, "libvio360_oracle_ba.so")
    src = os.path.join(_HERE, "ba_ref.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libvio360_oracle_ba.so"],
                              stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.vo_ba_chi2.restype = C.c_double
        _LIB.vo_pnp_chi2.restype = C.c_double
        _LIB.vo_apply_loss.restype = C.c_double
    return _LIB


def _d(a):
    return a.ctypes.data_as(_dp)


class Window:
    """Keeps the numpy arrays alive for the lifetime of the C struct."""

    def __init__(self, win):
        self.win = win
        self.keep = dict(
            pose_idx=np.ascontiguousarray(win["pose_idx"], np.int32),
            point_idx=np.ascontiguousarray(win["point_idx"], np.int32),
            obs_u=np.ascontiguousarray(win["obs_u"], np.float64),
            obs_v=np.ascontiguousarray(win["obs_v"], np.float64),
            outlier=np.ascontiguousarray(win.get("outlier", np.zeros(win["O"], np.uint8)), np.uint8),
            Twb=np.ascontiguousarray(win["Twb_init"], np.float64),
            pose_fixed=np.ascontiguousarray(win["pose_fixed"], np.uint8),
            point_fixed=np.ascontiguousarray(win["point_fixed"], np.uint8))
        k = self.keep
        w = _Window()
        w.K, w.M, w.O = int(win["K"]), int(win["M"]), int(win["O"])
        w.pose_idx = k["pose_idx"].ctypes.data_as(_ip)
        w.point_idx = k["point_idx"].ctypes.data_as(_ip)
        w.obs_u = _d(k["obs_u"]); w.obs_v = _d(k["obs_v"])
        w.outlier = k["outlier"].ctypes.data_as(_up)
        w.Twb_init = _d(k["Twb"])
        w.Tcb = (C.c_double * 16)(*np.asarray(win["Tcb"], np.float64).reshape(-1))
        w.cols, w.rows = float(win["cols"]), float(win["rows"])
        w.info = (C.c_double * 4)(*np.asarray(win["info"], np.float64).reshape(-1))
        w.huber_a = float(win["huber_a"])
        w.pose_fixed = k["pose_fixed"].ctypes.data_as(_up)
        w.point_fixed = k["point_fixed"].ctypes.data_as(_up)
        self.c = w

    def evaluate(self, pose_delta, points, apply_loss=True, nthreads=1, hoisted=False,
                 outputs=True):
        O = self.c.O
        pd = np.ascontiguousarray(pose_delta, np.float64); pts = np.ascontiguousarray(points, np.float64)
        cost = C.c_double(0.0)
        fn = lib().vo_window_evaluate_hoisted if hoisted else lib().vo_window_evaluate
        if outputs:
            r = np.empty((O, 2)); Jp = np.empty((O, 2, 6)); Jl = np.empty((O, 2, 3))
            fn(C.byref(self.c), _d(pd), _d(pts), int(apply_loss), _d(r), _d(Jp), _d(Jl),
               C.byref(cost), int(nthreads))
            return r, Jp, Jl, cost.value
        fn(C.byref(self.c), _d(pd), _d(pts), int(apply_loss), None, None, None,
           C.byref(cost), int(nthreads))
        return cost.value

    def normal_blocks(self, r, Jp, Jl):
        K, M, O = self.c.K, self.c.M, self.c.O
        Hpp = np.empty((K, 6, 6)); bp = np.empty((K, 6)); Hll = np.empty((M, 3, 3))
        bl = np.empty((M, 3)); Hpl = np.empty((O, 6, 3))
        lib().vo_normal_blocks(C.byref(self.c), _d(r), _d(Jp), _d(Jl), _d(Hpp), _d(bp),
                               _d(Hll), _d(bl), _d(Hpl))
        return Hpp, bp, Hll, bl, Hpl

    def schur(self, r, Jp, Jl, D=None):
        K = self.c.K
        nf = int(K - np.count_nonzero(self.keep["pose_fixed"]))
        n = 6 * nf
        S = np.empty((n, n)); rhs = np.empty(n); fm = np.empty(K, np.int32)
        Dp = None if D is None else _d(np.ascontiguousarray(D, np.float64))
        lib().vo_schur_eliminate(C.byref(self.c), _d(r), _d(Jp), _d(Jl), Dp, _d(S), _d(rhs),
                                 fm.ctypes.data_as(_ip))
        return S, rhs, fm

    def solve_step(self, r, Jp, Jl, D, S, rhs, fm):
        """Reduced-system solve + back substitution as Ceres does after Eliminate:
        x = S^-1 rhs (DenseSchurComplementSolver::SolveReducedLinearSystem, schur_complement_solver.cc;
        LAPACK here instead of Eigen LLT), then per landmark
        y = (E'E + D^2)^-1 (E'r - E'F x)  (SchurEliminator::BackSubstitute, schur_eliminator_impl.h:310-380).
        r/Jp/Jl are the loss-corrected rows; constant poses/points get zeros.  returns (dp K x 6, dl M x 3)."""
        K, M, O = self.c.K, self.c.M, self.c.O
        x = np.linalg.solve(S, rhs) if S.size else np.zeros(0)
        dp = np.zeros((K, 6))
        for p in range(K):
            if fm[p] >= 0:
                dp[p] = x[6 * fm[p]:6 * fm[p] + 6]
        pose_idx = np.ctypeslib.as_array(self.c.pose_idx, (O,)); point_idx = np.ctypeslib.as_array(self.c.point_idx, (O,))
        r = np.asarray(r).reshape(O, 2); Jp = np.asarray(Jp).reshape(O, 2, 6); Jl = np.asarray(Jl).reshape(O, 2, 3)
        s = r - np.einsum("oij,oj->oi", Jp, dp[pose_idx])          # F x is zero for constant poses (dp = 0)
        EtE = np.zeros((M, 3, 3)); g = np.zeros((M, 3))
        np.add.at(EtE, point_idx, np.einsum("oki,okj->oij", Jl, Jl))
        np.add.at(g, point_idx, np.einsum("oki,ok->oi", Jl, s))
        dl = np.zeros((M, 3))
        pf = self.keep["point_fixed"]
        Dl = np.zeros((M, 3)) if D is None else np.asarray(D)[6 * K:].reshape(M, 3)
        for l in range(M):
            if pf is not None and pf[l]:
                continue
            A = EtE[l] + np.diag(Dl[l] ** 2)
            if np.linalg.matrix_rank(A) == 3:
                dl[l] = np.linalg.solve(A, g[l])
        return dp, dl

    def lm_diagonal(self, Jp, Jl, radius=1e4):
        D = np.empty(6 * self.c.K + 3 * self.c.M)
        lib().vo_lm_diagonal(C.byref(self.c), _d(Jp), _d(Jl), C.c_double(radius), _d(D))
        return D

    def lm_solve(self, pose_delta, points, max_num_iterations=50, initial_radius=1e4, max_radius=1e16, min_radius=1e-32,
                 min_relative_decrease=1e-3, min_diag=1e-6, max_diag=1e32, function_tolerance=1e-6,
                 gradient_tolerance=1e-10, parameter_tolerance=1e-8, jacobi_scaling=True, max_invalid=5, apply_loss=True):
        """ceres::Solve as Optimizer::RunLocalBA configures it (Optimizer.cpp:73-81, 883-886), restated:
        TrustRegionMinimizer::Minimize (trust_region_minimizer.cc:67-139) with LevenbergMarquardtStrategy
        (levenberg_marquardt_strategy.cc:66-161), monotonic steps (trust_region_step_evaluator.cc:52-67), Jacobi
        scaling (trust_region_minimizer.cc:262-276), SPARSE_SCHUR = vo_schur_eliminate + LAPACK + BackSubstitute.
        Parameter blocks are plain vectors (no LocalParameterization is attached, Optimizer.cpp:771-773).
        returns (pose_delta, points, summary dict, trace list of per-iteration tuples)."""
        K, M, O = self.c.K, self.c.M, self.c.O
        x_p = np.array(pose_delta, np.float64).reshape(K, 6).copy(); x_l = np.array(points, np.float64).reshape(M, 3).copy()
        pose_idx = np.ctypeslib.as_array(self.c.pose_idx, (O,)); point_idx = np.ctypeslib.as_array(self.c.point_idx, (O,))
        free_p = self.keep["pose_fixed"] == 0; free_l = self.keep["point_fixed"] == 0

        def linearize(xp, xl):
            r, Jp, Jl, cost = self.evaluate(xp, xl, apply_loss=apply_loss)
            cp = np.zeros((K, 6)); cl = np.zeros((M, 3)); gp = np.zeros((K, 6)); gl = np.zeros((M, 3))
            np.add.at(cp, pose_idx, np.einsum("oia,oia->oa", Jp, Jp)); np.add.at(cl, point_idx, np.einsum("oia,oia->oa", Jl, Jl))
            np.add.at(gp, pose_idx, np.einsum("oia,oi->oa", Jp, r)); np.add.at(gl, point_idx, np.einsum("oia,oi->oa", Jl, r))
            cp[~free_p] = 0; cl[~free_l] = 0
            gmax = max(np.abs(gp[free_p]).max() if free_p.any() else 0.0, np.abs(gl[free_l]).max() if free_l.any() else 0.0)
            xnorm = np.sqrt((xp[free_p] ** 2).sum() + (xl[free_l] ** 2).sum())
            return r, Jp, Jl, cost, np.concatenate([cp.reshape(-1), cl.reshape(-1)]), gmax, xnorm

        r, Jp, Jl, x_cost, colsq, gmax, xnorm = linearize(x_p, x_l)
        scale = 1.0 / (1.0 + np.sqrt(colsq)) if jacobi_scaling else np.ones_like(colsq)
        radius, decrease = initial_radius, 2.0
        it = n_ok = n_bad = n_invalid = 0
        termination = 0
        trace = [(x_cost, 0.0, gmax, 0.0, 0.0, radius, 1, 0.0)]
        initial_cost = x_cost
        n_ok = 1
        if gmax <= gradient_tolerance:
            termination = 1
        while termination == 0:
            it += 1
            D = np.sqrt(np.clip(colsq * scale * scale, min_diag, max_diag) / radius) / scale
            S, rhs, fm = self.schur(r, Jp, Jl, D)
            try:
                np.linalg.cholesky(S)
                dp, dl = self.solve_step(r, Jp, Jl, D, S, rhs, fm)
                solver_ok = True
            except np.linalg.LinAlgError:
                dp = np.zeros((K, 6)); dl = np.zeros((M, 3)); solver_ok = False
            sp, sl = -dp, -dl
            sp[~free_p] = 0; sl[~free_l] = 0
            Jd = np.einsum("oia,oa->oi", Jp, sp[pose_idx]) + np.einsum("oia,oa->oi", Jl, sl[point_idx])
            model = float(-(Jd * (r + 0.5 * Jd)).sum())
            step_norm = float(np.sqrt((sp ** 2).sum() + (sl ** 2).sum()))
            valid = solver_ok and model > 0.0
            accepted, rel, cost_change, cand_cost = 0, 0.0, 0.0, 0.0
            if not valid:
                n_invalid += 1; n_bad += 1
                if n_invalid >= max_invalid:
                    termination = -1
                else:
                    radius /= decrease; decrease *= 2.0
                accepted = -1
            else:
                n_invalid = 0
                cand_cost = self.evaluate(x_p + sp, x_l + sl, apply_loss=apply_loss, outputs=False)
                cost_change = x_cost - cand_cost
                if step_norm <= parameter_tolerance * (xnorm + parameter_tolerance):
                    termination = 3
                elif abs(cost_change) <= function_tolerance * x_cost:
                    termination = 2
                else:
                    rel = cost_change / model
                    if rel > min_relative_decrease:
                        accepted = 1; n_ok += 1
                        x_p = x_p + sp; x_l = x_l + sl
                        r, Jp, Jl, x_cost, colsq, gmax, xnorm = linearize(x_p, x_l)
                        radius = min(max_radius, radius / max(1.0 / 3.0, 1.0 - (2.0 * rel - 1.0) ** 3)); decrease = 2.0
                    else:
                        n_bad += 1
                        radius /= decrease; decrease *= 2.0
            trace.append((x_cost if accepted == 1 else cand_cost, cost_change, gmax, step_norm, rel, radius, accepted, model))
            if termination == 0:
                if it >= max_num_iterations:
                    termination = 5
                elif accepted == 1 and gmax <= gradient_tolerance:
                    termination = 1
                elif radius <= min_radius:
                    termination = 4
        summary = dict(iterations=it, num_successful_steps=n_ok, num_unsuccessful_steps=n_bad, termination=termination,
                       initial_cost=initial_cost, final_cost=x_cost, final_radius=radius, final_gradient_max_norm=gmax)
        return x_p, x_l, summary, trace

    def chi2(self, pose_delta, points, thr=5.991):
        O = self.c.O
        pd = np.ascontiguousarray(pose_delta, np.float64); pts = np.ascontiguousarray(points, np.float64)
        chi2 = np.empty(O); flag = np.empty(O, np.uint8)
        lib().vo_window_chi2(C.byref(self.c), _d(pd), _d(pts), C.c_double(thr), _d(chi2),
                             flag.ctypes.data_as(_up))
        return chi2, flag


def make_factor(obs, cols, rows, Tcb, Twb_init, info, is_outlier=False, world_point=(0, 0, 0)):
    f = _Factor()
    f.obs = (C.c_double * 2)(*obs)
    f.cols, f.rows = float(cols), float(rows)
    f.Tcb = (C.c_double * 16)(*np.asarray(Tcb, np.float64).reshape(-1))
    f.Twb_init = (C.c_double * 16)(*np.asarray(Twb_init, np.float64).reshape(-1))
    f.info = (C.c_double * 4)(*np.asarray(info, np.float64).reshape(-1))
    f.is_outlier = int(bool(is_outlier))
    f.world_point = (C.c_double * 3)(*world_point)
    return f


def ba_evaluate(f, pose6, pt, want_jac=True, want_pose=True, want_point=True):
    r = np.full(2, np.nan); Jp = np.full((2, 6), np.nan); Jl = np.full((2, 3), np.nan)
    p6 = np.ascontiguousarray(pose6, np.float64); p3 = np.ascontiguousarray(pt, np.float64)
    ok = lib().vo_ba_evaluate(C.byref(f), _d(p6), _d(p3), _d(r), int(want_jac),
                              _d(Jp) if want_pose else None, _d(Jl) if want_point else None)
    return ok, r, Jp, Jl


def pnp_evaluate(f, pose6, want_jac=True):
    r = np.full(2, np.nan); Jp = np.full((2, 6), np.nan)
    p6 = np.ascontiguousarray(pose6, np.float64)
    ok = lib().vo_pnp_evaluate(C.byref(f), _d(p6), _d(r), _d(Jp) if want_jac else None)
    return ok, r, Jp


def ba_chi2(f, pose6, pt):
    p6 = np.ascontiguousarray(pose6, np.float64); p3 = np.ascontiguousarray(pt, np.float64)
    return lib().vo_ba_chi2(C.byref(f), _d(p6), _d(p3))


def pnp_chi2(f, pose6):
    p6 = np.ascontiguousarray(pose6, np.float64)
    return lib().vo_pnp_chi2(C.byref(f), _d(p6))


def so3_project(M):
    R = np.empty((3, 3)); Mc = np.ascontiguousarray(M, np.float64)
    lib().vo_so3_project(_d(Mc), _d(R))
    return R


def triangulate(b1, b2, T1w, T2w):
    """Estimator::TriangulateSinglePoint for n matches (fp32).  returns (points n x 3, valid n)."""
    b1 = np.ascontiguousarray(b1, np.float32).reshape(-1, 3); b2 = np.ascontiguousarray(b2, np.float32).reshape(-1, 3)
    T1 = np.ascontiguousarray(T1w, np.float32).reshape(16); T2 = np.ascontiguousarray(T2w, np.float32).reshape(16)
    n = b1.shape[0]
    pts = np.zeros((n, 3), np.float32); valid = np.zeros(n, np.uint8)
    fp = C.POINTER(C.c_float)
    f = lib().vo_triangulate
    f.restype = None
    f(b1.ctypes.data_as(fp), b2.ctypes.data_as(fp), T1.ctypes.data_as(fp), T2.ctypes.data_as(fp), n,
      pts.ctypes.data_as(fp), valid.ctypes.data_as(_up))
    return pts, valid
