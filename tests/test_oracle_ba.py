"""CPU tests pinning oracle/ba_ref.c (no reference golden vectors exist: parity unpinned).
Cross-checks: numpy/LAPACK mirror, finite differences over T*exp(eps), mpmath closed
forms, hand-built branch cases, and the host-compiled device arithmetic (tests/emu)."""
import ctypes as C

import numpy as np
import pytest

import ba_oracle
import ba_ref_np
from conftest import rel_err


@pytest.fixture(scope="module")
def win(synthetic):
    return synthetic.make_ba_window(K=6, M=120, seed=7, mid_iteration=True)


def test_so3_project_matches_lapack():
    rng = np.random.default_rng(0)
    for _ in range(200):
        M = rng.normal(size=(3, 3))
        R = ba_oracle.so3_project(M)
        assert np.allclose(R, ba_ref_np.so3_project(M), atol=1e-13)
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-13) and np.linalg.det(R) > 0
    # det < 0 input with distinct singular values: third column of U is flipped
    M = np.diag([3.0, 2.0, -1.0]) @ ba_ref_np.so3_exp(np.array([0.3, -0.2, 0.5]))
    assert np.allclose(ba_oracle.so3_project(M), ba_ref_np.so3_project(M), atol=1e-13)


def test_window_matches_numpy_mirror(win):
    W = ba_oracle.Window(win)
    for apply_loss in (True, False):
        r, Jp, Jl, cost = W.evaluate(win["pose_delta"], win["points"], apply_loss=apply_loss)
        r2, Jp2, Jl2, cost2 = ba_ref_np.window_evaluate(win, win["pose_delta"], win["points"], apply_loss)
        assert rel_err(r, r2) < 1e-12 and rel_err(Jp, Jp2) < 1e-12 and rel_err(Jl, Jl2) < 1e-12
        assert abs(cost - cost2) <= 1e-12 * cost2


def test_hoisted_and_threaded_variants_agree(win):
    W = ba_oracle.Window(win)
    a = W.evaluate(win["pose_delta"], win["points"], nthreads=1)
    b = W.evaluate(win["pose_delta"], win["points"], nthreads=4, hoisted=True)
    for x, y in zip(a[:3], b[:3]):
        assert np.array_equal(x, y)
    assert abs(a[3] - b[3]) < 1e-9 * a[3]


def test_schur_matches_dense_elimination(win):
    W = ba_oracle.Window(win)
    r, Jp, Jl, _ = W.evaluate(win["pose_delta"], win["points"])
    D = W.lm_diagonal(Jp, Jl, radius=1e4)
    S, rhs, fm = W.schur(r, Jp, Jl, D)
    S2, rhs2 = ba_ref_np.schur_dense(win, r, Jp, Jl, D)
    assert rel_err(S, S2) < 1e-11 and rel_err(rhs, rhs2) < 1e-11
    assert fm[0] == -1 and list(fm[1:]) == list(range(win["K"] - 1))


def test_normal_blocks_match_dense(win):
    W = ba_oracle.Window(win)
    r, Jp, Jl, _ = W.evaluate(win["pose_delta"], win["points"])
    Hpp, bp, Hll, bl, Hpl = W.normal_blocks(r, Jp, Jl)
    k = 3
    sel = win["pose_idx"] == k
    assert np.allclose(Hpp[k], np.einsum("oia,oib->ab", Jp[sel], Jp[sel]), rtol=1e-12)
    assert np.allclose(bp[k], np.einsum("oia,oi->a", Jp[sel], r[sel]), rtol=1e-12, atol=1e-12)
    assert np.all(Hpp[0] == 0)  # pose 0 is fixed


def _factor(win, i, **kw):
    k = win["pose_idx"][i]
    return ba_oracle.make_factor((win["obs_u"][i], win["obs_v"][i]), win["cols"], win["rows"], win["Tcb"],
                                 win["Twb_init"][k], win["info"], **kw)


def test_jacobians_by_finite_differences(win):
    """The pose Jacobian is the right-perturbation one: FD must perturb T <- T*exp(eps)
    (about delta = 0 this coincides with perturbing the parameter additively)."""
    rng = np.random.default_rng(1)
    zero = np.zeros(6)
    checked = 0
    for i in rng.choice(win["O"], 40, replace=False):
        f = _factor(win, i)
        pt = win["points"][win["point_idx"][i]]
        ok, r0, Jp, Jl = ba_oracle.ba_evaluate(f, zero, pt)
        if np.all(Jp == 0):
            continue
        h = 1e-6
        for a in range(6):
            e = np.zeros(6); e[a] = h
            rp = ba_oracle.ba_evaluate(f, e, pt)[1]; rm = ba_oracle.ba_evaluate(f, -e, pt)[1]
            assert np.allclose((rp - rm) / (2 * h), Jp[:, a], rtol=2e-5, atol=2e-5 * np.abs(Jp).max())
        for a in range(3):
            e = np.zeros(3); e[a] = h
            rp = ba_oracle.ba_evaluate(f, zero, pt + e)[1]; rm = ba_oracle.ba_evaluate(f, zero, pt - e)[1]
            assert np.allclose((rp - rm) / (2 * h), Jl[:, a], rtol=2e-5, atol=2e-5 * np.abs(Jl).max())
        checked += 1
    assert checked >= 30


def test_projection_closed_form_mpmath(win):
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 50
    rng = np.random.default_rng(2)
    cols, rows = win["cols"], win["rows"]
    for i in rng.choice(win["O"], 100, replace=False):
        k = win["pose_idx"][i]
        Rcw, tcw, _, _ = ba_ref_np.pose_chain(win["Twb_init"][k], win["Tcb"], win["pose_delta"][k])
        pt = win["points"][win["point_idx"][i]]
        Pc = [sum(mp.mpf(float(Rcw[a, b])) * mp.mpf(float(pt[b])) for b in range(3)) + mp.mpf(float(tcw[a])) for a in range(3)]
        L = mp.sqrt(Pc[0] ** 2 + Pc[1] ** 2 + Pc[2] ** 2)
        u = cols * (mp.mpf(0.5) + mp.atan2(Pc[0], Pc[2]) / (2 * mp.pi))
        v = rows * (mp.mpf(0.5) + mp.asin(Pc[1] / L) / mp.pi)
        du = mp.mpf(float(win["obs_u"][i])) - u; dv = mp.mpf(float(win["obs_v"][i])) - v
        if du > cols / 2: du -= cols
        elif du < -cols / 2: du += cols
        f = _factor(win, i)
        _, r, _, _ = ba_oracle.ba_evaluate(f, win["pose_delta"][k], pt, want_jac=False)
        if abs(du) > 100 or abs(dv) > 100:
            assert tuple(r) == (100.0, 100.0)
        else:
            assert abs(float(du) - r[0]) < 1e-9 and abs(float(dv) - r[1]) < 1e-9


def test_branch_cases(win):
    i = 5
    k = win["pose_idx"][i]
    pt = win["points"][win["point_idx"][i]]
    xi = win["pose_delta"][k]
    # (i) outlier flag
    f = _factor(win, i, is_outlier=True)
    ok, r, Jp, Jl = ba_oracle.ba_evaluate(f, xi, pt)
    assert ok == 1 and tuple(r) == (640.0, 480.0) and not Jp.any() and not Jl.any()
    assert ba_oracle.ba_chi2(f, xi, pt) == 0.0
    # (ii) point at the camera centre: L < 1e-10
    Rcw, tcw, _, _ = ba_ref_np.pose_chain(win["Twb_init"][k], win["Tcb"], xi)
    centre = -Rcw.T @ tcw
    f = _factor(win, i)
    ok, r, Jp, Jl = ba_oracle.ba_evaluate(f, xi, centre)
    assert ok == 1 and tuple(r) == (640.0, 360.0) and not Jp.any() and not Jl.any()
    assert ba_oracle.ba_chi2(f, xi, centre) == 1000.0
    fp = _factor(win, i, world_point=tuple(centre))
    ok, r, Jp = ba_oracle.pnp_evaluate(fp, xi)
    assert ok == 0 and np.isnan(r).all() and not Jp.any()          # residual untouched, returns false
    assert ba_oracle.pnp_chi2(fp, xi) == np.finfo(np.float64).max
    # (iii) > 100 px error
    f = ba_oracle.make_factor((win["obs_u"][i] + 150.0, win["obs_v"][i]), win["cols"], win["rows"], win["Tcb"],
                              win["Twb_init"][k], win["info"])
    ok, r, Jp, Jl = ba_oracle.ba_evaluate(f, xi, pt)
    assert tuple(r) == (100.0, 100.0) and not Jp.any() and not Jl.any()
    assert ba_oracle.ba_chi2(f, xi, pt) > 100.0 ** 2            # chi-square has no clamp
    # seam wrap: observation at u=5, projection near u=W-5
    Pc = np.array([np.sin(-np.pi + 2 * np.pi * 5.0 / win["cols"]) , 0.0, np.cos(-np.pi + 2 * np.pi * 5.0 / win["cols"])]) * 4.0
    Pw = Rcw.T @ (Pc - tcw)
    f = ba_oracle.make_factor((win["cols"] - 5.0, win["rows"] / 2), win["cols"], win["rows"], win["Tcb"],
                              win["Twb_init"][k], win["info"])
    ok, r, _, _ = ba_oracle.ba_evaluate(f, xi, Pw)
    assert abs(r[0] - (-10.0)) < 1e-6 and abs(r[1]) < 1e-6
    # (iv) point on the +Y axis: xz^2 < 1e-10 -> zero Jacobian, residual written
    Pw = Rcw.T @ (np.array([0.0, 3.0, 1e-6]) - tcw)   # z dominates the rounding noise in x
    f = ba_oracle.make_factor((win["cols"] / 2, win["rows"] - 1.0), win["cols"], win["rows"], win["Tcb"],
                              win["Twb_init"][k], win["info"])
    ok, r, Jp, Jl = ba_oracle.ba_evaluate(f, xi, Pw)
    assert not Jp.any() and not Jl.any() and np.isfinite(r).all() and abs(r[1]) < 100
    # theta < 1e-10 in exp
    R, t = ba_ref_np.se3_exp(np.array([0.1, 0.2, 0.3, 1e-12, 0, 0]))
    assert np.allclose(R, np.eye(3), atol=1e-11) and np.allclose(t, [0.1, 0.2, 0.3])


def test_pnp_matches_ba_pose_block(win):
    rng = np.random.default_rng(3)
    for i in rng.choice(win["O"], 30, replace=False):
        k = win["pose_idx"][i]
        pt = win["points"][win["point_idx"][i]]
        f = _factor(win, i, world_point=tuple(pt))
        ok1, r1, Jp1, _ = ba_oracle.ba_evaluate(f, win["pose_delta"][k], pt)
        ok2, r2, Jp2 = ba_oracle.pnp_evaluate(f, win["pose_delta"][k])
        assert ok2 == 1 and np.array_equal(r1, r2) and rel_err(Jp2, Jp1) < 1e-12
        assert abs(ba_oracle.ba_chi2(f, win["pose_delta"][k], pt) - ba_oracle.pnp_chi2(f, win["pose_delta"][k])) < 1e-12


def test_huber_and_corrector():
    lib = ba_oracle.lib()
    rho = (C.c_double * 3)()
    lib.vo_huber(C.c_double(1.0), C.c_double(0.25), rho)
    assert tuple(rho) == (0.25, 1.0, 0.0)
    lib.vo_huber(C.c_double(1.0), C.c_double(4.0), rho)
    assert tuple(rho) == (3.0, 0.5, -0.0625)
    r = np.array([3.0, 4.0]); J = np.ones((2, 6)); E = np.ones((2, 3))
    dp = C.POINTER(C.c_double)
    cost = lib.vo_apply_loss(C.c_double(1.0), r.ctypes.data_as(dp), 2, J.ctypes.data_as(dp), 6, E.ctypes.data_as(dp), 3)
    assert abs(cost - 0.5 * (2 * 5 - 1)) < 1e-15
    assert np.allclose(r, np.array([3.0, 4.0]) * np.sqrt(0.2)) and np.allclose(J, np.sqrt(0.2))


def test_device_arithmetic_on_host(win, emu_lib):
    """csrc/ba_math.cuh (what the kernels execute) compiled for the host == oracle."""
    dp = C.POINTER(C.c_double); ip = C.POINTER(C.c_int); up = C.POINTER(C.c_ubyte)
    K, O = win["K"], win["O"]
    posebuf = np.zeros((K, 33))
    Twb = np.ascontiguousarray(win["Twb_init"]); Tcb = np.ascontiguousarray(win["Tcb"]); pd = np.ascontiguousarray(win["pose_delta"])
    emu_lib.emu_pose_prepare(K, Twb.ctypes.data_as(dp), Tcb.ctypes.data_as(dp), pd.ctypes.data_as(dp), posebuf.ctypes.data_as(dp))
    for k in range(K):
        Rcw, tcw, Rbw, tbw = ba_ref_np.pose_chain(win["Twb_init"][k], win["Tcb"], win["pose_delta"][k])
        assert np.allclose(posebuf[k, :9].reshape(3, 3), Rcw, atol=1e-14) and np.allclose(posebuf[k, 9:12], tcw, atol=1e-14)
        assert np.allclose(posebuf[k, 12:21].reshape(3, 3), Rbw, atol=1e-14) and np.allclose(posebuf[k, 21:24], tbw, atol=1e-14)
    r = np.zeros((O, 2)); Jp = np.zeros((O, 2, 6)); Jl = np.zeros((O, 2, 3)); chi2 = np.zeros(O)
    pts = np.ascontiguousarray(win["points"]); info = np.ascontiguousarray(win["info"])
    cost = emu_lib.emu_evaluate(O, win["pose_idx"].ctypes.data_as(ip), win["point_idx"].ctypes.data_as(ip),
                                win["obs_u"].ctypes.data_as(dp), win["obs_v"].ctypes.data_as(dp),
                                win["outlier"].ctypes.data_as(up), posebuf.ctypes.data_as(dp), pts.ctypes.data_as(dp),
                                C.c_double(win["cols"]), C.c_double(win["rows"]), info.ctypes.data_as(dp),
                                Tcb.ctypes.data_as(dp), C.c_double(win["huber_a"]), 1, r.ctypes.data_as(dp),
                                Jp.ctypes.data_as(dp), Jl.ctypes.data_as(dp), chi2.ctypes.data_as(dp))
    W = ba_oracle.Window(win)
    r0, Jp0, Jl0, cost0 = W.evaluate(win["pose_delta"], win["points"])
    c0, _ = W.chi2(win["pose_delta"], win["points"])
    assert rel_err(r, r0) < 1e-12 and rel_err(Jp, Jp0) < 1e-12 and rel_err(Jl, Jl0) < 1e-12
    assert abs(cost - cost0) < 1e-12 * cost0 and rel_err(chi2, c0) < 1e-12
    # the optional fp32 mode (vio360_ba_evaluate_f32): same math in float, north-star bar 1e-4 relative
    r32 = np.zeros((O, 2)); Jp32 = np.zeros((O, 2, 6)); Jl32 = np.zeros((O, 2, 3))
    emu_lib.emu_evaluate_f32.restype = C.c_double
    cost32 = emu_lib.emu_evaluate_f32(O, win["pose_idx"].ctypes.data_as(ip), win["point_idx"].ctypes.data_as(ip),
                                      win["obs_u"].ctypes.data_as(dp), win["obs_v"].ctypes.data_as(dp),
                                      win["outlier"].ctypes.data_as(up), posebuf.ctypes.data_as(dp), pts.ctypes.data_as(dp),
                                      C.c_double(win["cols"]), C.c_double(win["rows"]), info.ctypes.data_as(dp),
                                      Tcb.ctypes.data_as(dp), C.c_double(win["huber_a"]), 1, r32.ctypes.data_as(dp),
                                      Jp32.ctypes.data_as(dp), Jl32.ctypes.data_as(dp))
    special = np.abs(Jp0).reshape(O, -1).max(axis=1) == 0
    # a float-rounded observation can cross the 100 px clamp the other way: compare the rows both precisions treat alike
    same = special == (np.abs(Jp32).reshape(O, -1).max(axis=1) == 0)
    assert same.mean() > 0.999
    assert rel_err(r32[same], r0[same]) < 1e-4 and rel_err(Jp32[same], Jp0[same]) < 1e-4 and rel_err(Jl32[same], Jl0[same]) < 1e-4
    assert abs(cost32 - cost0) < 1e-4 * cost0
    assert not np.array_equal(r32[~special], r0[~special])          # it really ran in float


def _mp_polar(mp, M):
    """SO3d(R) (LieUtils.cpp:275-288) at 50 digits: U V^T of the SVD, third column of U negated when det < 0."""
    U, S, Vt = mp.svd_r(M)
    R = U * Vt
    if mp.det(R) < 0:
        for i in range(3):
            U[i, 2] = -U[i, 2]
        R = U * Vt
    return R


def _mp_hat(mp, w):
    return mp.matrix([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])


def _mp_ba_factor(mp, win, i, xi, pt):
    """Independent 50-digit transcription of BAFactor::Evaluate (Factors.cpp:349-534) with the pose chain of
    LieUtils.cpp:203-219, 275-333 / LieUtils.h:207-283 -- written from the reference source, not from
    oracle/ba_ref.c or its numpy mirror.  Returns (r, J_pose 2x6, J_point 2x3) or None on a special branch."""
    f = lambda a: mp.mpf(float(a))
    k = win["pose_idx"][i]
    T0 = win["Twb_init"][k]
    R0 = _mp_polar(mp, mp.matrix([[f(T0[a, b]) for b in range(3)] for a in range(3)]))      # SE3d(m_T_wb_init)
    t0 = mp.matrix([f(T0[a, 3]) for a in range(3)])
    rho = mp.matrix([f(x) for x in xi[:3]]); phi = mp.matrix([f(x) for x in xi[3:]])
    th = mp.sqrt(sum(p * p for p in phi))
    I3 = mp.eye(3)
    if th < mp.mpf("1e-10"):                                                               # LieUtils.cpp:206-209, 316-318
        Rd = _mp_polar(mp, I3 + _mp_hat(mp, phi)); td = rho
    else:
        Kh = _mp_hat(mp, phi / th)
        Rd = _mp_polar(mp, I3 + mp.sin(th) * Kh + (1 - mp.cos(th)) * Kh * Kh)              # :211-218
        P = _mp_hat(mp, phi)
        V = I3 + (1 - mp.cos(th)) / th ** 2 * P + (th - mp.sin(th)) / th ** 3 * P * P       # :320-329
        td = V * rho
    Rwb = _mp_polar(mp, R0 * Rd); twb = t0 + R0 * td                                         # LieUtils.h:268-271
    Rbw = _mp_polar(mp, Rwb.T); tbw = Rbw * (-twb)                                           # :279-283
    Tcb = win["Tcb"]
    Rcb_raw = mp.matrix([[f(Tcb[a, b]) for b in range(3)] for a in range(3)])
    Rcb = _mp_polar(mp, Rcb_raw); tcb = mp.matrix([f(Tcb[a, 3]) for a in range(3)])         # SE3d Tcb(m_Tcb)
    Rcw = _mp_polar(mp, Rcb * Rbw); tcw = tcb + Rcb * tbw                                    # Tcb * Tbw
    Pw = mp.matrix([f(x) for x in pt])
    Pc = Rcw * Pw + tcw
    x, y, z = Pc[0], Pc[1], Pc[2]
    L = mp.sqrt(x * x + y * y + z * z)
    if L < mp.mpf("1e-10"):
        return None
    cols, rows = f(win["cols"]), f(win["rows"])
    u = cols * (mp.mpf(0.5) + mp.atan2(x, z) / (2 * mp.pi))
    v = rows * (mp.mpf(0.5) + mp.asin(y / L) / mp.pi)                                      # phi = -asin(y/L); v = rows(0.5 - phi/pi)
    du = f(win["obs_u"][i]) - u; dv = f(win["obs_v"][i]) - v
    if du > cols / 2: du -= cols
    elif du < -cols / 2: du += cols
    if abs(du) > 100 or abs(dv) > 100:
        return None
    info = np.asarray(win["info"], np.float64).reshape(2, 2)
    Lc = mp.cholesky(mp.matrix([[f(info[a, b]) for b in range(2)] for a in range(2)]))      # Eigen::LLT matrixL()
    r = Lc * mp.matrix([du, dv])
    xz2 = x * x + z * z; L2 = L * L; xz = mp.sqrt(xz2)
    if xz2 < mp.mpf("1e-10") or L2 < mp.mpf("1e-10"):
        return None
    Jc = mp.matrix([[-cols / (2 * mp.pi) * z / xz2, 0, cols / (2 * mp.pi) * x / xz2],       # Factors.cpp:486-492
                    [rows / mp.pi * (x * y) / (L2 * xz), -rows / mp.pi * xz / L2, rows / mp.pi * (y * z) / (L2 * xz)]])
    Jw = Lc * Jc
    Pb = Rbw * Pw + tbw                                                                     # :503
    Jcp = mp.zeros(3, 6)
    Jrot = Rcb_raw * _mp_hat(mp, Pb)                                                        # :509 (raw m_Tcb block)
    for a in range(3):
        for b in range(3):
            Jcp[a, b] = -Rcb_raw[a, b]                                                      # :506
            Jcp[a, 3 + b] = Jrot[a, b]
    return r, Jw * Jcp, Jw * (Rcb_raw * Rbw)                                               # :517, :527-530


def test_jacobians_closed_form_mpmath(win, synthetic):
    """the 1e-9 bar on J_pose / J_point anchored independently of the numpy mirror --
    oracle Jacobians against a 50-digit mpmath transcription of Factors.cpp:349-534 + the LieUtils pose chain
    (polar factor from mp.svd_r), on >= 100 factors whose T_wb_init went through fp32 (Optimizer.cpp:777) at
    a mid-iteration delta, plus factors at delta = 0 and a small-angle delta."""
    mp = pytest.importorskip("mpmath")
    mp.mp.dps = 50
    rng = np.random.default_rng(11)
    wins = [(win, win["pose_delta"]), (win, np.zeros_like(win["pose_delta"])),
            (win, np.tile(np.array([1e-3, -2e-3, 5e-4, 3e-12, -2e-12, 1e-12]), (win["K"], 1)))]
    big = synthetic.make_ba_window(K=12, M=60, degree=6, seed=77, mid_iteration=True, loop_radius=3.0)
    wins.append((big, big["pose_delta"]))
    checked = 0
    worst = 0.0
    for w, deltas in wins:
        n = 60 if w is win and deltas is win["pose_delta"] else 25
        for i in rng.choice(w["O"], n, replace=False):
            k = w["pose_idx"][i]
            pt = w["points"][w["point_idx"][i]]
            ref = _mp_ba_factor(mp, w, i, deltas[k], pt)
            fct = ba_oracle.make_factor((w["obs_u"][i], w["obs_v"][i]), w["cols"], w["rows"], w["Tcb"],
                                        w["Twb_init"][k], w["info"])
            ok, r, Jp, Jl = ba_oracle.ba_evaluate(fct, deltas[k], pt)
            if ref is None:
                assert not Jp.any() and not Jl.any()
                continue
            rm, Jpm, Jlm = ref
            Jp_ref = np.array([[float(Jpm[a, b]) for b in range(6)] for a in range(2)])
            Jl_ref = np.array([[float(Jlm[a, b]) for b in range(3)] for a in range(2)])
            r_ref = np.array([float(rm[0]), float(rm[1])])
            assert np.abs(r - r_ref).max() < 1e-9                      # pixels of magnitude ~1e3: absolute bar
            ep = rel_err(Jp, Jp_ref); el = rel_err(Jl, Jl_ref)
            worst = max(worst, ep, el)
            assert ep < 1e-9 and el < 1e-9, (i, ep, el)
            checked += 1
    assert checked >= 100, checked
    assert worst < 1e-10        # observed ~1e-13; the bar itself is 1e-9
