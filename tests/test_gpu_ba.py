"""GPU parity tests for path B through the C ABI (include/vio360.h) against the oracle.
Tolerance: 1e-9 relative (max-norm) in fp64, BASELINE.json north_star."""
import os

import numpy as np
import pytest

import ba_oracle
from conftest import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-9
GOLD = os.path.join(os.path.dirname(__file__), "golden", "ba_golden.npz")


def _oracle_all(win, radius=1e4):
    W = ba_oracle.Window(win)
    r, Jp, Jl, cost = W.evaluate(win["pose_delta"], win["points"])
    D = W.lm_diagonal(Jp, Jl, radius)
    S, rhs, fm = W.schur(r, Jp, Jl, D)
    blocks = W.normal_blocks(r, Jp, Jl)
    chi2, flag = W.chi2(win["pose_delta"], win["points"], 5.991)
    return dict(r=r, Jp=Jp, Jl=Jl, cost=cost, D=D, S=S, rhs=rhs, blocks=blocks, chi2=chi2, flag=flag)


def _check_window(bindings, ctx, win, ref=None):
    ref = ref or _oracle_all(win)
    ba = bindings.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    r, Jp, Jl, cost = ba.evaluate(True, True)
    assert rel_err(r, ref["r"]) < TOL and rel_err(Jp, ref["Jp"]) < TOL and rel_err(Jl, ref["Jl"]) < TOL
    assert abs(cost - ref["cost"]) < TOL * ref["cost"]
    # special-case rows must be reproduced exactly
    special = (np.abs(ref["Jp"]).reshape(len(r), -1).max(axis=1) == 0)
    assert np.array_equal(r[special], ref["r"][special]) and not Jp[special].any() and not Jl[special].any()
    r2, _, _, cost2 = ba.evaluate(True, False)
    Wo = ba_oracle.Window(win)
    ro, _, _, co = Wo.evaluate(win["pose_delta"], win["points"], apply_loss=False)
    assert rel_err(r2, ro) < TOL and abs(cost2 - co) < TOL * co
    colsq = ba.column_sqnorm(True)
    D = np.sqrt(np.clip(colsq, 1e-6, 1e32) / 1e4)
    free = np.concatenate([np.repeat(win["pose_fixed"] == 0, 6), np.repeat(win["point_fixed"] == 0, 3)])
    assert rel_err(D[free], ref["D"][free]) < TOL
    S, rhs, cost3 = ba.build_reduced_system(ref["D"], True)
    assert S.shape == ref["S"].shape
    assert rel_err(S, ref["S"]) < TOL and rel_err(rhs, ref["rhs"]) < TOL and abs(cost3 - ref["cost"]) < TOL * ref["cost"]
    assert np.array_equal(S, S.T)
    Hpp, bp, Hll, bl, Hpl = ba.normal_blocks(True)
    for a, b in zip((Hpp, bp, Hll, bl, Hpl), ref["blocks"]):
        assert rel_err(a, b) < TOL
    chi2, flag = ba.chi2(5.991)
    assert rel_err(chi2, ref["chi2"]) < TOL
    amb = np.abs(ref["chi2"] - 5.991) < 1e-6
    assert np.array_equal(flag[~amb], ref["flag"][~amb])
    ba.close()
    return S, rhs


def test_solve_reduced_and_back_substitution(bindings, ctx, synthetic):
    """block-envelope Cholesky of S on the device + landmark back substitution
    against LAPACK / numpy on the oracle's system.  A different factorisation order than Eigen's LLT,
    so the tolerance is 1e-7 relative (cond(S) ~ 1e6 with the LM diagonal), written here."""
    for win in (synthetic.make_ba_window(K=10, M=300, seed=20001),
                synthetic.make_ba_window(K=40, M=1500, degree=8, seed=20002),
                synthetic.make_ba_window(K=36, M=400, degree=30, seed=20003)):   # envelope wider than the shared-memory window
        W = ba_oracle.Window(win)
        r, Jp, Jl, _ = W.evaluate(win["pose_delta"], win["points"])
        D = W.lm_diagonal(Jp, Jl, 1e4)
        S, rhs, fm = W.schur(r, Jp, Jl, D)
        dp_ref, dl_ref = W.solve_step(r, Jp, Jl, D, S, rhs, fm)
        ba = bindings.BAWindow(ctx, win)
        ba.set_parameters(win["pose_delta"], win["points"])
        ba.build_reduced_system(D, True, fetch=False)
        dp, dl = ba.solve_reduced(True)
        assert rel_err(dp, dp_ref) < 1e-7 and rel_err(dl, dl_ref) < 1e-7
        assert not dp[win["pose_fixed"] != 0].any() and not dl[win["point_fixed"] != 0].any()
        # the step reduces the linearised cost: |r - J d|^2 + |D d|^2 < |r|^2
        dp2, _ = ba.solve_reduced(False)
        assert np.array_equal(dp, dp2)
        ba.close()


def test_triangulate_vs_oracle(ctx, synthetic):
    """fp32 DLT triangulation; same algorithm as the oracle, different FMA contraction:
    tolerance 2e-3 of the point's distance (fp32 with a 0.5 m baseline and up to 10 m depth)."""
    b1, b2, T1, T2, P = synthetic.make_triangulation_pairs(1000, seed=6)
    pts, valid = ctx.triangulate(b1, b2, T1, T2)
    opts, ovalid = ba_oracle.triangulate(b1, b2, T1, T2)
    assert np.array_equal(valid, ovalid) and valid.all()
    scale = np.linalg.norm(opts, axis=1)
    assert (np.linalg.norm(pts - opts, axis=1) / scale).max() < 2e-3
    assert np.median(np.linalg.norm(pts - P, axis=1) / scale) < 1e-3
    assert ctx.triangulate(b1[:0], b2[:0], T1, T2)[0].shape == (0, 3)


def test_golden_window(bindings, ctx):
    g = np.load(GOLD)
    K, M, O, cols, rows, huber = g["scalars"]
    win = {k[3:]: g[k] for k in g.files if k.startswith("in_")}
    win.update(K=int(K), M=int(M), O=int(O), cols=float(cols), rows=float(rows), huber_a=float(huber))
    ref = dict(r=g["r"], Jp=g["Jp"], Jl=g["Jl"], cost=float(g["cost"]), D=g["D"], S=g["S"], rhs=g["rhs"],
               blocks=(g["Hpp"], g["bp"], g["Hll"], g["bl"], g["Hpl"]), chi2=g["chi2"], flag=g["flag"])
    _check_window(bindings, ctx, win, ref)


def test_evaluate_fp32_mode(bindings, ctx, synthetic):
    """The optional fp32 mode of the factor evaluation (north star: 1e-4 relative): k_evaluate_f32 against the fp64
    oracle on the local window, at the linearisation point and mid-iteration, with and without the loss."""
    for mid in (False, True):
        win = synthetic.make_ba_window(K=10, M=2000, seed=20000, mid_iteration=mid)
        Wo = ba_oracle.Window(win)
        ba = bindings.BAWindow(ctx, win)
        ba.set_parameters(win["pose_delta"], win["points"])
        for loss in (True, False):
            r0, Jp0, Jl0, cost0 = Wo.evaluate(win["pose_delta"], win["points"], apply_loss=loss)
            r, Jp, Jl, cost = ba.evaluate(True, loss, f32=True)
            O = len(r0)
            same = (np.abs(Jp0).reshape(O, -1).max(axis=1) == 0) == (np.abs(Jp).reshape(O, -1).max(axis=1) == 0)
            assert same.mean() > 0.999          # a float-rounded projection may cross the 100 px clamp the other way
            assert rel_err(r[same], r0[same]) < 1e-4 and rel_err(Jp[same], Jp0[same]) < 1e-4 and rel_err(Jl[same], Jl0[same]) < 1e-4
            if same.all():
                assert abs(cost - cost0) < 1e-4 * cost0
            assert not np.array_equal(r, r0)    # it ran in float
            r64, _, _, _ = ba.evaluate(True, loss)          # the fp64 path is untouched by the mode
            assert rel_err(r64, r0) < TOL
        ba.close()


def test_local_window_c3(bindings, ctx, synthetic):
    """BASELINE config 3: 10 keyframes, ~20k observations, mid-iteration state."""
    win = synthetic.make_ba_window(K=10, M=2000, seed=20000, mid_iteration=True)
    assert 18000 < win["O"] <= 20000
    _check_window(bindings, ctx, win)
    win0 = synthetic.make_ba_window(K=10, M=2000, seed=20000, mid_iteration=False)
    _check_window(bindings, ctx, win0)


def test_unsorted_observations_and_ragged_structure(bindings, ctx, synthetic):
    win = synthetic.make_ba_window(K=12, M=300, degree=5, seed=5, mid_iteration=True, sort_by_landmark=False)
    # ragged: a landmark with no observation, one with a single observation, duplicates of a (pose, landmark) pair
    keep = ~((win["point_idx"] == 7) | ((win["point_idx"] == 9) & (np.arange(win["O"]) % 5 != 0)))
    for k in ("pose_idx", "point_idx", "obs_u", "obs_v", "outlier"):
        win[k] = np.ascontiguousarray(win[k][keep])
    dup = np.where(win["point_idx"] == 20)[0][:2]
    for k in ("pose_idx", "point_idx", "obs_u", "obs_v", "outlier"):
        win[k] = np.ascontiguousarray(np.concatenate([win[k], win[k][dup]]))
    win["obs_u"][-1] += 0.5
    win["O"] = int(win["obs_u"].shape[0])
    _check_window(bindings, ctx, win)


def test_high_degree_landmarks(bindings, ctx, synthetic):
    """landmarks seen by more than 32 poses exercise the multi-chunk path"""
    win = synthetic.make_ba_window(K=48, M=60, seed=9, mid_iteration=True, yaw_spread_deg=340.0)
    deg = np.bincount(win["point_idx"], minlength=win["M"])
    assert deg.max() > 32
    _check_window(bindings, ctx, win)


def test_wide_landmarks_without_atomics(bindings, ctx, synthetic):
    """landmarks whose observing poses span more than the 24-pose chunk window (a loop closure, a
    degree-50 landmark) are eliminated by the chunk kernel too -- pass 1 (k_wide_prepare) factorises E'E over all their
    observations, group-pair jobs contribute the blocks -- so the fp64-atomic kernel k_schur_warp is not launched, the
    build is deterministic (bit-identical when repeated) and still within 1e-9 of the oracle; mixed with regular
    landmarks, constant points, constant poses, outlier flags, with and without the LM diagonal."""
    rng = np.random.default_rng(4)
    # (a) every landmark wide: degree 50 on 120 poses; (b) a regular band plus loop closures
    wa = synthetic.make_ba_window(K=120, M=300, degree=50, seed=31, mid_iteration=True, loop_radius=5.0)
    wb = synthetic.make_ba_window(K=60, M=400, degree=8, seed=32, mid_iteration=True, loop_radius=4.0)
    # loop closure: 40 landmarks of the first poses are also seen from the last ones
    import ba_ref_np
    extra_p, extra_l, extra_u, extra_v = [], [], [], []
    for l in range(0, 40):
        for p in (55, 57, 59):     # the loop closes: pose 59 stands next to pose 0
            Rcw, tcw, _, _ = ba_ref_np.pose_chain(wb["Twb_init"][p], wb["Tcb"], wb["pose_delta"][p])
            u, v = synthetic.erp_project((Rcw @ wb["points"][l] + tcw)[None, :], wb["cols"], wb["rows"])
            u, v = float(u[0]) + rng.normal(), float(v[0]) + rng.normal()
            if 20 < u < wb["cols"] - 20 and 20 < v < wb["rows"] - 20:
                extra_p.append(p); extra_l.append(l); extra_u.append(np.float64(np.float32(u))); extra_v.append(np.float64(np.float32(v)))
    assert len(extra_p) > 60
    wb = dict(wb)
    wb["pose_idx"] = np.concatenate([wb["pose_idx"], np.array(extra_p, np.int32)]); wb["point_idx"] = np.concatenate([wb["point_idx"], np.array(extra_l, np.int32)])
    wb["obs_u"] = np.concatenate([wb["obs_u"], np.array(extra_u)]); wb["obs_v"] = np.concatenate([wb["obs_v"], np.array(extra_v)])
    wb["outlier"] = np.concatenate([wb["outlier"], np.zeros(len(extra_p), np.uint8)]); wb["O"] = len(wb["obs_u"])
    wb["pose_fixed"] = wb["pose_fixed"].copy(); wb["pose_fixed"][30] = 1
    for win in (wa, wb):
        ref = _oracle_all(win)
        ba = bindings.BAWindow(ctx, win)
        ba.set_parameters(win["pose_delta"], win["points"])
        ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
        S, rhs, cost = ba.build_reduced_system(ref["D"], True)
        S2, rhs2, cost2 = ba.build_reduced_system(ref["D"], True)
        assert ctx.kernel_time("k_schur_warp")[1] == 0 and ctx.kernel_time("k_wide_prepare")[1] == 2
        ctx.enable_kernel_timing(False)
        assert rel_err(S, ref["S"]) < TOL and rel_err(rhs, ref["rhs"]) < TOL and abs(cost - ref["cost"]) < TOL * ref["cost"]
        assert np.array_equal(S, S2) and np.array_equal(rhs, rhs2) and cost == cost2          # no atomics: bit-identical
        assert np.array_equal(S, S.T)
        ba.set_lm_diagonal(None)
        S0, rhs0, _ = ba.build_reduced_system(None, True)                                       # without an LM diagonal
        W = ba_oracle.Window(win)
        So, rhso, _ = W.schur(ref["r"], ref["Jp"], ref["Jl"], None)
        # without the diagonal a constant-free landmark seen from few poses can be ill conditioned: compare on the scale of S
        assert rel_err(S0, So) < 1e-8 and rel_err(rhs0, rhso) < 1e-8
        # solve + back substitution through the general envelope factorisation
        ba.build_reduced_system(ref["D"], True)
        dp, dl = ba.solve_reduced(True)
        odp, odl = W.solve_step(ref["r"], ref["Jp"], ref["Jl"], ref["D"], ref["S"], ref["rhs"], W.schur(ref["r"], ref["Jp"], ref["Jl"], ref["D"])[2])
        assert rel_err(dp, odp) < 1e-7 and rel_err(dl, odl) < 1e-7
        # sharded: the partial systems of three ranks sum to the full one (wide landmarks in every shard)
        Ssum = np.zeros_like(S); rsum = np.zeros_like(rhs)
        for rnk in range(3):
            ba.set_shard(rnk, 3)
            Sp, rp, _ = ba.build_reduced_system(ref["D"], True)
            Ssum += Sp; rsum += rp
        assert rel_err(Ssum, ref["S"]) < TOL and rel_err(rsum, ref["rhs"]) < TOL
        ba.close()


def test_outlier_flags_and_reuse(bindings, ctx, synthetic):
    win = synthetic.make_ba_window(K=6, M=150, seed=3, mid_iteration=True)
    ba = bindings.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    chi2, flag = ba.chi2(5.991)
    win2 = dict(win); win2["outlier"] = flag.copy()
    ba.set_outliers(flag)                                  # Optimizer.cpp:903-906
    r, Jp, Jl, cost = ba.evaluate(True, True)
    ref = ba_oracle.Window(win2).evaluate(win["pose_delta"], win["points"])
    assert rel_err(r, ref[0]) < TOL and rel_err(Jp, ref[1]) < TOL and abs(cost - ref[3]) < TOL * ref[3]
    r_raw = ba.evaluate(False, False)[0]                   # without the Huber correction
    assert np.all(r_raw[flag == 1] == (640.0, 480.0))
    ba.set_outliers(None)
    r, _, _, _ = ba.evaluate(False, True)
    ref = ba_oracle.Window(dict(win, outlier=np.zeros(win["O"], np.uint8))).evaluate(win["pose_delta"], win["points"])
    assert rel_err(r, ref[0]) < TOL
    ba.close()


def test_sharded_partials_sum_to_full_system(bindings, ctx, synthetic):
    """multi-GPU decomposition emulated on one device: the landmark shards'
    partial reduced systems add up to the single-GPU one."""
    win = synthetic.make_ba_window(K=40, M=1500, degree=8, seed=77, mid_iteration=True)
    ref = _oracle_all(win)
    ba = bindings.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    for world in (2, 4, 8):
        S = 0; rhs = 0; cost = 0
        for rank in range(world):
            ba.set_shard(rank, world)
            s, b, c = ba.build_reduced_system(ref["D"], True)
            S = S + s; rhs = rhs + b; cost += c
        assert rel_err(S, ref["S"]) < TOL and rel_err(rhs, ref["rhs"]) < TOL and abs(cost - ref["cost"]) < TOL * ref["cost"]
    ba.close()


def test_rebuild_after_shard_and_parameter_changes(bindings, ctx, synthetic):
    """The unsharded build only rewrites S inside its block envelope (zeros outside are kept from
    the first build): rebuilding after a sharded build, and at new parameters, must still match."""
    win = synthetic.make_ba_window(K=60, M=1500, degree=6, seed=91, mid_iteration=True)
    ref = _oracle_all(win)
    ba = bindings.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    S, rhs, _ = ba.build_reduced_system(ref["D"], True)
    assert rel_err(S, ref["S"]) < TOL and np.count_nonzero(ref["S"]) < 0.6 * S.size      # banded case
    ba.set_shard(1, 2)
    ba.build_reduced_system(ref["D"], True)               # fills a different envelope
    ba.set_shard(0, 1)
    S2, rhs2, _ = ba.build_reduced_system(ref["D"], True)
    assert np.array_equal(S2, S) and np.array_equal(rhs2, rhs)
    rng = np.random.default_rng(5)
    win2 = dict(win, pose_delta=win["pose_delta"] + 1e-3 * rng.standard_normal(win["pose_delta"].shape),
                points=win["points"] + 1e-3 * rng.standard_normal(win["points"].shape))
    ref2 = _oracle_all(win2)
    ba.set_parameters(win2["pose_delta"], win2["points"])
    S3, rhs3, _ = ba.build_reduced_system(ref2["D"], True)
    assert rel_err(S3, ref2["S"]) < TOL and rel_err(rhs3, ref2["rhs"]) < TOL
    assert not S3[ref2["S"] == 0].any()                   # nothing stale outside the envelope
    ba.close()


def test_band_exchange_buffer_sums_to_full_system(bindings, ctx, synthetic):
    """The compact block-band payload of sharded builds: summing the ranks' band buffers
    on the device and expanding gives the full reduced system, and the solver sees the sum."""
    import torch
    win = synthetic.make_ba_window(K=40, M=1500, degree=8, seed=78, mid_iteration=True)
    ref = _oracle_all(win)
    world = 3
    bas = []
    for rank in range(world):
        ba = bindings.BAWindow(ctx, win)
        ba.set_shard(rank, world)
        ba.set_parameters(win["pose_delta"], win["points"])
        ba.build_reduced_system(ref["D"], True, fetch=False)
        bas.append(ba)
    ctx.synchronize()
    views = []
    for ba in bas:
        ptr, nd = ba.reduced_band_dev()
        assert nd == ba.reduced_dim // 6 * min(ba.reduced_dim // 6, 24) * 36 + ba.reduced_dim + 1

        class _Buf:
            __cuda_array_interface__ = {"shape": (nd,), "typestr": "<f8", "data": (ptr, False), "version": 2}
        views.append(torch.as_tensor(_Buf(), device="cuda:0"))
    part0 = bas[0].fetch_reduced_system()                  # rank 0's partial through the lazy dense view
    views[0] += views[1] + views[2]                        # what the allreduce does
    torch.cuda.synchronize()
    bas[0].expand_band()
    S, rhs, cost = bas[0].fetch_reduced_system()
    assert rel_err(S, ref["S"]) < TOL and rel_err(rhs, ref["rhs"]) < TOL and abs(cost - ref["cost"]) < TOL * ref["cost"]
    assert np.array_equal(S, S.T) and not np.array_equal(part0[0], S)
    dp, _ = bas[0].solve_reduced(False)
    Wo = ba_oracle.Window(win)
    So, rhso, fm = Wo.schur(ref["r"], ref["Jp"], ref["Jl"], ref["D"])
    dp_ref, _ = Wo.solve_step(ref["r"], ref["Jp"], ref["Jl"], ref["D"], So, rhso, fm)
    assert rel_err(dp, dp_ref) < 1e-7
    # an unsharded build has no band buffer
    bas[1].set_shard(0, 1)
    bas[1].build_reduced_system(ref["D"], True, fetch=False)
    assert bas[1].reduced_band_dev() is None
    for ba in bas:
        ba.close()


def test_peer_exchange_same_process(bindings, synthetic):
    """vio360_ba_peer_exchange (one kernel: announce, wait, pull the covering bands, sum in rank order,
    LM diagonal, dense [S | rhs | cost]) with three ranks emulated by three contexts (= three streams) of one
    process: every rank must end with the full reduced system, bit-identical across ranks, over several
    epochs with changing parameters (the two band buffers alternate)."""
    win = synthetic.make_ba_window(K=40, M=1500, degree=8, seed=78, mid_iteration=True)
    ref = _oracle_all(win)
    world = 3
    ctxs = [bindings.Context(0) for _ in range(world)]
    bas = []
    for rank in range(world):
        ba = bindings.BAWindow(ctxs[rank], win)
        ba.set_shard(rank, world)
        ba.set_parameters(win["pose_delta"], win["points"])
        ba.set_lm_diagonal(ref["D"])
        bas.append(ba)
    blobs = [ba.peer_export() for ba in bas]
    assert all(len(b) == bindings.PEER_BLOB_BYTES for b in blobs)
    for ba in bas:
        ba.peer_attach(blobs)
    with pytest.raises(bindings.Vio360Error):
        bas[0].peer_exchange()                      # nothing built yet
    for epoch in range(3):
        if epoch == 2:                              # a new linearisation point: the other parity buffer again
            win2 = synthetic.make_ba_window(K=40, M=1500, degree=8, seed=78, mid_iteration=True)
            rng = np.random.default_rng(5)
            pd = win["pose_delta"] + 1e-3 * rng.standard_normal(win["pose_delta"].shape)
            pd[win["pose_fixed"] != 0] = 0.0
            win2["pose_delta"] = pd
            ref = _oracle_all(win2)
            for ba in bas:
                ba.set_parameters(pd, win["points"])
                ba.set_lm_diagonal(ref["D"])
        for ba in bas:
            ba.build_reduced_system(None, True, fetch=False)
        with pytest.raises(bindings.Vio360Error):
            bas[0].fetch_reduced_system()           # built but not exchanged: no silent partial result
        assert bas[0].reduced_band_dev() is None    # the band must not be all-reduced behind the exchange
        for c in ctxs:
            c.synchronize()                         # ranks sharing ONE device: the exchange kernels wait for each other and
                                                    # must not hold SM slots another rank's tile kernel still needs
        for ba in bas:
            ba.peer_exchange()                      # asynchronous: the three kernels meet on the device
        with pytest.raises(bindings.Vio360Error):
            bas[0].peer_exchange()                  # a second exchange without a new build
        outs = [ba.fetch_reduced_system() for ba in bas]
        S, rhs, cost = outs[0]
        assert rel_err(S, ref["S"]) < TOL and rel_err(rhs, ref["rhs"]) < TOL and abs(cost - ref["cost"]) < TOL * ref["cost"]
        assert np.array_equal(S, S.T)
        for S2, rhs2, cost2 in outs[1:]:
            assert np.array_equal(S2, S) and np.array_equal(rhs2, rhs) and cost2 == cost
    dp, _ = bas[1].solve_reduced(False)
    Wo = ba_oracle.Window(win2)
    So, rhso, fm = Wo.schur(ref["r"], ref["Jp"], ref["Jl"], ref["D"])
    dp_ref, _ = Wo.solve_step(ref["r"], ref["Jp"], ref["Jl"], ref["D"], So, rhso, fm)
    assert rel_err(dp, dp_ref) < 1e-7
    for ba in bas:
        ba.peer_detach()
    bas[0].build_reduced_system(None, True, fetch=False)      # back to the NCCL-style band
    assert bas[0].reduced_band_dev() is not None
    for ba in bas:
        ba.close()
    for c in ctxs:
        c.close()


def test_peer_exchange_two_devices_one_process(bindings, synthetic):
    """One process driving two GPUs (the reference is a single process): the ranks' buffers are mapped by plain peer
    access (cudaDeviceEnablePeerAccess) instead of cudaIpc handles; needs two GPUs."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    win = synthetic.make_ba_window(K=60, M=3000, degree=10, seed=91, mid_iteration=True)
    ref = _oracle_all(win)
    ctxs = [bindings.Context(0), bindings.Context(1)]
    bas = []
    for rank in range(2):
        ba = bindings.BAWindow(ctxs[rank], win)
        ba.set_shard(rank, 2)
        ba.set_parameters(win["pose_delta"], win["points"])
        ba.set_lm_diagonal(ref["D"])
        bas.append(ba)
    blobs = [ba.peer_export() for ba in bas]
    for ba in bas:
        ba.peer_attach(blobs)
    for _ in range(3):
        for ba in bas:
            ba.build_reduced_system(None, True, fetch=False)
        for ba in bas:
            ba.peer_exchange()
    outs = [ba.fetch_reduced_system() for ba in bas]
    S, rhs, cost = outs[0]
    assert rel_err(S, ref["S"]) < TOL and rel_err(rhs, ref["rhs"]) < TOL and abs(cost - ref["cost"]) < TOL * ref["cost"]
    assert np.array_equal(outs[1][0], S) and np.array_equal(outs[1][1], rhs) and outs[1][2] == cost
    for c in ctxs:
        c.synchronize()
    for ba in bas:
        ba.peer_detach()
    for ba in bas:
        ba.close()
    for c in ctxs:
        c.close()


def test_peer_exchange_two_processes():
    """The same exchange across PROCESSES (one per GPU, cudaIpc handles gathered through torch.distributed),
    checked against the NCCL allreduce of the band and the unsharded build; needs two GPUs."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29611", os.path.join(here, "dist_peer_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "PEER_OK" in out.stdout


def test_large_window_properties(bindings, ctx, synthetic):
    """BASELINE config 5 size (200 keyframes, 1M observations): size-independent checks --
    S symmetric PSD-ish, cost equals the evaluate() cost, linearity of the shard sum, and
    a sampled subset of residual rows against the oracle."""
    win = synthetic.make_ba_window(K=200, M=50000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
    assert win["O"] > 900000
    ba = bindings.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    r, Jp, Jl, cost = ba.evaluate(True, True)
    colsq = ba.column_sqnorm(True)
    D = np.sqrt(np.clip(colsq, 1e-6, 1e32) / 1e4)
    S, rhs, cost2 = ba.build_reduced_system(D, True)
    assert abs(cost - cost2) < 1e-12 * cost and np.array_equal(S, S.T)
    assert np.linalg.eigvalsh(S).min() > -1e-6 * np.abs(S).max()
    # sampled rows vs oracle
    rng = np.random.default_rng(0)
    idx = np.sort(rng.choice(win["O"], 3000, replace=False))
    sub = dict(win, O=len(idx), pose_idx=win["pose_idx"][idx], point_idx=win["point_idx"][idx],
               obs_u=np.ascontiguousarray(win["obs_u"][idx]), obs_v=np.ascontiguousarray(win["obs_v"][idx]),
               outlier=win["outlier"][idx])
    ro, Jpo, Jlo, _ = ba_oracle.Window(sub).evaluate(win["pose_delta"], win["points"])
    assert rel_err(r[idx], ro) < TOL and rel_err(Jp[idx], Jpo) < TOL and rel_err(Jl[idx], Jlo) < TOL
    # full Schur parity on this size too (the oracle needs a few seconds)
    Wo = ba_oracle.Window(win)
    So, rhso, fm = Wo.schur(r, Jp, Jl, D)
    assert rel_err(S, So) < TOL and rel_err(rhs, rhso) < TOL
    # and the solve at this size: 199 dependent block columns through the register-window band Cholesky
    dp_ref, dl_ref = Wo.solve_step(r, Jp, Jl, D, So, rhso, fm)
    dp, dl = ba.solve_reduced(True)
    assert rel_err(dp, dp_ref) < 1e-7 and rel_err(dl, dl_ref) < 1e-7
    ba.close()


def test_pnp(bindings, ctx, synthetic):
    win = synthetic.make_ba_window(K=3, M=400, seed=31, mid_iteration=True)
    k = 1
    sel = np.where(win["pose_idx"] == k)[0]
    obs = np.stack([win["obs_u"][sel], win["obs_v"][sel]], axis=1)
    pts = win["points"][win["point_idx"][sel]].copy()
    outl = win["outlier"][sel].copy(); outl[:3] = 1
    # a point at the camera centre -> Evaluate returns false
    import ba_ref_np
    Rcw, tcw, _, _ = ba_ref_np.pose_chain(win["Twb_init"][k], win["Tcb"], win["pose_delta"][k])
    pts[5] = -Rcw.T @ tcw
    out = ctx.pnp_evaluate(obs, pts, win["Twb_init"][k], win["Tcb"], win["cols"], win["rows"], win["info"], 1.0,
                           win["pose_delta"][k], outlier=outl)
    H = np.zeros((6, 6)); b = np.zeros(6); cost = 0.0
    for i in range(len(sel)):
        f = ba_oracle.make_factor(obs[i], win["cols"], win["rows"], win["Tcb"], win["Twb_init"][k], win["info"],
                                  is_outlier=bool(outl[i]), world_point=tuple(pts[i]))
        ok, r, Jp = ba_oracle.pnp_evaluate(f, win["pose_delta"][k])
        assert out["valid"][i] == ok
        c2 = ba_oracle.pnp_chi2(f, win["pose_delta"][k])
        assert out["chi2"][i] == c2 or abs(out["chi2"][i] - c2) < TOL * max(1.0, abs(c2))
        if not ok:
            continue
        import ctypes as C
        rr = r.copy(); JJ = Jp.copy()
        dp = C.POINTER(C.c_double)
        cost += ba_oracle.lib().vo_apply_loss(C.c_double(1.0), rr.ctypes.data_as(dp), 2, JJ.ctypes.data_as(dp), 6, None, 0)
        assert np.allclose(out["r"][i], rr, rtol=TOL, atol=TOL * 100) and rel_err(out["Jp"][i], JJ) < TOL or not JJ.any()
        H += JJ.T @ JJ; b += JJ.T @ rr
    assert rel_err(out["H"], H) < TOL and rel_err(out["b"], b) < TOL and abs(out["cost"] - cost) < TOL * cost


@pytest.mark.parametrize("K,M,degree,seed", [(10, 2000, None, 20000), (40, 4000, 12, 777)])
def test_lm_solve_vs_oracle(bindings, ctx, synthetic, K, M, degree, seed):
    """the device LM loop (vio360_ba_lm_solve) against the restated ceres::Solve of
    Optimizer::RunLocalBA (oracle Window.lm_solve: TrustRegionMinimizer + LevenbergMarquardtStrategy + Jacobi
    scaling + Schur + LAPACK) on config C3 and on a 40-keyframe co-visibility window: the same accept / reject
    sequence, the same termination, per-iteration cost and radius, final cost to 1e-9 and final parameters to
    1e-7 (relative to the largest parameter), then the chi-square outlier pass of Optimizer.cpp:898-915."""
    win = synthetic.make_ba_window(K=K, M=M, degree=degree, seed=seed, mid_iteration=True, loop_radius=1.0 if degree is None else 4.0)
    W = ba_oracle.Window(win)
    xp, xl, sm, tr = W.lm_solve(win["pose_delta"], win["points"])
    ba = bindings.BAWindow(ctx, win)
    gp, gl, gs, gt = ba.lm_solve(win["pose_delta"], win["points"])
    tr = np.array(tr)
    assert gs["termination"] == sm["termination"] and gs["iterations"] == sm["iterations"]
    assert gs["num_successful_steps"] == sm["num_successful_steps"] and gs["num_unsuccessful_steps"] == sm["num_unsuccessful_steps"]
    assert len(gt) == len(tr) and np.array_equal(gt[:, 6], tr[:, 6])                      # accept / reject / invalid sequence
    assert np.allclose(gt[:, 0], tr[:, 0], rtol=1e-9) and np.allclose(gt[:, 5], tr[:, 5], rtol=1e-6)   # cost, radius
    assert abs(gs["final_cost"] - sm["final_cost"]) < 1e-9 * sm["final_cost"]
    assert abs(gs["initial_cost"] - sm["initial_cost"]) < 1e-9 * sm["initial_cost"]
    assert sm["final_cost"] < 0.6 * sm["initial_cost"] and sm["iterations"] >= 3
    scale_p = max(np.abs(xp).max(), 1e-3); scale_l = np.abs(xl).max()
    assert np.abs(gp - xp).max() / scale_p < 1e-7 and np.abs(gl - xl).max() / scale_l < 1e-7
    assert not gp[0].any()                                                                # pose 0 is constant
    fixed = win["point_fixed"] != 0
    assert np.array_equal(gl[fixed], win["points"][fixed])
    # the window object now sits at the solution: the chi-square pass of Optimizer.cpp:898-915 on the device
    chi2, flag = ba.chi2(5.991)
    c2, f2 = W.chi2(xp, xl, 5.991)
    band = np.abs(c2 - 5.991) > 1e-5
    assert np.allclose(chi2, c2, rtol=1e-6, atol=1e-7) and np.array_equal(flag[band], f2[band])
    # options: no Jacobi scaling, a small radius and two iterations -> max-iterations termination, same trace
    xp2, xl2, sm2, tr2 = W.lm_solve(win["pose_delta"], win["points"], max_num_iterations=2, initial_radius=10.0, jacobi_scaling=False)
    gp2, gl2, gs2, gt2 = ba.lm_solve(win["pose_delta"], win["points"], max_num_iterations=2, initial_trust_region_radius=10.0, jacobi_scaling=0)
    assert gs2["termination"] == sm2["termination"] == 5 and np.array_equal(gt2[:, 6], np.array(tr2)[:, 6])
    assert np.abs(gp2 - xp2).max() / scale_p < 1e-8 and abs(gs2["final_cost"] - sm2["final_cost"]) < 1e-9 * sm2["final_cost"]
    ba.close()


def test_lm_solve_gradient_tolerance(bindings, ctx, synthetic):
    """GradientToleranceReached (trust_region_minimizer.cc:674-692).  The device loop makes this test in
    k_lm_linearized_finish and does not read the state back after a re-linearisation: the step that follows is a no-op
    and the read after k_lm_decide reports the termination.  Tolerances between the gradient norms of two consecutive
    accepted steps, and one above the initial gradient (no iteration at all)."""
    win = synthetic.make_ba_window(K=10, M=600, seed=11, mid_iteration=True)
    W = ba_oracle.Window(win)
    _, _, sm0, tr0 = W.lm_solve(win["pose_delta"], win["points"])
    g = np.array(tr0)[:, 2]
    assert sm0["iterations"] >= 3
    ba = bindings.BAWindow(ctx, win)
    picked = 0
    for i in range(1, len(g) - 1):
        if not (np.array(tr0)[i, 6] == 1 and g[i] < g[:i].min()):
            continue
        tol = 0.5 * (g[i] + g[:i].min())                       # first reached after iteration i
        xp, xl, sm, tr = W.lm_solve(win["pose_delta"], win["points"], gradient_tolerance=tol)
        gp, gl, gs, gt = ba.lm_solve(win["pose_delta"], win["points"], gradient_tolerance=tol)
        assert sm["termination"] == 1 and sm["iterations"] == i
        assert gs["termination"] == 1 and gs["iterations"] == i and len(gt) == len(tr)
        assert np.array_equal(gt[:, 6], np.array(tr)[:, 6]) and abs(gs["final_cost"] - sm["final_cost"]) < 1e-9 * sm["final_cost"]
        assert np.abs(gp - xp).max() < 1e-7 * max(np.abs(xp).max(), 1e-3) and np.abs(gl - xl).max() < 1e-7 * np.abs(xl).max()
        picked += 1
        if picked == 2:
            break
    assert picked >= 1
    xp, xl, sm, tr = W.lm_solve(win["pose_delta"], win["points"], gradient_tolerance=2.0 * g[0])
    gp, gl, gs, gt = ba.lm_solve(win["pose_delta"], win["points"], gradient_tolerance=2.0 * g[0])
    assert sm["termination"] == gs["termination"] == 1 and gs["iterations"] == sm["iterations"] == 0 and len(gt) == len(tr) == 1
    assert np.array_equal(gp, np.asarray(win["pose_delta"], np.float64).reshape(gp.shape))
    ba.close()


def test_lm_solve_rejected_steps(bindings, ctx, synthetic):
    """A start far from the optimum, no loss function and a huge initial radius: the 100-pixel clamp of
    BAFactor::Evaluate (Factors.cpp:424-441) makes the model over-predict, steps are REJECTED and the radius shrinks by
    2, 4, 8, ... (levenberg_marquardt_strategy.cc:156-160) until one is accepted again -- the same sequence on both sides."""
    win = synthetic.make_ba_window(K=10, M=600, seed=5, mid_iteration=True)
    rng = np.random.default_rng(1)
    pd = win["pose_delta"] + rng.normal(0, 0.05, win["pose_delta"].shape); pd[0] = 0
    W = ba_oracle.Window(win)
    xp, xl, sm, tr = W.lm_solve(pd, win["points"], initial_radius=1e10, apply_loss=False, max_num_iterations=15)
    tr = np.array(tr)
    assert (tr[:, 6] == 0).sum() >= 3 and (tr[1:, 6] == 1).sum() >= 3
    ba = bindings.BAWindow(ctx, win)
    gp, gl, gs, gt = ba.lm_solve(pd, win["points"], initial_trust_region_radius=1e10, apply_loss=0, max_num_iterations=15)
    assert np.array_equal(gt[:, 6], tr[:, 6]) and gs["termination"] == sm["termination"] and gs["iterations"] == sm["iterations"]
    assert np.allclose(gt[:, 5], tr[:, 5], rtol=1e-6) and np.allclose(gt[:, 0], tr[:, 0], rtol=1e-8)
    assert abs(gs["final_cost"] - sm["final_cost"]) < 1e-8 * sm["final_cost"]
    assert np.abs(gp - xp).max() < 1e-6 * max(np.abs(xp).max(), 1e-3) and np.abs(gl - xl).max() < 1e-6 * np.abs(xl).max()
    ba.close()


def test_fused_paths_fp32_mode(bindings, ctx, synthetic):
    """The north star's optional fp32 mode on the FUSED paths (vio360_ba_set_eval_precision): per-observation math in
    float, sums / Cholesky / Schur products in double.  Hessian blocks within 1e-4 (max-norm relative) of the fp64
    oracle; the reduced system, whose Schur complement amplifies the Jacobian rounding by the conditioning of E'E,
    within 1e-3; and it is really a different evaluation (not bit-equal to the fp64 build)."""
    win = synthetic.make_ba_window(K=10, M=2000, seed=20000, mid_iteration=True)
    ref = _oracle_all(win)
    ba = bindings.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    S64, rhs64, c64 = ba.build_reduced_system(ref["D"], True)
    ba.set_eval_precision(True)
    Hpp, bp, Hll, bl, Hpl = ba.normal_blocks(True)
    oHpp, obp, oHll, obl, oHpl = ref["blocks"]
    assert rel_err(Hpp, oHpp) < 1e-4 and rel_err(bp, obp) < 1e-4 and rel_err(Hll, oHll) < 1e-4 and rel_err(bl, obl) < 1e-4
    assert rel_err(Hpl, oHpl) < 1e-4
    S32, rhs32, c32 = ba.build_reduced_system(ref["D"], True)
    assert rel_err(S32, ref["S"]) < 1e-3 and rel_err(rhs32, ref["rhs"]) < 1e-3 and abs(c32 - ref["cost"]) < 1e-4 * ref["cost"]
    assert not np.array_equal(S32, S64)
    ba.set_eval_precision(False)
    S64b, _, _ = ba.build_reduced_system(ref["D"], True)
    assert np.array_equal(S64b, S64)
    # the LM loop in fp32 evaluation mode lands at the same minimum (cost within 1e-4)
    W = ba_oracle.Window(win)
    xp, xl, sm, tr = W.lm_solve(win["pose_delta"], win["points"])
    ba.set_eval_precision(True)
    gp, gl, gs, gt = ba.lm_solve(win["pose_delta"], win["points"])
    assert abs(gs["final_cost"] - sm["final_cost"]) < 1e-3 * sm["final_cost"]
    ba.close()
