"""Prints parity numbers of every kernel against the oracle (no asserts): used to debug
on the GPU box with one call.  Not part of the product."""
import importlib
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
b = importlib.import_module("360_visual_inertial_odometry_b200.bindings")
syn = importlib.import_module("360_visual_inertial_odometry_b200.synthetic")
import ba_oracle  # noqa
import fe_oracle as fo  # noqa


def rel(a, c):
    return float(np.max(np.abs(a - c)) / max(np.max(np.abs(c)), 1e-300))


def section(name, fn):
    print("==== " + name, flush=True)
    t = time.time()
    try:
        fn()
    except Exception:
        traceback.print_exc()
    print("     (%.2fs)" % (time.time() - t), flush=True)


ctx = b.Context(0)
print("fma peak (fp64, fp32) TFLOP/s:", ctx.probe_fma_peak())


def ba_small():
    for (K, M, deg, seed) in ((6, 200, None, 1), (10, 2000, None, 20000), (48, 60, None, 9), (40, 1500, 8, 77)):
        win = syn.make_ba_window(K=K, M=M, degree=deg, seed=seed, mid_iteration=True, yaw_spread_deg=340.0 if K == 48 else 60.0)
        W = ba_oracle.Window(win)
        ro, Jpo, Jlo, co = W.evaluate(win["pose_delta"], win["points"])
        D = W.lm_diagonal(Jpo, Jlo, 1e4)
        So, rhso, _ = W.schur(ro, Jpo, Jlo, D)
        blocks = W.normal_blocks(ro, Jpo, Jlo)
        c2o, fo_ = W.chi2(win["pose_delta"], win["points"])
        ba = b.BAWindow(ctx, win)
        ba.set_parameters(win["pose_delta"], win["points"])
        r, Jp, Jl, c = ba.evaluate(True, True)
        S, rhs, c3 = ba.build_reduced_system(D, True)
        got = ba.normal_blocks(True)
        c2, f = ba.chi2()
        print("K=%d M=%d O=%d: r %.2e Jp %.2e Jl %.2e cost %.2e | S %.2e rhs %.2e cost %.2e | blocks %s | chi2 %.2e flags %d"
              % (K, M, win["O"], rel(r, ro), rel(Jp, Jpo), rel(Jl, Jlo), abs(c - co) / co, rel(S, So), rel(rhs, rhso),
                 abs(c3 - co) / co, ["%.1e" % rel(x, y) for x, y in zip(got, blocks)], rel(c2, c2o), int((f != fo_).sum())))
        ba.close()


def fe_small():
    g = np.load(os.path.join(ROOT, "tests", "golden", "fe_golden.npz"))
    f0, f1, pts = g["f0"], g["f1"], g["pts"]
    H, W = f0.shape
    ctx.fe_configure(W, H, max_points=len(pts), max_batch=2)
    pyr = ctx.fe_pyramid(f0, 4)
    for l, name in ((1, "pyr1"), (2, "pyr2"), (3, "pyr3")):
        gg = g[name]
        d = pyr[l][:gg.shape[0], :gg.shape[1]].astype(int) - gg.astype(int)
        print("pyr level", l, "mismatch px:", int((d != 0).sum()), "max", int(np.abs(d).max()))
    xy, st, err, nt = ctx.fe_track(f0, f1, pts)
    print("LK golden: status mismatches", int((st != g["cv_status"]).sum()), "of", len(st), "tracked", nt, int(g["cv_status"].sum()))
    ok = (st == 1) & (g["cv_status"] == 1)
    dd = np.abs(xy - g["cv_next"])[ok].max(axis=1)
    print("   pos diff max %.5f  p99 %.5f  err diff %.5f" % (dd.max(), np.percentile(dd, 99), np.abs(err - g["cv_err"])[ok].max()))
    bad = np.where(st != g["cv_status"])[0]
    for i in bad[:10]:
        print("   mismatch idx", i, pts[i], "gpu", xy[i], st[i], "cv", g["cv_next"][i], g["cv_status"][i])
    worst = np.argsort(-np.where(ok, np.abs(xy - g["cv_next"]).max(axis=1), 0))[:5]
    for i in worst:
        print("   worst idx", i, pts[i], "gpu", xy[i], "cv", g["cv_next"][i])
    eig = ctx.fe_min_eigen_val(f0)
    eo = fo.min_eigen_val(f0)
    print("eig mismatches", int((eig != eo).sum()), "max abs", float(np.abs(eig - eo).max()))
    if (eig != eo).any():
        yy, xx = np.where(eig != eo)
        print("   first mismatches (y,x):", list(zip(yy[:8].tolist(), xx[:8].tolist())), "rows uniq", np.unique(yy)[:8], "cols uniq", np.unique(xx)[:8])
    for name, mask, mc in (("base", None, 1000), ("occ", g["occ_mask"], 1000), ("top25", None, 25)):
        got = ctx.fe_detect(f0, mask, mc)
        ref = g["gftt_" + name]
        same = got.shape == ref.shape and np.array_equal(got, ref)
        print("gftt", name, "n", len(got), len(ref), "equal", same)
        if not same:
            m = min(len(got), len(ref))
            first = next((i for i in range(m) if not np.array_equal(got[i], ref[i])), m)
            print("   first diff at", first, got[first:first + 3].tolist(), ref[first:first + 3].tolist())


def fe_big():
    W, H = 3840, 1920
    import torch
    f0 = syn.render_erp_frame(W, H, 0, seed=360, n_waves=64, sigma_omega=120.0, xp=torch, device="cuda").cpu().numpy()
    f1 = syn.render_erp_frame(W, H, 1, seed=360, n_waves=64, sigma_omega=120.0, xp=torch, device="cuda").cpu().numpy()
    pts = syn.make_track_points(W, H, 1000, seed=1000)
    ctx.fe_configure(W, H, max_points=1000, max_batch=1)
    t = time.time(); xy, st, err, nt = ctx.fe_track(f0, f1, pts); t1 = time.time() - t
    t = time.time(); xy, st, err, nt = ctx.fe_track(f0, f1, pts); t2 = time.time() - t
    t = time.time(); oxy, ost, oerr, _ = fo.lk_track(f0, f1, pts); t3 = time.time() - t
    t = time.time(); cxy, cst, cerr = fo.cv2_lk_track(f0, f1, pts); t4 = time.time() - t
    ok = (st == 1) & (ost == 1)
    print("C2 LK: status mism vs oracle %d vs cv2 %d ; pos max vs oracle %.5f vs cv2 %.5f ; tracked %d ; times gpu %.4f/%.4f oracle %.3f cv2 %.3f"
          % ((st != ost).sum(), (st != cst).sum(), np.abs(xy - oxy)[ok].max(), np.abs(xy - cxy)[(st == 1) & (cst == 1)].max(), nt, t1, t2, t3, t4))
    t = time.time(); got = ctx.fe_detect(f0, None, 1000); t1 = time.time() - t
    t = time.time(); got = ctx.fe_detect(f0, None, 1000); t2 = time.time() - t
    t = time.time(); ref = fo.good_features(f0, fo.base_mask(W, H), 1000); t3 = time.time() - t
    t = time.time(); cref = fo.cv2_good_features(f0, fo.base_mask(W, H), 1000); t4 = time.time() - t
    print("C2 GFTT: n %d/%d/%d equal oracle %s equal cv2 %s ; times gpu %.4f/%.4f oracle %.3f cv2 %.3f"
          % (len(got), len(ref), len(cref), np.array_equal(got, ref), np.array_equal(got, cref), t1, t2, t3, t4))


section("BA", ba_small)
section("FE golden", fe_small)
section("FE C2 size", fe_big)
print("launches", ctx.launch_count)
