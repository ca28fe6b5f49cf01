"""Regenerates the committed golden fixtures.  Run in the build container:

    python tests/golden/make_golden.py

fe_golden.npz   inputs + outputs of cv2 (the OpenCV the reference's front-end delegates
                to) called with the reference's exact arguments
                (src/processing/FeatureTracker.cpp:222-223, 238-240).
ba_golden.npz   inputs + outputs of oracle/ba_ref.c for a small window (the reference
                itself cannot be built here: no Eigen/glog/OpenCV C++; parity unpinned).
The cv2 version used is stored inside the file."""
import importlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import cv2  # noqa: E402

import ba_oracle  # noqa: E402
import fe_oracle  # noqa: E402

syn = importlib.import_module("360_visual_inertial_odometry_b200.synthetic")


def make_fe():
    W, H = 640, 320
    f0 = syn.render_erp_frame(W, H, 0, seed=11, n_waves=96, sigma_omega=40.0)
    f1 = syn.render_erp_frame(W, H, 3, seed=11, n_waves=96, sigma_omega=40.0)
    # a flat patch (min-eigenvalue rejects) present in both frames
    f0[150:200, 300:360] = 90; f1[150:200, 300:360] = 90
    rng = np.random.default_rng(5)
    pts = syn.make_track_points(W, H, 260, seed=1001, seam_frac=0.1)
    extra = np.array([[2.0, 100.0], [W - 2.5, 160.0], [320.0, 1.5], [330.0, H - 1.25], [0.0, 0.0],
                      [W - 1.0, H - 1.0], [325.5, 170.25], [310.0, 160.0], [-30.0, 50.0], [W + 25.0, 60.0],
                      [100.0, -40.0], [5.125, 5.875]], np.float32)
    pts = np.concatenate([pts, extra, rng.uniform([0, 0], [W, H], (28, 2)).astype(np.float32)]).astype(np.float32)
    nxt, st, err = fe_oracle.cv2_lk_track(f0, f1, pts)
    mask = fe_oracle.base_mask(W, H, 0.15, 20)
    c_base = fe_oracle.cv2_good_features(f0, mask, 1000)
    occ = mask.copy()
    for p in c_base[::3]:
        cv2.circle(occ, (int(p[0]), int(p[1])), 30, 0, -1)
    c_occ = fe_oracle.cv2_good_features(f0, occ, 1000)
    c_top = fe_oracle.cv2_good_features(f0, mask, 25)
    eig = cv2.cornerMinEigenVal(f0, 3, ksize=3)
    pyr = [f0]
    for _ in range(3):
        pyr.append(cv2.pyrDown(pyr[-1]))
    np.savez_compressed(os.path.join(HERE, "fe_golden.npz"), f0=f0, f1=f1, pts=pts, cv_next=nxt, cv_status=st,
                        cv_err=err, occ_mask=occ, gftt_base=c_base, gftt_occ=c_occ, gftt_top25=c_top,
                        eig_rows=eig[::8].copy(), eig_bits_sum=np.array([eig.view(np.uint32).astype(np.uint64).sum()], np.uint64), pyr1=pyr[1], pyr2=pyr[2], pyr3=pyr[3], cv2_version=cv2.__version__)
    print("fe_golden: tracked %d/%d, corners %d/%d/%d" % (st.sum(), len(st), len(c_base), len(c_occ), len(c_top)))


def make_ba():
    win = syn.make_ba_window(K=5, M=40, seed=123, mid_iteration=True, flagged_outlier_frac=0.03)
    W = ba_oracle.Window(win)
    r, Jp, Jl, cost = W.evaluate(win["pose_delta"], win["points"])
    D = W.lm_diagonal(Jp, Jl, 1e4)
    S, rhs, fm = W.schur(r, Jp, Jl, D)
    Hpp, bp, Hll, bl, Hpl = W.normal_blocks(r, Jp, Jl)
    chi2, flag = W.chi2(win["pose_delta"], win["points"], 5.991)
    arrays = {k: v for k, v in win.items() if isinstance(v, np.ndarray)}
    scal = {k: v for k, v in win.items() if not isinstance(v, np.ndarray)}
    np.savez_compressed(os.path.join(HERE, "ba_golden.npz"), r=r, Jp=Jp, Jl=Jl, cost=cost, D=D, S=S, rhs=rhs,
                        Hpp=Hpp, bp=bp, Hll=Hll, bl=bl, Hpl=Hpl, chi2=chi2, flag=flag,
                        scalars=np.array([scal["K"], scal["M"], scal["O"], scal["cols"], scal["rows"], scal["huber_a"]]),
                        **{"in_" + k: v for k, v in arrays.items()})
    print("ba_golden: O=%d cost=%.6f" % (win["O"], cost))


if __name__ == "__main__":
    make_fe()
    make_ba()
