"""Rotation-only RANSAC: the fp32 restatement in oracle/fe_ref.c against an
independent float64 numpy evaluation, plus the reference's degenerate-input behaviour
(FeatureTracker.cpp:258-261, 321-325)."""
import importlib

import numpy as np

from oracle import fe_oracle as fo

syn = importlib.import_module("360_visual_inertial_odometry_b200.synthetic")
W, H = 3840, 1920


def _np_counts(prev, cur, samples, thr_deg):
    def px2b(xy):
        xy = xy.astype(np.float64)
        lon = (xy[:, 0] / W - 0.5) * 2 * np.pi
        lat = -(xy[:, 1] / H - 0.5) * np.pi
        return np.stack([np.cos(lat) * np.sin(lon), -np.sin(lat), np.cos(lat) * np.cos(lon)], 1)
    b1, b2 = px2b(prev), px2b(cur)
    thr = np.deg2rad(thr_deg)
    out = []
    for s in samples:
        Hm = sum(np.outer(b2[i], b1[i]) for i in s)
        U, _, Vt = np.linalg.svd(Hm)
        R = U @ Vt
        if np.linalg.det(R) < 0:
            U[:, 2] *= -1
            R = U @ Vt
        c = np.clip(np.sum((b1 @ R.T) * b2, 1), -1, 1)
        out.append(int((np.arccos(c) < thr).sum()))
    return np.array(out)


def test_oracle_matches_float64_model():
    prev, cur, is_out = syn.make_rotation_tracks(W, H, 600, seed=7)
    samples = syn.draw_ransac_samples(600, 120, seed=8)
    best, mask, counts, bi, errs = fo.ransac_rotation(prev, cur, W, H, samples, 1.0)
    ref = _np_counts(prev, cur, samples, 1.0)
    assert (counts >= 0).all()
    # fp32 acos near 0 is coarse (step ~3e-4 rad), so allow a few borderline points per hypothesis
    assert np.abs(counts - ref).max() <= 0.02 * 600
    assert best == counts.max() and bi == int(np.argmax(counts))      # first maximum wins (strict >)
    assert mask.sum() == best
    assert mask[~is_out].mean() > 0.99 and mask[is_out].mean() < 0.15
    assert np.array_equal(mask, (errs < np.float32(np.deg2rad(1.0))).astype(np.uint8))


def test_degenerate_inputs():
    prev, cur, _ = syn.make_rotation_tracks(W, H, 2, seed=1)
    best, mask, _, bi, _ = fo.ransac_rotation(prev, cur, W, H, np.zeros((5, 3), np.int32), 1.0)
    assert best == 2 and mask.tolist() == [1, 1] and bi == -1
    # no hypothesis gathers an inlier -> everything kept, 0 returned
    rng = np.random.default_rng(3)
    prev = np.stack([rng.uniform(20, W - 20, 50), rng.uniform(300, 1600, 50)], 1).astype(np.float32)
    cur = np.stack([rng.uniform(20, W - 20, 50), rng.uniform(300, 1600, 50)], 1).astype(np.float32)
    samples = syn.draw_ransac_samples(50, 20, seed=4)
    best, mask, counts, bi, _ = fo.ransac_rotation(prev, cur, W, H, samples, 1e-4)
    assert best == 0 and bi == -1 and mask.all() and (counts <= 0).all()
