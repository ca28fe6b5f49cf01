"""Batched DLT triangulation: the fp32 restatement of
Estimator::TriangulateSinglePoint against a float64 numpy SVD and the known points."""
import importlib

import numpy as np

import ba_oracle

syn = importlib.import_module("360_visual_inertial_odometry_b200.synthetic")


def test_oracle_triangulation_matches_svd_and_truth():
    b1, b2, T1, T2, P = syn.make_triangulation_pairs(400, seed=5)
    pts, valid = ba_oracle.triangulate(b1, b2, T1, T2)
    assert valid.all()
    ref = np.zeros_like(P)
    for i in range(len(P)):
        A = np.stack([b1[i, 0] * T1[2] - b1[i, 2] * T1[0], b1[i, 1] * T1[2] - b1[i, 2] * T1[1],
                      b2[i, 0] * T2[2] - b2[i, 2] * T2[0], b2[i, 1] * T2[2] - b2[i, 2] * T2[1]]).astype(np.float64)
        h = np.linalg.svd(A)[2][3]
        ref[i] = h[:3] / h[3]
    scale = np.linalg.norm(ref, axis=1)
    # fp32 DLT: error grows with depth / baseline; 0.5 m baseline, <= 10 m depth -> well below 1 %
    assert (np.linalg.norm(pts - ref, axis=1) / scale).max() < 5e-3
    assert np.median(np.linalg.norm(pts - P, axis=1) / scale) < 1e-3


def test_degenerate_pair_is_rejected():
    # identical bearings and poses: the null space is a line through the camera, w can vanish or the point is non-finite
    b = np.array([[0.0, 0.0, 1.0]], np.float32)
    T = np.eye(4, dtype=np.float32)
    pts, valid = ba_oracle.triangulate(b, b, T, T)
    assert valid[0] in (0, 1) and np.isfinite(pts).all()
