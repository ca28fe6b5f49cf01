"""Oracle of the feature-list filters (oracle/fe_ref.c: vf_cluster_filter, vf_grid_cap) against an independent,
literal Python transcription of FeatureTracker::RemoveClusteredFeatures (FeatureTracker.cpp:404-497) and
Frame::AssignFeaturesToGrid / LimitFeaturesPerGrid (Frame.cpp:108-202) in numpy float32 scalars.  The reference
ships no test or golden vector for them (parity unpinned by the reference itself)."""
import numpy as np

import fe_oracle as fo

F = np.float32


def make_features(n, w, h, seed, clusters=3, outside=0.03):
    """uniform features + a few tight clusters (so that some cells are 'clustered') + points outside the image"""
    rng = np.random.default_rng(seed)
    pts = rng.uniform([0, 0], [w, h], (n, 2))
    per = max(4, n // 20)
    if clusters * per > n:
        clusters = 0
    for c in range(clusters):
        ctr = rng.uniform([0.1 * w, 0.1 * h], [0.9 * w, 0.9 * h])
        pts[c * per:(c + 1) * per] = ctr + rng.normal(0, 2.0 + 6.0 * c, (per, 2))
    k = int(outside * n)
    if k:
        pts[-k:] = rng.uniform([-40, -40], [w + 40, h + 40], (k, 2))
    pts = np.round(pts * 8) / 8
    track = rng.integers(1, 6, n).astype(np.int32)      # few distinct values: many ties
    return pts.astype(np.float32), track


def py_cluster_filter(xy, w, h, cols, rows, ratio):
    n = len(xy)
    keep = np.ones(n, np.uint8)
    if n < 4:
        return keep
    cw, ch = F(w) / F(cols), F(h) / F(rows)
    thr = np.sqrt(cw * cw + ch * ch, dtype=F) * F(ratio)
    cells = {}
    for i, (x, y) in enumerate(xy):
        col = max(min(int(F(x) / cw), cols - 1), 0); row = max(min(int(F(y) / ch), rows - 1), 0)
        cells.setdefault((row, col), []).append(i)
    for idx in cells.values():
        if len(idx) < 4:
            continue
        mx = F(0); my = F(0)
        for i in idx:
            mx = F(mx + xy[i, 0]); my = F(my + xy[i, 1])
        mx = F(mx / F(len(idx))); my = F(my / F(len(idx)))
        var = F(0)
        for i in idx:
            dx = F(xy[i, 0] - mx); dy = F(xy[i, 1] - my)
            var = F(var + F(F(dx * dx) + F(dy * dy)))
        var = F(var / F(len(idx)))
        if np.sqrt(var, dtype=F) < thr:
            keep[idx] = 0
    return keep


def py_grid_cap(xy, track, w, h, cols, rows, cap):
    n = len(xy)
    keep = np.zeros(n, np.uint8)
    cw, ch = F(w) / F(cols), F(h) / F(rows)
    cells = {}
    for i, (x, y) in enumerate(xy):
        if x < 0 or x >= w or y < 0 or y >= h:
            continue
        col = min(int(F(x) / cw), cols - 1); row = min(int(F(y) / ch), rows - 1)
        cells.setdefault((row, col), []).append(i)
    for idx in cells.values():
        if len(idx) > cap:
            idx = sorted(idx, key=lambda i: -int(track[i]))[:cap]     # Python's sort is stable
        keep[idx] = 1
    return keep


def test_cluster_filter_oracle_vs_transcription():
    removed = 0
    for seed, (w, h, n) in enumerate([(1920, 960, 600), (3840, 1920, 1000), (640, 320, 300), (963, 481, 257)]):
        xy, _ = make_features(n, w, h, seed)
        for ratio in (0.15, 0.25):
            got = fo.cluster_filter(xy, w, h, 20, 10, ratio)
            ref = py_cluster_filter(xy, w, h, 20, 10, ratio)
            assert np.array_equal(got, ref)
            removed += int((got == 0).sum())
    assert removed > 0
    # fewer than four features: untouched; empty list
    xy, _ = make_features(3, 640, 320, 9, clusters=0, outside=0)
    assert fo.cluster_filter(xy, 640, 320).all()
    assert len(fo.cluster_filter(np.zeros((0, 2), np.float32), 640, 320)) == 0
    # four coincident points in one cell: std 0 -> removed
    xy = np.array([[100, 100]] * 4 + [[500, 200]], np.float32)
    assert np.array_equal(fo.cluster_filter(xy, 640, 320), [0, 0, 0, 0, 1])


def test_grid_cap_oracle_vs_transcription():
    dropped = 0
    for seed, (w, h, n, cap) in enumerate([(1920, 960, 1500, 4), (3840, 1920, 3000, 10), (640, 320, 900, 4), (963, 481, 500, 1)]):
        xy, track = make_features(n, w, h, 20 + seed)
        got = fo.grid_cap(xy, track, w, h, 20, 10, cap)
        ref = py_grid_cap(xy, track, w, h, 20, 10, cap)
        assert np.array_equal(got, ref)
        dropped += int((got == 0).sum())
    assert dropped > 0
    # points on / outside the image border are dropped (Frame.cpp:125-128); x == w is outside, x == 0 is inside
    xy = np.array([[0, 0], [640, 10], [-0.125, 5], [639.875, 319.875], [10, 320]], np.float32)
    assert np.array_equal(fo.grid_cap(xy, np.ones(5, np.int32), 640, 320), [1, 0, 0, 1, 0])
    assert len(fo.grid_cap(np.zeros((0, 2), np.float32), np.zeros(0, np.int32), 640, 320)) == 0
