"""BASELINE configs[0] (headless): FeatureTracker::TrackFeatures (src/processing/FeatureTracker.cpp:61-206) driven frame
after frame on the synthetic 1920x960 textured-sphere sequence, as ONE device-resident call per frame
(vio360_fe_track_frame_batch), against the same chain on the CPU: cv2 for the two OpenCV calls the reference makes,
oracle/fe_ref.c for the reference's own list logic (post filter, rotation RANSAC with the injected sample list,
RemoveClusteredFeatures, LimitFeaturesPerGrid, CreateFeatureMask).  Per frame: identical feature ids and track
counts in the reference's order, pixel coordinates within 0.05 px (new corners exact), identical
m_num_tracked / m_num_detected."""
import numpy as np
import pytest

import fe_oracle as fo

pytestmark = pytest.mark.gpu
POS_TOL = 0.05


def _frames(synthetic, W, H, n, seed, sigma):
    import torch
    return [synthetic.render_erp_frame(W, H, t, seed=seed, n_waves=24, sigma_omega=sigma, xp=torch, device="cuda").cpu().numpy()
            for t in range(n)]


def _run_sequence(ctx, frames, W, H, cfg, use_cv2, seed, n_seq=1):
    rng = np.random.default_rng(seed)
    seqs = [ctx.seq_create(**cfg) for _ in range(n_seq)]
    refs = [fo.SequenceTracker(W, H, max_features=cfg["max_features"], min_distance=cfg["min_distance"], quality=cfg["quality"],
                               grid_cols=cfg["grid_cols"], grid_rows=cfg["grid_rows"], max_per_cell=cfg["max_per_cell"],
                               remove_clustered=cfg["remove_clustered"], std_ratio=cfg["std_ratio"],
                               ransac_thr_deg=cfg["ransac_thr_deg"], use_cv2=use_cv2) for _ in range(n_seq)]
    stats = []
    try:
        for t in range(len(frames[0])):
            words = [rng.integers(0, 2 ** 32, (cfg["ransac_iters"], 3), dtype=np.uint64).astype(np.uint32) for _ in range(n_seq)]
            outs = ctx.fe_track_frame_batch(seqs, [frames[q][t] for q in range(n_seq)], words)
            for q in range(n_seq):
                nt, nd = refs[q].step(frames[q][t], words[q])
                ids, xy, tc, gnt, gnd = outs[q]
                assert (gnt, gnd) == (nt, nd), (t, q, gnt, gnd, nt, nd)
                assert np.array_equal(ids, refs[q].ids), (t, q)
                assert np.array_equal(tc, refs[q].tc), (t, q)
                d = np.abs(xy - refs[q].xy)
                assert d.max() < POS_TOL, (t, q, float(d.max()))
                assert np.array_equal(xy[tc == 0], refs[q].xy[tc == 0])          # this frame's corners: integer, exact
                # the CPU chain continues from ITS positions: a 1e-4 px difference must not be re-injected every frame
                stats.append((t, len(ids), nt, nd, float(d.max())))
    finally:
        for h in seqs:
            ctx.seq_destroy(h)
    return stats


def test_c1_sequence_chain_vs_cv2(ctx, synthetic):
    """configs[0]: 1920x960, default_config.yaml feature_detection (1000 / 0.01 / 30 / 20x10x10), clustered-cell removal
    on (ratio 0.25), RANSAC 1000 hypotheses at 2 degrees (FeatureTracker.cpp:36-37), 30 frames."""
    W, H, T = 1920, 960, 30
    frames = _frames(synthetic, W, H, T, 360, 60.0)
    ctx.fe_configure(W, H, max_points=2000, max_batch=2)
    cfg = dict(max_features=1000, min_distance=30.0, quality=0.01, grid_cols=20, grid_rows=10, max_per_cell=10,
               remove_clustered=True, std_ratio=0.25, ransac_iters=1000, ransac_thr_deg=2.0)
    stats = _run_sequence(ctx, [frames], W, H, cfg, True, 1)
    n = np.array([s[1] for s in stats]); nt = np.array([s[2] for s in stats]); nd = np.array([s[3] for s in stats])
    assert n.min() > 300 and nt[1:].min() > 200 and nd[0] > 300          # a real sequence: features survive and are replenished
    assert max(s[4] for s in stats) < POS_TOL


def test_sequence_chain_batch_and_variants(ctx, synthetic):
    """two independent sequences per call (configs[3] shape at a small size), tight caps so that every branch fires:
    grid cap 4 (FeatureTracker.cpp:41), few features (detection skipped once the list is full), no cluster removal,
    a tiny RANSAC; the CPU chain runs on oracle/fe_ref.c alone here."""
    W, H, T = 640, 320, 8
    fa = _frames(synthetic, W, H, T, 11, 40.0)
    fb = _frames(synthetic, W, H, T, 12, 40.0)
    ctx.fe_configure(W, H, max_points=400, max_batch=2)
    for cfg in (dict(max_features=200, min_distance=12.0, quality=0.01, grid_cols=20, grid_rows=10, max_per_cell=4,
                     remove_clustered=True, std_ratio=0.15, ransac_iters=64, ransac_thr_deg=1.0),
                dict(max_features=60, min_distance=20.0, quality=0.05, grid_cols=8, grid_rows=4, max_per_cell=50,
                     remove_clustered=False, std_ratio=0.25, ransac_iters=16, ransac_thr_deg=5.0)):
        stats = _run_sequence(ctx, [fa, fb], W, H, cfg, False, 7, n_seq=2)
        assert min(s[1] for s in stats) > 10
