"""GPU parity tests for path A through the C ABI against the oracle (oracle/fe_ref.c,
itself pinned to cv2) and the committed cv2 golden vectors.
Bars (BASELINE.json north_star): pyramid / status / keypoint indices bit exact,
tracked positions within 0.05 px."""
import os

import numpy as np
import pytest

import fe_oracle as fo

pytestmark = pytest.mark.gpu
POS_TOL = 0.05
GOLD = os.path.join(os.path.dirname(__file__), "golden", "fe_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _bits_sum(a):
    return np.array([a.view(np.uint32).astype(np.uint64).sum()], np.uint64)


def test_golden_pyramid_lk_gftt(ctx, gold):
    f0, f1, pts = gold["f0"], gold["f1"], gold["pts"]
    H, W = f0.shape
    ctx.fe_configure(W, H, max_points=len(pts), max_batch=2)
    pyr = ctx.fe_pyramid(f0, 4)
    for l, name in ((1, "pyr1"), (2, "pyr2"), (3, "pyr3")):
        g = gold[name]
        assert np.array_equal(pyr[l][:g.shape[0], :g.shape[1]], g)
    xy, st, err, nt = ctx.fe_track(f0, f1, pts)
    assert np.array_equal(st, gold["cv_status"]) and nt == int(gold["cv_status"].sum())
    ok = st == 1
    assert np.abs(xy - gold["cv_next"])[ok].max() < POS_TOL
    assert np.abs(err - gold["cv_err"])[ok].max() < 0.05
    eig = ctx.fe_min_eigen_val(f0)
    assert np.array_equal(eig[::8], gold["eig_rows"]) and np.array_equal(_bits_sum(eig), gold["eig_bits_sum"])
    assert np.array_equal(ctx.fe_detect(f0, None, 1000), gold["gftt_base"])
    assert np.array_equal(ctx.fe_detect(f0, gold["occ_mask"], 1000), gold["gftt_occ"])
    assert np.array_equal(ctx.fe_detect(f0, None, 25), gold["gftt_top25"])


@pytest.mark.parametrize("shape", [(481, 963), (240, 320), (67, 91), (1920, 3840)])
def test_pyramid_bit_exact(ctx, shape):
    H, W = shape
    rng = np.random.default_rng(H)
    img = rng.integers(0, 256, (H, W), dtype=np.uint8)
    ctx.fe_configure(W, H, max_points=8, max_batch=1)
    lv = img
    got = ctx.fe_pyramid(img, 4 if min(H, W) > 200 else 2)
    for l in range(1, len(got)):
        lv = fo.pyr_down(lv)
        assert np.array_equal(got[l][:lv.shape[0], :lv.shape[1]], lv), "level %d" % l


@pytest.mark.parametrize("size,seed,n", [((640, 320), 3, 500), ((1280, 640), 4, 1000), ((963, 481), 6, 300)])
def test_lk_vs_oracle(ctx, synthetic, size, seed, n):
    W, H = size
    f0 = synthetic.render_erp_frame(W, H, 0, seed=seed, n_waves=48, sigma_omega=40.0)
    f1 = synthetic.render_erp_frame(W, H, 2, seed=seed, n_waves=48, sigma_omega=40.0)
    f0[H // 2:H // 2 + 40, W // 2:W // 2 + 60] = 77; f1[H // 2:H // 2 + 40, W // 2:W // 2 + 60] = 77
    pts = synthetic.make_track_points(W, H, n, seed=seed, seam_frac=0.1)
    rng = np.random.default_rng(seed)
    pts[:20] = rng.uniform([-30, -30], [W + 30, H + 30], (20, 2)).astype(np.float32)   # border / out-of-image
    ctx.fe_configure(W, H, max_points=n, max_batch=2)
    xy, st, err, nt = ctx.fe_track(f0, f1, pts)
    oxy, ost, oerr, ocnt = fo.lk_track(f0, f1, pts)
    assert np.array_equal(st, ost) and nt == ocnt
    ok = ost == 1
    assert np.abs(xy - oxy)[ok].max() < POS_TOL
    assert np.abs(err - oerr)[ok].max() < 0.05
    assert (ost == 0).sum() > 0
    # untracked points still report OpenCV's last estimate
    assert np.abs(xy - oxy)[~ok].max() < POS_TOL if (~ok).any() else True


def test_track_batch_cache_and_keep(ctx, synthetic):
    W, H = 640, 320
    frames = [synthetic.render_erp_frame(W, H, t, seed=21, n_waves=32, sigma_omega=40.0) for t in range(3)]
    pts = synthetic.make_track_points(W, H, 200, seed=2)
    ctx.fe_configure(W, H, max_points=256, max_batch=4)
    # two "sequences" in one batch; keys let the second call re-use frame 1's pyramid
    out1 = ctx.fe_track_batch_host([(frames[0], frames[1], pts, 100, 101), (frames[1], frames[2], pts[:150], 101, 102)])
    out2 = ctx.fe_track_batch_host([(frames[1], frames[2], pts[:150], 101, 102)])
    assert np.array_equal(out1[1][0], out2[0][0]) and np.array_equal(out1[1][1], out2[0][1])
    for (prev, cur, p), (xy, st, err, keep) in zip(((frames[0], frames[1], pts), (frames[1], frames[2], pts[:150])), out1):
        oxy, ost, _, _ = fo.lk_track(prev, cur, p)
        assert np.array_equal(st, ost) and np.abs(xy - oxy)[ost == 1].max() < POS_TOL
        assert np.array_equal(keep, fo.post_filter(xy, st, W, H, 0.15, 20.0))
    # empty input: returns 0 tracked and touches nothing (FeatureTracker.cpp:233-235)
    xy, st, err, nt = ctx.fe_track(frames[0], frames[1], np.zeros((0, 2), np.float32))
    assert nt == 0 and len(st) == 0


def test_cluster_filter_and_grid_cap_vs_oracle(ctx):
    """RemoveClusteredFeatures and LimitFeaturesPerGrid on the device, bit exact against the
    oracle over ragged multi-frame batches (an empty frame, a frame with fewer than four features, ties, points
    outside the image, the reference's grid 20 x 10 with caps 4 (FeatureTracker.cpp:41) and 10 (default_config.yaml:16))."""
    from test_grid_cpu import make_features
    W, H = 3840, 1920
    ctx.fe_configure(W, H, max_points=3000, max_batch=2)
    sizes = [1000, 0, 3, 3000, 257, 4]
    frames, tracks = [], []
    for f, n in enumerate(sizes):
        xy, tr = make_features(n, W, H, 40 + f) if n else (np.zeros((0, 2), np.float32), np.zeros(0, np.int32))
        frames.append(xy); tracks.append(tr)
    for ratio in (0.15, 0.25):
        got = ctx.fe_cluster_filter(frames, W, H, 20, 10, ratio)
        for f in range(len(sizes)):
            assert np.array_equal(got[f], fo.cluster_filter(frames[f], W, H, 20, 10, ratio)), (f, ratio)
    assert any((g == 0).any() for g in got)
    for cap in (4, 10, 1):
        got = ctx.fe_grid_cap(frames, tracks, W, H, 20, 10, cap)
        for f in range(len(sizes)):
            assert np.array_equal(got[f], fo.grid_cap(frames[f], tracks[f], W, H, 20, 10, cap)), (f, cap)
    assert any((g == 0).any() for g in got)


def test_track_batch_host_upload_groups(ctx, synthetic, monkeypatch):
    """Host-space batches larger than the upload group are cut into groups whose frames travel on a copy stream
    while the previous group's kernels run: the results must be the ones of the same jobs issued one by one, also
    when a frame uploaded for one group is the cached previous frame of a job in a later group, and for ragged
    and empty jobs."""
    W, H = 640, 320
    frames = [synthetic.render_erp_frame(W, H, t, seed=33, n_waves=32, sigma_omega=40.0) for t in range(12)]
    pts = synthetic.make_track_points(W, H, 200, seed=7)
    monkeypatch.setenv("VIO360_FE_H2D_GROUP", "3")
    ctx.fe_configure(W, H, max_points=256, max_batch=16)
    monkeypatch.delenv("VIO360_FE_H2D_GROUP")
    ns = [200, 150, 0, 77, 200, 1, 200, 33, 200, 120, 64]
    # a chain: job i tracks frame i -> i + 1; keys make frame i + 1 of job i the cached prev of job i + 1 (next group or same)
    pairs = [(frames[i], frames[i + 1], pts[:ns[i]], 500 + i, 501 + i) for i in range(11)]
    l0 = ctx.launch_count
    outs = ctx.fe_track_batch_host(pairs)
    assert ctx.launch_count - l0 == 4 * (3 + 1)          # four groups: three pyramid levels + one LK launch each
    ctx.fe_configure(W, H, max_points=256, max_batch=16)
    for i, (xy, st, err, keep) in enumerate(outs):
        if ns[i] == 0:
            continue
        (xy1, st1, err1, keep1), = ctx.fe_track_batch_host([(frames[i], frames[i + 1], pts[:ns[i]], 0, 0)])
        assert np.array_equal(xy, xy1) and np.array_equal(st, st1) and np.array_equal(err, err1) and np.array_equal(keep, keep1)
    oxy, ost, _, _ = fo.lk_track(frames[4], frames[5], pts[:ns[4]])
    assert np.array_equal(outs[4][1], ost) and np.abs(outs[4][0] - oxy)[ost == 1].max() < POS_TOL
    # unkeyed host frames in a grouped batch (every frame uploaded, nothing cached)
    ctx.fe_configure(W, H, max_points=256, max_batch=16)
    outs2 = ctx.fe_track_batch_host([(frames[i], frames[i + 1], pts[:ns[i]], 0, 0) for i in range(11)])
    for a, b2 in zip(outs, outs2):
        assert np.array_equal(a[0], b2[0]) and np.array_equal(a[1], b2[1])


@pytest.mark.parametrize("size,seed", [((960, 480), 3), ((1920, 960), 4), ((3840, 1920), 5)])
def test_detect_vs_oracle(ctx, synthetic, size, seed):
    W, H = size
    f0 = synthetic.render_erp_frame(W, H, 0, seed=seed, n_waves=48, sigma_omega=40.0 * W / 960)
    ctx.fe_configure(W, H, max_points=8, max_batch=1)
    eig = ctx.fe_min_eigen_val(f0)
    assert np.array_equal(eig, fo.min_eigen_val(f0))
    mask = fo.base_mask(W, H)
    ref = fo.good_features(f0, mask, 1000)
    got = ctx.fe_detect(f0, None, 1000)
    assert np.array_equal(got, ref) and len(ref) > 50
    occ = mask.copy()
    yy, xx = np.mgrid[0:H, 0:W]
    for p in ref[::2]:
        occ[(xx - p[0]) ** 2 + (yy - p[1]) ** 2 <= 900] = 0
    assert np.array_equal(ctx.fe_detect(f0, occ, 1000), fo.good_features(f0, occ, 1000))
    assert np.array_equal(ctx.fe_detect(f0, None, 10), ref[:10])
    # empty mask -> no corners
    assert len(ctx.fe_detect(f0, np.zeros_like(mask), 1000)) == 0


@pytest.mark.parametrize("narrow", ["0", "1"])
def test_min_eigen_strip_seams(ctx, monkeypatch, narrow):
    """Every position of the right image border inside a strip of k_min_eig (60 columns per warp) and of
    k_min_eig_narrow (28): a last strip of one column puts the reflected column into the halo lane of the strip before."""
    monkeypatch.setenv("VIO360_EIG_NARROW", narrow)
    rng = np.random.default_rng(17)
    H = 37
    for W in list(range(56, 64)) + list(range(84, 92)) + list(range(118, 124)) + [180, 181, 182, 481]:
        img = rng.integers(0, 256, (H, W), dtype=np.uint8)
        ctx.fe_configure(W, H, max_points=8, max_batch=1)
        assert np.array_equal(ctx.fe_min_eigen_val(img), fo.min_eigen_val(img)), W


def test_detect_batch_and_grid_variants(ctx, synthetic, monkeypatch):
    """One batched call over different images (single merged sort), odd image size (byte loads /
    scalar stores), a small minDistance (cell grid in global memory), minDistance < 1 (no grid),
    no corner limit, and a candidate buffer that starts too small and is regrown."""
    W, H = 963, 481
    frames = [synthetic.render_erp_frame(W, H, t, seed=21 + t, n_waves=40, sigma_omega=45.0) for t in range(3)]
    mask = fo.base_mask(W, H)
    occ = mask.copy()
    occ[:, 300:500] = 0
    monkeypatch.setenv("VIO360_DET_CAND", "64")
    ctx.fe_configure(W, H, max_points=8, max_batch=4)
    assert np.array_equal(ctx.fe_min_eigen_val(frames[0]), fo.min_eigen_val(frames[0]))
    for md, mc in ((30.0, 1000), (3.0, 400), (0.5, 300), (12.0, 0)):
        got = ctx.fe_detect_batch(frames, [None, occ, None], mc, min_dist=md)
        for b in range(3):
            ref = fo.good_features(frames[b], occ if b == 1 else mask, mc, min_dist=md)
            assert np.array_equal(got[b], ref), (md, mc, b)
            assert len(ref) > 20


@pytest.mark.parametrize("size", [(963, 481), (1920, 960), (61, 45)])
def test_detect_fused_variant(ctx, synthetic, monkeypatch, size):
    """VIO360_DET_FUSED=1: no response map; the march applies the 3x3 test to its strip's interior columns against a
    threshold that is still rising, k_seam_candidates judges the strip seams, k_sort_select cuts at the final
    threshold.  Same corners as the default path: masks, no corner limit (every candidate comes back), a candidate
    segment that starts too small, a width of 1 mod 60."""
    W, H = size
    monkeypatch.setenv("VIO360_DET_FUSED", "1")
    monkeypatch.setenv("VIO360_DET_CAND", "256")
    frames = [synthetic.render_erp_frame(W, H, t, seed=51 + t, n_waves=40, sigma_omega=45.0 * max(1.0, W / 960)) for t in range(2)]
    mask = fo.base_mask(W, H)
    occ = mask.copy()
    occ[:, W // 3:W // 2] = 0
    ctx.fe_configure(W, H, max_points=8, max_batch=2)
    for md, mc in ((30.0, 1000), (3.0, 0), (12.0, 40)):
        got = ctx.fe_detect_batch(frames, [None, occ], mc, min_dist=md)
        for b in range(2):
            ref = fo.good_features(frames[b], occ if b == 1 else mask, mc, min_dist=md)
            assert np.array_equal(got[b], ref), (md, mc, b)
    assert len(ctx.fe_detect(frames[0], np.zeros_like(mask), 1000)) == 0


@pytest.mark.parametrize("fused", ["0", "1"])
def test_detect_dense_candidates(ctx, monkeypatch, fused):
    """A 3 x 3 periodic texture: every 3x3 box sees each phase once, so the response is one constant plateau and EVERY
    pixel of that half is a candidate (ties broken by raster index, as cv2's pointer comparison does) -- more than a
    block of k_candidates4 compacts in shared memory (1024 keys per 512 x 16 pixels; the rest is appended one by one),
    more than a warp of the fused march buffers between two flushes, and more than the first candidate segment holds."""
    monkeypatch.setenv("VIO360_DET_FUSED", fused)
    monkeypatch.setenv("VIO360_DET_CAND", "4096")
    W, H = 1024, 96
    rng = np.random.default_rng(77)
    tile = rng.integers(0, 256, (3, 3), dtype=np.uint8)
    img = np.tile(tile, (H // 3, W // 3 + 1))[:, :W].copy()
    img[:, W // 2:] = np.maximum(img[:, W // 2:] // 4, rng.integers(0, 60, (H, W // 2), dtype=np.uint8))   # half plateau, half noise
    ctx.fe_configure(W, H, max_points=8, max_batch=1)
    for md, mc in ((1.0, 0), (5.0, 3000)):
        ref = fo.good_features(img, fo.base_mask(W, H), mc, min_dist=md)
        got = ctx.fe_detect(img, None, mc, min_dist=md)
        assert np.array_equal(got, ref), (md, mc)
        assert len(ref) > 2000


def test_detect_with_occupied_features(ctx, synthetic):
    """the occupancy mask of CreateFeatureMask (FeatureTracker.cpp:386-402) built on
    the device from the feature positions equals detection with the rasterised cv::circle mask."""
    W, H = 960, 480
    f0 = synthetic.render_erp_frame(W, H, 0, seed=31, n_waves=48, sigma_omega=40.0)
    f1 = synthetic.render_erp_frame(W, H, 1, seed=32, n_waves=48, sigma_omega=40.0)
    ctx.fe_configure(W, H, max_points=8, max_batch=2)
    base = fo.base_mask(W, H)
    first = fo.good_features(f0, base, 120)
    occ_pts = first[::2] + np.float32(0.5)              # half-integer centres round to even
    occ_pts = np.concatenate([occ_pts, np.array([[3.0, 100.0], [W - 2.0, 300.0]], np.float32)])   # clipped discs
    for r in (30, 7):
        mask = fo.feature_mask(base, occ_pts, r)
        ref0 = fo.good_features(f0, mask, 200)
        ref1 = fo.good_features(f1, base, 200)
        got = ctx.fe_detect_batch([f0, f1], None, 200, occupied=[occ_pts, None], occupied_radius=r)
        assert np.array_equal(got[0], ref0) and np.array_equal(got[1], ref1) and len(ref0) > 20
        # an explicit mask and discs together: discs are drawn on top of the caller's mask
        extra = base.copy(); extra[:, 400:520] = 0
        got = ctx.fe_detect_batch([f0], [extra], 200, occupied=[occ_pts], occupied_radius=r)
        assert np.array_equal(got[0], fo.good_features(f0, fo.feature_mask(extra, occ_pts, r), 200))


def test_ransac_rotation_vs_oracle(ctx, synthetic):
    """rotation RANSAC with an injected sample list.  fp32 trigonometry differs
    by an ulp between libm and the device, so counts may differ on borderline points; the
    tolerance is written out: <= 1 % of the points per hypothesis, and mask differences only
    where the oracle's angular error is within 2e-4 rad of the threshold."""
    W, H = 3840, 1920
    ctx.fe_configure(W, H, max_points=8, max_batch=2)
    for n, iters, thr_deg, seed in ((1000, 200, 1.0, 11), (300, 64, 0.1, 12)):
        prev, cur, is_out = synthetic.make_rotation_tracks(W, H, n, seed=seed)
        samples = synthetic.draw_ransac_samples(n, iters, seed=seed + 100)
        best, mask, counts = ctx.fe_ransac_rotation(prev, cur, samples, thr_deg)
        obest, omask, ocounts, obi, oerrs = fo.ransac_rotation(prev, cur, W, H, samples, thr_deg)
        assert np.array_equal(counts < 0, ocounts < 0)
        assert np.abs(counts - ocounts).max() <= max(2, 0.01 * n)
        assert abs(best - obest) <= max(2, 0.01 * n) and mask.sum() == best == counts.max()
        if int(np.argmax(counts)) == obi:
            thr = np.float32(np.deg2rad(thr_deg))
            diff = mask != omask
            assert (np.abs(oerrs[diff] - thr) < 2e-4).all()
    # batch of two + degenerate inputs (FeatureTracker.cpp:258-261)
    p1, c1, _ = synthetic.make_rotation_tracks(W, H, 500, seed=21)
    p2, c2, _ = synthetic.make_rotation_tracks(W, H, 2, seed=22)
    s1 = synthetic.draw_ransac_samples(500, 50, seed=23)
    (b1, m1, k1), (b2, m2, k2) = ctx.fe_ransac_rotation_batch([(p1, c1, s1), (p2, c2, np.zeros((50, 3), np.int32))], 1.0)
    ob1 = fo.ransac_rotation(p1, c1, W, H, s1, 1.0)[0]
    assert abs(b1 - ob1) <= 5 and b2 == 2 and m2.tolist() == [1, 1]


def test_full_size_properties_c2(ctx, synthetic):
    """BASELINE config 2 size: a frame tracked onto itself does not move; status all ones."""
    W, H = 3840, 1920
    f0 = synthetic.render_erp_frame(W, H, 0, seed=360, n_waves=32, sigma_omega=120.0)
    pts = synthetic.make_track_points(W, H, 1000, seed=1000)
    ctx.fe_configure(W, H, max_points=1000, max_batch=1)
    xy, st, err, nt = ctx.fe_track(f0, f0, pts)
    ost = fo.lk_track(f0, f0, pts)[1]
    assert np.array_equal(st, ost)
    assert np.abs(xy - pts)[st == 1].max() < 1e-3 and err[st == 1].max() == 0.0
    f1 = np.roll(f0, 3, axis=1)             # pure horizontal shift by 3 px
    xy, st, err, nt = ctx.fe_track(f0, f1, pts)
    oxy, ost, _, _ = fo.lk_track(f0, f1, pts)
    assert np.array_equal(st, ost) and np.abs(xy - oxy)[ost == 1].max() < POS_TOL
    inner = (ost == 1) & (pts[:, 0] > 200) & (pts[:, 0] < W - 200)
    assert np.median(np.abs(xy[inner, 0] - pts[inner, 0] - 3.0)) < 0.05


def test_full_size_properties_c4(ctx, synthetic):
    """BASELINE config 4 size (7680x3840, 3000 features per frame): a two-sequence batch against the
    oracle on a shifted frame pair, a frame tracked onto itself, and the detector on the 8K frame."""
    W, H = 7680, 3840
    f0 = synthetic.render_erp_frame(W, H, 0, seed=361, n_waves=16, sigma_omega=120.0)
    f1 = np.roll(np.roll(f0, 2, axis=1), 1, axis=0)
    pts = synthetic.make_track_points(W, H, 3000, seed=1001)
    ctx.fe_configure(W, H, max_points=3000, max_batch=2)
    out = ctx.fe_track_batch_host([(f0, f1, pts, 1, 2), (f0, f0, pts[:2500], 1, 1)])
    oxy, ost, oerr, _ = fo.lk_track(f0, f1, pts)
    xy, st, err, keep = out[0]
    assert np.array_equal(st, ost) and np.abs(xy - oxy)[ost == 1].max() < POS_TOL
    assert np.array_equal(keep, fo.post_filter(xy, st, W, H, 0.15, 20.0))
    inner = (ost == 1) & (pts[:, 0] > 200) & (pts[:, 0] < W - 200)
    assert np.median(np.abs(xy[inner, 0] - pts[inner, 0] - 2.0)) < 0.05
    xy0, st0, err0, _ = out[1]
    assert np.array_equal(st0, fo.lk_track(f0, f0, pts[:2500])[1])
    assert np.abs(xy0 - pts[:2500])[st0 == 1].max() < 1e-3 and err0[st0 == 1].max() == 0.0
    ocorners = fo.good_features(f0, fo.base_mask(W, H), 3000)
    assert np.array_equal(ctx.fe_detect(f0, None, 3000), ocorners) and len(ocorners) > 1000


def _render_pair(synthetic, W, H, seed, sigma, n_waves=24, dt=1):
    """consecutive rendered frames of the moving camera (camera_pose: translation + yaw + pitch), on the GPU"""
    import torch
    f0 = synthetic.render_erp_frame(W, H, 0, seed=seed, n_waves=n_waves, sigma_omega=sigma, xp=torch, device="cuda").cpu().numpy()
    f1 = synthetic.render_erp_frame(W, H, dt, seed=seed, n_waves=n_waves, sigma_omega=sigma, xp=torch, device="cuda").cpu().numpy()
    return f0, f1


@pytest.mark.parametrize("size,n,seed,sigma", [((3840, 1920), 1000, 360, 120.0), ((7680, 3840), 3000, 361, 240.0)])
def test_track_and_detect_vs_cv2_live_full_size(ctx, synthetic, size, n, seed, sigma):
    """the CUDA path against cv2 ITSELF (the dependency the reference calls,
    FeatureTracker.cpp:222-223, 238-240) at BASELINE configs 2 and 4, on CONSECUTIVE RENDERED frames of the
    moving camera (not np.roll shifts): status bytes equal, positions within 0.05 px, corners identical."""
    cv2 = pytest.importorskip("cv2")
    W, H = size
    f0, f1 = _render_pair(synthetic, W, H, seed, sigma)
    pts = synthetic.make_track_points(W, H, n, seed=1000 + seed)
    ctx.fe_configure(W, H, max_points=n, max_batch=1)
    xy, st, err, nt = ctx.fe_track(f0, f1, pts)
    cxy, cst, cerr = fo.cv2_lk_track(f0, f1, pts)
    assert np.array_equal(st, cst), "status differs on %d of %d points" % (int((st != cst).sum()), n)
    ok = cst == 1
    assert ok.sum() > 0.8 * n
    assert np.abs(xy - cxy)[ok].max() < POS_TOL
    assert np.abs(err - cerr)[ok].max() < 0.05
    moved = np.linalg.norm(cxy - pts, axis=1)[ok]
    assert np.median(moved) > 0.5           # real sub-pixel motion, not a frame tracked onto itself
    # detection: cv2.goodFeaturesToTrack with the reference's arguments and its polar & boundary mask
    mask = fo.base_mask(W, H)
    cc = fo.cv2_good_features(f0, mask, n)
    got = ctx.fe_detect(f0, None, n)
    assert np.array_equal(got, cc) and len(cc) > 0.5 * n
    # re-detection as TrackFeatures does it: discs of radius min_distance at the tracked features
    occ = xy[ok][: n // 2]
    m2 = fo.feature_mask(mask, occ, 30)
    cc2 = fo.cv2_good_features(f0, m2, n)
    got2 = ctx.fe_detect_batch([f0], None, n, occupied=[occ], occupied_radius=30)[0]
    assert np.array_equal(got2, cc2)


def test_device_frames_alias_lifetime_and_failed_batch(ctx, synthetic, bindings):
    """(1) level 0 of a keyed device-space frame is aliased -- a cache hit must read it through the
    pointer of the CURRENT call, not the one remembered from the call that built the pyramid; (2) a batch
    rejected at validation (or failing before its uploads are enqueued) must not leave cache slots that a
    retry with the same keys takes for loaded frames."""
    import torch
    W, H, n = 640, 320, 200
    frames = [synthetic.render_erp_frame(W, H, t, seed=44, n_waves=32, sigma_omega=40.0) for t in range(3)]
    pts = synthetic.make_track_points(W, H, n, seed=5)
    ctx.fe_configure(W, H, max_points=n, max_batch=2)
    dev = [torch.from_numpy(f).cuda() for f in frames]
    dpts = torch.from_numpy(pts).cuda()
    oxy = torch.zeros((n, 2), dtype=torch.float32, device="cuda"); ost = torch.zeros(n, dtype=torch.uint8, device="cuda")

    def run(prev, cur, kp, kc):
        jobs = (bindings.TrackJob * 1)()
        j = jobs[0]
        j.prev, j.prev_pitch, j.prev_key, j.cur, j.cur_pitch, j.cur_key = prev.data_ptr(), W, kp, cur.data_ptr(), W, kc
        j.prev_xy, j.n, j.cur_xy, j.status, j.err, j.keep = dpts.data_ptr(), n, oxy.data_ptr(), ost.data_ptr(), None, None
        ctx.fe_track_batch_raw(jobs, 1, bindings.MEM_DEVICE)
        ctx.synchronize()
        return oxy.cpu().numpy().copy(), ost.cpu().numpy().copy()

    run(dev[0], dev[1], 9001, 9002)                      # builds and caches the pyramid of frame 1 under key 9002
    moved = dev[1].clone()                               # the caller moves frame 1 and recycles the old buffer
    dev[1].fill_(0)
    xy, st = run(moved, dev[2], 9002, 9003)              # cache hit on 9002: level 0 must come from `moved`
    rxy, rst, _, _ = fo.lk_track(frames[1], frames[2], pts)
    assert np.array_equal(st, rst) and np.abs(xy - rxy)[rst == 1].max() < POS_TOL

    # (2) host-space batch whose second job is invalid: nothing may be cached under keys 7001 / 7002
    ctx.fe_configure(W, H, max_points=n, max_batch=2)
    ctx.fe_track(frames[2], frames[0], pts, 0, 0)        # fill the slot pool with other pixels
    bad = np.zeros((n + 1, 2), np.float32)
    with pytest.raises(bindings.Vio360Error):
        ctx.fe_track_batch_host([(frames[0], frames[1], pts, 7001, 7002), (frames[1], frames[2], bad, 7002, 7003)])
    (xy, st, err, keep), = ctx.fe_track_batch_host([(frames[0], frames[1], pts, 7001, 7002)])
    rxy, rst, _, _ = fo.lk_track(frames[0], frames[1], pts)
    assert np.array_equal(st, rst) and np.abs(xy - rxy)[rst == 1].max() < POS_TOL


def test_host_register_in_place(ctx, synthetic, bindings):
    """vio360_host_register: a pageable numpy frame page-locked in place gives the same tracking result; registering
    twice is fine, unregistering an unknown range is an error."""
    W, H = 1280, 640
    ctx.fe_configure(W, H, max_points=300, max_batch=2)
    f0 = synthetic.render_erp_frame(W, H, 0, seed=5, n_waves=32, sigma_omega=60.0)
    f1 = synthetic.render_erp_frame(W, H, 1, seed=5, n_waves=32, sigma_omega=60.0)
    pts = synthetic.make_track_points(W, H, 300, seed=6)
    ref = ctx.fe_track(f0, f1, pts)
    ctx.host_register(f0); ctx.host_register(f1); ctx.host_register(f1)
    got = ctx.fe_track(f0, f1, pts)
    ctx.host_unregister(f0); ctx.host_unregister(f1)
    assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1]) and got[3] == ref[3]
    with pytest.raises(bindings.Vio360Error):
        ctx.host_unregister(np.zeros(16, np.uint8))
