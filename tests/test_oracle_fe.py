"""CPU tests pinning oracle/fe_ref.c against (a) the committed cv2 golden vectors and
(b) cv2 itself, live, with the reference's exact arguments."""
import os

import numpy as np
import pytest

import fe_oracle as fo

GOLD = os.path.join(os.path.dirname(__file__), "golden", "fe_golden.npz")
POS_TOL = 0.05      # px, BASELINE.json north_star


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _bits_sum(a):
    return np.array([a.view(np.uint32).astype(np.uint64).sum()], np.uint64)


def test_pyramid_bit_exact_golden(gold):
    lv = gold["f0"]
    for name in ("pyr1", "pyr2", "pyr3"):
        lv = fo.pyr_down(lv)
        assert np.array_equal(lv, gold[name])


def test_lk_against_golden(gold):
    xy, st, err, cnt = fo.lk_track(gold["f0"], gold["f1"], gold["pts"])
    assert np.array_equal(st, gold["cv_status"]) and cnt == int(gold["cv_status"].sum())
    ok = st == 1
    assert np.abs(xy - gold["cv_next"])[ok].max() < POS_TOL
    assert np.abs(err - gold["cv_err"])[ok].max() < 0.05
    assert (st == 0).sum() >= 10          # the fixture exercises the status=0 branches


def test_gftt_against_golden(gold):
    f0 = gold["f0"]
    h, w = f0.shape
    mask = fo.base_mask(w, h, 0.15, 20)
    assert np.array_equal(fo.good_features(f0, mask, 1000), gold["gftt_base"])
    assert np.array_equal(fo.good_features(f0, gold["occ_mask"], 1000), gold["gftt_occ"])
    assert np.array_equal(fo.good_features(f0, mask, 25), gold["gftt_top25"])
    eig = fo.min_eigen_val(f0)
    assert np.array_equal(eig[::8], gold["eig_rows"]) and np.array_equal(_bits_sum(eig), gold["eig_bits_sum"])


def test_post_filter_and_masks():
    m = fo.base_mask(960, 480, 0.15, 20)
    assert m[:72].max() == 0 and m[72, 20] == 255 and m[72, 19] == 0 and m[407, 939] == 255
    assert m[408:].max() == 0 and m[300, 940] == 0       # int(480*(1-0.15f)) == 408
    xy = np.array([[100, 100], [100, 71.9], [19.9, 200], [940.1, 200], [940, 200], [100, 408.1], [100, 200]], np.float32)
    st = np.array([1, 1, 1, 1, 1, 1, 0], np.uint8)
    assert list(fo.post_filter(xy, st, 960, 480)) == [1, 0, 0, 0, 1, 0, 0]


# ---- live comparisons with the dependency itself -------------------------------------
cv2 = pytest.importorskip("cv2")


@pytest.mark.parametrize("shape", [(481, 963), (240, 320), (67, 45)])
def test_pyr_scharr_vs_cv2(shape):
    rng = np.random.default_rng(shape[0])
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    assert np.array_equal(fo.pyr_down(img), cv2.pyrDown(img))
    s = fo.scharr(img)
    assert np.array_equal(s[..., 0], cv2.Scharr(img, cv2.CV_16S, 1, 0))
    assert np.array_equal(s[..., 1], cv2.Scharr(img, cv2.CV_16S, 0, 1))


@pytest.mark.parametrize("size,seed", [((640, 320), 3), ((1280, 640), 4)])
def test_lk_vs_cv2_live(synthetic, size, seed):
    W, H = size
    f0 = synthetic.render_erp_frame(W, H, 0, seed=seed, n_waves=48, sigma_omega=40.0)
    f1 = synthetic.render_erp_frame(W, H, 2, seed=seed, n_waves=48, sigma_omega=40.0)
    pts = synthetic.make_track_points(W, H, 400, seed=seed, seam_frac=0.1)
    a, sa, ea, _ = fo.lk_track(f0, f1, pts)
    b, sb, eb = fo.cv2_lk_track(f0, f1, pts)
    assert np.array_equal(sa, sb)
    ok = sb == 1
    assert np.abs(a - b)[ok].max() < POS_TOL
    for nt in (1, 4):                      # thread-count invariance of the dependency
        cv2.setNumThreads(nt)
        b2, sb2, _ = fo.cv2_lk_track(f0, f1, pts)
        assert np.array_equal(sb2, sb) and np.array_equal(b2, b)
    cv2.setNumThreads(-1)


@pytest.mark.parametrize("size,seed", [((960, 480), 3), ((1920, 960), 4)])
def test_gftt_vs_cv2_live(synthetic, size, seed):
    W, H = size
    f0 = synthetic.render_erp_frame(W, H, 0, seed=seed, n_waves=48, sigma_omega=50.0)
    mask = fo.base_mask(W, H)
    assert np.array_equal(cv2.cornerMinEigenVal(f0, 3, ksize=3), fo.min_eigen_val(f0))
    a = fo.cv2_good_features(f0, mask)
    b = fo.good_features(f0, mask)
    assert np.array_equal(a, b) and len(a) > 50


def test_feature_mask_matches_cv2_circle():
    """CreateFeatureMask (FeatureTracker.cpp:386-402): the restated midpoint-circle fill equals
    cv2.circle(mask, (cvRound(x), cvRound(y)), r, 0, -1) including clipped and half-integer centres."""
    cv2 = pytest.importorskip("cv2")
    w, h = 321, 200
    rng = np.random.default_rng(9)
    pts = np.concatenate([rng.uniform(-20, [w + 20, h + 20], (60, 2)),
                          np.array([[0.5, 0.5], [1.5, 2.5], [w - 1, h - 1], [10.5, 20.5], [160.49, 99.51]])]).astype(np.float32)
    for r in (0, 1, 2, 7, 30, 64):
        ref = np.full((h, w), 255, np.uint8)
        for x, y in pts:
            cv2.circle(ref, (int(np.rint(x)), int(np.rint(y))), r, 0, -1)
        got = fo.feature_mask(np.full((h, w), 255, np.uint8), pts, r)
        assert np.array_equal(got, ref), r


def test_sequence_chain_cv2_vs_restatement(synthetic):
    """The CPU statement of FeatureTracker::TrackFeatures (fe_oracle.SequenceTracker) gives the same per-frame feature
    ids / track counts whether the two OpenCV calls are cv2 itself or oracle/fe_ref.c (positions within 0.05 px),
    and it behaves like a tracker: ids persist, track counts grow, lost features are replenished."""
    W, H, T = 640, 320, 6
    frames = [synthetic.render_erp_frame(W, H, t, seed=11, n_waves=24, sigma_omega=40.0) for t in range(T)]
    rng = np.random.default_rng(5)
    kw = dict(max_features=200, min_distance=12.0, grid_cols=20, grid_rows=10, max_per_cell=4, std_ratio=0.15, ransac_thr_deg=1.0)
    a = fo.SequenceTracker(W, H, use_cv2=True, **kw); b = fo.SequenceTracker(W, H, use_cv2=False, **kw)
    seen_old = False
    for t in range(T):
        words = rng.integers(0, 2 ** 32, (64, 3), dtype=np.uint64).astype(np.uint32)
        ra = a.step(frames[t], words); rb = b.step(frames[t], words)
        assert ra == rb and np.array_equal(a.ids, b.ids) and np.array_equal(a.tc, b.tc)
        assert np.abs(a.xy - b.xy).max() < POS_TOL
        b.xy = a.xy.copy()                       # keep the two chains on identical inputs
        if t == 0:
            assert ra[0] == 0 and ra[1] > 50 and (a.tc == 0).all()
        else:
            assert ra[0] > 20 and a.tc.max() == t
            seen_old = seen_old or (a.ids[a.tc == t] < 200).any()
        assert len(np.unique(a.ids)) == len(a.ids)
    assert seen_old
    s = fo.ransac_samples_from_words(np.array([[0, 0, 0], [2 ** 32 - 1, 2 ** 32 - 1, 0], [2 ** 31, 2 ** 31, 2 ** 31]], np.uint32), 5)
    assert s.tolist() == [[0, 1, 2], [4, 0, 1], [2, 3, 4]]
