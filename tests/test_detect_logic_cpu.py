"""CPU statements of two exactness arguments the detection kernels rest on (csrc/fe_detect.cuh), checked on random data
with numpy only:

1. the fused march (k_min_eig_cand) tests a pixel against its RAW neighbours and against a threshold that is still
   rising; cv::goodFeaturesToTrack (featureselect.cpp) thresholds the map first (THRESH_TOZERO), dilates it 3x3 and
   keeps the pixels equal to the dilated value.  For values above the final threshold the two tests select the same
   pixels, whatever lower threshold was in force when a pixel was emitted;
2. k_sort_select drops, between two tranches and between two chunks of a tranche, every remaining candidate that
   already lies within minDistance of an accepted corner.  The greedy walk over the candidates in descending order
   accepts the same corners with and without that pruning.
"""
import numpy as np
import pytest


def _cv_candidates(eig, thr):
    """featureselect.cpp: threshold(TOZERO), dilate 3x3, val != 0 && val == dilated, interior pixels only."""
    H, W = eig.shape
    t = np.where(eig > thr, eig, np.float32(0))
    pad = np.pad(t, 1, mode="constant", constant_values=-np.inf)
    dil = np.max([pad[dy:dy + H, dx:dx + W] for dy in range(3) for dx in range(3)], axis=0)
    c = (t != 0) & (t == dil)
    c[0, :] = c[-1, :] = False
    c[:, 0] = c[:, -1] = False
    return c


def _raw_candidates(eig, thr_map):
    """The march's test: above the (per pixel, possibly lower) threshold in force, positive, not below any raw neighbour."""
    H, W = eig.shape
    pad = np.pad(eig, 1, mode="constant", constant_values=-np.inf)
    nb = np.max([pad[dy:dy + H, dx:dx + W] for dy in range(3) for dx in range(3)], axis=0)
    c = (eig > thr_map) & (eig > 0) & (eig == nb)
    c[0, :] = c[-1, :] = False
    c[:, 0] = c[:, -1] = False
    return c


@pytest.mark.parametrize("seed", [0, 1, 2, 3])
def test_raw_neighbour_test_equals_thresholded_dilate(seed):
    rng = np.random.default_rng(seed)
    H, W = 60, 90
    eig = rng.random((H, W)).astype(np.float32) ** 4
    eig[rng.random((H, W)) < 0.2] = 0                                   # flat regions
    eig[10:14, 20:30] = eig[10, 20]                                       # a plateau of equal values
    q = np.repeat(np.repeat(rng.random((H // 6, W // 6)).astype(np.float32), 6, 0), 6, 1)
    eig = np.maximum(eig, (q > 0.9) * np.float32(0.5))                    # blocks of ties
    thr = np.float32(0.01 * eig.max())
    ref = _cv_candidates(eig, thr)
    # thresholds in force while the march runs: anything from zero up to the final one, different per pixel
    rising = (rng.random((H, W)).astype(np.float32) * thr).astype(np.float32)
    emitted = _raw_candidates(eig, rising)
    kept = emitted & (eig > thr)                                          # k_sort_select's floor
    assert np.array_equal(kept, ref)
    assert ref.sum() > 20 and emitted.sum() >= ref.sum()


def _greedy(keys_xy, min_dist, max_corners, prune_every=None):
    """goodFeaturesToTrack's walk over candidates sorted by descending value; prune_every = n: after every n walked
    candidates all not yet walked ones within min_dist of an accepted corner are dropped (k_sort_select)."""
    acc = []
    live = list(range(len(keys_xy)))
    md2 = min_dist * min_dist
    walked = 0
    while live:
        i = live.pop(0)
        x, y = keys_xy[i]
        if all((x - ax) ** 2 + (y - ay) ** 2 >= md2 for ax, ay in acc):
            acc.append((x, y))
            if max_corners and len(acc) >= max_corners:
                break
        walked += 1
        if prune_every and walked % prune_every == 0:
            live = [j for j in live if all((keys_xy[j][0] - ax) ** 2 + (keys_xy[j][1] - ay) ** 2 >= md2 for ax, ay in acc)]
    return acc


@pytest.mark.parametrize("seed,min_dist,max_corners", [(0, 8.0, 0), (1, 12.0, 40), (2, 5.0, 0), (3, 20.0, 25)])
def test_pruning_covered_candidates_keeps_the_greedy_result(seed, min_dist, max_corners):
    rng = np.random.default_rng(seed)
    n = 1500
    xy = np.stack([rng.integers(0, 160, n), rng.integers(0, 120, n)], 1)
    xy = np.unique(xy, axis=0)
    rng.shuffle(xy)                                                       # the order stands for descending value
    pts = [tuple(map(int, p)) for p in xy]
    ref = _greedy(pts, min_dist, max_corners)
    for every in (1, 7, 64, 512):
        assert _greedy(pts, min_dist, max_corners, prune_every=every) == ref
    assert len(ref) > 10


def test_integer_widening_of_squares_is_exact():
    """widen_sq (fe_detect.cuh): float -> double on the integer pipe for non-negative NORMAL floats and zero --
    hi = (bits >> 3) + 0x38000000, lo = bits << 29.  The march feeds it squares of Sobel gradients scaled by
    float(1/3060): zero or >= 2^-70, never subnormal."""
    rng = np.random.default_rng(5)
    f = np.concatenate([
        np.zeros(4, np.float32),
        np.float32(2.0) ** rng.integers(-70, 20, 20000).astype(np.float32) * (1 + rng.random(20000).astype(np.float32)),
        np.array([np.finfo(np.float32).tiny, 1.0, 3060.0 ** -2, np.finfo(np.float32).max], np.float32)])
    b = f.view(np.uint32).astype(np.uint64)
    hi = np.where(b != 0, (b >> np.uint64(3)) + np.uint64(0x38000000), np.uint64(0))
    lo = (b << np.uint64(29)) & np.uint64(0xffffffff)
    d = ((hi << np.uint64(32)) | lo).view(np.float64)
    assert np.array_equal(d, f.astype(np.float64))
