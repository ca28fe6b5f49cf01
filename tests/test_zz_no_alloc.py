"""vio360_fe_ransac_rotation_noalloc (include/vio360.h): the rotation RANSAC with its intermediates in a static
device array, for callers that may not allocate.  No counterpart in the reference; the check is against the
existing entry point (vio360_fe_ransac_rotation_batch), which launches the same kernels on pooled scratch, so
the bar is equality: mask bytes, inlier count and winning hypothesis.

Collected last on purpose.  Uses a context of its own, so it does not depend
on what earlier tests configured."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H = 3840, 1920
MAX_POINTS, MAX_ITERS = 4096, 1024     # VIO360_RANSAC_NOALLOC_MAX_POINTS / _MAX_ITERS


@pytest.fixture(scope="module")
def own_ctx(bindings):
    c = bindings.Context(0)
    c.fe_configure(W, H, max_points=8, max_batch=2)
    yield c
    c.close()


def _noalloc(c, torch, prev, cur, samples, thr_deg, check_memory=False):
    dev = "cuda:0"
    n, iters = len(prev), len(samples)
    tp = torch.from_numpy(np.ascontiguousarray(prev, np.float32)).to(dev)
    tc = torch.from_numpy(np.ascontiguousarray(cur, np.float32)).to(dev)
    ts = torch.from_numpy(np.ascontiguousarray(samples, np.int32)).to(dev)
    mask = torch.full((max(n, 1),), 7, dtype=torch.uint8, device=dev)       # poisoned outputs
    best = torch.full((2,), -5, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    c.fe_ransac_rotation_noalloc(tp.data_ptr(), tc.data_ptr(), n, ts.data_ptr(), iters, thr_deg, mask.data_ptr(), best.data_ptr())
    c.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    if check_memory:
        assert free1 == free0, "the call changed the device's free memory by %d bytes" % (free0 - free1)
    return best.cpu().numpy(), mask.cpu().numpy()[:n]


def test_noalloc_ransac_equals_pooled_entry_point(own_ctx, synthetic, bindings):
    import torch
    c = own_ctx
    # two calls in a row (the second smaller, so the array still holds the first call's bearings and counts
    # beyond the second's ranges), then the maxima, then a degenerate list
    cases = ((1000, 200, 1.0, 11), (300, 64, 0.1, 12), (MAX_POINTS, MAX_ITERS, 0.5, 13), (700, 128, 2.0, 14))
    for k, (n, iters, thr_deg, seed) in enumerate(cases):
        prev, cur, _ = synthetic.make_rotation_tracks(W, H, n, seed=seed)
        samples = synthetic.draw_ransac_samples(n, iters, seed=seed + 100)
        ref_best, ref_mask, ref_counts = c.fe_ransac_rotation(prev, cur, samples, thr_deg)
        best, mask = _noalloc(c, torch, prev, cur, samples, thr_deg, check_memory=k > 0)
        assert int(best[0]) == ref_best and np.array_equal(mask, ref_mask), (n, iters)
        assert ref_best > 0.3 * n and int(best[1]) == int(np.argmax(ref_counts))
        # and once more on the same inputs: the array's previous contents do not matter
        best2, mask2 = _noalloc(c, torch, prev, cur, samples, thr_deg, check_memory=True)
        assert np.array_equal(best2, best) and np.array_equal(mask2, mask)
    # FeatureTracker.cpp:258-261: fewer than three points -> everything is an inlier
    prev, cur, _ = synthetic.make_rotation_tracks(W, H, 2, seed=22)
    best, mask = _noalloc(c, torch, prev, cur, np.zeros((16, 3), np.int32), 1.0)
    assert best.tolist() == [2, -1] and mask.tolist() == [1, 1]
    # no hypothesis at all: iters = 0 -> count 0, all ones (FeatureTracker.cpp:321-322)
    prev, cur, _ = synthetic.make_rotation_tracks(W, H, 50, seed=23)
    best, mask = _noalloc(c, torch, prev, cur, np.zeros((0, 3), np.int32), 1.0)
    assert best.tolist() == [0, -1] and mask.tolist() == [1] * 50


def test_noalloc_ransac_rejects_sizes_beyond_the_static_array(own_ctx, bindings):
    import torch
    c = own_ctx
    buf = torch.zeros(4 * (MAX_POINTS + 8), dtype=torch.float32, device="cuda:0")
    out = torch.zeros(MAX_POINTS + 8, dtype=torch.int32, device="cuda:0")
    p = buf.data_ptr()
    for n, iters in ((MAX_POINTS + 1, 16), (100, MAX_ITERS + 1)):
        with pytest.raises(bindings.Vio360Error, match="static scratch"):
            c.fe_ransac_rotation_noalloc(p, p, n, out.data_ptr(), iters, 1.0, out.data_ptr(), out.data_ptr())
    with pytest.raises(bindings.Vio360Error):
        c.fe_ransac_rotation_noalloc(p, p, 10, out.data_ptr(), 4, 1.0, out.data_ptr(), 0)     # no output array
