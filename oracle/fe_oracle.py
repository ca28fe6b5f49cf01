"""oracle/fe_oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes binding of oracle/fe_ref.c plus thin wrappers that call cv2 (the third-party
dependency the reference's front-end delegates to) with EXACTLY the reference's
arguments (FeatureTracker.cpp:33-38, 222-223, 238-240).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

# The reference builds cv::TermCriteria(EPS|COUNT, 30, 0.01f): the float literal widens
# to this double; the same value is re-used as minEigThreshold (FeatureTracker.cpp:35,240)
LK_EPS = float(np.float32(0.01))
LK_WIN = 21
LK_MAX_LEVEL = 3
LK_MAX_ITER = 30
GFTT_QUALITY = float(np.float32(0.01))   # float m_quality_level widened to double
GFTT_MIN_DIST = float(np.float32(30.0))


def build(force=False):
    so = os.path.join(_HERE, "libvio360_oracle_fe.so")
    src = os.path.join(_HERE, "fe_ref.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libvio360_oracle_fe.so"],
                              stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def pyr_down(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty(((h + 1) // 2, (w + 1) // 2), np.uint8)
    lib().vf_pyr_down(_p(img, C.c_ubyte), w, h, C.c_size_t(img.strides[0]), _p(out, C.c_ubyte),
                      C.c_size_t(out.strides[0]))
    return out


def scharr(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty((h, w, 2), np.int16)
    lib().vf_scharr(_p(img, C.c_ubyte), w, h, C.c_size_t(img.strides[0]), _p(out, C.c_short))
    return out


def lk_track(prev, cur, pts, max_level=LK_MAX_LEVEL, win=LK_WIN, max_iter=LK_MAX_ITER,
             eps=LK_EPS, min_eig=LK_EPS, want_err=True):
    prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
    n = pts.shape[0]
    h, w = prev.shape
    out = np.zeros((n, 2), np.float32); st = np.zeros(n, np.uint8); err = np.zeros(n, np.float32)
    cnt = lib().vf_lk_track(_p(prev, C.c_ubyte), C.c_size_t(prev.strides[0]), _p(cur, C.c_ubyte),
                            C.c_size_t(cur.strides[0]), w, h, max_level, win, max_iter,
                            C.c_double(eps), C.c_double(min_eig), _p(pts, C.c_float), n,
                            _p(out, C.c_float), _p(st, C.c_ubyte),
                            _p(err, C.c_float) if want_err else None)
    return out, st, err, cnt


def min_eigen_val(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty((h, w), np.float32)
    lib().vf_min_eigen_val(_p(img, C.c_ubyte), w, h, C.c_size_t(img.strides[0]), _p(out, C.c_float))
    return out


def good_features(img, mask=None, max_corners=1000, quality=GFTT_QUALITY, min_dist=GFTT_MIN_DIST):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    cap = max_corners if max_corners > 0 else w * h
    out = np.zeros((cap, 2), np.float32)
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
        n = lib().vf_good_features(_p(img, C.c_ubyte), w, h, C.c_size_t(img.strides[0]),
                                   _p(mask, C.c_ubyte), C.c_size_t(mask.strides[0]), max_corners,
                                   C.c_double(quality), C.c_double(min_dist), _p(out, C.c_float))
    else:
        n = lib().vf_good_features(_p(img, C.c_ubyte), w, h, C.c_size_t(img.strides[0]), None,
                                   C.c_size_t(0), max_corners, C.c_double(quality),
                                   C.c_double(min_dist), _p(out, C.c_float))
    return out[:n].copy()


def base_mask(w, h, polar_ratio=0.15, boundary_margin=20):
    m = np.empty((h, w), np.uint8)
    lib().vf_base_mask(w, h, C.c_float(polar_ratio), boundary_margin, _p(m, C.c_ubyte))
    return m


def post_filter(xy, status, w, h, polar_thr=0.15, margin=20.0):
    xy = np.ascontiguousarray(xy, np.float32); status = np.ascontiguousarray(status, np.uint8)
    keep = np.empty(status.shape[0], np.uint8)
    lib().vf_post_filter(_p(xy, C.c_float), _p(status, C.c_ubyte), status.shape[0], w, h,
                         C.c_float(polar_thr), C.c_float(margin), _p(keep, C.c_ubyte))
    return keep


def feature_mask(mask, pts, radius):
    """FeatureTracker::CreateFeatureMask: filled cv::circle discs zeroed in a copy of `mask`."""
    out = np.ascontiguousarray(mask, np.uint8).copy()
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
    h, w = out.shape
    f = lib().vf_feature_mask
    f.restype = None
    f(_p(out, C.c_ubyte), w, h, _p(pts, C.c_float), pts.shape[0], int(radius))
    return out


def ransac_rotation(prev_xy, cur_xy, w, h, samples, threshold_deg):
    """FeatureTracker::RejectOutliersRotationRANSAC with an injected sample list (iters x 3).
    returns (best_inliers, mask, counts, best_iter, errs)."""
    prev_xy = np.ascontiguousarray(prev_xy, np.float32); cur_xy = np.ascontiguousarray(cur_xy, np.float32)
    samples = np.ascontiguousarray(samples, np.int32).reshape(-1, 3)
    n, iters = prev_xy.shape[0], samples.shape[0]
    mask = np.zeros(n, np.uint8); counts = np.zeros(max(iters, 1), np.int32); errs = np.zeros(max(n, 1), np.float32)
    bi = C.c_int(-1)
    f = lib().vf_ransac_rotation
    f.restype = C.c_int
    best = f(_p(prev_xy, C.c_float), _p(cur_xy, C.c_float), n, w, h, _p(samples, C.c_int), iters,
             C.c_float(threshold_deg), _p(mask, C.c_ubyte), _p(counts, C.c_int), C.byref(bi), _p(errs, C.c_float))
    return best, mask, counts[:iters], bi.value, errs[:n]


def cluster_filter(xy, w, h, cols=20, rows=10, std_ratio=0.25):
    """FeatureTracker::RemoveClusteredFeatures -> keep mask."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    keep = np.zeros(max(xy.shape[0], 1), np.uint8)
    f = lib().vf_cluster_filter
    f.restype = None
    f(_p(xy, C.c_float), xy.shape[0], w, h, cols, rows, C.c_float(std_ratio), _p(keep, C.c_ubyte))
    return keep[:xy.shape[0]]


def grid_cap(xy, track, w, h, cols=20, rows=10, max_per_cell=4):
    """Frame::AssignFeaturesToGrid + LimitFeaturesPerGrid -> keep mask."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2); track = np.ascontiguousarray(track, np.int32)
    keep = np.zeros(max(xy.shape[0], 1), np.uint8)
    f = lib().vf_grid_cap
    f.restype = None
    f(_p(xy, C.c_float), _p(track, C.c_int), xy.shape[0], w, h, cols, rows, max_per_cell, _p(keep, C.c_ubyte))
    return keep[:xy.shape[0]]


# ------------------------------------------------------------------ the real dependency
def cv2_lk_track(prev, cur, pts):
    """cv::calcOpticalFlowPyrLK exactly as FeatureTracker.cpp:238-240 calls it."""
    import cv2
    crit = (cv2.TERM_CRITERIA_EPS | cv2.TERM_CRITERIA_COUNT, LK_MAX_ITER, LK_EPS)
    nxt, st, err = cv2.calcOpticalFlowPyrLK(prev, cur, np.ascontiguousarray(pts, np.float32).reshape(-1, 1, 2),
                                            None, winSize=(LK_WIN, LK_WIN), maxLevel=LK_MAX_LEVEL,
                                            criteria=crit, flags=0, minEigThreshold=LK_EPS)
    return nxt.reshape(-1, 2), st.reshape(-1), err.reshape(-1)


def cv2_good_features(img, mask=None, max_corners=1000, quality=GFTT_QUALITY, min_dist=GFTT_MIN_DIST):
    """cv::goodFeaturesToTrack exactly as FeatureTracker.cpp:222-223 calls it."""
    import cv2
    c = cv2.goodFeaturesToTrack(img, max_corners, quality, min_dist, mask=mask,
                                blockSize=3, useHarrisDetector=False, k=0.04)
    return np.zeros((0, 2), np.float32) if c is None else c.reshape(-1, 2)


# ------------------------------------------------------------------ the per-frame chain
def ransac_samples_from_words(words, n):
    """The sample rule of vio360_fe_track_frame_batch (include/vio360.h): word r -> (r * n) >> 32, an index
    equal to an earlier one of its triple moves to (index + 1) % n.  The reference draws from
    std::random_device (FeatureTracker.cpp:273-289): any three distinct uniform indices are a faithful run."""
    words = np.ascontiguousarray(words, np.uint32).reshape(-1, 3)
    out = np.zeros(words.shape, np.int32)
    for it in range(words.shape[0]):
        idx = []
        for s in range(3):
            v = (int(words[it, s]) * n) >> 32
            while v in idx:
                v = (v + 1) % n
            idx.append(v)
        out[it] = idx
    return out


class SequenceTracker:
    """FeatureTracker::TrackFeatures (FeatureTracker.cpp:61-206) frame after frame, on the CPU: cv2 for the two
    OpenCV calls (or oracle/fe_ref.c with use_cv2=False), fe_ref.c for the reference's own list logic.
    State = what Frame / Feature carry between frames: ids, pixel coordinates, track counts."""

    def __init__(self, w, h, max_features=1000, min_distance=30.0, quality=0.01, grid_cols=20, grid_rows=10,
                 max_per_cell=10, remove_clustered=True, std_ratio=0.25, ransac_thr_deg=2.0, use_cv2=True,
                 polar_ratio=0.15, margin=20):
        self.w, self.h = w, h
        self.max_features, self.min_distance, self.quality = max_features, float(np.float32(min_distance)), float(np.float32(quality))
        self.grid = (grid_cols, grid_rows, max_per_cell)
        self.remove_clustered, self.std_ratio, self.ransac_thr_deg = remove_clustered, std_ratio, ransac_thr_deg
        self.use_cv2, self.polar_ratio, self.margin = use_cv2, polar_ratio, margin
        self.base = base_mask(w, h, polar_ratio, margin)
        self.polar = base_mask(w, h, polar_ratio, 0)              # Camera::CreatePolarMask: rows only
        self.ids = np.zeros(0, np.int32); self.xy = np.zeros((0, 2), np.float32); self.tc = np.zeros(0, np.int32)
        self.next_id = 0
        self.prev = None

    def _detect(self, img, mask):
        if self.use_cv2:
            return cv2_good_features(img, mask, self.max_features, self.quality, self.min_distance).astype(np.float32)
        return good_features(img, mask, self.max_features, self.quality, self.min_distance)

    def _grid_cap(self):
        keep = grid_cap(self.xy, self.tc, self.w, self.h, *self.grid) if len(self.xy) else np.zeros(0, np.uint8)
        k = keep.astype(bool)
        self.ids, self.xy, self.tc = self.ids[k], self.xy[k], self.tc[k]

    def step(self, img, words):
        """returns (n_tracked, n_detected); the frame's features are self.ids / self.xy / self.tc"""
        n_tracked = 0
        if self.prev is not None and len(self.xy):
            if self.use_cv2:
                cur, st, _ = cv2_lk_track(self.prev, img, self.xy)
            else:
                cur, st, _, _ = lk_track(self.prev, img, self.xy)
            keep = post_filter(cur, st, self.w, self.h, self.polar_ratio, float(self.margin)).astype(bool)   # :118-126
            gp, gc, gi = self.xy[keep], cur[keep], np.nonzero(keep)[0]
            if len(gc) >= 3:                                                                              # :131
                samples = ransac_samples_from_words(words, len(gc))
                _, mask, _, _, _ = ransac_rotation(gp, gc, self.w, self.h, samples, self.ransac_thr_deg)
            else:
                mask = np.ones(len(gc), np.uint8)
            m = mask.astype(bool)
            self.ids, self.xy, self.tc = self.ids[gi[m]], gc[m].copy(), self.tc[gi[m]] + 1               # :139-163
            n_tracked = len(self.ids)
            if self.remove_clustered and len(self.xy) >= 4:                                               # :168
                k = cluster_filter(self.xy, self.w, self.h, self.grid[0], self.grid[1], self.std_ratio).astype(bool)
                self.ids, self.xy, self.tc = self.ids[k], self.xy[k], self.tc[k]
            self._grid_cap()                                                                               # :171-172
        else:
            self.ids = np.zeros(0, np.int32); self.xy = np.zeros((0, 2), np.float32); self.tc = np.zeros(0, np.int32)
        n_detected = 0
        if len(self.ids) < self.max_features:                                                              # :176
            # CreateFeatureMask starts from the POLAR mask (:389); DetectNewFeatures ands it with polar & boundary (:212-219)
            mask = feature_mask(self.polar, self.xy, int(self.min_distance)) & self.base
            new = self._detect(img, mask)
            n_detected = len(new)
            self.ids = np.concatenate([self.ids, self.next_id + np.arange(n_detected, dtype=np.int32)])
            self.xy = np.concatenate([self.xy, new.reshape(-1, 2)]).astype(np.float32)
            self.tc = np.concatenate([self.tc, np.zeros(n_detected, np.int32)])
            self.next_id += n_detected
            self._grid_cap()                                                                               # :199-200
        self.prev = img
        return n_tracked, n_detected
