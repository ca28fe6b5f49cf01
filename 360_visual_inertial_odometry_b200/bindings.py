"""ctypes binding of libvio360_cuda.so (include/vio360.h).

This is the Python-side twin of the C++ host shims in host/: thin, no arithmetic, no
fallback.  Loading fails loudly when the library has not been built, and every call
raises Vio360Error when the C ABI reports an error (e.g. no CUDA device).
"""
import ctypes as C
import weakref
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvio360_cuda.so")

MEM_HOST, MEM_DEVICE = 0, 1
PEER_BLOB_BYTES = 128   # VIO360_PEER_BLOB_BYTES

EXPORTS = [
    "vio360_version", "vio360_create", "vio360_destroy", "vio360_last_error", "vio360_synchronize",
    "vio360_stream", "vio360_launch_count", "vio360_probe_fma_peak", "vio360_host_register", "vio360_host_unregister",
    "vio360_enable_kernel_timing", "vio360_kernel_time", "vio360_reset_kernel_times",
    "vio360_fe_configure", "vio360_fe_track", "vio360_fe_track_batch", "vio360_fe_detect",
    "vio360_fe_detect_batch", "vio360_fe_pyramid", "vio360_fe_min_eigen_val",
    "vio360_fe_ransac_rotation", "vio360_fe_ransac_rotation_batch", "vio360_fe_ransac_rotation_noalloc", "vio360_fe_cluster_filter", "vio360_fe_grid_cap",
    "vio360_seq_create", "vio360_seq_destroy", "vio360_seq_reset", "vio360_seq_capacity", "vio360_fe_track_frame_batch",
    "vio360_ba_create", "vio360_ba_destroy", "vio360_ba_set_structure", "vio360_ba_set_outliers",
    "vio360_ba_set_shard", "vio360_ba_set_parameters", "vio360_ba_evaluate", "vio360_ba_evaluate_f32",
    "vio360_ba_get_residuals_jacobians", "vio360_ba_normal_blocks", "vio360_ba_reduced_dim",
    "vio360_ba_build_reduced_system", "vio360_ba_set_lm_diagonal", "vio360_ba_reduced_system_dev", "vio360_ba_fetch_reduced_system",
    "vio360_ba_reduced_band_dev", "vio360_ba_expand_band",
    "vio360_ba_set_eval_precision", "vio360_lm_default_options", "vio360_ba_lm_solve", "vio360_ba_jacobian_right_multiply",
    "vio360_ba_peer_export", "vio360_ba_peer_attach", "vio360_ba_peer_exchange", "vio360_ba_peer_detach",
    "vio360_ba_column_sqnorm", "vio360_ba_chi2", "vio360_ba_solve_reduced", "vio360_pnp_evaluate", "vio360_triangulate",
]


class Vio360Error(RuntimeError):
    pass


class TrackJob(C.Structure):
    _fields_ = [("prev", C.c_void_p), ("prev_pitch", C.c_size_t), ("prev_key", C.c_uint64),
                ("cur", C.c_void_p), ("cur_pitch", C.c_size_t), ("cur_key", C.c_uint64),
                ("prev_xy", C.c_void_p), ("n", C.c_int),
                ("cur_xy", C.c_void_p), ("status", C.c_void_p), ("err", C.c_void_p), ("keep", C.c_void_p)]


class RansacJob(C.Structure):
    _fields_ = [("prev_xy", C.c_void_p), ("cur_xy", C.c_void_p), ("n", C.c_int), ("samples", C.c_void_p),
                ("inlier_mask", C.c_void_p), ("best_inliers", C.c_void_p), ("counts", C.c_void_p)]


class DetectJob(C.Structure):
    _fields_ = [("img", C.c_void_p), ("pitch", C.c_size_t), ("key", C.c_uint64),
                ("mask", C.c_void_p), ("mask_pitch", C.c_size_t),
                ("out_xy", C.c_void_p), ("out_n", C.c_void_p),
                ("occupied_xy", C.c_void_p), ("n_occupied", C.c_int), ("occupied_radius", C.c_int)]


class LmOptions(C.Structure):
    _fields_ = [("max_num_iterations", C.c_int), ("initial_trust_region_radius", C.c_double), ("max_trust_region_radius", C.c_double),
                ("min_trust_region_radius", C.c_double), ("min_relative_decrease", C.c_double), ("min_lm_diagonal", C.c_double),
                ("max_lm_diagonal", C.c_double), ("function_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
                ("parameter_tolerance", C.c_double), ("jacobi_scaling", C.c_int), ("max_num_consecutive_invalid_steps", C.c_int),
                ("apply_loss", C.c_int)]


class LmSummary(C.Structure):
    _fields_ = [("iterations", C.c_int), ("num_successful_steps", C.c_int), ("num_unsuccessful_steps", C.c_int), ("termination", C.c_int),
                ("initial_cost", C.c_double), ("final_cost", C.c_double), ("final_radius", C.c_double),
                ("final_gradient_max_norm", C.c_double), ("n_trace", C.c_int)]


class SeqConfig(C.Structure):
    _fields_ = [("max_features", C.c_int), ("min_distance", C.c_float), ("quality_level", C.c_float),
                ("grid_cols", C.c_int), ("grid_rows", C.c_int), ("max_features_per_grid", C.c_int),
                ("remove_clustered", C.c_int), ("clustered_std_ratio", C.c_float),
                ("ransac_iterations", C.c_int), ("ransac_threshold_deg", C.c_float)]


class FrameJob(C.Structure):
    _fields_ = [("seq", C.c_void_p), ("image", C.c_void_p), ("pitch", C.c_size_t), ("ransac_words", C.c_void_p),
                ("ids", C.c_void_p), ("xy", C.c_void_p), ("track_count", C.c_void_p),
                ("n_features", C.c_void_p), ("n_tracked", C.c_void_p), ("n_detected", C.c_void_p)]


_lib = None


def load():
    """dlopen the in-tree library; no fallback of any kind."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Vio360Error("%s is missing: run `make` (or __graft_entry__.build()) first; "
                          "there is no CPU fallback for the CUDA paths" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.vio360_last_error.restype = C.c_char_p
    lib.vio360_last_error.argtypes = [C.c_void_p]
    lib.vio360_stream.restype = C.c_void_p
    lib.vio360_stream.argtypes = [C.c_void_p]
    lib.vio360_launch_count.restype = C.c_longlong
    lib.vio360_launch_count.argtypes = [C.c_void_p]
    lib.vio360_destroy.restype = None
    lib.vio360_destroy.argtypes = [C.c_void_p]
    lib.vio360_ba_destroy.restype = None
    lib.vio360_ba_destroy.argtypes = [C.c_void_p]
    lib.vio360_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.vio360_synchronize.argtypes = [C.c_void_p]
    lib.vio360_probe_fma_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.vio360_host_register.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    lib.vio360_host_unregister.argtypes = [C.c_void_p, C.c_void_p]
    lib.vio360_enable_kernel_timing.argtypes = [C.c_void_p, C.c_int]
    lib.vio360_kernel_time.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
    lib.vio360_reset_kernel_times.argtypes = [C.c_void_p]
    lib.vio360_fe_configure.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double,
                                        C.c_double, C.c_float, C.c_int, C.c_int, C.c_int]
    lib.vio360_fe_track.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p, C.c_size_t,
                                    C.c_uint64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.POINTER(C.c_int)]
    lib.vio360_fe_track_batch.argtypes = [C.c_void_p, C.POINTER(TrackJob), C.c_int, C.c_int]
    lib.vio360_fe_detect.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint64, C.c_void_p, C.c_size_t,
                                     C.c_int, C.c_double, C.c_double, C.c_void_p, C.POINTER(C.c_int)]
    lib.vio360_fe_ransac_rotation.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float,
                                              C.c_void_p, C.c_void_p]
    lib.vio360_fe_ransac_rotation_batch.argtypes = [C.c_void_p, C.POINTER(RansacJob), C.c_int, C.c_int, C.c_float, C.c_int]
    lib.vio360_fe_ransac_rotation_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float,
                                                      C.c_void_p, C.c_void_p]
    lib.vio360_fe_detect_batch.argtypes = [C.c_void_p, C.POINTER(DetectJob), C.c_int, C.c_int, C.c_double,
                                           C.c_double, C.c_int]
    lib.vio360_fe_pyramid.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p),
                                      C.POINTER(C.c_size_t)]
    lib.vio360_fe_min_eigen_val.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    lib.vio360_fe_cluster_filter.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 6 + [C.c_float, C.c_void_p, C.c_int]
    lib.vio360_fe_grid_cap.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_int]
    lib.vio360_seq_create.argtypes = [C.c_void_p, C.POINTER(SeqConfig), C.POINTER(C.c_void_p)]
    lib.vio360_seq_destroy.restype = None
    lib.vio360_seq_destroy.argtypes = [C.c_void_p]
    lib.vio360_seq_reset.argtypes = [C.c_void_p]
    lib.vio360_seq_capacity.argtypes = [C.c_void_p]
    lib.vio360_fe_track_frame_batch.argtypes = [C.c_void_p, C.POINTER(FrameJob), C.c_int, C.c_int]
    lib.vio360_ba_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    lib.vio360_ba_set_structure.argtypes = [C.c_void_p] + [C.c_void_p] * 6 + [C.c_double, C.c_double, C.c_void_p,
                                                                               C.c_double, C.c_void_p, C.c_void_p]
    lib.vio360_ba_set_outliers.argtypes = [C.c_void_p, C.c_void_p]
    lib.vio360_ba_set_shard.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.vio360_ba_set_parameters.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.vio360_ba_evaluate.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.vio360_ba_evaluate_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double)]
    lib.vio360_ba_get_residuals_jacobians.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.vio360_ba_normal_blocks.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5
    lib.vio360_ba_reduced_dim.argtypes = [C.c_void_p]
    lib.vio360_ba_build_reduced_system.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                   C.POINTER(C.c_double)]
    lib.vio360_ba_set_lm_diagonal.argtypes = [C.c_void_p, C.c_void_p]
    lib.vio360_ba_reduced_system_dev.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    lib.vio360_ba_reduced_band_dev.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    lib.vio360_ba_expand_band.argtypes = [C.c_void_p]
    lib.vio360_ba_peer_export.argtypes = [C.c_void_p, C.c_void_p]
    lib.vio360_ba_peer_attach.argtypes = [C.c_void_p, C.c_void_p]
    lib.vio360_ba_peer_exchange.argtypes = [C.c_void_p]
    lib.vio360_ba_peer_detach.argtypes = [C.c_void_p]
    lib.vio360_ba_fetch_reduced_system.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
    lib.vio360_ba_column_sqnorm.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.vio360_ba_chi2.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p]
    lib.vio360_triangulate.argtypes = [C.c_void_p] * 5 + [C.c_int, C.c_void_p, C.c_void_p]
    lib.vio360_ba_solve_reduced.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.vio360_ba_set_eval_precision.argtypes = [C.c_void_p, C.c_int]
    lib.vio360_ba_jacobian_right_multiply.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    lib.vio360_lm_default_options.argtypes = [C.POINTER(LmOptions)]
    lib.vio360_ba_lm_solve.argtypes = [C.c_void_p, C.POINTER(LmOptions), C.c_void_p, C.c_void_p, C.POINTER(LmSummary), C.c_void_p, C.c_int]
    lib.vio360_pnp_evaluate.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_double, C.c_double, C.c_void_p,
                                                                                   C.c_double] + [C.c_void_p] * 8
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


class Context:
    """vio360_ctx: one device, one stream."""

    def __init__(self, device=0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.vio360_create(int(device), C.byref(h))
        if rc != 0:
            raise Vio360Error("vio360_create failed (%d): %s" % (rc, self.lib.vio360_last_error(None).decode()))
        self.h = h
        self.device = device
        self._children = weakref.WeakSet()    # BA windows of this context: destroyed before it, whatever the GC order

    def check(self, rc):
        if rc != 0:
            raise Vio360Error("vio360 error %d: %s" % (rc, self.lib.vio360_last_error(self.h).decode()))

    def close(self):
        if getattr(self, "h", None):
            for c in list(getattr(self, "_children", ())):
                c.close()
            self.lib.vio360_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def host_register(self, arr):
        """page-locks a numpy array's buffer in place (vio360_host_register); keep the array alive until host_unregister"""
        self.check(self.lib.vio360_host_register(self.h, arr.ctypes.data, arr.nbytes))

    def host_unregister(self, arr):
        self.check(self.lib.vio360_host_unregister(self.h, arr.ctypes.data))

    def synchronize(self):
        self.check(self.lib.vio360_synchronize(self.h))

    @property
    def stream(self):
        return self.lib.vio360_stream(self.h)

    @property
    def launch_count(self):
        return int(self.lib.vio360_launch_count(self.h))

    def enable_kernel_timing(self, on=True):
        self.check(self.lib.vio360_enable_kernel_timing(self.h, int(on)))

    def reset_kernel_times(self):
        self.check(self.lib.vio360_reset_kernel_times(self.h))

    def kernel_time(self, name):
        ms = C.c_double(0); n = C.c_longlong(0)
        self.check(self.lib.vio360_kernel_time(self.h, name.encode(), C.byref(ms), C.byref(n)))
        return ms.value, int(n.value)

    def probe_fma_peak(self):
        a, b = C.c_double(0), C.c_double(0)
        self.check(self.lib.vio360_probe_fma_peak(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # ---------------------------------------------------------------- front-end
    def fe_configure(self, width, height, max_level=3, win=21, max_iters=30,
                     eps=float(np.float32(0.01)), min_eig_thr=float(np.float32(0.01)),
                     polar_ratio=0.15, boundary_margin=20, max_points=1000, max_batch=1):
        self.fe_cfg = dict(W=width, H=height, max_points=max_points, max_batch=max_batch)
        self.check(self.lib.vio360_fe_configure(self.h, width, height, max_level, win, max_iters, eps, min_eig_thr,
                                                polar_ratio, boundary_margin, max_points, max_batch))

    def fe_track(self, prev, cur, pts, prev_key=0, cur_key=0, want_err=True):
        prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        n = pts.shape[0]
        out = np.zeros((n, 2), np.float32); st = np.zeros(n, np.uint8); err = np.zeros(n, np.float32)
        nt = C.c_int(0)
        self.check(self.lib.vio360_fe_track(self.h, _ptr(prev), prev.strides[0], prev_key, _ptr(cur), cur.strides[0],
                                            cur_key, _ptr(pts), n, _ptr(out), _ptr(st),
                                            _ptr(err) if want_err else None, C.byref(nt)))
        return out, st, err, nt.value

    def fe_track_batch_host(self, pairs, want_keep=True):
        """pairs: list of (prev, cur, pts, prev_key, cur_key) host arrays -> list of (xy, status, err, keep)."""
        jobs = (TrackJob * len(pairs))()
        keep_alive, outs = [], []
        for i, (prev, cur, pts, pk, ck) in enumerate(pairs):
            prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
            pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
            n = pts.shape[0]
            xy = np.zeros((n, 2), np.float32); st = np.zeros(n, np.uint8); err = np.zeros(n, np.float32)
            keep = np.zeros(n, np.uint8)
            keep_alive.append((prev, cur, pts))
            outs.append((xy, st, err, keep))
            j = jobs[i]
            j.prev, j.prev_pitch, j.prev_key = prev.ctypes.data, prev.strides[0], pk
            j.cur, j.cur_pitch, j.cur_key = cur.ctypes.data, cur.strides[0], ck
            j.prev_xy, j.n = pts.ctypes.data, n
            j.cur_xy, j.status, j.err = xy.ctypes.data, st.ctypes.data, err.ctypes.data
            j.keep = keep.ctypes.data if want_keep else None
        self.check(self.lib.vio360_fe_track_batch(self.h, jobs, len(pairs), MEM_HOST))
        return outs

    # ---- device-resident per-frame chain (FeatureTracker::TrackFeatures, FeatureTracker.cpp:61-206)
    def seq_create(self, max_features=1000, min_distance=30.0, quality=0.01, grid_cols=20, grid_rows=10, max_per_cell=10,
                   remove_clustered=True, std_ratio=0.25, ransac_iters=1000, ransac_thr_deg=2.0):
        cfg = SeqConfig(max_features, min_distance, quality, grid_cols, grid_rows, max_per_cell, int(remove_clustered),
                        std_ratio, ransac_iters, ransac_thr_deg)
        h = C.c_void_p()
        self.check(self.lib.vio360_seq_create(self.h, C.byref(cfg), C.byref(h)))
        return h

    def seq_destroy(self, seq):
        self.lib.vio360_seq_destroy(seq)

    def seq_reset(self, seq):
        self.check(self.lib.vio360_seq_reset(seq))

    def fe_track_frame_batch(self, seqs, images, words):
        """one frame for each sequence (host arrays) -> list of (ids, xy, track_count, n_tracked, n_detected)"""
        nj = len(seqs)
        jobs = (FrameJob * nj)()
        keep, outs = [], []
        for b in range(nj):
            cap = self.lib.vio360_seq_capacity(seqs[b])
            img = np.ascontiguousarray(images[b], np.uint8)
            w = np.ascontiguousarray(words[b], np.uint32)
            ids = np.zeros(cap, np.int32); xy = np.zeros((cap, 2), np.float32); tc = np.zeros(cap, np.int32)
            cnt = (C.c_int * 3)()
            keep.append((img, w, cnt)); outs.append((ids, xy, tc, cnt))
            j = jobs[b]
            j.seq, j.image, j.pitch, j.ransac_words = seqs[b], img.ctypes.data, img.strides[0], w.ctypes.data
            j.ids, j.xy, j.track_count = ids.ctypes.data, xy.ctypes.data, tc.ctypes.data
            j.n_features, j.n_tracked, j.n_detected = C.addressof(cnt), C.addressof(cnt) + 4, C.addressof(cnt) + 8
        self.check(self.lib.vio360_fe_track_frame_batch(self.h, jobs, nj, MEM_HOST))
        return [(ids[:cnt[0]].copy(), xy[:cnt[0]].copy(), tc[:cnt[0]].copy(), cnt[1], cnt[2]) for ids, xy, tc, cnt in outs]

    def fe_track_frame_batch_raw(self, jobs, n_jobs, mem_space):
        self.check(self.lib.vio360_fe_track_frame_batch(self.h, jobs, n_jobs, mem_space))

    def fe_track_batch_raw(self, jobs, n_jobs, mem_space):
        self.check(self.lib.vio360_fe_track_batch(self.h, jobs, n_jobs, mem_space))

    def _grid_pack(self, frames, track=None):
        nf = len(frames)
        ns = np.array([len(f) for f in frames], np.int32)
        stride = max(int(ns.max()) if nf else 0, 1)
        xy = np.zeros((nf, stride, 2), np.float32); tr = np.zeros((nf, stride), np.int32)
        for f in range(nf):
            xy[f, :ns[f]] = np.asarray(frames[f], np.float32).reshape(-1, 2)
            if track is not None:
                tr[f, :ns[f]] = np.asarray(track[f], np.int32)
        return nf, ns, stride, xy, tr, np.zeros((nf, stride), np.uint8)

    def fe_cluster_filter(self, frames, width, height, cols=20, rows=10, std_ratio=0.25):
        """FeatureTracker::RemoveClusteredFeatures over a list of per-frame (n, 2) point arrays -> list of keep masks."""
        nf, ns, stride, xy, _, keep = self._grid_pack(frames)
        self.check(self.lib.vio360_fe_cluster_filter(self.h, _ptr(xy), _ptr(ns), nf, stride, width, height, cols, rows,
                                                     std_ratio, _ptr(keep), MEM_HOST))
        return [keep[f, :ns[f]].copy() for f in range(nf)]

    def fe_grid_cap(self, frames, track, width, height, cols=20, rows=10, max_per_cell=4):
        """Frame::LimitFeaturesPerGrid over a list of per-frame point arrays and track counts -> list of keep masks."""
        nf, ns, stride, xy, tr, keep = self._grid_pack(frames, track)
        self.check(self.lib.vio360_fe_grid_cap(self.h, _ptr(xy), _ptr(tr), _ptr(ns), nf, stride, width, height, cols, rows,
                                               max_per_cell, _ptr(keep), MEM_HOST))
        return [keep[f, :ns[f]].copy() for f in range(nf)]

    def fe_detect(self, img, mask=None, max_corners=1000, quality=float(np.float32(0.01)),
                  min_dist=float(np.float32(30.0)), key=0):
        img = np.ascontiguousarray(img, np.uint8)
        if mask is not None:
            mask = np.ascontiguousarray(mask, np.uint8)
        cap = max_corners if max_corners > 0 else 1 << 18
        out = np.zeros((cap, 2), np.float32)
        n = C.c_int(0)
        self.check(self.lib.vio360_fe_detect(self.h, _ptr(img), img.strides[0], key, _ptr(mask),
                                             mask.strides[0] if mask is not None else 0, max_corners, quality,
                                             min_dist, _ptr(out), C.byref(n)))
        return out[:n.value].copy()

    def fe_detect_batch(self, imgs, masks=None, max_corners=1000, quality=float(np.float32(0.01)),
                        min_dist=float(np.float32(30.0)), out_cap=None, occupied=None, occupied_radius=30):
        """One vio360_fe_detect_batch call over independent images (host memory); list of (n, 2) arrays."""
        nb = len(imgs)
        imgs = [np.ascontiguousarray(i, np.uint8) for i in imgs]
        masks = [None if m is None else np.ascontiguousarray(m, np.uint8) for m in (masks or [None] * nb)]
        cap = max_corners if max_corners > 0 else (out_cap or imgs[0].size)
        outs = [np.zeros((cap, 2), np.float32) for _ in range(nb)]
        ns = np.zeros(nb, np.int32)
        jobs = (DetectJob * nb)()
        for b in range(nb):
            j = jobs[b]
            j.img, j.pitch, j.key = imgs[b].ctypes.data, imgs[b].strides[0], 0
            j.mask = masks[b].ctypes.data if masks[b] is not None else None
            j.mask_pitch = masks[b].strides[0] if masks[b] is not None else 0
            j.out_xy, j.out_n = outs[b].ctypes.data, ns.ctypes.data + 4 * b
            if occupied is not None and occupied[b] is not None and len(occupied[b]):
                occ = np.ascontiguousarray(occupied[b], np.float32).reshape(-1, 2)
                imgs.append(occ)   # keep alive
                j.occupied_xy, j.n_occupied, j.occupied_radius = occ.ctypes.data, occ.shape[0], occupied_radius
        self.check(self.lib.vio360_fe_detect_batch(self.h, jobs, nb, max_corners, quality, min_dist, MEM_HOST))
        return [outs[b][:ns[b]].copy() for b in range(nb)]

    def triangulate(self, b1, b2, T1w, T2w):
        """Estimator::TriangulateSinglePoint for all matches of a keyframe pair: (points n x 3, valid n)."""
        b1 = np.ascontiguousarray(b1, np.float32).reshape(-1, 3); b2 = np.ascontiguousarray(b2, np.float32).reshape(-1, 3)
        T1 = np.ascontiguousarray(T1w, np.float32).reshape(16); T2 = np.ascontiguousarray(T2w, np.float32).reshape(16)
        n = b1.shape[0]
        pts = np.zeros((n, 3), np.float32); valid = np.zeros(n, np.uint8)
        self.check(self.lib.vio360_triangulate(self.h, _ptr(b1), _ptr(b2), _ptr(T1), _ptr(T2), n, _ptr(pts), _ptr(valid)))
        return pts, valid

    def fe_ransac_rotation(self, prev_xy, cur_xy, samples, threshold_deg):
        """FeatureTracker::RejectOutliersRotationRANSAC with an injected (iters, 3) sample list.
        returns (best_inliers, mask, counts)."""
        return self.fe_ransac_rotation_batch([(prev_xy, cur_xy, samples)], threshold_deg)[0]

    def fe_ransac_rotation_batch(self, items, threshold_deg):
        nb = len(items)
        jobs = (RansacJob * nb)()
        keep, outs = [], []
        iters = None
        for b, (prev_xy, cur_xy, samples) in enumerate(items):
            prev_xy = np.ascontiguousarray(prev_xy, np.float32).reshape(-1, 2)
            cur_xy = np.ascontiguousarray(cur_xy, np.float32).reshape(-1, 2)
            samples = np.ascontiguousarray(samples, np.int32).reshape(-1, 3)
            iters = samples.shape[0] if iters is None else iters
            assert samples.shape[0] == iters
            n = prev_xy.shape[0]
            mask = np.zeros(n, np.uint8); best = np.zeros(1, np.int32); counts = np.zeros(max(iters, 1), np.int32)
            keep.append((prev_xy, cur_xy, samples))
            outs.append((best, mask, counts))
            j = jobs[b]
            j.prev_xy, j.cur_xy, j.n, j.samples = _ptr(prev_xy), _ptr(cur_xy), n, _ptr(samples)
            j.inlier_mask, j.best_inliers, j.counts = _ptr(mask), best.ctypes.data, counts.ctypes.data
        self.check(self.lib.vio360_fe_ransac_rotation_batch(self.h, jobs, nb, iters or 0, C.c_float(threshold_deg), MEM_HOST))
        return [(int(b[0]), m, c[:iters]) for b, m, c in outs]

    def fe_ransac_rotation_noalloc(self, prev_xy_dev, cur_xy_dev, n, samples_dev, iters, threshold_deg, mask_dev, best_dev):
        """vio360_fe_ransac_rotation_noalloc: raw device addresses in, kernels enqueued on the context's stream,
        nothing allocated and nothing synchronised (call synchronize() before reading the outputs)."""
        self.check(self.lib.vio360_fe_ransac_rotation_noalloc(self.h, prev_xy_dev, cur_xy_dev, n, samples_dev, iters,
                                                              C.c_float(threshold_deg), mask_dev, best_dev))

    def fe_detect_batch_raw(self, jobs, n_jobs, max_corners, quality, min_dist, mem_space):
        self.check(self.lib.vio360_fe_detect_batch(self.h, jobs, n_jobs, max_corners, quality, min_dist, mem_space))

    def fe_pyramid(self, img, n_levels):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        outs = []
        for _ in range(n_levels):
            outs.append(np.zeros((h, w), np.uint8))
            w, h = (w + 1) // 2, (h + 1) // 2
        ptrs = (C.c_void_p * n_levels)(*[o.ctypes.data for o in outs])
        pitches = (C.c_size_t * n_levels)(*[o.strides[0] for o in outs])
        self.check(self.lib.vio360_fe_pyramid(self.h, _ptr(img), img.strides[0], n_levels, ptrs, pitches))
        return outs

    def fe_min_eigen_val(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        out = np.zeros(img.shape, np.float32)
        self.check(self.lib.vio360_fe_min_eigen_val(self.h, _ptr(img), img.strides[0], _ptr(out)))
        return out

    # ---------------------------------------------------------------- PnP
    def pnp_evaluate(self, obs_uv, world_pts, Twb_init, Tcb, cols, rows, info, huber_delta, pose_delta,
                     outlier=None, want_jac=True, want_chi2=True):
        obs_uv = np.ascontiguousarray(obs_uv, np.float64); world_pts = np.ascontiguousarray(world_pts, np.float64)
        N = obs_uv.shape[0]
        Twb_init = np.ascontiguousarray(Twb_init, np.float64); Tcb = np.ascontiguousarray(Tcb, np.float64)
        info = np.ascontiguousarray(info, np.float64); pose_delta = np.ascontiguousarray(pose_delta, np.float64)
        if outlier is not None:
            outlier = np.ascontiguousarray(outlier, np.uint8)
        r = np.zeros((N, 2)); Jp = np.zeros((N, 2, 6)) if want_jac else None
        valid = np.zeros(N, np.uint8); H = np.zeros((6, 6)); b = np.zeros(6); cost = np.zeros(1)
        chi2 = np.zeros(N) if want_chi2 else None
        self.check(self.lib.vio360_pnp_evaluate(self.h, N, _ptr(obs_uv), _ptr(world_pts), _ptr(outlier), _ptr(Twb_init),
                                                _ptr(Tcb), float(cols), float(rows), _ptr(info), float(huber_delta),
                                                _ptr(pose_delta), _ptr(r), _ptr(Jp), _ptr(valid), _ptr(H), _ptr(b),
                                                _ptr(cost), _ptr(chi2)))
        return dict(r=r, Jp=Jp, valid=valid, H=H, b=b, cost=float(cost[0]), chi2=chi2)


class BAWindow:
    """vio360_ba: the device twin of one Optimizer::RunLocalBA problem."""

    def __init__(self, ctx, win):
        self.ctx = ctx
        self.lib = ctx.lib
        self.K, self.M, self.O = int(win["K"]), int(win["M"]), int(win["O"])
        h = C.c_void_p()
        ctx.check(self.lib.vio360_ba_create(ctx.h, self.K, self.M, self.O, C.byref(h)))
        self.h = h
        ctx._children.add(self)
        a = dict(pose_idx=np.ascontiguousarray(win["pose_idx"], np.int32),
                 point_idx=np.ascontiguousarray(win["point_idx"], np.int32),
                 u=np.ascontiguousarray(win["obs_u"], np.float64), v=np.ascontiguousarray(win["obs_v"], np.float64),
                 Twb=np.ascontiguousarray(win["Twb_init"], np.float64), Tcb=np.ascontiguousarray(win["Tcb"], np.float64),
                 info=np.ascontiguousarray(win["info"], np.float64),
                 pf=np.ascontiguousarray(win["pose_fixed"], np.uint8), lf=np.ascontiguousarray(win["point_fixed"], np.uint8))
        ctx.check(self.lib.vio360_ba_set_structure(self.h, _ptr(a["pose_idx"]), _ptr(a["point_idx"]), _ptr(a["u"]),
                                                   _ptr(a["v"]), _ptr(a["Twb"]), _ptr(a["Tcb"]), float(win["cols"]),
                                                   float(win["rows"]), _ptr(a["info"]), float(win["huber_a"]),
                                                   _ptr(a["pf"]), _ptr(a["lf"])))
        if "outlier" in win and win["outlier"] is not None:
            self.set_outliers(win["outlier"])

    def close(self):
        if getattr(self, "h", None):
            if getattr(self.ctx, "h", None):      # the native window points into its context
                self.lib.vio360_ba_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_outliers(self, outlier):
        o = None if outlier is None else np.ascontiguousarray(outlier, np.uint8)
        self.ctx.check(self.lib.vio360_ba_set_outliers(self.h, _ptr(o)))

    def set_shard(self, rank, world):
        self.ctx.check(self.lib.vio360_ba_set_shard(self.h, rank, world))

    def set_parameters(self, pose_delta, points):
        pd = np.ascontiguousarray(pose_delta, np.float64); pts = np.ascontiguousarray(points, np.float64)
        self.ctx.check(self.lib.vio360_ba_set_parameters(self.h, _ptr(pd), _ptr(pts)))

    def evaluate(self, want_jacobians=True, apply_loss=True, fetch=True, f32=False):
        cost = C.c_double(0)
        fn = self.lib.vio360_ba_evaluate_f32 if f32 else self.lib.vio360_ba_evaluate
        self.ctx.check(fn(self.h, int(want_jacobians), int(apply_loss), C.byref(cost)))
        if not fetch:
            return cost.value
        r = np.zeros((self.O, 2)); Jp = np.zeros((self.O, 2, 6)) if want_jacobians else None
        Jl = np.zeros((self.O, 2, 3)) if want_jacobians else None
        self.ctx.check(self.lib.vio360_ba_get_residuals_jacobians(self.h, _ptr(r), _ptr(Jp), _ptr(Jl)))
        return r, Jp, Jl, cost.value

    def normal_blocks(self, apply_loss=True):
        Hpp = np.zeros((self.K, 6, 6)); bp = np.zeros((self.K, 6)); Hll = np.zeros((self.M, 3, 3))
        bl = np.zeros((self.M, 3)); Hpl = np.zeros((self.O, 6, 3))
        self.ctx.check(self.lib.vio360_ba_normal_blocks(self.h, int(apply_loss), _ptr(Hpp), _ptr(bp), _ptr(Hll),
                                                        _ptr(bl), _ptr(Hpl)))
        return Hpp, bp, Hll, bl, Hpl

    @property
    def reduced_dim(self):
        return int(self.lib.vio360_ba_reduced_dim(self.h))

    def build_reduced_system(self, lm_diag=None, apply_loss=True, fetch=True):
        n = self.reduced_dim
        D = None if lm_diag is None else np.ascontiguousarray(lm_diag, np.float64)
        if not fetch:
            self.ctx.check(self.lib.vio360_ba_build_reduced_system(self.h, _ptr(D), int(apply_loss), None, None, None))
            return None
        S = np.zeros((n, n)); rhs = np.zeros(n); cost = C.c_double(0)
        self.ctx.check(self.lib.vio360_ba_build_reduced_system(self.h, _ptr(D), int(apply_loss), _ptr(S), _ptr(rhs),
                                                               C.byref(cost)))
        return S, rhs, cost.value

    def set_lm_diagonal(self, lm_diag):
        D = None if lm_diag is None else np.ascontiguousarray(lm_diag, np.float64)
        self.ctx.check(self.lib.vio360_ba_set_lm_diagonal(self.h, _ptr(D)))

    def reduced_system_dev(self):
        p = C.c_void_p(); n = C.c_size_t(0)
        self.ctx.check(self.lib.vio360_ba_reduced_system_dev(self.h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def reduced_band_dev(self):
        """(device pointer, doubles) of the block-band exchange buffer of a sharded build, or None when the
        last build has none (not sharded / landmarks outside the chunk windows)."""
        p = C.c_void_p(); n = C.c_size_t(0)
        if self.lib.vio360_ba_reduced_band_dev(self.h, C.byref(p), C.byref(n)) != 0:
            return None
        return p.value, n.value

    def expand_band(self):
        self.ctx.check(self.lib.vio360_ba_expand_band(self.h))

    def peer_export(self):
        """This rank's exchange-buffer descriptor (bytes) for vio360_ba_peer_attach on every rank."""
        blob = C.create_string_buffer(PEER_BLOB_BYTES)
        self.ctx.check(self.lib.vio360_ba_peer_export(self.h, blob))
        return blob.raw

    def peer_attach(self, blobs):
        """blobs: the world's descriptors in rank order (list of bytes)."""
        raw = b"".join(blobs)
        assert len(raw) % PEER_BLOB_BYTES == 0
        self.ctx.check(self.lib.vio360_ba_peer_attach(self.h, C.c_char_p(raw)))

    def peer_exchange(self):
        self.ctx.check(self.lib.vio360_ba_peer_exchange(self.h))

    def peer_detach(self):
        self.ctx.check(self.lib.vio360_ba_peer_detach(self.h))

    def fetch_reduced_system(self):
        n = self.reduced_dim
        S = np.zeros((n, n)); rhs = np.zeros(n); cost = C.c_double(0)
        self.ctx.check(self.lib.vio360_ba_fetch_reduced_system(self.h, _ptr(S), _ptr(rhs), C.byref(cost)))
        return S, rhs, cost.value

    def solve_reduced(self, want_points=True):
        """x of S x = rhs (K x 6, zeros for constant poses) and the back-substituted landmark step (M x 3)."""
        dp = np.zeros((self.K, 6)); dl = np.zeros((self.M, 3)) if want_points else None
        self.ctx.check(self.lib.vio360_ba_solve_reduced(self.h, _ptr(dp), _ptr(dl)))
        return dp, dl

    def set_eval_precision(self, fp32):
        self.ctx.check(self.lib.vio360_ba_set_eval_precision(self.h, int(bool(fp32))))

    def jacobian_right_multiply(self, x, apply_loss=True):
        x = np.ascontiguousarray(x, np.float64).reshape(-1)
        y = np.zeros((self.O, 2))
        self.ctx.check(self.lib.vio360_ba_jacobian_right_multiply(self.h, int(apply_loss), _ptr(x), _ptr(y)))
        return y

    def lm_solve(self, pose_delta, points, **overrides):
        """vio360_ba_lm_solve: returns (pose_delta K x 6, points M x 3, summary dict, trace n x 8)."""
        opt = LmOptions()
        self.ctx.check(self.lib.vio360_lm_default_options(C.byref(opt)))
        for k, v in overrides.items():
            setattr(opt, k, v)
        pd = np.array(pose_delta, np.float64).reshape(self.K, 6).copy(); pts = np.array(points, np.float64).reshape(self.M, 3).copy()
        cap = opt.max_num_iterations + 2
        trace = np.zeros((cap, 8)); sm = LmSummary()
        self.ctx.check(self.lib.vio360_ba_lm_solve(self.h, C.byref(opt), _ptr(pd), _ptr(pts), C.byref(sm), _ptr(trace), cap))
        summary = {f: getattr(sm, f) for f, _ in LmSummary._fields_}
        return pd, pts, summary, trace[:sm.n_trace].copy()

    def column_sqnorm(self, apply_loss=True):
        out = np.zeros(6 * self.K + 3 * self.M)
        self.ctx.check(self.lib.vio360_ba_column_sqnorm(self.h, int(apply_loss), _ptr(out)))
        return out

    def chi2(self, thr=5.991):
        c = np.zeros(self.O); f = np.zeros(self.O, np.uint8)
        self.ctx.check(self.lib.vio360_ba_chi2(self.h, float(thr), _ptr(c), _ptr(f)))
        return c, f
