"""tools/chain_perf.py -- per-kernel split of vio360_fe_track_frame_batch on device-resident 1920x960 frames (developer
probe).  usage: python tools/chain_perf.py [sequences]"""
import ctypes as C
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "360_visual_inertial_odometry_b200"
NAMES = ("k_pyr_down", "k_pyr_down8", "k_lk_track", "k_seq_post_lk", "k_ransac_bearings", "k_ransac_hypotheses", "k_ransac_pick",
         "k_seq_post_ransac", "k_occupancy_discs", "k_min_eig", "k_candidates", "k_sort_select", "k_seq_append", "k_cluster_filter", "k_grid_cap")


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    Ws, Hs, T = 1920, 960, 10
    b = importlib.import_module(PKG + ".bindings")
    syn = importlib.import_module(PKG + ".synthetic")
    ctx = b.Context(0)
    ctx.fe_configure(Ws, Hs, max_points=2000, max_batch=S)
    dev = torch.device("cuda:0")
    frames = [[syn.render_erp_frame(Ws, Hs, t, seed=360 + q, n_waves=24, sigma_omega=60.0, xp=torch, device=dev) for t in range(T)] for q in range(S)]
    rng = np.random.default_rng(42)
    words = torch.from_numpy(rng.integers(0, 2 ** 32, (T, S, 1000, 3), dtype=np.uint64).astype(np.uint32).view(np.int32)).to(dev)
    cfg = dict(max_features=1000, min_distance=30.0, quality=0.01, grid_cols=20, grid_rows=10, max_per_cell=10,
               remove_clustered=True, std_ratio=0.25, ransac_iters=1000, ransac_thr_deg=2.0)
    cap = 2000
    d_ids = torch.zeros((S, cap), dtype=torch.int32, device=dev); d_xy = torch.zeros((S, cap, 2), dtype=torch.float32, device=dev)
    d_tc = torch.zeros((S, cap), dtype=torch.int32, device=dev)
    cnt = (C.c_int * (3 * S))()
    for timing in (False, True):
        seqs = [ctx.seq_create(**cfg) for _ in range(S)]
        jobs = (b.FrameJob * S)()

        def step(t):
            for q in range(S):
                j = jobs[q]
                j.seq = seqs[q]; j.pitch = Ws
                j.image, j.ransac_words = frames[q][t].data_ptr(), words[t, q].data_ptr()
                j.ids, j.xy, j.track_count = d_ids[q].data_ptr(), d_xy[q].data_ptr(), d_tc[q].data_ptr()
                j.n_features, j.n_tracked, j.n_detected = C.addressof(cnt) + 12 * q, C.addressof(cnt) + 12 * q + 4, C.addressof(cnt) + 12 * q + 8
            ctx.fe_track_frame_batch_raw(jobs, S, b.MEM_DEVICE)

        step(0); step(1)
        ctx.enable_kernel_timing(timing); ctx.reset_kernel_times()
        t0 = time.perf_counter()
        for t in range(2, T):
            step(t)
        dt = (time.perf_counter() - t0) / (T - 2)
        out = {"sequences": S, "kernel_events": timing, "ms_per_call": dt * 1e3, "detected_last": [cnt[3 * q + 2] for q in range(min(S, 4))]}
        if timing:
            tot = 0.0
            for name in NAMES:
                ms, n = ctx.kernel_time(name)
                if n:
                    out[name] = round(ms / (T - 2), 4)
                    tot += ms / (T - 2)
            out["kernel_ms_per_call"] = round(tot, 4)
        print(json.dumps(out))
        ctx.enable_kernel_timing(False)
        for h in seqs:
            ctx.seq_destroy(h)
    ctx.close()


if __name__ == "__main__":
    main()
