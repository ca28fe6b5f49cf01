"""tools/fe_e2e_probe.py -- host-space vio360_fe_track_batch (64 frame pairs, 3840x1920): wall time per call and the
kernels' own times, for the current VIO360_FE_H2D_GROUP (developer probe)."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    b = importlib.import_module(bench.PKG + ".bindings")
    syn = importlib.import_module(bench.PKG + ".synthetic")
    W, H, B, N = bench.W, bench.H, 64, bench.NFEAT
    ctx = b.Context(0)
    ctx.fe_configure(W, H, max_points=N, max_batch=B)
    base = [syn.render_erp_frame(W, H, t, seed=360, n_waves=8, sigma_omega=120.0, xp=torch, device="cuda").cpu() for t in range(2)]
    hring = [[base[t].clone().pin_memory() for _ in range(B)] for t in range(2)]
    pts = torch.from_numpy(np.stack([syn.make_track_points(W, H, N, seed=1000 + s) for s in range(B)])).pin_memory()
    hxy = torch.zeros((B, N, 2), dtype=torch.float32).pin_memory()
    hst = torch.zeros((B, N), dtype=torch.uint8).pin_memory()
    herr = torch.zeros((B, N), dtype=torch.float32).pin_memory()
    tabs = []
    for par in range(2):
        jobs = (b.TrackJob * B)()
        for s in range(B):
            j = jobs[s]
            j.prev, j.prev_pitch, j.cur, j.cur_pitch = hring[par][s].data_ptr(), W, hring[1 - par][s].data_ptr(), W
            j.prev_xy, j.n = pts[s].data_ptr(), N
            j.cur_xy, j.status, j.err, j.keep = hxy[s].data_ptr(), hst[s].data_ptr(), herr[s].data_ptr(), None
        tabs.append(jobs)

    def step(i):
        jobs = tabs[i % 2]
        for s in range(B):
            jobs[s].prev_key = (1 << 40) + i * B + s
            jobs[s].cur_key = (1 << 40) + (i + 1) * B + s
        ctx.fe_track_batch_raw(jobs, B, b.MEM_HOST)
    step(0); step(1)
    t0 = time.perf_counter()
    for i in range(2, 8):
        step(i)
    dt = (time.perf_counter() - t0) / 6
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    step(8); step(9)
    k = {n: (round(ctx.kernel_time(n)[0] / 2, 3), ctx.kernel_time(n)[1] // 2) for n in ("k_lk_track", "k_pyr_down")}
    print("group %s: %.2f ms per call = %.0f frames/s; kernel ms per call (launches): %s" % (os.environ.get("VIO360_FE_H2D_GROUP", "default"), dt * 1e3, B / dt, k))
    ctx.close()


if __name__ == "__main__":
    main()
