"""tools/det_perf.py -- per-kernel timing of the detector on a batch of device-resident synthetic frames (developer
probe; the judged numbers come from bench.py).  usage: python tools/det_perf.py [frames] [W] [H]"""
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


def main():
    nb = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
    H = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
    b = importlib.import_module(PKG + ".bindings")
    syn = importlib.import_module(PKG + ".synthetic")
    ctx = b.Context(0)
    ctx.fe_configure(W, H, max_points=8, max_batch=nb)
    base = [torch.from_numpy(syn.render_erp_frame(W, H, t, seed=360, n_waves=48, sigma_omega=40.0 * W / 960)).cuda() for t in range(2)]
    frames = [base[s & 1].roll(7 * s, 1).contiguous() for s in range(nb)]
    out_xy = torch.zeros((nb, 1000, 2), dtype=torch.float32, device="cuda")
    out_n = (C.c_int * nb)()

    def step():
        jobs = (b.DetectJob * nb)()
        for s in range(nb):
            j = jobs[s]
            j.img, j.pitch, j.key = frames[s].data_ptr(), W, 0
            j.mask, j.mask_pitch = None, 0
            j.out_xy, j.out_n = out_xy[s].data_ptr(), C.addressof(out_n) + 4 * s
        ctx.fe_detect_batch_raw(jobs, nb, 1000, float(np.float32(0.01)), float(np.float32(30.0)), b.MEM_DEVICE)

    for _ in range(2):
        step()
    t0 = time.perf_counter()
    for _ in range(5):
        step()
    dt = (time.perf_counter() - t0) / 5
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    for _ in range(3):
        step()
    out = {"frames": nb, "W": W, "H": H, "ms_per_step": dt * 1e3, "frames_per_s": nb / dt, "corners0": int(out_n[0])}
    for name in ("k_min_eig", "k_candidates", "k_sort_select"):
        ms, n = ctx.kernel_time(name)
        out[name] = ms / max(n, 1)
    print(json.dumps(out))
    ctx.close()


if __name__ == "__main__":
    main()
