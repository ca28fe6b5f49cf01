"""tools/ba_perf.py -- per-kernel timing of the BA reduced-system build (developer probe; the
judged numbers come from bench.py).  usage: python tools/ba_perf.py [c3|c5|dense] [steps]"""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "360_visual_inertial_odometry_b200"


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c5"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    b = importlib.import_module(PKG + ".bindings")
    syn = importlib.import_module(PKG + ".synthetic")
    ctx = b.Context(0)
    if which == "c3":
        win = syn.make_ba_window(K=10, M=2000, seed=20000, mid_iteration=True)
    elif which == "dense":
        win = syn.make_ba_window(K=200, M=20000, degree=50, seed=1000001, mid_iteration=True, loop_radius=6.0)
    else:
        win = syn.make_ba_window(K=200, M=50000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
    ba = b.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    colsq = ba.column_sqnorm(True)
    D = np.sqrt(np.clip(colsq, 1e-6, 1e32) / 1e4)
    ba.set_lm_diagonal(D)
    for _ in range(2):
        ba.build_reduced_system(None, True, fetch=False)
    ctx.synchronize()
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    for _ in range(steps):
        ba.build_reduced_system(None, True, fetch=False)
    ctx.synchronize()
    out = {"window": which, "O": int(win["O"]), "steps": steps}
    for name in ("k_schur_tile", "k_schur_reduce", "k_schur_warp", "k_finish_S", "k_sum_partials"):
        ms, n = ctx.kernel_time(name)
        if n:
            out[name + "_ms"] = ms / n
    deg = np.bincount(win["point_idx"], minlength=win["M"]).astype(np.float64)
    out["flop"] = 451.0 * win["O"] + float(np.sum(50 + 130 * deg + 216 * deg * (deg + 1) / 2))
    if "k_schur_tile_ms" in out:
        out["tile_tflops"] = out["flop"] / (out["k_schur_tile_ms"] * 1e-3) / 1e12
    print(json.dumps(out))
    ba.close(); ctx.close()


if __name__ == "__main__":
    main()
