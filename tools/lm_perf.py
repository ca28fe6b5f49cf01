"""tools/lm_perf.py -- per-kernel split of vio360_ba_lm_solve (developer probe).  usage: python tools/lm_perf.py [c3|c5]"""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "360_visual_inertial_odometry_b200"
NAMES = ("k_lm_cost", "k_blocks_pose", "k_blocks_landmark", "k_lm_linearized", "k_lm_diagonal", "k_wide_prepare", "k_schur_tile",
         "k_schur_reduce", "k_chol_band", "k_chol_band_back", "k_back_substitute", "k_lm_candidate", "k_lm_model_change",
         "k_lm_decide", "k_expand_band", "k_chol_skyline", "k_finish_S", "k_sum_partials")


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c5"
    b = importlib.import_module(PKG + ".bindings")
    syn = importlib.import_module(PKG + ".synthetic")
    ctx = b.Context(0)
    if which == "c3":
        win = syn.make_ba_window(K=10, M=2000, seed=20000, mid_iteration=True)
    else:
        win = syn.make_ba_window(K=200, M=50000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
    ba = b.BAWindow(ctx, win)
    for timing in (False, True):
        ctx.enable_kernel_timing(timing); ctx.reset_kernel_times()
        ba.lm_solve(win["pose_delta"], win["points"])
        t0 = time.perf_counter()
        _, _, sm, _ = ba.lm_solve(win["pose_delta"], win["points"])
        ms = (time.perf_counter() - t0) * 1e3
        out = {"window": which, "kernel_events": timing, "ms": ms, "iterations": sm["iterations"]}
        if timing:
            tot = 0.0
            for name in NAMES:
                t, n = ctx.kernel_time(name)
                if n:
                    out[name] = {"ms_per_launch": t / n, "launches_per_solve": n / 2, "ms_per_solve": t / 2}
                    tot += t / 2
            out["kernel_ms_per_solve"] = tot
        print(json.dumps(out))
    ba.close(); ctx.close()


if __name__ == "__main__":
    main()
