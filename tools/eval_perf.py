"""tools/eval_perf.py -- k_evaluate (fp64) against k_evaluate_f32 on the 1M-observation window (developer probe)."""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "360_visual_inertial_odometry_b200"


def main():
    b = importlib.import_module(PKG + ".bindings")
    syn = importlib.import_module(PKG + ".synthetic")
    ctx = b.Context(0)
    win = syn.make_ba_window(K=200, M=50000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
    ba = b.BAWindow(ctx, win)
    ba.set_parameters(win["pose_delta"], win["points"])
    out = {"O": int(win["O"])}
    for f32 in (False, True):
        for _ in range(2):
            ba.evaluate(True, True, fetch=False, f32=f32)
        ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
        for _ in range(10):
            ba.evaluate(True, True, fetch=False, f32=f32)
        name = "k_evaluate_f32" if f32 else "k_evaluate"
        ms, n = ctx.kernel_time(name)
        ctx.enable_kernel_timing(False)
        out[name + "_ms"] = ms / max(n, 1)
        out[name + "_gflops"] = 235.0 * win["O"] / (ms / max(n, 1) * 1e-3) / 1e9
        out[name + "_out_GBps"] = 160.0 * win["O"] / (ms / max(n, 1) * 1e-3) / 1e9
    print(json.dumps(out))
    ba.close(); ctx.close()


if __name__ == "__main__":
    main()
