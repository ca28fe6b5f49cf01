import sys, importlib, numpy as np
sys.path.insert(0, "/root/repo")
PKG = "360_visual_inertial_odometry_b200"
b = importlib.import_module(PKG + ".bindings"); syn = importlib.import_module(PKG + ".synthetic")
ctx = b.Context(0)
win = syn.make_ba_window(K=200, M=5000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
ba = b.BAWindow(ctx, win); ba.set_parameters(win["pose_delta"], win["points"])
colsq = ba.column_sqnorm(True); D = np.sqrt(np.clip(colsq, 1e-6, 1e32) / 1e4)
ba.build_reduced_system(D, True, fetch=False)
for _ in range(2): ba.solve_reduced(False)
ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
for _ in range(5): ba.solve_reduced(False)
print("k_chol_band ms", ctx.kernel_time("k_chol_band"), "back", ctx.kernel_time("k_chol_band_back"))
ba.close(); ctx.close()
