"""NVLink bytes of the reduced-camera-system exchange: ONE process drives two GPUs (plain peer access),
so that ncu can profile k_band_pull without a multi-rank launch.  Run with VIO360_PEER_PUBLISH=2 (the reduce publishes,
the pull only reads flags): under ncu kernels are serialised, and the default protocol (each rank's pull announces its
band) would wait for a kernel that has not been launched yet.
  VIO360_PEER_PUBLISH=2 ncu --metrics nvlrx__bytes.sum,nvltx__bytes.sum -k regex:k_band_pull python tools/nvlink_probe.py"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "360_visual_inertial_odometry_b200"
b = importlib.import_module(PKG + ".bindings")
syn = importlib.import_module(PKG + ".synthetic")


def main():
    win = syn.make_ba_window(K=200, M=50000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
    ctxs = [b.Context(0), b.Context(1)]
    bas = []
    for rank in range(2):
        ba = b.BAWindow(ctxs[rank], win)
        ba.set_parameters(win["pose_delta"], win["points"])
        if rank == 0:
            colsq = ba.column_sqnorm(True)
            D = np.sqrt(np.clip(colsq, 1e-6, 1e32) / 1e4)
        ba.set_shard(rank, 2)
        ba.set_parameters(win["pose_delta"], win["points"])
        ba.set_lm_diagonal(D)
        bas.append(ba)
    blobs = [ba.peer_export() for ba in bas]
    for ba in bas:
        ba.peer_attach(blobs)
    for _ in range(3):
        for ba in bas:
            ba.build_reduced_system(None, True, fetch=False)
        for ba in bas:
            ba.peer_exchange()
    outs = [ba.fetch_reduced_system() for ba in bas]
    print("exchange ok: ranks identical", bool(np.array_equal(outs[0][0], outs[1][0])), "n", outs[0][0].shape[0])
    for c in ctxs:
        c.synchronize()
    for ba in bas:
        ba.peer_detach()
    for ba in bas:
        ba.close()
    for c in ctxs:
        c.close()


if __name__ == "__main__":
    main()
