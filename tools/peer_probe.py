"""tools/peer_probe.py -- per-kernel times of build + vio360_ba_peer_exchange in an LM loop on every rank, developer probe
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/peer_probe.py"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    rank, local_rank, world = bench.dist_env()
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist.init_process_group("nccl", device_id=device)
    b = importlib.import_module(bench.PKG + ".bindings")
    syn = importlib.import_module(bench.PKG + ".synthetic")
    ctx = b.Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    win = syn.make_ba_window(K=200, M=50000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
    ba = b.BAWindow(ctx, win)
    ba.set_shard(rank, world)
    ba.set_parameters(win["pose_delta"], win["points"])
    blobs = [None] * world
    dist.all_gather_object(blobs, ba.peer_export())
    ba.peer_attach(blobs)
    dist.barrier()
    ba.build_reduced_system(None, True, fetch=False)
    # skew probe: build + exchange, events around each part
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    for _ in range(30):
        ba.build_reduced_system(None, True, fetch=False)
        ba.peer_exchange()
    ctx.synchronize()
    print("rank %d: in the LM loop %s" % (rank, {k: round(ctx.kernel_time(k)[0] / max(ctx.kernel_time(k)[1], 1) * 1e3, 2) for k in ("k_schur_tile", "k_schur_reduce", "k_band_pull")}), flush=True)
    dist.barrier()
    ba.peer_detach(); ba.close(); ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
