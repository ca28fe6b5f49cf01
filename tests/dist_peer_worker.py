"""Worker of tests/test_gpu_ba.py::test_peer_exchange_two_processes (launched by torch.distributed.run, one
rank per GPU): the peer-memory exchange of the sharded reduced system must equal the NCCL allreduce of the band
(up to summation order) and the unsharded build, and be bit-identical on all ranks."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PKG = "360_visual_inertial_odometry_b200"


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=device)
    b = importlib.import_module(PKG + ".bindings")
    syn = importlib.import_module(PKG + ".synthetic")
    ctx = b.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    for (K, M, deg, seed) in ((40, 1500, 8, 78), (200, 6000, 20, 79)):
        win = syn.make_ba_window(K=K, M=M, degree=deg, seed=seed, mid_iteration=True, loop_radius=6.0)
        full = b.BAWindow(ctx, win)
        full.set_parameters(win["pose_delta"], win["points"])
        D = np.sqrt(np.clip(full.column_sqnorm(True), 1e-6, 1e32) / 1e4)
        S0, rhs0, cost0 = full.build_reduced_system(D, True)
        full.close()
        ba = b.BAWindow(ctx, win)
        ba.set_shard(rank, world)
        ba.set_parameters(win["pose_delta"], win["points"])
        ba.set_lm_diagonal(D)
        # NCCL path: allreduce of the band, expand
        ba.build_reduced_system(None, True, fetch=False)
        ptr, nd = ba.reduced_band_dev()

        class _Buf:
            __cuda_array_interface__ = {"shape": (nd,), "typestr": "<f8", "data": (ptr, False), "version": 2}
        packed = torch.as_tensor(_Buf(), device=device)
        with torch.cuda.stream(stream):
            dist.all_reduce(packed)
        ba.expand_band()
        S1, rhs1, cost1 = ba.fetch_reduced_system()
        assert rel(S1, S0) < 1e-11 and rel(rhs1, rhs0) < 1e-11 and abs(cost1 - cost0) < 1e-11 * cost0
        del packed
        # peer-memory path
        blobs = [None] * world
        dist.all_gather_object(blobs, ba.peer_export())
        ba.peer_attach(blobs)
        dist.barrier()
        for it in range(4):
            ba.build_reduced_system(None, True, fetch=False)
            ba.peer_exchange()
        S2, rhs2, cost2 = ba.fetch_reduced_system()
        assert rel(S2, S0) < 1e-11 and rel(rhs2, rhs0) < 1e-11 and abs(cost2 - cost0) < 1e-11 * cost0, (rel(S2, S0), rel(rhs2, rhs0))
        assert np.array_equal(S2, S2.T)
        # bit-identical on all ranks (rank-ordered sum)
        h = torch.tensor([float(np.sum(S2)), float(np.sum(rhs2)), cost2], dtype=torch.float64, device=device)
        hs = [torch.zeros_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        assert all(torch.equal(x, hs[0]) for x in hs)
        dist.barrier()
        ba.peer_detach()
        ba.close()
    dist.barrier()
    if rank == 0:
        print("PEER_OK")
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
