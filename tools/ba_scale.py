"""tools/ba_scale.py -- the BA section of bench.py alone (developer probe for the multi-GPU exchange).
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/ba_scale.py [steps]"""
import argparse
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    rank, local_rank, world = bench.dist_env()
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    b = importlib.import_module(bench.PKG + ".bindings")
    syn = importlib.import_module(bench.PKG + ".synthetic")
    ctx = b.Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.synchronize()
    args = argparse.Namespace(ba_steps=steps, skip_cpu=True)
    res = bench.bench_ba(ctx, b, syn, torch, dist, rank, world, device, barrier, stream, args)
    if rank == 0:
        print(json.dumps(res))
    barrier()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
