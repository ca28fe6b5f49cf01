"""Host-to-device copy rate with 1..N GPUs copying at once (the host-fabric ceiling as a number).
Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/h2d_fabric_probe.py
Every rank copies a pinned 472 MB buffer (64 frames of 3840x1920) to its GPU `reps` times between two barriers, as one
cudaMemcpyAsync per frame on one stream -- the copy pattern of vio360_fe_track_batch.  Rank 0 prints one JSON line."""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    x = torch.empty(64, 1920, 3840, dtype=torch.uint8).pin_memory()
    y = torch.empty_like(x, device="cuda")
    xs = [x[i] for i in range(64)]
    reps = 10

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(2):
        y.copy_(x, non_blocking=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        for i in range(64):
            y[i].copy_(xs[i], non_blocking=True)
    torch.cuda.synchronize()
    mine = time.perf_counter() - t0
    barrier()
    total = time.perf_counter() - t0
    g = torch.tensor([x.numel() * reps / mine / 1e9], dtype=torch.float64, device="cuda")
    allg = [torch.zeros_like(g) for _ in range(world)]
    if world > 1:
        dist.all_gather(allg, g)
    else:
        allg = [g]
    if rank == 0:
        per = [round(float(v.item()), 2) for v in allg]
        print(json.dumps({"gpus": world, "bytes_per_rank_per_rep": x.numel(), "reps": reps, "h2d_gbs_per_rank": per,
                          "aggregate_gbs": round(x.numel() * reps * world / total / 1e9, 2),
                          "cpu_count": os.cpu_count()}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
