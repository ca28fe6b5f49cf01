#!/usr/bin/env python
"""bench.py -- headline benchmark of the two hot paths on N B200s of one node.

  python bench.py --gpus N --steps K --warmup W            (this repo's CUDA path)
  python bench.py --impl reference --gpus N ...             (the reference's CPU path)

Metric (BASELINE.json): ERP frames/s @3840x1920, 1000 features (config 2), with the BA
factor-evaluation figure (observations/s, config 5 = 200 keyframes / 1M observations,
plus config 3 at N=1) reported in the "ba" object of the same JSON line.

One step of the headline path = one tracking pass over a batch of independent frame pairs
(one per sequence, `--batch` sequences per GPU): pyramid of the new frame + pyramidal LK of
1000 features; the previous frame's pyramid is cached on the device exactly as a sequence
would have it.  Inputs (2 x batch frames per GPU = 0.94 GB) are far larger than L2.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = "360_visual_inertial_odometry_b200"
sys.path.insert(0, ROOT)

W, H, NFEAT, LEVELS, WIN = 3840, 1920, 1000, 4, 21
METRIC = "ERP tracking frames/s @3840x1920, 1000 features, 4-level pyramid, 21x21 KLT"


def algorithmic_bytes():
    pyr = 1.640625 * W * H                                   # SURVEY 8-D, K1 per new frame
    lk = NFEAT * LEVELS * ((WIN + 2) ** 2 + (WIN + 1) ** 2) + 21 * NFEAT   # K3 per frame pair
    return pyr, lk


class ClockSampler(threading.Thread):
    """samples nvidia-smi during the timed region (B200_PROFILING.md clocks line)"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index; self.stop_flag = False; self.rows = []

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def ncu_traffic(kernel, batch):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full`
    capture (profiles/ncu_traffic.json, written by tools/ncu_summary.py --traffic).  Only reported when the
    capture was taken on this workload AND on this version of the kernel source (sha1 of the .cuh), else None:
    a profiler figure cannot be measured inside a timed run, so it is tied to what it was measured on."""
    import hashlib
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[kernel]
        src = os.path.join(ROOT, PKG, "csrc", rec["source"])
        sha = hashlib.sha1(open(src, "rb").read()).hexdigest()
        wl = rec["workload"]
        if sha != rec["source_sha1"] or (wl["W"], wl["H"], wl["batch"], wl["points"]) != (W, H, batch, NFEAT):
            return None, "stale: capture %s was taken on another kernel version or workload" % rec.get("capture")
        return float(rec["bytes_per_launch"]), "dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, %s" % rec["capture"]
    except Exception as e:   # no capture committed
        return None, "no capture (%s)" % type(e).__name__


def measured_peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def make_frames(torch, syn, batch, rank, device):
    """2 x batch synthetic textured-sphere frames on the device (consecutive time steps)."""
    ring = [[], []]
    for s in range(batch):
        seed = 360 + rank * batch + s
        for t in (0, 1):
            ring[t].append(syn.render_erp_frame(W, H, t, seed=seed, n_waves=24, sigma_omega=120.0, xp=torch, device=device))
    return ring


# =========================================================================== our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    rank, local_rank, world = dist_env()
    if args.gpus != world and world > 1:
        pass
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)
    b = importlib.import_module(PKG + ".bindings")
    syn = importlib.import_module(PKG + ".synthetic")
    ctx = b.Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.stream, device=device)
    B = args.batch
    ctx.fe_configure(W, H, max_points=NFEAT, max_batch=B)
    ring = make_frames(torch, syn, B, rank, device)
    pts_h = np.stack([syn.make_track_points(W, H, NFEAT, seed=1000 + rank * B + s) for s in range(B)])
    pts_d = torch.from_numpy(pts_h).to(device)
    out_xy = torch.zeros((B, NFEAT, 2), dtype=torch.float32, device=device)
    out_st = torch.zeros((B, NFEAT), dtype=torch.uint8, device=device)
    out_err = torch.zeros((B, NFEAT), dtype=torch.float32, device=device)
    out_keep = torch.zeros((B, NFEAT), dtype=torch.uint8, device=device)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ctx.synchronize()

    key_base = [1 << 32]

    dev_jobs = []
    for par in range(2):      # job tables for even / odd steps; only the keys change per step
        jobs = (b.TrackJob * B)()
        for s in range(B):
            j = jobs[s]
            prev, cur = ring[par][s], ring[1 - par][s]
            j.prev, j.prev_pitch, j.cur, j.cur_pitch = prev.data_ptr(), W, cur.data_ptr(), W
            j.prev_xy, j.n = pts_d[s].data_ptr(), NFEAT
            j.cur_xy, j.status, j.err, j.keep = out_xy[s].data_ptr(), out_st[s].data_ptr(), out_err[s].data_ptr(), out_keep[s].data_ptr()
        dev_jobs.append(jobs)

    def dev_step(i):
        jobs = dev_jobs[i % 2]
        for s in range(B):
            jobs[s].prev_key = key_base[0] + i * B + s
            jobs[s].cur_key = key_base[0] + (i + 1) * B + s
        ctx.fe_track_batch_raw(jobs, B, b.MEM_DEVICE)

    # ---- device-resident throughput ("value")
    for i in range(args.warmup):
        dev_step(i)
    barrier()
    l0 = ctx.launch_count
    sampler = ClockSampler(local_rank); sampler.start()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(args.warmup, args.warmup + args.steps):
        dev_step(i)
    e1.record(stream)
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    launches = ctx.launch_count - l0
    # second pass over the same K steps with a CUDA event pair around every kernel launch (on the context's
    # stream): the per-kernel durations of the roofline object.  Kept out of the headline region above.
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    k0 = torch.cuda.Event(enable_timing=True); k1 = torch.cuda.Event(enable_timing=True)
    k0.record(stream)
    for i in range(args.warmup + args.steps, args.warmup + 2 * args.steps):
        dev_step(i)
    k1.record(stream)
    barrier()
    ms_events_total = k0.elapsed_time(k1)
    lk_ms, lk_n = ctx.kernel_time("k_lk_track")
    pyr_ms, pyr_n = ctx.kernel_time("k_pyr_down")
    ctx.enable_kernel_timing(False)
    tracked = int(out_st.sum().item())
    # the timed region is ~10 ms, shorter than one nvidia-smi query: keep the same steps running for ~1.5 s more so
    # that the sampler sees clocks and throttle reasons UNDER THIS LOAD (reported beside `value`, not part of it)
    s0 = torch.cuda.Event(enable_timing=True); s1 = torch.cuda.Event(enable_timing=True)
    sus_steps, t_sus, i_sus = 0, time.perf_counter(), args.warmup + 2 * args.steps
    s0.record(stream)
    while time.perf_counter() - t_sus < 1.5:
        for _ in range(50):
            dev_step(i_sus + sus_steps); sus_steps += 1
        ctx.synchronize()
    s1.record(stream)
    barrier()
    sampler.stop_flag = True
    sustained_ms = s0.elapsed_time(s1) / max(sus_steps, 1)
    last_dev = i_sus + sus_steps - 1        # the frame pair out_st now belongs to
    frames = B * args.steps * world
    value = frames / (ms_total * 1e-3)

    # ---- end to end through the C ABI with pinned HOST buffers
    e2e_steps = max(20, args.steps)
    hring = [[torch.empty((H, W), dtype=torch.uint8).pin_memory() for _ in range(B)] for _ in range(2)]
    for t in range(2):
        for s in range(B):
            hring[t][s].copy_(ring[t][s])
    hpts = torch.from_numpy(pts_h).pin_memory()
    hxy = torch.zeros((B, NFEAT, 2), dtype=torch.float32).pin_memory()
    hst = torch.zeros((B, NFEAT), dtype=torch.uint8).pin_memory()
    herr = torch.zeros((B, NFEAT), dtype=torch.float32).pin_memory()
    key_base[0] = 1 << 40

    host_jobs = []
    for par in range(2):      # job tables for even / odd steps; only the keys change per step
        jobs = (b.TrackJob * B)()
        for s in range(B):
            j = jobs[s]
            j.prev, j.prev_pitch = hring[par][s].data_ptr(), W
            j.cur, j.cur_pitch = hring[1 - par][s].data_ptr(), W
            j.prev_xy, j.n = hpts[s].data_ptr(), NFEAT
            j.cur_xy, j.status, j.err, j.keep = hxy[s].data_ptr(), hst[s].data_ptr(), herr[s].data_ptr(), None
        host_jobs.append(jobs)

    def host_step(i):
        jobs = host_jobs[i % 2]
        for s in range(B):
            jobs[s].prev_key = key_base[0] + i * B + s
            jobs[s].cur_key = key_base[0] + (i + 1) * B + s
        ctx.fe_track_batch_raw(jobs, B, b.MEM_HOST)

    host_step(0); host_step(1)
    barrier()
    t0 = time.perf_counter()
    for i in range(2, 2 + e2e_steps):
        host_step(i)
    my_s = time.perf_counter() - t0          # this rank's own elapsed time (the call returns with the results on the host)
    barrier()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_value = B * e2e_steps * world / float(e2e_s.item())
    # host-to-device rate every rank reached inside its own timed region (the host fabric limit as
    # a measured number): bytes this rank uploaded / this rank's own elapsed time
    my_gbs = torch.tensor([B * e2e_steps * (W * H + NFEAT * 8) / my_s / 1e9], dtype=torch.float64, device=device)
    if world > 1:
        all_gbs = [torch.zeros_like(my_gbs) for _ in range(world)]
        dist.all_gather(all_gbs, my_gbs)
        h2d_gbs = [float(x.item()) for x in all_gbs]
    else:
        h2d_gbs = [float(my_gbs.item())]
    # The same call from memory the caller did NOT allocate pinned: plain numpy buffers (pageable, what a cv::Mat is),
    # then the same buffers page-locked in place through vio360_host_register (what the shim's RegisterFrameBuffer /
    # SetAutoRegisterFrames do).  Fewer sequences, so that the pageable leg stays short.
    host_kinds = {}
    if rank == 0 or world == 1:
        Bk = min(B, 16)
        npring = [[np.empty((H, W), np.uint8) for _ in range(Bk)] for _ in range(2)]
        for t in range(2):
            for s in range(Bk):
                npring[t][s][...] = hring[t][s].numpy()
        np_xy = np.zeros((Bk, NFEAT, 2), np.float32); np_st = np.zeros((Bk, NFEAT), np.uint8); np_err = np.zeros((Bk, NFEAT), np.float32)
        np_pts = np.ascontiguousarray(pts_h[:Bk])
        kjobs = []
        for par in range(2):
            jobs = (b.TrackJob * Bk)()
            for s in range(Bk):
                j = jobs[s]
                j.prev, j.prev_pitch = npring[par][s].ctypes.data, W
                j.cur, j.cur_pitch = npring[1 - par][s].ctypes.data, W
                j.prev_xy, j.n = np_pts[s].ctypes.data, NFEAT
                j.cur_xy, j.status, j.err, j.keep = np_xy[s].ctypes.data, np_st[s].ctypes.data, np_err[s].ctypes.data, None
            kjobs.append(jobs)

        def kind_step(i, base):
            jobs = kjobs[i % 2]
            for s in range(Bk):
                jobs[s].prev_key = base + i * Bk + s
                jobs[s].cur_key = base + (i + 1) * Bk + s
            ctx.fe_track_batch_raw(jobs, Bk, b.MEM_HOST)

        def time_kind(base, n):
            kind_step(0, base); kind_step(1, base)
            t1 = time.perf_counter()
            for i in range(2, 2 + n):
                kind_step(i, base)
            return Bk * n / (time.perf_counter() - t1)
        host_kinds["pageable"] = time_kind(1 << 44, 6)
        t1 = time.perf_counter()
        for t in range(2):
            for s in range(Bk):
                ctx.host_register(npring[t][s])
        reg_s = time.perf_counter() - t1
        host_kinds["registered_in_place"] = time_kind(1 << 45, 12)
        for t in range(2):
            for s in range(Bk):
                ctx.host_unregister(npring[t][s])
        host_kinds.update(sequences=Bk, unit="frames/s", register_ms_per_frame=reg_s * 1e3 / (2 * Bk),
                          note="vio360_fe_track_batch from numpy buffers: pageable as allocated, then page-locked in place with vio360_host_register (one call per buffer, not per step)")
    # same frame pair as the last device-resident step (one more untimed call when the ring parity differs):
    # the status bytes of the host-buffer path must equal the device path's
    last_host = 1 + e2e_steps
    if last_host % 2 != last_dev % 2:
        host_step(last_host + 1)
    e2e_status_equal = bool((hst.to(device) == out_st).all().item())

    # ---- roofline of the dominant kernel
    hbm_peak, peak_src = measured_peaks()
    pyr_b, lk_b = algorithmic_bytes()
    dominant = "k_lk_track" if lk_ms >= pyr_ms else "k_pyr_down"
    if dominant == "k_lk_track":
        per_launch_bytes = lk_b * B; dur_ms = lk_ms / max(lk_n, 1)
    else:
        per_launch_bytes = pyr_b * B; dur_ms = pyr_ms / max(args.steps, 1)   # three level launches per step
    achieved = per_launch_bytes / (dur_ms * 1e-3) / 1e9
    frame_bytes = (pyr_b + lk_b) * B
    step_ms = ms_total / args.steps
    traffic, traffic_src = ncu_traffic(dominant, B)
    roofline = {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src, "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": per_launch_bytes, "avg_launch_ms": dur_ms,
                "kernel_share_of_step": {"k_lk_track": lk_ms / ms_events_total, "k_pyr_down": pyr_ms / ms_events_total},
                "kernel_timing": "CUDA event pair around every launch on the context's stream, in a second pass over the same %d steps (%.4f ms/step with the event pairs, %.4f without)" % (args.steps, ms_events_total / args.steps, ms_total / args.steps),
                "tracking_frame": {"algorithmic_bytes_per_step": frame_bytes,
                                   "achieved_gbs": frame_bytes / (step_ms * 1e-3) / 1e9,
                                   "frac": frame_bytes / (step_ms * 1e-3) / 1e9 / hbm_peak},
                "k_pyr_down": {"achieved_gbs": pyr_b * B / (pyr_ms / max(args.steps, 1) * 1e-3) / 1e9 if pyr_ms > 0 else None}}

    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic",
            "config": {"workload": "configs[1]: ERP tracking 3840x1920, 4-level pyramid, 1000 features/frame, 21x21 KLT",
                       "frames_per_step_per_gpu": B, "sequences": B * world, "sharding": "independent sequences, no collective",
                       "l2": "inputs (2 x %d frames = %.2f GB per GPU) exceed the 126 MB L2" % (B, 2 * B * W * H / 1e9),
                       "prev_pyramid": "cached on device (sequence mode)"},
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": B * (W * H + NFEAT * 8),
                    "d2h_bytes_per_step": B * NFEAT * (8 + 4 + 1 + 1), "steps": e2e_steps,
                    "h2d_gbs_per_rank": h2d_gbs, "host_memory": host_kinds,
                    "note": "vio360_fe_track_batch with pinned host buffers; H2D of each new frame and D2H of results inside the timed region; frames are uploaded on a copy stream in groups of 4 jobs and each group's kernels run behind the next groups' copies"},
            "gpu_launches": launches, "roofline": roofline,
            "clocks": dict(sampler.summary(), sustained_ms_per_step=sustained_ms, sustained_steps=sus_steps,
                           note="sampled over the timed region plus the same steps repeated for 1.5 s"),
            "tracked_points_last_step": tracked, "e2e_status_equals_device_path": e2e_status_equal}

    # ---- detection (reported separately)
    if not args.skip_detect:
        line["detect"] = bench_detect(ctx, b, torch, ring, B, world, device, barrier, stream)
        line["ransac"] = bench_ransac(ctx, syn)
        line["pnp_triangulate"] = bench_pnp_triangulate(ctx, syn, args, rank)
        line["sequence_chain"] = bench_sequence(ctx, b, syn, torch, device, barrier, args, rank)
    # ---- BA factor evaluation
    if not args.skip_ba:
        line["ba"] = bench_ba(ctx, b, syn, torch, dist, rank, world, device, barrier, stream, args)
    # ---- configs[3]: 8K tracking, 64 sequences over the ranks (re-configures the front end: keep it last)
    if not args.skip_c4:
        line["tracking_c4"] = bench_c4(ctx, b, syn, torch, dist, rank, world, device, barrier, stream, args)
    if rank == 0 and world == 1 and not args.skip_cpu:
        # parity on the benchmark's own inputs: the frame pairs the CPU leg runs, once more through the CUDA path
        ctx.fe_configure(W, H, max_points=NFEAT, max_batch=B)
        jobs = dev_jobs[0]                       # prev = frame t, cur = frame t + 1 of every sequence
        for s in range(B):
            jobs[s].prev_key = 0; jobs[s].cur_key = 0
        ctx.fe_track_batch_raw(jobs, B, b.MEM_DEVICE)
        ctx.synchronize()
        gpu_out = (out_xy.cpu().numpy(), out_st.cpu().numpy(), out_err.cpu().numpy())
        line["cpu_baseline"], line["parity"] = cpu_baseline_tracking(ring, pts_h, torch, gpu_out)
        if "detect" in line:
            cb, par = cpu_baseline_detect(ring, line["detect"].pop("_corners"))
            line["detect"]["cpu_baseline"] = cb
            line["parity"]["gftt_vs_cv2"] = par
    if "detect" in line:
        line["detect"].pop("_corners", None)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line))


def bench_c4(ctx, b, syn, torch, dist, rank, world, device, barrier, stream, args):
    """BASELINE configs[3]: 7680x3840, 3000 features per frame, 64 independent sequences sharded over the
    ranks (sequence s -> rank s mod world, no collective): strong scaling of one 64-frame step."""
    W4, H4, N4, SEQ = 7680, 3840, 3000, 64
    mine = [s for s in range(SEQ) if s % world == rank]
    B4 = len(mine)
    ctx.fe_configure(W4, H4, max_points=N4, max_batch=B4)
    ring = [[], []]
    for s in mine:
        for t in (0, 1):
            ring[t].append(syn.render_erp_frame(W4, H4, t, seed=360 + s, n_waves=24, sigma_omega=240.0, xp=torch, device=device))
    pts_d = torch.from_numpy(np.stack([syn.make_track_points(W4, H4, N4, seed=3000 + s) for s in mine])).to(device)
    out_xy = torch.zeros((B4, N4, 2), dtype=torch.float32, device=device)
    out_st = torch.zeros((B4, N4), dtype=torch.uint8, device=device)
    out_err = torch.zeros((B4, N4), dtype=torch.float32, device=device)
    out_keep = torch.zeros((B4, N4), dtype=torch.uint8, device=device)
    tables = []
    for par in range(2):
        jobs = (b.TrackJob * B4)()
        for i in range(B4):
            j = jobs[i]
            j.prev, j.prev_pitch, j.cur, j.cur_pitch = ring[par][i].data_ptr(), W4, ring[1 - par][i].data_ptr(), W4
            j.prev_xy, j.n = pts_d[i].data_ptr(), N4
            j.cur_xy, j.status, j.err, j.keep = out_xy[i].data_ptr(), out_st[i].data_ptr(), out_err[i].data_ptr(), out_keep[i].data_ptr()
        tables.append(jobs)

    def step(i):
        jobs = tables[i % 2]
        for k in range(B4):
            jobs[k].prev_key = (1 << 44) + i * B4 + k
            jobs[k].cur_key = (1 << 44) + (i + 1) * B4 + k
        ctx.fe_track_batch_raw(jobs, B4, b.MEM_DEVICE)

    steps = max(2, min(args.steps, 6))
    for i in range(3):
        step(i)
    barrier()
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(3, 3 + steps):
        step(i)
    e1.record(stream)
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    lk_ms, lk_n = ctx.kernel_time("k_lk_track")
    pyr_ms, _ = ctx.kernel_time("k_pyr_down")
    ctx.enable_kernel_timing(False)
    step_ms = float(ms.item()) / steps
    frame_bytes = 1.640625 * W4 * H4 + N4 * 4 * 1013 + 21 * N4      # SURVEY 8-D tracking frame
    hbm_peak = measured_peaks()[0]
    fps = SEQ / (step_ms * 1e-3)
    tracked = torch.tensor([int(out_st.sum().item())], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(tracked)
    return {"metric": "ERP frames/s @7680x3840, 3000 feats (64 sequences over all ranks)", "value": fps, "unit": "frames/s",
            "scaling": "strong", "n_gpus": world, "sequences_per_gpu": B4, "steps": steps, "ms_per_step": step_ms,
            "kernel_ms_per_launch": {"k_lk_track": lk_ms / max(lk_n, 1), "k_pyr_down_per_step": pyr_ms / steps},
            "tracked_points_last_step": int(tracked.item()),
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_frame": frame_bytes, "achieved": fps / world * frame_bytes / 1e9,
                         "peak": hbm_peak, "unit": "GB/s", "frac": fps / world * frame_bytes / 1e9 / hbm_peak}}


def bench_sequence(ctx, b, syn, torch, device, barrier, args, rank):
    """BASELINE configs[0], headless: FeatureTracker::TrackFeatures (FeatureTracker.cpp:61-206) frame after frame at 1920x960 with
    default_config.yaml's detector settings, as ONE device-resident call per frame and batch of sequences
    (vio360_fe_track_frame_batch: track -> post filter -> rotation RANSAC -> cluster removal -> grid cap -> masked
    re-detection -> grid cap), host frames in, final feature lists out.  Beside it the same chain on the host cores:
    cv2 for the two OpenCV calls + the C restatement of the reference's list logic."""
    Ws, Hs, S, T = 1920, 960, 16, 10
    ctx.fe_configure(Ws, Hs, max_points=2000, max_batch=S)
    frames = [[syn.render_erp_frame(Ws, Hs, t, seed=360 + rank * S + q, n_waves=24, sigma_omega=60.0, xp=torch, device=device) for t in range(T)]
              for q in range(S)]
    hframes = [[f.cpu().pin_memory() for f in fs] for fs in frames]
    rng = np.random.default_rng(42)
    words = torch.from_numpy(rng.integers(0, 2 ** 32, (T, S, 1000, 3), dtype=np.uint64).astype(np.uint32).view(np.int32)).pin_memory()
    cfg = dict(max_features=1000, min_distance=30.0, quality=0.01, grid_cols=20, grid_rows=10, max_per_cell=10,
               remove_clustered=True, std_ratio=0.25, ransac_iters=1000, ransac_thr_deg=2.0)
    cap = 2000
    ids = torch.zeros((S, cap), dtype=torch.int32).pin_memory(); xy = torch.zeros((S, cap, 2), dtype=torch.float32).pin_memory()
    tc = torch.zeros((S, cap), dtype=torch.int32).pin_memory()
    cnt = (C.c_int * (3 * S))()
    out = {}
    for space, name in ((b.MEM_HOST, "host_frames"), (b.MEM_DEVICE, "device_frames")):
        seqs = [ctx.seq_create(**cfg) for _ in range(S)]
        if space == b.MEM_DEVICE:
            d_ids = torch.zeros((S, cap), dtype=torch.int32, device=device); d_xy = torch.zeros((S, cap, 2), dtype=torch.float32, device=device)
            d_tc = torch.zeros((S, cap), dtype=torch.int32, device=device); d_words = words.to(device)
        jobs = (b.FrameJob * S)()

        def step(t):
            for q in range(S):
                j = jobs[q]
                j.seq = seqs[q]; j.pitch = Ws
                if space == b.MEM_HOST:
                    j.image, j.ransac_words = hframes[q][t].data_ptr(), words[t, q].data_ptr()
                    j.ids, j.xy, j.track_count = ids[q].data_ptr(), xy[q].data_ptr(), tc[q].data_ptr()
                else:
                    j.image, j.ransac_words = frames[q][t].data_ptr(), d_words[t, q].data_ptr()
                    j.ids, j.xy, j.track_count = d_ids[q].data_ptr(), d_xy[q].data_ptr(), d_tc[q].data_ptr()
                j.n_features, j.n_tracked, j.n_detected = C.addressof(cnt) + 12 * q, C.addressof(cnt) + 12 * q + 4, C.addressof(cnt) + 12 * q + 8
            ctx.fe_track_frame_batch_raw(jobs, S, space)

        step(0); step(1)
        barrier()
        l0 = ctx.launch_count
        t0 = time.perf_counter()
        for t in range(2, T):
            step(t)
        barrier()
        dt = time.perf_counter() - t0
        n_feat = [cnt[3 * q] for q in range(S)]; n_trk = [cnt[3 * q + 1] for q in range(S)]
        out[name] = {"value": S * (T - 2) / dt, "unit": "frames/s", "ms_per_frame_batch": dt / (T - 2) * 1e3, "sequences": S,
                     "frames_timed": T - 2, "gpu_launches_per_call": (ctx.launch_count - l0) / (T - 2),
                     "features_last_frame_mean": float(np.mean(n_feat)), "tracked_last_frame_mean": float(np.mean(n_trk))}
        if space == b.MEM_HOST:
            last = (ids[0, :cnt[0]].numpy().copy(), xy[0, :cnt[0]].numpy().copy(), tc[0, :cnt[0]].numpy().copy())
            out[name]["h2d_bytes_per_call"] = S * (Ws * Hs + 12000); out[name]["d2h_bytes_per_call"] = S * (16 * cap + 80)
        for h in seqs:
            ctx.seq_destroy(h)
    # the reference's own situation: ONE sequence, frame after frame from host memory -- the latency of a frame
    seq1 = ctx.seq_create(**cfg)
    job1 = (b.FrameJob * 1)()

    def step1(t):
        j = job1[0]
        j.seq = seq1; j.pitch = Ws
        j.image, j.ransac_words = hframes[0][t].data_ptr(), words[t, 0].data_ptr()
        j.ids, j.xy, j.track_count = ids[0].data_ptr(), xy[0].data_ptr(), tc[0].data_ptr()
        j.n_features, j.n_tracked, j.n_detected = C.addressof(cnt), C.addressof(cnt) + 4, C.addressof(cnt) + 8
        ctx.fe_track_frame_batch_raw(job1, 1, b.MEM_HOST)
    step1(0); step1(1)
    t0 = time.perf_counter()
    for t in range(2, T):
        step1(t)
    dt1 = time.perf_counter() - t0
    ctx.seq_destroy(seq1)
    out["single_sequence"] = {"ms_per_frame": dt1 / (T - 2) * 1e3, "value": (T - 2) / dt1, "unit": "frames/s", "frames_timed": T - 2,
                              "note": "one sequence, host frames, one call per frame (H2D of the frame and D2H of the feature list inside)"}
    res = {"metric": "TrackFeatures chain frames/s @1920x960 (configs[0] headless): track + filter + RANSAC + cluster + grid cap + re-detect",
           "value": out["host_frames"]["value"], "unit": "frames/s", **out}
    if rank == 0 and not args.skip_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import fe_oracle
        ref = fe_oracle.SequenceTracker(Ws, Hs, max_features=1000, min_distance=30.0, quality=0.01, max_per_cell=10, use_cv2=True)
        f0 = [hframes[0][t].numpy() for t in range(T)]
        w0 = words[:, 0].numpy().view(np.uint32)
        ref.step(f0[0], w0[0]); ref.step(f0[1], w0[1])
        t0 = time.perf_counter()
        for t in range(2, T):
            ref.step(f0[t], w0[t])
        dtc = time.perf_counter() - t0
        import cv2
        res["cpu_baseline"] = {"value": (T - 2) / dtc, "unit": "frames/s", "cores": int(cv2.getNumThreads()), "kind": "reference",
                               "sample": "%d frames of one sequence: cv2.calcOpticalFlowPyrLK + cv2.goodFeaturesToTrack %s with the reference's arguments, the reference's list logic restated in C (oracle/fe_ref.c)" % (T - 2, cv2.__version__)}
        gi, gx, gt = last
        res["parity"] = {"sequence": 0, "frame": T - 1, "ids_equal": bool(np.array_equal(gi, ref.ids)), "track_counts_equal": bool(np.array_equal(gt, ref.tc)),
                         "max_abs_dxy_px": float(np.abs(gx - ref.xy).max()) if gx.shape == ref.xy.shape and len(gx) else None,
                         "features": int(len(gi)), "tolerance_px": 0.05}
    return res


def bench_ransac(ctx, syn):
    """rotation RANSAC, 1000 tracked points x 1000 hypotheses at 2 degrees (FeatureTracker.cpp:36-37),
    16 frames per call through the host-memory C ABI (copies inside the timed region)."""
    nb, n, iters = 16, 1000, 1000
    items = []
    for s in range(nb):
        prev, cur, _ = syn.make_rotation_tracks(W, H, n, seed=500 + s)
        items.append((prev, cur, syn.draw_ransac_samples(n, iters, seed=600 + s)))
    for _ in range(3):
        res = ctx.fe_ransac_rotation_batch(items, 2.0)
    reps = 10
    t0 = time.perf_counter()
    for _ in range(reps):
        res = ctx.fe_ransac_rotation_batch(items, 2.0)
    dt = (time.perf_counter() - t0) / reps
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    ctx.fe_ransac_rotation_batch(items, 2.0)
    kern = {}
    for name in ("k_ransac_bearings", "k_ransac_hypotheses", "k_ransac_pick"):
        ms, cnt = ctx.kernel_time(name)
        kern[name] = ms / max(cnt, 1)
    ctx.enable_kernel_timing(False)
    return {"metric": "rotation RANSAC frames/s (1000 points, 1000 hypotheses)", "value": nb / dt, "unit": "frames/s",
            "frames_per_call": nb, "ms_per_call": dt * 1e3, "kernel_ms_per_launch": kern,
            "inliers_first_frame": int(res[0][0]),
            "parity_note": "unpinned by the reference (Eigen JacobiSVD absent here): tolerance-based against the fp32 C restatement, <= 1 % inlier-count difference per hypothesis (tests/test_gpu_fe.py)"}


def bench_pnp_triangulate(ctx, syn, args, rank):
    """batched PnP evaluation (PnPFactor::Evaluate + compute_chi_square for all matches of a frame,
    Optimizer.cpp:180-267) and batched DLT triangulation (Estimator::TriangulateSinglePoint, Estimator.cpp:1082-1139),
    through the host-memory C ABI (copies inside the timed region), beside the C restatement on one host core."""
    out = {}
    win = syn.make_ba_window(K=3, M=1000, degree=3, seed=77, mid_iteration=True)
    k = 1
    sel = np.where(win["pose_idx"] == k)[0]
    obs = np.stack([win["obs_u"][sel], win["obs_v"][sel]], axis=1)
    pts = win["points"][win["point_idx"][sel]].copy()
    call = lambda: ctx.pnp_evaluate(obs, pts, win["Twb_init"][k], win["Tcb"], win["cols"], win["rows"], win["info"], 1.0, win["pose_delta"][k])
    for _ in range(3):
        call()
    reps = 50
    t0 = time.perf_counter()
    for _ in range(reps):
        call()
    dt = (time.perf_counter() - t0) / reps
    out["pnp"] = {"metric": "PnP factor evaluations/s (residual, Jacobian, Huber, H = J'J, b = J'r, chi2 of one frame's matches)",
                  "value": len(sel) / dt, "unit": "obs/s", "observations_per_call": int(len(sel)), "ms_per_call": dt * 1e3}
    b1, b2, T1, T2, _ = syn.make_triangulation_pairs(20000, seed=6)
    for _ in range(3):
        ctx.triangulate(b1, b2, T1, T2)
    t0 = time.perf_counter()
    for _ in range(reps):
        ctx.triangulate(b1, b2, T1, T2)
    dtt = (time.perf_counter() - t0) / reps
    out["triangulate"] = {"metric": "DLT triangulations/s", "value": len(b1) / dtt, "unit": "points/s", "points_per_call": int(len(b1)), "ms_per_call": dtt * 1e3,
                          "parity_note": "unpinned: 2e-3 of the point's distance against the fp32 C restatement (tests/test_gpu_ba.py)"}
    out["pnp"]["parity_note"] = "unpinned by the reference (no vectors): 1e-9 against the C restatement of PnPFactor (tests/test_gpu_ba.py::test_pnp)"
    if rank == 0 and not args.skip_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import ba_oracle
        n = min(len(sel), 300)
        t0 = time.perf_counter()
        for i in range(n):
            f = ba_oracle.make_factor(obs[i], win["cols"], win["rows"], win["Tcb"], win["Twb_init"][k], win["info"], world_point=tuple(pts[i]))
            ba_oracle.pnp_evaluate(f, win["pose_delta"][k]); ba_oracle.pnp_chi2(f, win["pose_delta"][k])
        out["pnp"]["cpu_baseline"] = {"value": n / (time.perf_counter() - t0), "unit": "obs/s", "cores": 1, "kind": "port",
                                      "sample": "%d factors, one call per factor as Ceres does (C restatement behind ctypes; the per-call overhead of the binding is included)" % n}
        t0 = time.perf_counter()
        ba_oracle.triangulate(b1[:5000], b2[:5000], T1, T2)
        out["triangulate"]["cpu_baseline"] = {"value": 5000 / (time.perf_counter() - t0), "unit": "points/s", "cores": 1, "kind": "port", "sample": "5000 pairs"}
    return out


def bench_detect(ctx, b, torch, ring, B, world, device, barrier, stream):
    nb = min(B, 64)
    out_xy = torch.zeros((nb, 1000, 2), dtype=torch.float32, device=device)
    out_n = (C.c_int * nb)()

    def step():
        jobs = (b.DetectJob * nb)()
        for s in range(nb):
            j = jobs[s]
            j.img, j.pitch, j.key = ring[0][s].data_ptr(), W, 0
            j.mask, j.mask_pitch = None, 0
            j.out_xy, j.out_n = out_xy[s].data_ptr(), C.addressof(out_n) + 4 * s
        ctx.fe_detect_batch_raw(jobs, nb, 1000, float(np.float32(0.01)), float(np.float32(30.0)), b.MEM_DEVICE)

    for _ in range(3):
        step()
    barrier()
    reps = 10
    t0 = time.perf_counter()
    for _ in range(reps):
        step()
    barrier()
    dt = (time.perf_counter() - t0) / reps
    ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
    for _ in range(3):
        step()
    barrier()
    kern = {}
    for name in ("k_min_eig", "k_candidates", "k_sort_select"):
        ms, n = ctx.kernel_time(name)
        kern[name] = ms / max(n, 1)
    ctx.enable_kernel_timing(False)
    fps = nb / dt
    alg_bytes = float(W * H + 1000 * 8)   # SURVEY 8-D K4+K5: image in, corners out; the response map is internal
    corners = [out_xy[s, :out_n[s]].cpu().numpy() for s in range(min(nb, 2))]
    return {"_corners": corners,
            "metric": "ERP detection frames/s @3840x1920 (goodFeaturesToTrack, 1000 corners)", "value": fps * world,
            "unit": "frames/s", "frames_per_step_per_gpu": nb, "corners_first_frame": int(out_n[0]),
            "ms_per_step": dt * 1e3, "kernel_ms_per_launch": kern, "k_min_eig_ms_per_launch": kern["k_min_eig"],
            "timing": "host clock around the C-ABI call (it ends with a stream synchronize); no library kernel, no intermediate read-back",
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_frame": alg_bytes, "achieved": fps * alg_bytes / 1e9,
                         "peak": measured_peaks()[0], "frac": fps * alg_bytes / 1e9 / measured_peaks()[0], "unit": "GB/s"}}


def bench_ba(ctx, b, syn, torch, dist, rank, world, device, barrier, stream, args):
    fp64_peak, fp32_peak = ctx.probe_fma_peak()
    res = {"metric": "BA factor evals (observations/s), fused residual+Jacobian+Huber+Schur blocks", "unit": "obs/s",
           "fp64_peak_tflops_measured": fp64_peak, "dtype": "f64"}

    def run(win, steps, warmup, sharded, exchange="nccl"):
        ba = b.BAWindow(ctx, win)
        if sharded:
            ba.set_shard(rank, world)
        ba.set_parameters(win["pose_delta"], win["points"])
        colsq = ba.column_sqnorm(True)
        D = np.sqrt(np.clip(colsq, 1e-6, 1e32) / 1e4)
        ba.set_lm_diagonal(D)
        multi = sharded and world > 1
        peer = multi and exchange == "peer"
        if peer:
            # SURVEY 8-E through peer memory: one kernel per rank pulls the covering bands over NVLink, sums, expands
            blobs = [None] * world
            dist.all_gather_object(blobs, ba.peer_export())
            ba.peer_attach(blobs)
            dist.barrier()
            band, packed, nd = None, None, 0
        else:
            ba.build_reduced_system(None, True, fetch=False)
            ptr, nd = ba.reduced_system_dev()
            band = ba.reduced_band_dev() if multi else None
            if band is not None:      # block-band payload [blocks | rhs | cost] (1.4 MB instead of 11.4 MB at 200 keyframes)
                ptr, nd = band

            class _Buf:   # zero-copy torch view of the exchange buffer
                __cuda_array_interface__ = {"shape": (nd,), "typestr": "<f8", "data": (ptr, False), "version": 2}
            packed = torch.as_tensor(_Buf(), device=device)

        def exchange_step():
            if peer:
                ba.peer_exchange()
            elif multi:
                with torch.cuda.stream(stream):
                    dist.all_reduce(packed)     # one allreduce of the exchange buffer
                if band is not None:
                    ba.expand_band()            # dense [S | rhs | cost] for the solver, on the same stream

        def step():
            ba.build_reduced_system(None, True, fetch=False)
            exchange_step()
        def timed(nsteps):
            barrier()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(nsteps):
                step()
            e1.record(stream)
            barrier()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t
        for _ in range(warmup):
            step()
        ms = timed(steps)                          # the step as a caller runs it
        ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
        ms_events = timed(steps)                   # again with an event pair around every kernel (per-kernel times; ~2 us each)
        ktimes = {name: ctx.kernel_time(name) for name in ("k_wide_prepare", "k_schur_tile", "k_schur_reduce", "k_schur_warp", "k_finish_S", "k_sum_partials", "k_expand_band", "k_band_pull")}
        per_rank = None
        if multi:   # every rank's kernel times: the exchange kernel's time includes waiting for the slowest rank
            per_rank = [None] * world
            dist.all_gather_object(per_rank, {k: round(v[0] / max(v[1], 1), 5) for k, v in ktimes.items() if v[1] > 0})
        dom = max((k for k in ktimes if k != "k_band_pull"), key=lambda k: ktimes[k][0])
        k_ms, k_n = ktimes[dom]
        ctx.enable_kernel_timing(False)
        # end to end: parameters from the host, reduced system back to the host
        t0 = time.perf_counter()
        for _ in range(max(2, steps // 2)):
            ba.set_parameters(win["pose_delta"], win["points"])
            ba.build_reduced_system(D, True, fetch=not multi)
            if multi:
                exchange_step()
                ba.fetch_reduced_system()
        barrier()
        e2e_s = (time.perf_counter() - t0) / max(2, steps // 2)
        S, rhs, cost = ba.fetch_reduced_system()
        deg = np.bincount(win["point_idx"], minlength=win["M"]).astype(np.float64)
        flops = 451.0 * win["O"] + float(np.sum(50 + 130 * deg + 216 * deg * (deg + 1) / 2))
        out = {"O": int(win["O"]), "K": int(win["K"]), "M": int(win["M"]), "ms_per_step": float(ms.item()) / steps,
               "ms_per_step_with_kernel_events": float(ms_events.item()) / steps,
               "value": win["O"] / (float(ms.item()) / steps * 1e-3), "e2e_value": win["O"] / e2e_s,
               "algorithmic_flop_per_step": flops, "allreduce_doubles": int(nd) if (multi and not peer) else 0,
               "exchange": ("peer memory (k_band_pull behind each rank's k_schur_reduce: announce the band, wait for the row flags of the rows a CTA needs, pull covering bands over NVLink, rank-ordered sum, LM diagonal, dense expand in one kernel; no grid-wide wait)"
                            if peer else ("NCCL allreduce of the block band + k_expand_band" if multi else "none")),
               "roofline": {"bound": "fp64", "kernel": dom, "achieved": flops / world / (k_ms / max(k_n, 1) * 1e-3) / 1e12 if k_ms > 0 else None,
                            "peak": fp64_peak, "unit": "TFLOP/s", "avg_launch_ms": k_ms / max(k_n, 1),
                            "kernel_ms_per_step": {k: v[0] / steps for k, v in ktimes.items() if v[1] > 0}},
               "cost": cost, "S_trace": float(np.trace(S))}
        if out["roofline"]["achieved"]:
            out["roofline"]["frac"] = out["roofline"]["achieved"] / fp64_peak
        if per_rank is not None:
            out["kernel_ms_per_launch_by_rank"] = per_rank
        if multi:
            # cross-device correctness carried by the scaling run itself: EVERY rank fetches the exchanged system
            # and compares it with its own unsharded build of the same window (max-norm relative error)
            step(); ctx.synchronize()
            S_x, rhs_x, cost_x = ba.fetch_reduced_system()
            full = b.BAWindow(ctx, win)
            full.set_parameters(win["pose_delta"], win["points"])
            full.set_lm_diagonal(D)
            S_f, rhs_f, cost_f = full.build_reduced_system(None, True)
            full.close()
            mine = {"rank": rank, "S": float(np.max(np.abs(S_x - S_f)) / np.max(np.abs(S_f))),
                    "rhs": float(np.max(np.abs(rhs_x - rhs_f)) / np.max(np.abs(rhs_f))),
                    "cost": float(abs(cost_x - cost_f) / abs(cost_f))}
            allp = [None] * world
            dist.all_gather_object(allp, mine)
            out["exchange_parity"] = {"per_rank": allp, "max": max(max(a["S"], a["rhs"], a["cost"]) for a in allp),
                                      "tolerance": 1e-9, "what": "max-norm relative error of the exchanged [S | rhs | cost] on each rank against that rank's own unsharded build"}
        if world == 1:
            # the optional fp32 mode of the fused path: float per-observation math on the FP32 pipe, double sums and SYRK
            ba.set_eval_precision(True)
            for _ in range(3):
                step()
            ms32 = timed(steps)
            ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
            timed(steps)
            k32_ms, k32_n = ctx.kernel_time("k_schur_tile_f32")
            ctx.enable_kernel_timing(False)
            S32, rhs32, cost32 = ba.fetch_reduced_system()
            ba.set_eval_precision(False)
            step(); ctx.synchronize()
            S64, rhs64, cost64 = ba.fetch_reduced_system()
            a32 = flops / (k32_ms / max(k32_n, 1) * 1e-3) / 1e12 if k32_ms > 0 else None
            out["fp32_eval_mode"] = {"ms_per_step": float(ms32.item()) / steps, "value": win["O"] / (float(ms32.item()) / steps * 1e-3),
                                     "kernel": "k_schur_tile<float>", "avg_launch_ms": k32_ms / max(k32_n, 1),
                                     "achieved_tflops": a32, "frac_of_fp64_peak": a32 / fp64_peak if a32 else None,
                                     "frac_of_fp32_peak": a32 / fp32_peak if a32 else None, "fp32_peak_tflops_measured": fp32_peak,
                                     "note": "80 % of the algorithmic flops are the double-precision SYRK on the FP64 tensor pipe, so the FP64 peak is the bound that applies; the float factor evaluation runs beside it on the FP32 pipe",
                                     "S_rel_err_vs_fp64_build": float(np.max(np.abs(S32 - S64)) / np.max(np.abs(S64))),
                                     "rhs_rel_err_vs_fp64_build": float(np.max(np.abs(rhs32 - rhs64)) / np.max(np.abs(rhs64))),
                                     "tolerance": {"blocks": 1e-4, "reduced_system": 1e-3}}
            # reduced-system solve + landmark back substitution on the device
            ba.solve_reduced(True)
            ctx.enable_kernel_timing(True); ctx.reset_kernel_times()
            t0 = time.perf_counter()
            for _ in range(3):
                ba.solve_reduced(True)
            solve_s = (time.perf_counter() - t0) / 3
            out["solve"] = {"ms_per_call_host_clock": solve_s * 1e3,
                            "kernel_ms_per_launch": {k: ctx.kernel_time(k)[0] / max(ctx.kernel_time(k)[1], 1)
                                                     for k in ("k_chol_band", "k_chol_band_back", "k_chol_skyline", "k_back_substitute")},
                            "parity_note": "unpinned: 1e-7 relative against LAPACK on the restated system (a different factorisation order than Eigen's LLT)"}
            ctx.enable_kernel_timing(False)
            # one whole LM linearisation on the device: parameters in, build + solve, (dx_pose, dx_landmark) back;
            # the reduced system itself never crosses PCIe
            t0 = time.perf_counter()
            for _ in range(5):
                ba.set_parameters(win["pose_delta"], win["points"])
                ba.build_reduced_system(D, True, fetch=False)
                ba.solve_reduced(True)
            lm_s = (time.perf_counter() - t0) / 5
            out["e2e_lm_step"] = {"ms_per_step_host_clock": lm_s * 1e3, "value": win["O"] / lm_s, "unit": "obs/s",
                                  "h2d_bytes_per_step": int(8 * (6 * win["K"] + 3 * win["M"])),
                                  "d2h_bytes_per_step": int(8 * (6 * win["K"] + 3 * win["M"]))}
            # the whole ceres::Solve of Optimizer::RunLocalBA as one device-resident call (vio360_ba_lm_solve)
            ba.lm_solve(win["pose_delta"], win["points"])
            t0 = time.perf_counter()
            gp, gl, sm, tr = ba.lm_solve(win["pose_delta"], win["points"])
            lm_total = time.perf_counter() - t0
            evals = (sm["iterations"] + sm["num_successful_steps"]) * win["O"]      # one candidate cost per iteration + one linearisation per accepted step
            out["e2e_lm_solve"] = {"ms_host_clock": lm_total * 1e3, "iterations": sm["iterations"], "successful_steps": sm["num_successful_steps"],
                                   "termination": sm["termination"], "initial_cost": sm["initial_cost"], "final_cost": sm["final_cost"],
                                   "ms_per_iteration": lm_total * 1e3 / max(sm["iterations"], 1), "factor_evals_per_s": evals / lm_total,
                                   "h2d_bytes": int(8 * (6 * win["K"] + 3 * win["M"])), "d2h_bytes": int(8 * (6 * win["K"] + 3 * win["M"]) + 128 * (sm["iterations"] + sm["num_successful_steps"]))}
            if rank == 0 and not args.skip_cpu and win["O"] <= 50000:
                # the restated ceres::Solve on the host (oracle, numpy + C): same iteration count, accept/reject sequence and final cost
                sys.path.insert(0, os.path.join(ROOT, "oracle"))
                import ba_oracle
                t0 = time.perf_counter()
                xp, xl, osm, otr = ba_oracle.Window(win).lm_solve(win["pose_delta"], win["points"])
                o_s = time.perf_counter() - t0
                out["e2e_lm_solve"]["oracle"] = {"ms": o_s * 1e3, "kind": "port (numpy + C restatement, 1 thread)", "iterations": osm["iterations"], "final_cost": osm["final_cost"],
                                                 "same_accept_reject_sequence": bool(len(tr) == len(otr) and all(int(a[6]) == int(b_[6]) for a, b_ in zip(tr, otr))),
                                                 "final_cost_rel_diff": abs(sm["final_cost"] - osm["final_cost"]) / osm["final_cost"],
                                                 "max_param_diff": float(max(np.abs(gp - xp).max(), np.abs(gl - xl).max()))}
        if peer:
            barrier()
            ba.peer_detach()
        ba.close()
        return out

    if world == 1:
        win3 = syn.make_ba_window(K=10, M=2000, seed=20000, mid_iteration=True)
        res["c3_local_window"] = run(win3, args.ba_steps * 4, 3, False)
    win5 = syn.make_ba_window(K=200, M=50000, degree=20, seed=1000000, mid_iteration=True, loop_radius=6.0)
    r5 = run(win5, args.ba_steps, 3, True, "peer")
    if world > 1:
        # the library-collective variant beside it (same shards, ncclAllReduce of the band + expand kernel)
        rn = run(win5, args.ba_steps, 3, True, "nccl")
        r5["nccl_variant"] = {k: rn[k] for k in ("ms_per_step", "value", "e2e_value", "allreduce_doubles", "exchange")}
        r5["nccl_variant"]["kernel_ms_per_step"] = rn["roofline"]["kernel_ms_per_step"]
        r5["nccl_variant"]["S_trace"] = rn["S_trace"]
    res["c5_large_window"] = r5
    # SURVEY 8-D / 8-E dense variant: 20 000 landmarks of degree 50 (the Schur-dominated regime).  Every landmark spans
    # more than the 24-pose chunk window: pass 1 (k_wide_prepare) + group-pair jobs of the same chunk kernel, no atomics
    wind = syn.make_ba_window(K=200, M=20000, degree=50, seed=1000001, mid_iteration=True, loop_radius=6.0)
    res["dense_window_degree50"] = run(wind, max(2, args.ba_steps // 2), 2, True, "peer")
    res["value"] = r5["value"]; res["n_gpus"] = world
    res["config"] = {"workload": "configs[4]: 200 keyframes, 1M observations, landmark-range sharding, reduced-camera-system blocks summed over NVLink peer memory (NCCL allreduce variant reported beside it)"}
    if rank == 0 and world == 1 and not args.skip_cpu:
        res["cpu_baseline"] = cpu_baseline_ba(syn, win5)
    return res


# ===================================================================== CPU baselines
def cpu_baseline_tracking(ring, pts_h, torch, gpu_out=None, max_pairs=24, budget_s=20.0):
    """cv2.calcOpticalFlowPyrLK = the reference's own CPU implementation of the path
    (FeatureTracker.cpp:238-240), exact arguments, all host threads (and once more on one thread).  The outputs
    of the first calls are kept and compared with the CUDA path's on the same frame pairs -> `parity`."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cv2
    import fe_oracle
    n = min(len(ring[0]), 4)
    pairs = [(ring[0][s].cpu().numpy(), ring[1][s].cpu().numpy(), pts_h[s]) for s in range(n)]
    cv_out = [fe_oracle.cv2_lk_track(*p) for p in pairs]
    t0 = time.perf_counter(); done = 0
    while done < max_pairs and time.perf_counter() - t0 < budget_s:
        fe_oracle.cv2_lk_track(*pairs[done % n]); done += 1
    dt = time.perf_counter() - t0
    cv2.setNumThreads(1)
    t1 = time.perf_counter(); done1 = 0
    while done1 < 8 and time.perf_counter() - t1 < 8.0:
        fe_oracle.cv2_lk_track(*pairs[done1 % n]); done1 += 1
    dt1 = time.perf_counter() - t1
    cv2.setNumThreads(-1)
    base = {"value": done / dt, "unit": "frames/s", "cores": int(cv2.getNumThreads()), "kind": "reference",
            "sample": "%d calls of cv2.calcOpticalFlowPyrLK %s (both pyramids rebuilt per call, as the reference does) on the same synthetic frames" % (done, cv2.__version__),
            "host_cpus": os.cpu_count(), "single_thread_value": done1 / dt1}
    parity = None
    if gpu_out is not None:
        gxy, gst, gerr = gpu_out
        mism, maxd, maxe, ntr, moved = 0, 0.0, 0.0, 0, []
        for s, (cxy, cst, cerr) in enumerate(cv_out):
            mism += int((gst[s] != cst).sum())
            ok = (cst == 1) & (gst[s] == 1)
            ntr += int(ok.sum())
            if ok.any():
                maxd = max(maxd, float(np.abs(gxy[s] - cxy)[ok].max()))
                maxe = max(maxe, float(np.abs(gerr[s][ok] - cerr[ok]).max()))   # err of untracked points is undefined on both sides
                moved.append(float(np.median(np.linalg.norm(cxy - pts_h[s], axis=1)[ok])))
        parity = {"lk_vs_cv2": {"pairs": n, "points": int(n * pts_h.shape[1]), "frames": "consecutive rendered frames t=0,1 at %dx%d (the timed inputs)" % (W, H),
                                "status_mismatches": mism, "status_equal": mism == 0, "tracked_both": ntr,
                                "max_abs_dxy_px": maxd, "tolerance_px": 0.05, "within_tolerance": maxd < 0.05,
                                "max_abs_derr": maxe, "median_motion_px": float(np.median(moved)) if moved else None}}
    return base, parity


def cpu_baseline_detect(ring, gpu_corners):
    """cv2.goodFeaturesToTrack with the reference's arguments (FeatureTracker.cpp:212-223) on the frames
    bench_detect ran: timing (all host threads) and corner-by-corner comparison with the CUDA path."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cv2
    import fe_oracle
    frames = [ring[0][s].cpu().numpy() for s in range(len(gpu_corners))]
    mask = fe_oracle.base_mask(W, H)
    cv_c = [fe_oracle.cv2_good_features(f, mask, 1000) for f in frames]
    t0 = time.perf_counter(); done = 0
    while done < 6 and time.perf_counter() - t0 < 10.0:
        fe_oracle.cv2_good_features(frames[done % len(frames)], mask, 1000); done += 1
    dt = time.perf_counter() - t0
    eq = [bool(np.array_equal(a, c)) for a, c in zip(gpu_corners, cv_c)]
    base = {"value": done / dt, "unit": "frames/s", "cores": int(cv2.getNumThreads()), "kind": "reference",
            "sample": "%d calls of cv2.goodFeaturesToTrack %s (1000, 0.01, 30, polar & boundary mask, block 3) on the benchmark frames" % (done, cv2.__version__)}
    par = {"frames": len(frames), "corners": [int(len(c)) for c in cv_c], "corners_identical": all(eq)}
    return base, par


def cpu_baseline_ba(syn, win5):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ba_oracle
    nthreads = min(4, os.cpu_count() or 1)     # the reference's options.num_threads = 4 (Optimizer.cpp:79)
    sub_M = 5000                                # bounded sample: the first 5000 landmarks (100k observations)
    keep = win5["point_idx"] < sub_M
    sub = dict(win5, M=sub_M, O=int(keep.sum()), pose_idx=np.ascontiguousarray(win5["pose_idx"][keep]),
               point_idx=np.ascontiguousarray(win5["point_idx"][keep]), obs_u=np.ascontiguousarray(win5["obs_u"][keep]),
               obs_v=np.ascontiguousarray(win5["obs_v"][keep]), outlier=np.ascontiguousarray(win5["outlier"][keep]),
               points=np.ascontiguousarray(win5["points"][:sub_M]), point_fixed=np.ascontiguousarray(win5["point_fixed"][:sub_M]))
    Wo = ba_oracle.Window(sub)
    r, Jp, Jl, _ = Wo.evaluate(sub["pose_delta"], sub["points"], nthreads=nthreads)
    D = Wo.lm_diagonal(Jp, Jl)
    t0 = time.perf_counter(); reps = 0
    while reps < 20 and time.perf_counter() - t0 < 20.0:
        r, Jp, Jl, _ = Wo.evaluate(sub["pose_delta"], sub["points"], nthreads=nthreads)
        Wo.schur(r, Jp, Jl, D); reps += 1
    dt = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for _ in range(3):
        r, Jp, Jl, _ = Wo.evaluate(sub["pose_delta"], sub["points"], nthreads=nthreads, hoisted=True)
        Wo.schur(r, Jp, Jl, D)
    dth = (time.perf_counter() - t0) / 3
    base = {"value": sub["O"] / dt, "unit": "obs/s", "cores": nthreads, "kind": "port",
            "sample": "%d observations (first %d landmarks of the 1M window), per-factor 6-SVD pose chain as the reference executes it + Huber + Schur elimination (C restatement, Schur single-threaded)" % (sub["O"], sub_M),
            "hoisted_pose_chain_value": sub["O"] / dth}
    # All host cores, evaluation AND elimination threaded: the landmark range is cut into one contiguous part per
    # core, every part evaluates its factors and eliminates its landmarks into a private S / rhs (what Ceres'
    # SchurEliminator does per thread with its chunk loop, schur_eliminator_impl.h:190-262), then the parts are added.
    from concurrent.futures import ThreadPoolExecutor
    ncores = os.cpu_count() or 1
    bounds = np.linspace(0, sub_M, ncores + 1).astype(np.int64)
    parts = []
    for c in range(ncores):
        lo, hi = int(bounds[c]), int(bounds[c + 1])
        k = (sub["point_idx"] >= lo) & (sub["point_idx"] < hi)
        w = dict(sub, M=hi - lo, O=int(k.sum()), pose_idx=np.ascontiguousarray(sub["pose_idx"][k]),
                 point_idx=np.ascontiguousarray(sub["point_idx"][k] - lo), obs_u=np.ascontiguousarray(sub["obs_u"][k]),
                 obs_v=np.ascontiguousarray(sub["obs_v"][k]), outlier=np.ascontiguousarray(sub["outlier"][k]),
                 points=np.ascontiguousarray(sub["points"][lo:hi]), point_fixed=np.ascontiguousarray(sub["point_fixed"][lo:hi]))
        Wc = ba_oracle.Window(w)
        rc, Jpc, Jlc, _ = Wc.evaluate(w["pose_delta"], w["points"])
        parts.append((Wc, w, Wc.lm_diagonal(Jpc, Jlc)))

    def one(p):
        Wc, w, Dc = p
        rc, Jpc, Jlc, _ = Wc.evaluate(w["pose_delta"], w["points"], nthreads=1)
        return Wc.schur(rc, Jpc, Jlc, Dc)[:2]
    with ThreadPoolExecutor(ncores) as ex:
        list(ex.map(one, parts))
        t0 = time.perf_counter(); reps = 0
        while reps < 20 and time.perf_counter() - t0 < 10.0:
            res = list(ex.map(one, parts))
            S = sum(x[0] for x in res); rhs = sum(x[1] for x in res); reps += 1
        dta = (time.perf_counter() - t0) / reps
    base.update(all_cores_value=sub["O"] / dta, all_cores=ncores,
                all_cores_note="same sample, one landmark range per host core (evaluation + elimination + the sum of the per-core systems)")
    return base


# ================================================================= reference arm
def run_reference(args):
    rank, local_rank, world = dist_env()
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cv2
    import fe_oracle
    syn = importlib.import_module(PKG + ".synthetic")
    n_pairs = 4
    try:
        import torch
        use_cuda = torch.cuda.is_available()
    except Exception:
        use_cuda = False
    frames = []
    for s in range(n_pairs):
        for t in (0, 1):
            if use_cuda:
                frames.append(syn.render_erp_frame(W, H, t, seed=360 + s, n_waves=24, sigma_omega=120.0, xp=torch, device="cuda").cpu().numpy())
            else:
                frames.append(syn.render_erp_frame(W, H, t, seed=360 + s, n_waves=8, sigma_omega=120.0))
    pts = [syn.make_track_points(W, H, NFEAT, seed=1000 + s) for s in range(n_pairs)]

    def step():
        for s in range(n_pairs):
            fe_oracle.cv2_lk_track(frames[2 * s], frames[2 * s + 1], pts[s])
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = n_pairs * args.steps / dt
    sample = "each step = %d calls of cv2.calcOpticalFlowPyrLK %s with the reference's arguments (FeatureTracker.cpp:238-240)" % (n_pairs, cv2.__version__)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "configs[1]: ERP tracking 3840x1920, 4-level pyramid, 1000 features/frame, 21x21 KLT",
                       "frames_per_step": n_pairs, "note": "rank 0 only; the reference has no multi-GPU or multi-process path"},
            "cpu_baseline": {"value": value, "unit": "frames/s", "cores": int(cv2.getNumThreads()), "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="independent sequences (frame pairs) per GPU per step")
    ap.add_argument("--ba-steps", type=int, default=20)
    ap.add_argument("--skip-ba", action="store_true")
    ap.add_argument("--skip-detect", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-c4", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
