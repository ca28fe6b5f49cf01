"""Turns `ncu -i X.ncu-rep --page raw --csv` output into the short per-launch summary that
is committed under profiles/ (the .ncu-rep itself stays in gpurun_out/, which is scratch)."""
import csv
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__maximum_warps_per_active_cycle_pct", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "lts__t_sector_hit_rate.pct", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct",
        "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct",
        "smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct", "smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct",
        "smsp__warp_issue_stalled_wait_per_warp_active.pct", "smsp__warp_issue_stalled_not_selected_per_warp_active.pct"]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    for r in rows[2:]:
        print("kernel: " + r[ki])
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print("  %-78s %18s %s" % (w, r[i], units[i]))
        print()


def to_bytes(value, unit):
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(value.replace(",", "")) * mult[unit]


def traffic(path, kernel, source, capture, w, h, batch, points, out):
    """ncu_summary.py raw.csv --traffic <kernel> <source under csrc/> <capture name> W H batch points out.json:
    records dram__bytes_read.sum + dram__bytes_write.sum of the kernel's captured launch, tied to the sha1 of the
    kernel's source file and to the workload, in the file bench.py's ncu_traffic() reads."""
    import hashlib
    import json
    import os
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    r = [x for x in rows[2:] if kernel in x[ki]][0]
    ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    total = to_bytes(r[ir], units[ir]) + to_bytes(r[iw], units[iw])
    root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
    src = os.path.join(root, "360_visual_inertial_odometry_b200", "csrc", source)
    rec = json.load(open(out)) if os.path.exists(out) else {}
    rec[kernel] = {"bytes_per_launch": total, "source": source, "source_sha1": hashlib.sha1(open(src, "rb").read()).hexdigest(),
                   "capture": capture, "workload": {"W": int(w), "H": int(h), "batch": int(batch), "points": int(points)}}
    json.dump(rec, open(out, "w"), indent=1, sort_keys=True)
    print(kernel, total, "bytes per launch ->", out)


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "--traffic":
        traffic(sys.argv[1], *sys.argv[3:])
    else:
        main(sys.argv[1])
