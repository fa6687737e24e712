#!/usr/bin/env python3
"""Transposed per-launch summary of an `ncu -i X.ncu-rep --page raw --csv` dump: one row per metric, one column per launch.
python tools/ncu_summary.py raw.csv > profiles/ncu_<what>_summary.txt"""
import csv
import sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread", "launch__grid_size",
           "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_active",
           "lts__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio", "memory_l2_theoretical_sectors_local",
           "l1tex__t_bytes.sum", "l1tex__t_sectors.sum", "lts__t_bytes.sum", "lts__t_sectors.sum", "lts__t_sectors_srcunit_tex.sum",
           "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "sm__cycles_elapsed.max", "gpc__cycles_elapsed.max"]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]


def col(name):
    """exact name first: the raw page repeats some metrics under a section prefix, with empty cells"""
    c = [i for i, h in enumerate(hdr) if h == name] or [i for i, h in enumerate(hdr) if h.endswith("." + name)]
    return c[0] if c else -1


def num(r, i):
    try:
        return float(r[i].replace(",", ""))
    except (ValueError, IndexError):
        return float("nan")


def to_bytes(v, unit):
    return v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)


def to_seconds(v, unit):
    return v * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(unit, 1e-9)


k = col("Kernel Name")
print("%-90s %-16s %s" % ("Kernel Name", "", " | ".join(r[k][:24] for r in data)))
for m in METRICS:
    i = col(m)
    if i < 0:
        continue
    print("%-90s %-16s %s" % (m, units[i], " | ".join(r[i] for r in data)))


# achieved bandwidths per launch: bytes through L1 / L2 / DRAM over the launch's duration
t = col("gpu__time_duration.sum")
if t >= 0:
    for label, names in (("achieved L1 traffic", ["l1tex__t_bytes.sum"]), ("achieved L2 traffic", ["lts__t_bytes.sum"]),
                         ("achieved DRAM traffic", ["dram__bytes_read.sum", "dram__bytes_write.sum"])):
        cols = [col(n) for n in names]
        if min(cols) < 0:
            continue
        out = []
        for r in data:
            secs = to_seconds(num(r, t), units[t])
            b = sum(to_bytes(num(r, c), units[c]) for c in cols)
            out.append("%.1f" % (b / secs / 1e9) if secs > 0 else "nan")
        print("%-90s %-16s %s" % (label, "GB/s", " | ".join(out)))
