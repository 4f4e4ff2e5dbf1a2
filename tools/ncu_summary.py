#!/usr/bin/env python
"""Summarise ncu output for profiles/: `launches` (CSV of gpu__time_duration.sum per launch ->
per-kernel totals and shares) and `raw` (a .ncu-rep -> the handful of counters DESIGN.md cites).
Runs where ncu is installed (no GPU needed to read a report)."""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "smsp__average_warp_latency_issue_stalled", "smsp__average_warps_issue_stalled",
    "smsp__pcsamp_warps_issue_stalled", "derived__smsp__sass_thread_inst_executed_op",
    "smsp__sass_average_branch_targets_threads_uniform.pct", "sm__cycles_active.avg", "lts__t_sectors_srcunit_tex_op_read.sum",
    "lts__t_sectors_srcunit_tex_lookup_miss.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "smsp__warp_issue_stalled",
]


def launches(path):
    rows = [l for l in open(path) if not l.startswith("==")]
    r = list(csv.reader(io.StringIO("".join(rows))))
    h = r[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    d = collections.OrderedDict()
    for row in r[1:]:
        if len(row) <= vi:
            continue
        d.setdefault(row[ki], []).append(float(row[vi].replace(",", "")))
    tot = sum(sum(v) for v in d.values())
    print(f"# per-kernel device time from {path} ({r[1][ui]}; cold-cache, serialised under ncu: compare SHARES)")
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        print(f"{sum(v) / tot:7.4f}  n={len(v):4d}  total={sum(v):14.0f}  avg={sum(v) / len(v):12.0f}  {k[:110]}")


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    r = list(csv.reader(io.StringIO(out)))
    h, units = r[0], r[1]
    for row in r[2:]:
        name = row[h.index("Kernel Name")] if "Kernel Name" in h else "?"
        print(f"# {name[:120]}")
        for i, n in enumerate(h):
            if any(n.startswith(k) for k in KEYS):
                print(f"{n:90s} {units[i]:14s} {row[i]}")


if __name__ == "__main__":
    {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
