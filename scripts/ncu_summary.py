#!/usr/bin/env python
"""Summarise ncu output into profiles/ (tracked).

  python scripts/ncu_summary.py launches <launches.csv>         # per-kernel time shares
  python scripts/ncu_summary.py kernel <report.ncu-rep>          # key metrics + executed SASS mix
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        a = agg.setdefault(r[ki][:90], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    ours = sum(a[1] for k, a in agg.items() if ("alnr::" in k or "_kernel" in k) and "at::" not in k and "peak_" not in k)
    print("# per-kernel device time (ncu --metrics gpu__time_duration.sum, cold-cache, serialised)")
    print("# share_all = of every launch in the process; share_path = of the alnr:: hot-path kernels only")
    print("%-92s %5s %12s %9s %10s" % ("kernel", "n", "total_ms", "share_all", "share_path"))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        path = ("%9.1f%%" % (100 * a[1] / ours)) if (("alnr::" in k or "_kernel" in k) and "at::" not in k and "peak_" not in k) else "         -"
        print("%-92s %5d %12.3f %8.1f%% %s" % (k, a[0], a[1] / 1e6, 100 * a[1] / tot, path))


def kernel(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print("## kernel:", name)
        for i, n in enumerate(hdr):
            if n in KEYS:
                print("%-86s %-12s %s" % (n, units[i], r[i]))
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    if len(rows) < 3:
        return
    hdr = rows[1]
    ia, ie = hdr.index("Source"), hdr.index("Instructions Executed")
    agg, tot = collections.Counter(), 0
    for r in rows[2:]:
        if len(r) <= ie:
            continue
        m = re.match(r"^(@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)\b", r[ia].strip())
        if not m:
            continue
        try:
            n = int(r[ie])
        except ValueError:
            continue
        agg[m.group(2).split(".")[0]] += n
        tot += n
    print("## executed warp-instructions by opcode (source page), total %d" % tot)
    for k, v in agg.most_common(16):
        print("%-12s %14d %6.2f%%" % (k, v, 100.0 * v / tot))


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel}[sys.argv[1]](sys.argv[2])
