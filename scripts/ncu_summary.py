#!/usr/bin/env python3
"""Summarise a gpurun ncu capture into profiles/: launch-list shares + key metrics per kernel.
  python scripts/ncu_summary.py <launches.csv> <prof.ncu-rep> <out.md> [title]"""
import collections
import csv
import subprocess
import sys

launches, rep, out = sys.argv[1:4]
title = sys.argv[4] if len(sys.argv) > 4 else "ncu summary"
rows = list(csv.reader(l for l in open(launches) if l.startswith('"')))
h = rows[0]
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[1:]:
    n = r[ki].split("(")[0]
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += float(r[vi].replace(",", "")) / 1e3
tot = sum(a[1] for a in agg.values())
md = ["# %s\n" % title, "## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`; cold-cache, serialised: compare shares)\n",
      "| kernel | launches | total us | share |", "|---|---|---|---|"]
for n, a in sorted(agg.items(), key=lambda x: -x[1][1]):
    md.append("| %s | %d | %.1f | %.1f%% |" % (n, a[0], a[1], 100 * a[1] / tot))
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"]).decode()
rr = list(csv.reader(raw.splitlines()))
hh = rr[0]
want = [("gpu__time_duration.sum", "time"), ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("smsp__inst_executed.sum", "warp insts"), ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "global ld sectors"),
        ("l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "global ld requests"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_scoreboard"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall short_scoreboard"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall lg_throttle")]
idx = [(hh.index(w), lab) for w, lab in want if w in hh]
units = rr[1]
md += ["", "## `ncu --set full` (per launch)\n", "| kernel | " + " | ".join(l for _, l in idx) + " |", "|---|" + "---|" * len(idx)]
seen = collections.Counter()
for r in rr[2:]:
    name = r[hh.index("Kernel Name")].split("(")[0]
    seen[name] += 1
    if seen[name] > 1:
        name += " [launch %d]" % seen[name]
    md.append("| %s | " % name + " | ".join("%s %s" % (r[i][:12], units[i]) for i, _ in idx) + " |")
open(out, "w").write("\n".join(md) + "\n")
print("\n".join(md))
