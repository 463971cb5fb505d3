#!/usr/bin/env python
"""Compact per-launch table from an .ncu-rep:  ncu_brief.py <file.ncu-rep>"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h = rows[0]
cols = {"dur_us": "gpu__time_duration.sum", "inst": "smsp__inst_executed.sum", "sm%": "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dramR_MB": "dram__bytes_read.sum", "dramW_MB": "dram__bytes_write.sum", "occ%": "sm__warps_active.avg.pct_of_peak_sustained_active",
        "regs": "launch__registers_per_thread", "thr/inst": "smsp__thread_inst_executed_per_inst_executed.ratio",
        "ipc": "sm__inst_executed.avg.per_cycle_active",
        "st_long": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "st_lg": "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "st_bar": "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "st_short": "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "st_wait": "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "st_math": "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "st_mio": "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "st_branch": "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "st_noinst": "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "l2hit%": "lts__t_sector_hit_rate.pct", "l1hit%": "l1tex__t_sector_hit_rate.pct"}
for r in rows[2:]:
    name = r[h.index("Kernel Name")].split("(")[0]
    grid = r[h.index("Grid Size")]
    out = []
    for k, m in cols.items():
        if m in h:
            try: out.append(f"{k}={float(r[h.index(m)].replace(',', '')):.4g}")
            except ValueError: out.append(f"{k}={r[h.index(m)]}")
    print(name, grid, " ".join(out))
