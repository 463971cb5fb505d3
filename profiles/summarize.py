#!/usr/bin/env python
"""Turns gpurun_out/launches.csv (ncu --metrics gpu__time_duration.sum) and an .ncu-rep (ncu --set full) into the
markdown summaries committed under profiles/.   usage: summarize.py <launches.csv> <prof.ncu-rep> <out.md> [title]"""
import collections
import csv
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "sm__inst_executed.sum", "smsp__inst_executed.sum", "smsp__inst_executed.avg.per_cycle_active",
    # pipe-level evidence: instructions per pipe, IPC, issue-slot use
    "sm__inst_executed.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum",
    "sm__inst_executed_pipe_fmaheavy.sum", "sm__inst_executed_pipe_lsu.sum", "sm__inst_executed_pipe_xu.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
]


def launches(path):
    lines = open(path).read().splitlines()
    start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
    agg = collections.OrderedDict()
    for r in csv.DictReader(lines[start:]):
        agg.setdefault(r["Kernel Name"].split("(")[0], []).append(float(r["Metric Value"]))
    tot = sum(sum(v) for v in agg.values())
    out = ["| kernel | launches | avg us | share of listed time |", "|---|---:|---:|---:|"]
    for k, v in agg.items():
        out.append(f"| `{k}` | {len(v)} | {sum(v) / len(v) / 1000:.1f} | {100 * sum(v) / tot:.1f}% |")
    return out


def full(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0]
        out += [f"", f"**`{name}`** (launch id {r[0]})", "", "| metric | value | unit |", "|---|---:|---|"]
        for m in METRICS:
            if m in hdr:
                out.append(f"| {m} | {r[hdr.index(m)]} | {units[hdr.index(m)]} |")
    return out


if __name__ == "__main__":
    lc, rep, dst = sys.argv[1:4]
    title = sys.argv[4] if len(sys.argv) > 4 else dst
    md = [f"# {title}", "", "## Launch list (`ncu --metrics gpu__time_duration.sum --clock-control none`; cold-cache, serialised: compare shares)", ""]
    md += launches(lc)
    md += ["", "## `ncu --set full --clock-control none` of the top kernels (per launch)"]
    md += full(rep)
    open(dst, "w").write("\n".join(md) + "\n")
    print("\n".join(md))
