#!/usr/bin/env python
"""Summarises an .ncu-rep (raw page + source page) into a small text file for profiles/.

    python tools/ncu_summary.py gpurun_out/r1_thread.ncu-rep profiles/r1_thread_summary.txt
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_elapsed", "smsp__issue_active.avg.pct",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__inst_executed.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "smsp__average_warp_latency_per_inst_issued", "smsp__warps_eligible.avg",
        "smsp__average_warps_issue_stalled"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units = rows[0], rows[1]
    lines = []
    for data in rows[2:]:
        name = data[head.index("Kernel Name")] if "Kernel Name" in head else "?"
        lines.append("kernel: %s" % name)
        for n, u, v in zip(head, units, data):
            if any(k in n for k in KEYS) and v not in ("", "0"):
                lines.append("  %-95s %-14s %s" % (n, u, v))
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = rows[1]
    d = [r for r in rows[2:] if len(r) == len(h)]
    isamp, iex = h.index("# Samples"), h.index("Instructions Executed")
    tot = sum(int(r[isamp]) for r in d) or 1
    lines.append("")
    lines.append("SASS instructions: %d (%d KB); warp-level instructions executed: %d; samples: %d" % (len(d), len(d) * 16 // 1024, sum(int(r[iex]) for r in d), tot))
    stalls = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
    agg = {c: sum(int(r[h.index(c)] or 0) for r in d) for c in stalls}
    lines.append("warp-state samples by reason (all samples):")
    for c, v in sorted(agg.items(), key=lambda kv: -kv[1])[:10]:
        lines.append("  %-28s %6.2f %%" % (c, 100.0 * v / tot))
    hot = [r for r in d if int(r[iex]) > 0.05 * max(int(x[iex]) for x in d)]
    lines.append("instructions executed >= 5 %% as often as the hottest one: %d (%d KB of code)" % (len(hot), len(hot) * 16 // 1024))
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[-16:]))


if __name__ == "__main__":
    main()
