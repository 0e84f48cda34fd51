#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into the handful of metrics the
roofline discussion in DESIGN.md uses.  usage: ncu_summary.py <rep> [<rep>...]"""
import csv
import subprocess
import sys

WANT = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__maximum_warps_per_active_cycle_pct",
]


def main():
    for rep in sys.argv[1:]:
        out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True,
                             text=True).stdout
        rows = list(csv.reader(out.splitlines()))
        hdr, units, data = rows[0], rows[1], rows[2:]
        print("# " + rep)
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print("%-68s %-8s %s" % (w, units[i], " | ".join(r[i] for r in data)))
        # stall reasons (warp-state sampling), top few
        stalls = [(h, i) for i, h in enumerate(hdr)
                  if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")]
        vals = sorted(((float(data[0][i] or 0), h) for h, i in stalls), reverse=True)[:8]
        for v, h in vals:
            print("stall %-62s %.3f" % (h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))
        print()


if __name__ == "__main__":
    main()
