#!/usr/bin/env python
"""profiles/r02_traffic.json from the `ncu --set full` captures of a step's three kernels (read here, no GPU
needed): DRAM bytes per launch, issue-slot use, lanes per instruction -- what bench.py's `roofline.traffic`
quotes.  usage: make_traffic.py <update.ncu-rep> <rebuild.ncu-rep> <suffstats_tile.ncu-rep> > profiles/r02_traffic.json"""
import csv
import json
import subprocess
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def first_row(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2]

    def get(name, scale=True):
        i = hdr.index(name)
        v = float(data[i].replace(",", ""))
        return v * UNIT.get(units[i], 1.0) if scale else v

    return {
        "kernel": data[hdr.index("Kernel Name")],
        "grid": int(float(data[hdr.index("launch__grid_size")])),
        "registers": int(float(data[hdr.index("launch__registers_per_thread")])),
        "ms": get("gpu__time_duration.sum"),
        "dram_bytes_read": get("dram__bytes_read.sum"),
        "dram_bytes_write": get("dram__bytes_write.sum"),
        "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active", False),
        "lanes_per_inst": get("smsp__thread_inst_executed_per_inst_executed.ratio", False),
        "warps_active_pct": get("sm__warps_active.avg.pct_of_peak_sustained_active", False),
        "warp_instructions": get("smsp__inst_executed.sum", False),
        "l1_hit_pct": get("l1tex__t_sector_hit_rate.pct", False),
        "l2_hit_pct": get("lts__t_sector_hit_rate.pct", False),
    }


def main():
    upd, reb, jd = (first_row(r) for r in sys.argv[1:4])
    out = {
        "kernel": upd["kernel"], "workload": "bal8", "n_sites_per_gpu": 12500000,
        "dram_bytes_read": upd["dram_bytes_read"], "dram_bytes_write": upd["dram_bytes_write"],
        "traffic_bytes_per_launch": upd["dram_bytes_read"] + upd["dram_bytes_write"],
        "issue_active_pct": upd["issue_active_pct"], "lanes_per_inst": upd["lanes_per_inst"],
        "warps_active_pct": upd["warps_active_pct"],
        "source": "ncu --set full --clock-control none, one launch each (tools/sessions/run_gpu_final.sh, run_gpu_final2.sh); summaries in "
                  "profiles/r02_ncu_summary_final.txt",
        "update_kernel": upd,
        "other_kernels": {"csr_rebuild_kernel": reb, "suffstats_tile_kernel": jd},
    }
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
