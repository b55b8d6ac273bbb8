#!/usr/bin/env python
"""Text summary of an ncu report for profiles/: key raw metrics per profiled launch plus the
hottest source lines (needs -lineinfo).  usage: ncu_summary.py report.ncu-rep [n_lines] > profiles/x.txt"""
import csv
import os
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__grid_size",
    "launch__block_size", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
    "sm__maximum_warps_per_active_cycle_pct", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
]


def main():
    rep = sys.argv[1]
    ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    print(f"# {os.path.basename(rep)}: {len(rows) - 2} profiled launch(es)  (ncu --set full --clock-control none; cold cache, serialised)")
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print(f"\n## launch id {r[0]}: {name[:90]}")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:62s} {r[i]:>16s} {units[i]}")
    here = os.path.dirname(os.path.abspath(__file__))
    print("\n## hottest source lines (stall samples / executed warp instructions)")
    sys.stdout.flush()
    subprocess.run([sys.executable, os.path.join(here, "ncu_lines.py"), rep, str(ntop)])


if __name__ == "__main__":
    main()
