#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into a small text file for profiles/.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/out.txt"""
import csv
import io
import re
import subprocess
import sys

PATS = [r'^gpu__time_duration\.sum$', r'^launch__(grid_size|block_size|registers_per_thread|waves_per_multiprocessor|occupancy_limit_\w+)$',
        r'^sm__cycles_elapsed\.avg\.per_second$', r'^sm__inst_executed_pipe_(fma|xu|alu|lsu|uniform|fp64|tc)\.avg\.pct_of_peak_sustained_active$',
        r'^sm__pipe_(fma|alu|fp64|tensor)_cycles_active\.avg\.pct_of_peak_sustained_active$',
        r'^smsp__issue_active\.avg\.pct_of_peak_sustained_active$', r'^sm__inst_executed\.avg\.per_cycle_active$',
        r'^smsp__warps_(active|eligible)\.avg\.per_cycle_active$', r'^sm__warps_active\.avg\.pct_of_peak_sustained_active$',
        r'^smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio$', r'^dram__bytes_(read|write)\.sum$',
        r'^gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed$', r'^smsp__inst_executed\.sum$',
        r'^lts__t_sector_hit_rate\.pct$', r'^l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$']


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write(f"# ncu --set full --clock-control none summary of {rep}\n")
        for k, r in enumerate(rows[2:]):
            d = dict(zip(hdr, zip(units, r)))
            f.write(f"\n## launch {k}: {d.get('Kernel Name', ('', '?'))[1]}\n")
            for h in sorted(d):
                if any(re.search(p, h) for p in PATS):
                    f.write(f"{h:90s} {d[h][1]:>18s} {d[h][0]}\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
