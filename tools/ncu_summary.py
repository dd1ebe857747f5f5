#!/usr/bin/env python3
"""Summarise one kernel of an .ncu-rep (ncu --set full) into the plain-text metric list kept under profiles/:
usage: ncu_summary.py <file.ncu-rep> "<header comment>" > profiles/rNN_<kernel>_ncu_raw.txt
Keeps the metrics the design notes argue with: duration, DRAM bytes, cache hit rates, launch shape and occupancy
limits, issue-slot and pipe utilisation, warp-stall reasons per issued instruction, executed instructions."""
import csv
import re
import subprocess
import sys

KEEP = re.compile(r"^(dram__bytes_(read|write)\.sum(\.per_second|\.pct_of_peak_sustained_elapsed)?|gpu__time_duration\.sum|"
                  r"gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|l1tex__t_sector_hit_rate\.pct|lts__t_sector_hit_rate\.pct|"
                  r"launch__(block_size|grid_size|registers_per_thread|shared_mem_per_block_dynamic|waves_per_multiprocessor|"
                  r"occupancy_limit_\w+)|sm__icc_request_hit_rate\.pct|sm__inst_executed_pipe_(alu|fma|lsu|xu)\.avg\.pct_of_peak_sustained_active|"
                  r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
                  r"smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio|smsp__inst_executed\.sum|"
                  r"smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__thread_inst_executed_per_inst_executed\.ratio|"
                  r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum|smsp__inst_executed_op_shared_(ld|st)\.sum)$")


def main():
    rep, header = sys.argv[1], sys.argv[2]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    print("# " + header)
    for name, unit, val in sorted(zip(hdr, units, vals)):
        if KEEP.match(name) and val not in ("", "n/a"):
            print(name, unit, val.replace(",", ""))


if __name__ == "__main__":
    main()
