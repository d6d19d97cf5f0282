#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box with `ncu -i`) into the few counters DESIGN.md argues from.
usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rN_ncu_<name>.txt"""
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum|launch__(registers_per_thread|grid_size|block_size|"
    r"shared_mem_per_block_dynamic)|sm__cycles_(active|elapsed)\.avg|"
    r"l1tex__data_pipe_lsu_wavefronts(_mem_shared(_op_(ld|st))?|_mem_lgds)?\.(sum|avg)(\.pct_of_peak_sustained_elapsed)?|"
    r"l1tex__data_pipe_tex_wavefronts(_mem_\w+)?\.(sum|avg)(\.pct_of_peak_sustained_elapsed)?|l1tex__tex_writeback_active\.avg\.pct_of_peak_sustained_elapsed|"
    r"l1tex__texin_\w+\.(sum|avg)(\.pct_of_peak_sustained_elapsed)?|smsp__inst_executed_pipe_tex\.sum|sm__inst_executed_pipe_tex\.avg\.pct_of_peak_sustained_active|"
    r"l1tex__lsu_writeback_active\.avg\.pct_of_peak_sustained_elapsed|l1tex__data_bank_(reads|writes)\.avg\.pct_of_peak_sustained_elapsed|"
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared(_op_(ld|st))?\.sum|"
    r"sm__inst_executed_pipe_(alu|fma|lsu|xu|adu|uniform)\.avg\.pct_of_peak_sustained_active|"
    r"sm__issue_active\.avg\.pct_of_peak_sustained_elapsed|smsp__inst_executed\.sum|smsp__sass_inst_executed_op_(shared_ld|shared_st|global_ld)\.sum|"
    r"sm__warps_active\.avg\.per_cycle_active|dram__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"lts__throughput\.avg\.pct_of_peak_sustained_elapsed|sm__throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"smsp__average_warps_issue_stalled_\w+_per_issue_active\.ratio)$")


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print("== kernel %s  grid %s block %s" % (d.get("Kernel Name"), d.get("Grid Size"), d.get("Block Size")))
        for h, u in zip(hdr, units):
            if KEEP.match(h):
                print("%-95s %-10s %s" % (h, u, d[h]))


if __name__ == "__main__":
    main()
