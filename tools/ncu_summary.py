#!/usr/bin/env python
"""Summarise an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the handful of
metrics the roofline discussion needs.  Usage: tools/ncu_summary.py file.ncu-rep [regex]"""
import csv
import io
import re
import subprocess
import sys

KEYS = [r"gpu__time_duration\.sum", r"dram__bytes_read\.sum$", r"dram__bytes_write\.sum$",
        r"gpu__dram_throughput\.avg\.pct", r"dram__throughput\.avg\.pct_of_peak_sustained_elapsed$",
        r"launch__registers_per_thread", r"launch__occupancy_limit", r"sm__warps_active\.avg\.pct",
        r"sm__throughput\.avg\.pct", r"lts__throughput\.avg\.pct", r"l1tex__throughput\.avg\.pct",
        r"sm__inst_executed_pipe_fp64", r"sm__pipe_fp64_cycles_active", r"sm__inst_executed_pipe_lsu",
        r"smsp__inst_executed\.sum$", r"sm__cycles_elapsed\.max", r"smsp__cycles_active\.avg",
        r"smsp__issue_active\.avg\.pct", r"smsp__warp_issue_stalled.*_per_warp_active\.pct",
        r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$", r"smsp__inst_executed_op_shared",
        r"lts__t_sectors_op_red\.sum$", r"launch__shared_mem_per_block", r"launch__grid_size", r"launch__block_size",
        r"sm__sass_thread_inst_executed_op_d(fma|add|mul)_pred_on\.sum$", r"smsp__sass_thread_inst_executed_op_d"]


def main():
    rep = sys.argv[1]
    extra = sys.argv[2] if len(sys.argv) > 2 else None
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    pats = [re.compile(k) for k in KEYS] + ([re.compile(extra)] if extra else [])
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"== {d.get('Kernel Name')}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
        for h, u, v in zip(hdr, units, r):
            name = h.split(".", 2)[-1] if h.count(".") >= 2 and h.split(".")[1].startswith("Triage") else h
            if any(p.search(h) for p in pats):
                print(f"  {h} [{u}] = {v}")


if __name__ == "__main__":
    main()
