#!/usr/bin/env python
"""Summarise an ncu --set full report: python tools/ncu_summary.py report.ncu-rep [regex ...]"""
import csv, io, re, subprocess, sys
rep = sys.argv[1]
pats = sys.argv[2:] or [
    r"^gpu__time_duration\.sum$", r"^dram__bytes_(read|write)\.sum$", r"dram__throughput\.avg\.pct", r"^sm__inst_executed\.sum$",
    r"^smsp__inst_executed\.avg\.per_cycle_active$", r"^sm__inst_executed_pipe_(alu|fma|fmaheavy|lsu|xu|uniform|adu|cbu)\.sum$",
    r"^sm__warps_active\.avg\.pct_of_peak", r"issue_active\.avg\.pct", r"^sm__throughput\.avg\.pct", r"^launch__(registers_per_thread|grid_size|block_size|shared_mem_per_block_dynamic|occupancy_limit.*)$",
    r"^smsp__average_warps?_issue_stalled_.*ratio$", r"^smsp__average_warp_latency_issue_stalled.*ratio$", r"lts__t_sector_hit_rate\.pct", r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$",
    r"^lts__t_bytes\.sum$", r"^l1tex__t_bytes\.sum$", r"pipe_alu_cycles_active.*pct", r"pipe_fma.*cycles_active.*pct", r"inst_executed_pipe_lsu.*pct", r"smsp__pcsamp_warps_issue_stalled_.*(?<!not_issued)$"]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print("==", r[hdr.index("Kernel Name")][:90], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
    for i, h in enumerate(hdr):
        if any(re.search(p, h) for p in pats):
            v = r[i]
            if v not in ("", "0", "0.0"):
                print(f"   {h:85s} {v} {units[i]}")
