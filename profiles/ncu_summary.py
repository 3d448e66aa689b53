#!/usr/bin/env python
"""Distil an .ncu-rep into the text summary committed under profiles/: python profiles/ncu_summary.py <report> <kernel regex> [title]
(raw-page counters that the roofline discussion uses + the SASS-level stall/opcode summary of ncu_sass_summary.py)."""
import csv
import io
import subprocess
import sys

rep, regex = sys.argv[1], sys.argv[2]
title = sys.argv[3] if len(sys.argv) > 3 else regex
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum", "sm__cycles_active.avg"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--kernel-name", "regex:" + regex], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
print(title)
for r in rows[2:3]:
    print("kernel:", r[hdr.index("Kernel Name")])
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w:78s} {r[i]} {units[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + regex], capture_output=True, text=True).stdout
open("/tmp/_ncu_src.csv", "w").write(src)
print()
print(subprocess.run([sys.executable, __file__.replace("ncu_summary.py", "ncu_sass_summary.py"), "/tmp/_ncu_src.csv"], capture_output=True, text=True).stdout)
