#!/usr/bin/env python
"""Extract the metrics the notes quote from an .ncu-rep (raw page) into a small CSV: one row per metric, one column per
profiled launch.   python scripts/ncu_extract.py gpurun_out/x.ncu-rep > profiles/x_ncu_full.csv"""
import csv
import subprocess
import sys

WANT = ("gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
        "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed")
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
cols = [i for i, h in enumerate(hdr) if h in WANT or (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"))]
w = csv.writer(sys.stdout)
name_i = hdr.index("Kernel Name")
w.writerow(["metric", "unit"] + [r[name_i].split("(")[0] for r in data])
for i in cols:
    w.writerow([hdr[i], units[i]] + [r[i] for r in data])
