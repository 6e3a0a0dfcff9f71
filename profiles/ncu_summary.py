#!/usr/bin/env python3
"""Prints the handful of ncu metrics the roofline discussion needs from a .ncu-rep
(run here, on the CPU box: `python profiles/ncu_summary.py gpurun_out/x.ncu-rep`)."""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "launch__registers_per_thread ",
        "launch__occupancy_limit", "smsp__inst_executed.sum ", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__pipe_tensor_cycles_active",
        "smsp__inst_executed_pipe_lsu", "smsp__inst_executed_pipe_alu", "smsp__inst_executed_pipe_fma",
        "sm__inst_executed_pipe_tensor", "lts__t_bytes.sum ", "launch__grid_size", "launch__block_size",
        "smsp__average_warps_issue_stalled", "sm__cycles_elapsed.max ", "launch__shared_mem_per_block"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
    print("==", name[:100])
    for h, u, v in zip(hdr, units, r):
        if any((h + " ").startswith(w) or w.strip() in h and w.endswith("stalled") for w in WANT):
            try:
                if float(v.replace(",", "")) == 0 and "stalled" in h:
                    continue
            except ValueError:
                pass
            print(f"  {h:75s} {v} {u}")
