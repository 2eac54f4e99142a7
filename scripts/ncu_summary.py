#!/usr/bin/env python
"""Condenses an ncu report (ncu -i X.ncu-rep --page raw --csv) into a small TSV of the metrics the design argues
with, for profiles/.  Usage: ncu_summary.py report.ncu-rep out.tsv ["header comment"]"""
import csv
import io
import subprocess
import sys

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed",
        "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed",
        "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed",
        "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum.per_cycle_elapsed",
        "sm__sass_thread_inst_executed_op_dfma_pred_on.sum.peak_sustained",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg.per_second", "sm__cycles_elapsed.avg",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
        "l1tex__t_bytes.sum", "lts__t_bytes.sum", "smsp__average_warp_latency_issue_stalled", "local_load", "local_store",
        "smsp__pcsamp_warps_issue_stalled", "smsp__average_warps_issue_stalled")
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
hdr, units = rows[0], rows[1]
with open(out, "w") as f:
    if len(sys.argv) > 3:
        f.write("# " + sys.argv[3] + "\n")
    for vals in rows[2:]:
        f.write("Kernel Name\t\t" + vals[hdr.index("Kernel Name")] + "\n")
        for h, u, v in zip(hdr, units, vals):
            if any(k in h for k in KEEP) and v not in ("", "n/a"):
                f.write(f"{h}\t{u}\t{v}\n")
print(open(out).read()[:3000])
