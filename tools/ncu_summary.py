#!/usr/bin/env python
"""Summarises an .ncu-rep (read here on the CPU box): python tools/ncu_summary.py gpurun_out/x.ncu-rep [pattern ...]"""
import csv
import io
import subprocess
import sys

DEFAULT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct", "smsp__inst_executed.sum", "sm__inst_executed_pipe", "smsp__issue_active.avg.pct",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "warp_issue_stalled", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_op_read.sum", "sm__cycles_active.avg", "smsp__cycles_active.avg",
    "gpu__dram_throughput", "l1tex__throughput", "lts__throughput", "sm__inst_executed.sum", "sm__sass_inst_executed_op_shared",
    "smsp__pcsamp",
]


def main():
    rep = sys.argv[1]
    pats = sys.argv[2:] or DEFAULT
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    print("kernels:", [r[ki][:40] for r in data])
    for i, h in enumerate(hdr):
        if any(p in h for p in pats):
            vals = [r[i] for r in data]
            if all(v in ("0", "", "n/a") for v in vals):
                continue
            print("%-90s %-10s %s" % (h, units[i], vals))


if __name__ == "__main__":
    main()
