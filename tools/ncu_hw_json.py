#!/usr/bin/env python
"""Hardware-side figures of the dominant kernel out of an `ncu --set full` capture, as the JSON bench.py reads
(profiles/r2_hw.json):  python tools/ncu_hw_json.py gpurun_out/x.ncu-rep <kernel regex> <nodes> "<command that was profiled>" """
import csv
import io
import json
import re
import subprocess
import sys

WANT = {
    "duration_us_under_ncu": ("gpu__time_duration.sum", 1e-3),
    "dram_bytes_read": ("dram__bytes_read.sum", None),
    "dram_bytes_write": ("dram__bytes_write.sum", None),
    "warp_instructions": ("smsp__inst_executed.sum", 1),
    "issue_active_pct": ("smsp__issue_active.avg.pct_of_peak_sustained_active", 1),
    "inst_executed_pct_of_peak": ("sm__inst_executed.sum.pct_of_peak_sustained_elapsed", 1),
    "alu_pipe_pct": ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", 1),
    "fma_pipe_pct": ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", 1),
    "lsu_pipe_pct": ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", 1),
    "warps_active_pct": ("sm__warps_active.avg.pct_of_peak_sustained_active", 1),
    "l1tex_throughput_pct": ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    "dram_throughput_pct": ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
    "registers_per_thread": ("launch__registers_per_thread", 1),
}
UNIT_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
UNIT_TIME_NS = {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}


def main():
    rep, pattern, nodes, cmd = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    data = [r for r in data if re.search(pattern, r[ki])]
    if not data:
        raise SystemExit("no kernel matches %r" % pattern)
    res = {"kernel": data[0][ki].split("(")[0], "launches_averaged": len(data), "nodes": nodes,
           "source": "ncu --set full --clock-control none inside `%s` (cold caches, serialised launches)" % cmd}
    for key, (metric, scale) in WANT.items():
        if metric not in hdr:
            continue
        i = hdr.index(metric)
        vals = [float(r[i].replace(",", "")) for r in data if r[i] not in ("", "n/a")]
        if not vals:
            continue
        v = sum(vals) / len(vals)
        if key.startswith("dram_bytes"):
            v = int(v * UNIT_BYTES.get(units[i], 1))
        elif key == "duration_us_under_ncu":
            v = v * UNIT_TIME_NS.get(units[i], 1) / 1e3
        res[key] = v
    if "dram_bytes_read" in res and "duration_us_under_ncu" in res:
        res["dram_gbs_measured"] = (res["dram_bytes_read"] + res.get("dram_bytes_write", 0)) / (res["duration_us_under_ncu"] * 1e-6) / 1e9
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
