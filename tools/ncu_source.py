#!/usr/bin/env python
"""SASS-level view of an ncu report: python tools/ncu_source.py rep [kernel-regex]
Prints instruction counts by opcode, regions between barriers, and the top stall sites."""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    ins = []
    for r in rows[2:]:
        if len(r) < len(hdr) or not r[0].startswith("0x"):
            continue
        ins.append(r)
    tot_exec = sum(int(r[ix["Instructions Executed"]]) for r in ins)
    tot_samp = sum(int(r[ix["# Samples"]]) for r in ins)
    print("SASS instructions: %d, warp-instr executed: %d, samples: %d" % (len(ins), tot_exec, tot_samp))
    byop = collections.Counter()
    sop = collections.Counter()
    for r in ins:
        op = r[ix["Source"]].split()[0]
        if op.startswith("@"):
            op = r[ix["Source"]].split()[1]
        op = op.split(".")[0]
        byop[op] += int(r[ix["Instructions Executed"]])
        sop[op] += int(r[ix["# Samples"]])
    print("\nby opcode (executed, share, samples):")
    for op, n in byop.most_common(25):
        print("  %-10s %10d %5.1f%% %6d" % (op, n, 100.0 * n / tot_exec, sop[op]))
    # regions split at BAR / BRA-heavy points: print cumulative per 'BAR' boundaries
    print("\nregions between BAR.SYNC (static order):")
    acc_e = acc_s = 0
    start = 0
    for i, r in enumerate(ins):
        acc_e += int(r[ix["Instructions Executed"]])
        acc_s += int(r[ix["# Samples"]])
        if "BAR.SYNC" in r[ix["Source"]] or "EXIT" in r[ix["Source"]] or i == len(ins) - 1:
            if acc_e:
                print("  sass[%4d..%4d] executed %9d (%4.1f%%) samples %5d (%4.1f%%)  ends with %s" % (
                    start, i, acc_e, 100.0 * acc_e / tot_exec, acc_s, 100.0 * acc_s / max(tot_samp, 1), r[ix["Source"]].strip()[:40]))
            start = i + 1
            acc_e = acc_s = 0
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    print("\ntop stall sites:")
    top = sorted(ins, key=lambda r: -int(r[ix["# Samples"]]))[:25]
    for r in top:
        st = {h: int(r[ix[h]]) for h in stall_cols if int(r[ix[h]])}
        main_st = sorted(st.items(), key=lambda kv: -kv[1])[:3]
        print("  #%4d %-52s samp %4s exec %8s %s" % (ins.index(r), r[ix["Source"]].strip()[:52], r[ix["# Samples"]],
                                                  r[ix["Instructions Executed"]], main_st))


if __name__ == "__main__":
    main()
