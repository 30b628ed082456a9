#!/usr/bin/env python
"""One line per bench.py output file: value, us per step, e2e, kernel timings."""
import json
import sys
for f in sys.argv[1:]:
    try:
        txt = [l for l in open(f).read().splitlines() if l.startswith("{")][-1]
        d = json.loads(txt)
        e = d.get("e2e") or {}
        k = d.get("kernels_us") or {}
        print("%-32s n=%d value %.3e  %.2f us/step  e2e %.1f us (copying %.1f)  K1 cold %.1f warm %.1f  step warm %.1f  strong %s" % (
            f.split("/")[-1], d["n_gpus"], d["value"], d["ms_per_step"] * 1e3, e.get("ms_per_step", 0) * 1e3,
            (e.get("copying") or {}).get("ms_per_step", 0) * 1e3, k.get("score_cold_us", 0), k.get("score_l2warm_us", 0),
            k.get("step_l2warm_us", 0), (d.get("strong") or {}).get("ms_per_step")))
    except Exception as ex:
        print(f, "unreadable:", ex)
