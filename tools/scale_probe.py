#!/usr/bin/env python
"""How do the kernels scale with the number of nodes?  (run on the GPU box)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _pkg

gt = _pkg.load_package()
configs = (4, 2, 3)
sizes = (1000, 10000, 25000, 50000, 100000, 200000, 400000, 800000)
if len(sys.argv) > 1:
    configs = (int(sys.argv[1]),)
    sizes = tuple(int(x) for x in sys.argv[2:]) or sizes
for config in configs:
    for n in sizes:
        if config == 3 and n > 100000:
            continue
        lv, ng, fr, _ = gt.synth_cluster(config, 0x5EED0000 + config, 0, n)
        with gt.Allocator(n) as a:
            a.load_nodes(lv, ng, fr)
            subsets, adds = a.count_subsets()
            t = {k: a.time_kernel(k, 30) * 1e3 for k in ("step", "step_nograph", "bin", "score")}
            with gt.Allocator(n, flags=gt.FLAG_MEMO_PASS) as am:
                am.load_nodes(lv, ng, fr)
                am.score_all(fetch=False)
                t["memo"] = am.time_kernel("step", 30) * 1e3
            print("config %2d N %7d subsets %11d  step %8.2f us (no graph %6.2f)  bin %7.2f  score %8.2f  memo-pass %6.2f   %.3e subsets/s" % (
                config, n, subsets, t["step"], t["step_nograph"], t["bin"], t["score"], t["memo"], subsets / (t["step"] * 1e-6)))
