#!/usr/bin/env python
"""Is a pass bound by the host's launch rate or by the GPU?  Enqueue time vs completion time of T passes."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _pkg
gt = _pkg.load_package()
for N in (1000, 50000, 100000, 200000):
    lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004, 0, N)
    with gt.Allocator(N) as a:
        a.load_nodes(lv, ng, fr)
        for _ in range(50):
            a.score_all_async()
        a.sync()
        T = 2000
        t0 = time.perf_counter()
        for _ in range(T):
            a.score_all_async()
        t1 = time.perf_counter()
        a.sync()
        t2 = time.perf_counter()
        print("N %7d  enqueue %.2f us/pass   until done %.2f us/pass   time_kernel step %.2f" % (
            N, (t1 - t0) / T * 1e6, (t2 - t0) / T * 1e6, a.time_kernel("step", 200) * 1e3))
