"""How does the per-pass time of the bench loop depend on the number of timed passes?  (fixed cost of a short run)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import _pkg
gt = _pkg.load_package()
N, R = 100000, 8
stream = torch.cuda.Stream()
ctxs = []
for r in range(R):
    lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004 + r, 0, N)
    a = gt.Allocator(N, stream=stream.cuda_stream)
    a.load_nodes(lv, ng, fr)
    a.score_all(fetch=False)
    ctxs.append(a)
for steps in (1, 2, 5, 10, 20, 40, 80, 160, 640, 2000):
    res = []
    for rep in range(5):
        gt.score_all_async_many(ctxs, 5)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        gt.score_all_async_many(ctxs, steps)
        e1.record(stream)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        res.append((e0.elapsed_time(e1) * 1e3, (t1 - t0) * 1e6))
    dev = sorted(x[0] for x in res)[2]
    host = sorted(x[1] for x in res)[2]
    print("steps %5d  device %9.1f us  (%6.2f us/pass; minus 17.7/pass: %6.1f us)   host enqueue %8.1f us (%5.2f us/pass)" % (
        steps, dev, dev / steps, dev - 17.7 * steps, host, host / steps))
