"""Where does an end-to-end `sort` cycle spend its time?  (run on the GPU box)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _pkg, torch
gt = _pkg.load_package()
N = 100000
lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004, 0, N)
pin = torch.empty(N, dtype=torch.int16).pin_memory(); fnp = pin.numpy().view(np.uint16); fnp[:] = fr
sc = torch.empty(N, dtype=torch.uint8).pin_memory().numpy()
pl = np.zeros(16, np.uint64)
with gt.Allocator(N) as a:
    a.load_nodes(lv, ng, fr)
    for _ in range(20):
        a.update_free(fnp); a.ext_sort_into(2, sc, pl)
    T = 500
    t0 = time.perf_counter()
    for _ in range(T): a.update_free(fnp)
    t1 = time.perf_counter()
    for _ in range(T): a.update_free(fnp); a.ext_sort_into(2, sc, pl)
    t2 = time.perf_counter()
    for _ in range(T): a.ext_sort_into(2, sc, pl)       # nothing dirty: sort kernel + copies only
    t3 = time.perf_counter()
    for _ in range(T): a.score_all_async(); a.sync()
    t4 = time.perf_counter()
    print("update_free alone      %.1f us" % ((t1 - t0) / T * 1e6))
    print("update_free + ext_sort %.1f us" % ((t2 - t1) / T * 1e6))
    print("ext_sort, nothing dirty %.1f us" % ((t3 - t2) / T * 1e6))
    print("score_all_async + sync %.1f us" % ((t4 - t3) / T * 1e6))
