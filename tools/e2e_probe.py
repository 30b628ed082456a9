"""Where does an end-to-end `sort` cycle spend its time?  (run on the GPU box)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _pkg
gt = _pkg.load_package()
N = 100000
R = 8
ctxs = []
for r in range(R):                       # rotated replicas: cold L2 as in bench.py
    lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004 + r, 0, N)
    a = gt.Allocator(N)
    a.load_nodes(lv, ng, fr)
    a.host_free()[:N] = fr
    ctxs.append(a)
pl = np.zeros(16, np.uint64)
T = 400


def timed(f):
    for i in range(20):
        f(i)
    t0 = time.perf_counter()
    for i in range(T):
        f(i)
    return (time.perf_counter() - t0) / T * 1e6


def pass_only(i):
    a = ctxs[i % R]; a.score_all_async(); a.sync()


def upload_pass(i):
    a = ctxs[i % R]; a.host_free_changed(); a.score_all_async(); a.sync()


def sort_cycle(i):
    a = ctxs[i % R]; a.host_free_changed(); a.ext_sort_inplace(1 + i % 8, pl)


def sort_cycle_no_keys(i):
    a = ctxs[i % R]; a.host_free_changed(); a.ext_sort_inplace(1 + i % 8, None)


def sort_from_table(i):
    a = ctxs[i % R]; a.ext_sort_inplace(1 + i % 8, pl)      # nothing changed: score kernel + sync only


def sync_only(i):
    ctxs[i % R].sync()


print("sync alone                                  %6.1f us" % timed(sync_only))
print("pass (masks resident) + sync                %6.1f us" % timed(pass_only))
print("masks read in place by K0 + pass + sync     %6.1f us" % timed(upload_pass))
print("full sort cycle (scores left by the pass)   %6.1f us" % timed(sort_cycle))
print("full sort cycle without the placement keys  %6.1f us" % timed(sort_cycle_no_keys))
for a in ctxs:
    a.score_all(fetch=False)
print("sort from the table (nothing changed)       %6.1f us" % timed(sort_from_table))
for a in ctxs:
    a.close()
