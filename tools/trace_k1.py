#!/usr/bin/env python
"""Per-warp timeline of one k_score launch (needs a -DGT_TRACE build: GT_LIB=build/libgputopo_trace.so)."""
import ctypes as C
import os
import sys
from collections import defaultdict

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _pkg

gt = _pkg.load_package()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004, 0, N)
SL = 64
COLD = int(os.environ.get("GT_TRACE_COLD", "0"))     # this many other clusters pass in between (their tables evict ours from L2)
others = []
for r in range(COLD):
    lv2, ng2, fr2, _ = gt.synth_cluster(4, 0x5EED0104 + r, 0, N)
    o = gt.Allocator(N)
    o.load_nodes(lv2, ng2, fr2)
    others.append(o)
with gt.Allocator(N) as a:
    a.load_nodes(lv, ng, fr)
    for _ in range(3):
        a.score_all(fetch=False)
    L = gt._lib()
    L.gt_debug_trace.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.c_uint32]
    nw = 148 * 8 * 4
    buf = np.zeros(nw * SL, np.uint64)
    for o in others:
        o.score_all(fetch=False)
    L.gt_debug_trace(a._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)), len(buf))   # clear
    for o in others:
        o.score_all(fetch=False)
        L.gt_debug_trace(o._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)), len(buf))
    a.score_all(fetch=False)
    L.gt_debug_trace(a._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)), len(buf))
for o in others:
    o.close()
print("other clusters scored in between:", COLD)
tr = buf.reshape(nw, SL)
t = (tr >> np.uint64(8)).astype(np.int64)
ev = (tr & np.uint64(0xFF)).astype(np.int64)
used = (tr != 0).any(axis=1)
t0 = t[used][:, 0].min()
print("warps traced:", used.sum())
phase = defaultdict(list)
unit_dur = defaultdict(list)
units_per_warp = []
for w in np.nonzero(used)[0]:
    e, tt = ev[w], t[w]
    n = int((tr[w] != 0).sum())
    marks = {int(e[i]): int(tt[i]) for i in range(n) if e[i] in (1, 2, 3, 4, 5, 6, 7)}
    for k in (1, 2, 6, 7, 3, 4, 5):
        if k in marks:
            phase[k].append(marks[k] - t0)
    cnt = 0
    i = 0
    while i < n:
        if e[i] >= 32:
            p = int(e[i]) - 32
            start = int(tt[i])
            j = i + 1
            sub = {}
            while j < n and e[j] < 32 and e[j] not in (4, 5, 6, 7):
                sub[int(e[j])] = int(tt[j])
                j += 1
            end = sub.get(14, sub.get(13, start))
            rec = [sub.get(10, start) - start, sub.get(11, start) - sub.get(10, start), sub.get(12, start) - sub.get(11, start),
                   sub.get(13, start) - sub.get(12, sub.get(20, start)), sub.get(14, start) - sub.get(13, start), end - start]
            if p <= 8:   # columns reused: gathers-wait, tree
                rec = [sub.get(21, start) - start, sub.get(20, start) - sub.get(21, start), sub.get(22, start) - sub.get(20, start),
                       sub.get(13, start) - sub.get(22, start), sub.get(14, start) - sub.get(13, start), end - start]
            unit_dur[p].append(rec)
            cnt += 1
            i = j
        else:
            i += 1
    units_per_warp.append(cnt)
# the warps that finish last: what did they run?
fin = []
for w in np.nonzero(used)[0]:
    e, tt = ev[w], t[w]
    n = int((tr[w] != 0).sum())
    seq = [(int(e[i]) - 32, int(tt[i]) - t0) for i in range(n) if e[i] >= 32]
    end = [int(tt[i]) - t0 for i in range(n) if e[i] == 4]
    fin.append((end[0] if end else 0, w, seq))
fin.sort(reverse=True)
print("latest warps (finish ns, warp, [(p, unit start ns)...]):")
for f in fin[:12]:
    print("  ", f)
first = defaultdict(list)
for f in fin:
    if f[2]:
        first[f[2][0][0]].append(f[2][0][1])
print("first unit of a warp: p -> (count, median start ns, max start ns)")
for pp in sorted(first, reverse=True):
    v = np.array(first[pp])
    print("   p=%2d %5d %7d %7d" % (pp, len(v), np.median(v), v.max()))
hist = np.histogram([f[0] for f in fin], bins=12)
print("finish-time histogram:", list(zip(hist[1][:-1].astype(int), hist[0])))
names = {1: "kernel entry", 2: "after griddep wait", 6: "control block arrived", 7: "scan done", 3: "after the barrier", 4: "work loop done", 5: "after final barrier"}
for k in (1, 2, 6, 7, 3, 4, 5):
    if not phase[k]:
        continue
    v = np.array(phase[k])
    print("%-22s min %7d  median %7d  max %7d ns" % (names[k], v.min(), np.median(v), v.max()))
print("units per warp: mean %.2f max %d" % (np.mean(units_per_warp), max(units_per_warp)))
print("per-unit ns (median): p  count | list-wait  stage  setup+tree  writeout  ticket-wait | total   (p <= 8: gathers-wait, tree, row stores, fold)")
for p in sorted(unit_dur, reverse=True):
    r = np.array(unit_dur[p])
    med = np.median(r, axis=0)
    print("   p=%2d %6d | %8d %6d %10d %9d %11d | %6d   (max total %d)" % (p, len(r), med[0], med[1], med[2], med[3], med[4], med[5], r[:, 5].max()))
