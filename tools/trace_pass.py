#!/usr/bin/env python
"""Kernel-level timeline of consecutive passes (needs a -DGT_TRACE build: GT_LIB=build/trace_base.so)."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _pkg

gt = _pkg.load_package()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004, 0, N)
with gt.Allocator(N) as a:
    a.load_nodes(lv, ng, fr)
    L = gt._lib()
    L.gt_debug_ktrace.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
    buf = np.zeros((64, 8), np.uint64)
    ps = C.c_uint32(0)
    for _ in range(40):
        a.score_all_async()
    L.gt_debug_ktrace(a._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(ps))   # reset
    p0 = ps.value
    for _ in range(40):
        a.score_all_async()
    L.gt_debug_ktrace(a._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(ps))
    rows = [buf[(p0 + i) & 63].astype(np.int64) for i in range(40)]
    base = rows[0][0]
    print("pass | K0 entry  K0 go-entry  K0 end-go | K1 entry-K0end  K1 go-K0end  K1 len (go..last exit) | next K0 go - K1 end | pass period")
    per = []
    for i in range(5, 39):
        r, nx = rows[i], rows[i + 1]
        per.append(nx[5] - r[5])
        if i < 17:
            print("%4d | %8d %7d %7d | %8d %8d %8d | %8d | %8d" % (i, r[0] - base, r[5] - r[0], r[1] - r[5], r[2] - r[1],
                                                                   r[3] - r[1], r[4] - r[3], nx[5] - r[4], nx[5] - r[5]))
    print("median period %.0f ns" % np.median(per))
