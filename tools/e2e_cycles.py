"""Per-cycle times of the first `sort` cycles after a synchronisation (run on the GPU box)."""
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
    a.host_free()[:N] = fr
    ctxs.append(a)
pl = np.zeros(16, np.uint64)
for warm in (8, 24):
    for rep in range(3):
        for i in range(warm):
            c = ctxs[i % R]; c.host_free_changed(); c.ext_sort_inplace(1 + i % 8, pl)
        torch.cuda.synchronize()
        ts = []
        for i in range(40):
            t0 = time.perf_counter()
            c = ctxs[i % R]; c.host_free_changed(); c.ext_sort_inplace(1 + i % 8, pl)
            ts.append((time.perf_counter() - t0) * 1e6)
        print("warm-up %2d cycles: first 20 mean %.1f  next 20 mean %.1f   cycles: %s" % (
            warm, np.mean(ts[:20]), np.mean(ts[20:]), " ".join("%.0f" % t for t in ts[:24])))
        time.sleep(0.3)
for a in ctxs:
    a.close()
