import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import _pkg
gt = _pkg.load_package()
lv, ng, fr, _ = gt.synth_cluster(2, 1, 0, 1000)
with gt.Allocator(1000) as a:
    a.load_nodes(lv, ng, fr)
    for k in ("noop", "noop_grid", "noop_smem", "bin", "score", "step_nograph", "step"):
        print(k, "%.2f us" % (a.time_kernel(k, 200) * 1e3))
