#!/usr/bin/env python
"""Regenerates tests/golden/sweep_small.npz from the CPU oracle: for each small fixture topology the best table
(16 x u32 per free mask, default cost table and the design.md mark table) of every free mask, plus the placement
keys of the 1000-node config-2 cluster.  The oracle itself is pinned by known_answers.json; this file pins the
oracle (and, in the gpu tests, the CUDA path) against silent changes.  Run from the repo root."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol


def build():
    o = ol.load()
    out = {}
    for name in ("DGX1", "F1_RING", "F2_PCIE", "F3_TREE4"):
        lv, n, _, _ = o.class_topology(ol.CLS[name])
        N = 1 << n
        levels = np.tile(lv, (N, 1))
        free = np.arange(N, dtype=np.uint16)
        for preset in (0, 1):
            best, _ = o.exhaustive_cluster(levels, np.full(N, n, np.uint8), free, o.lut(preset), 1)
            out["%s_lut%d" % (name, preset)] = best
    lv, n, fr, _ = o.synth_cluster(2, 0x5EED0002, 0, 1000)
    best, _ = o.exhaustive_cluster(lv, n, fr, o.lut(0), 1)
    out["config2_free"] = fr
    out["config2_place"] = o.cluster_argmin(best)
    return out


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "sweep_small.npz"), **build())
    print("wrote", os.path.join(HERE, "sweep_small.npz"))
