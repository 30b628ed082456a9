"""ctypes view of oracle/libgt_oracle.so -- test infrastructure, never imported by product code."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")
INFEASIBLE = 0xFFFF0000
LV = dict(X=0, NVS=1, NV4=2, NV3=3, NV2=4, NV1=5, PSB=6, PIX=7, PXB=8, PHB=9, NODE=10, SYS=11)
CLS = dict(DGX1=0, F1_RING=1, F2_PCIE=2, F3_TREE4=3, NVS16=4, DUAL8=5)

u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)


class Tree(C.Structure):
    _fields_ = [("n_nodes", C.c_int), ("parent", C.c_int * 32), ("gpu_id", C.c_int * 32),
                ("access_cost", C.c_int * 32), ("level", C.c_int * 32)]


def _p(a, t):
    return a.ctypes.data_as(t)


class Oracle:
    def __init__(self, lib):
        self.lib = lib
        L = lib
        L.gto_score_int.restype = C.c_int
        L.gto_score_milli.restype = C.c_int
        L.gto_lut_cmax.restype = C.c_uint32
        L.gto_subset_cost.restype = C.c_uint32
        L.gto_exhaustive_node.restype = C.c_uint64
        L.gto_exhaustive_cluster.restype = C.c_uint64
        L.gto_greedy_node.restype = C.c_uint16
        L.gto_greedy_cluster.restype = C.c_uint64
        L.gto_gaia_singular.restype = C.c_int
        L.gto_gaia_link.restype = C.c_uint16
        L.gto_gaia_dispatch.restype = C.c_uint16
        L.gto_splitmix64_at.restype = C.c_uint64
        L.gto_splitmix64_at.argtypes = [C.c_uint64, C.c_uint64]
        L.gto_replay.restype = C.c_uint64

    def lut(self, preset=0):
        a = np.zeros(16, np.uint8)
        self.lib.gto_lut_preset(C.c_int(preset), _p(a, u8p))
        return a

    def cmax(self, lut):
        return int(self.lib.gto_lut_cmax(_p(np.ascontiguousarray(lut, np.uint8), u8p)))

    def score_int(self, s, l, cmax):
        return int(self.lib.gto_score_int(C.c_uint32(s), C.c_uint32(l), C.c_uint32(cmax)))

    def score_milli(self, s, l, cmax):
        return int(self.lib.gto_score_milli(C.c_uint32(s), C.c_uint32(l), C.c_uint32(cmax)))

    def subset_cost(self, levels, lut, mask):
        return int(self.lib.gto_subset_cost(_p(levels, u8p), _p(lut, u8p), C.c_uint16(mask)))

    def exhaustive_node(self, levels, n, free, lut):
        out = np.zeros(16, np.uint32)
        cnt = self.lib.gto_exhaustive_node(_p(levels, u8p), C.c_int(n), C.c_uint16(free), _p(lut, u8p), _p(out, u32p))
        return out, int(cnt)

    def exhaustive_cluster(self, levels, n_gpus, free, lut, threads=1):
        N = len(n_gpus)
        out = np.zeros((N, 16), np.uint32)
        cnt = self.lib.gto_exhaustive_cluster(_p(levels, u8p), _p(n_gpus, u8p), _p(free, u16p), C.c_uint32(N),
                                              _p(lut, u8p), _p(out, u32p), C.c_int(threads))
        return out, int(cnt)

    def greedy_cluster(self, levels, n_gpus, free, lut):
        N = len(n_gpus)
        out = np.zeros((N, 16), np.uint32)
        cnt = self.lib.gto_greedy_cluster(_p(levels, u8p), _p(n_gpus, u8p), _p(free, u16p), C.c_uint32(N), _p(lut, u8p),
                                          _p(out, u32p))
        return out, int(cnt)

    def cluster_argmin(self, best, first=0):
        best = np.ascontiguousarray(best, np.uint32)
        out = np.zeros(16, np.uint64)
        self.lib.gto_cluster_argmin(_p(best, u32p), C.c_uint32(best.shape[0]), C.c_uint32(first), _p(out, u64p))
        return out

    def greedy_node(self, levels, n, free, lut, m):
        cost = C.c_uint32(0)
        mask = self.lib.gto_greedy_node(_p(levels, u8p), C.c_int(n), C.c_uint16(free), _p(lut, u8p), C.c_int(m),
                                        C.byref(cost))
        return int(mask), int(cost.value)

    def gaia_singular(self, n, free, access_cost, sibling_mask):
        return int(self.lib.gto_gaia_singular(C.c_int(n), C.c_uint16(free), _p(access_cost, u8p), _p(sibling_mask, u16p)))

    def k1_farthest(self, levels, n, free, lut):
        return int(self.lib.gto_k1_farthest(_p(levels, u8p), C.c_int(n), C.c_uint16(int(free)), _p(lut, u8p)))

    def gaia_link(self, tree, free, m):
        return int(self.lib.gto_gaia_link(C.byref(tree), C.c_uint16(free), C.c_int(m)))

    def gaia_dispatch(self, tree, free, m):
        return int(self.lib.gto_gaia_dispatch(C.byref(tree), C.c_uint16(free), C.c_int(m)))

    def tree_to_tables(self, tree, n):
        lv = np.zeros(256, np.uint8); ac = np.zeros(16, np.uint8); sib = np.zeros(16, np.uint16)
        self.lib.gto_tree_to_tables(C.byref(tree), C.c_int(n), _p(lv, u8p), _p(ac, u8p), _p(sib, u16p))
        return lv, ac, sib

    def class_topology(self, cls):
        lv = np.zeros(256, np.uint8); ac = np.zeros(16, np.uint8); sib = np.zeros(16, np.uint16)
        n = C.c_uint8(0)
        self.lib.gto_class_topology(C.c_int(cls), _p(lv, u8p), C.byref(n), _p(ac, u8p), _p(sib, u16p))
        return lv, int(n.value), ac, sib

    def splitmix(self, seed, idx):
        return int(self.lib.gto_splitmix64_at(seed, idx))

    def synth_cluster(self, config, seed, first, count):
        lv = np.zeros((count, 256), np.uint8); n = np.zeros(count, np.uint8)
        fr = np.zeros(count, np.uint16); cl = np.zeros(count, np.uint8)
        self.lib.gto_synth_cluster(C.c_int(config), C.c_uint64(seed), C.c_uint32(first), C.c_uint32(count),
                                   _p(lv, u8p), _p(n, u8p), _p(fr, u16p), _p(cl, u8p))
        return lv, n, fr, cl

    def synth_events(self, config, seed, n_events):
        k = np.zeros(n_events, np.uint8); arg = np.zeros(n_events, np.uint32)
        self.lib.gto_synth_events(C.c_int(config), C.c_uint64(seed), C.c_uint32(n_events), _p(k, u8p), _p(arg, u32p))
        return k, arg

    def replay(self, levels, n_gpus, free, lut, ev_k, ev_arg):
        free = free.copy()
        E = len(ev_k)
        on = np.zeros(E, np.uint32); om = np.zeros(E, np.uint16)
        placed = self.lib.gto_replay(_p(levels, u8p), _p(n_gpus, u8p), _p(free, u16p), C.c_uint32(len(n_gpus)),
                                     _p(lut, u8p), _p(ev_k, u8p), _p(ev_arg, u32p), C.c_uint32(E),
                                     _p(on, u32p), _p(om, u16p))
        return int(placed), on, om, free


def make_tree(spec):
    """spec: nested tuples ('LEVEL', child, child, ...) with leaves (gpu_id, access_cost)."""
    t = Tree()
    nodes = []

    def add(s, parent):
        idx = len(nodes)
        if isinstance(s[0], str):
            nodes.append((parent, -1, 0, LV["SYS" if s[0] == "SOC" else s[0]]))
            for c in s[1:]:
                add(c, idx)
        else:
            nodes.append((parent, s[0], s[1], 0))
    add(spec, 0)
    t.n_nodes = len(nodes)
    for i, (p, g, a, l) in enumerate(nodes):
        t.parent[i] = p; t.gpu_id[i] = g; t.access_cost[i] = a; t.level[i] = l
    return t


# Gaia Fig. 4 (gpu_topology_tree.png) and Fig. 7 (paper p.257)
F2_TREE = ("SOC", ("PXB", (6, 4), (7, 4)),
           ("PHB", ("PXB", (4, 3), (5, 3)),
            ("PXB", ("PIX", (0, 1), (1, 1)), ("PXB", (2, 2), (3, 2)))))
F3_TREE = ("SOC", ("PXB", (0, 2), (1, 2)), ("PIX", (2, 1), (3, 1)))


def build():
    subprocess.run(["make", "-s", "-C", ODIR], check=True)


def load():
    so = os.path.join(ODIR, "libgt_oracle.so")
    src = [os.path.join(ODIR, f) for f in ("gt_oracle.c", "gt_oracle.h")]
    if (not os.path.exists(so)) or any(os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        build()
    return Oracle(C.CDLL(so))
