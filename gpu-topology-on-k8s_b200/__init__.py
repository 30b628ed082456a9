"""gputopo_b200 -- host-side mirror of the reference's scheduler-extender / device-plugin surface
for the topology-aware GPU allocation path, over the C-ABI of libgputopo.so (include/gputopo.h).

Python is the harness language here (the reference's Go toolchain is absent from this image); the
product is the shared library.  Nothing in this package scores subsets on the CPU and nothing here
imports oracle/: if the CUDA library is missing or no device is present, calls fail loudly.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
INFEASIBLE = 0xFFFF0000
NO_PLACE = 0xFFFFFFFFFFFFFFFF
LV = dict(X=0, NVS=1, NV4=2, NV3=3, NV2=4, NV1=5, PSB=6, PIX=7, PXB=8, PHB=9, NODE=10, SYS=11)
LUT_DEFAULT, LUT_DESIGN_MARKS = 0, 1
K1_FIRST_UNUSED, K1_GAIA_SINGULAR, K1_FARTHEST = 0, 1, 2
FLAG_GRAPH, FLAG_NO_MEMO, FLAG_MEMO_PASS = 2, 4, 8

EXPORTS = [
    "gt_create", "gt_destroy", "gt_strerror", "gt_last_error", "gt_lut_preset", "gt_set_cost_lut",
    "gt_score_int", "gt_score_milli", "gt_load_nodes", "gt_update_free", "gt_score_all", "gt_score_dirty",
    "gt_best_node", "gt_allocate", "gt_commit", "gt_release", "gt_score_all_async", "gt_sync", "gt_stream",
    "gt_device_best", "gt_num_nodes", "gt_launch_count", "gt_count_subsets", "gt_time_kernel",
    "gt_comm_unique_id", "gt_comm_init", "gt_synth_cluster", "gt_score_nodes_static", "gt_schedule", "gt_synth_events",
    "gt_parse_topo_matrix", "gt_annotation_encode", "gt_annotation_decode", "gt_group_encode", "gt_group_decode",
    "gt_bind_annotations", "gt_ext_sort", "gt_ext_bind", "gt_plugin_allocate",
    "gt_tree_tables", "gt_load_trees", "gt_gaia_select", "gt_comm_mode",
    "gt_wire_parse_extender_args", "gt_wire_host_priorities", "gt_wire_parse_binding_args", "gt_wire_binding_result",
    "gt_set_node_names", "gt_ext_sort_json", "gt_ext_bind_json", "gt_ext_pod_deleted",
    "gt_level_from_p2plink", "gt_levels_from_p2plink", "gt_lut_cmax",
    "gt_time_pass_rotated", "gt_best_node_local", "gt_comm_share",
    "gt_host_free", "gt_host_free_changed", "gt_host_scores", "gt_score_all_async_many", "gt_score_all_async_many_timed",
]
STATIC_MAX_NODES = 4096

u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)


class GtConfig(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("device", C.c_int32), ("max_nodes", C.c_uint32),
                ("first_node", C.c_uint32), ("k1_policy", C.c_uint32), ("flags", C.c_uint32),
                ("stream", C.c_void_p)]


class GtTree(C.Structure):
    """Gaia resource-access cost tree (include/gputopo.h gt_tree): index 0 = root, gpu_id >= 0 marks a leaf."""
    _fields_ = [("n_nodes", C.c_int32), ("parent", C.c_int32 * 32), ("gpu_id", C.c_int32 * 32),
                ("access_cost", C.c_int32 * 32), ("level", C.c_int32 * 32)]


def make_tree(spec):
    """spec: nested tuples (level_code, child, ...) for topology nodes and (gpu_id, access_cost) pairs for leaves,
    e.g. (LV_SYS, (LV_PXB, (0, 2), (1, 2)), (LV_PIX, (2, 1), (3, 1))); a node is a leaf when it has two ints."""
    t = GtTree()
    rows = []

    def add(s, parent):
        idx = len(rows)
        if len(s) == 2 and all(isinstance(x, int) for x in s):
            rows.append((parent, s[0], s[1], 0))
        else:
            rows.append((parent, -1, 0, int(s[0])))
            for ch in s[1:]:
                add(ch, idx)
    add(spec, 0)
    if len(rows) >= 32:
        raise ValueError("a tree has at most 31 nodes")
    t.n_nodes = len(rows)
    for i, (p, g, a, l) in enumerate(rows):
        t.parent[i], t.gpu_id[i], t.access_cost[i], t.level[i] = p, g, a, l
    return t


def tree_tables(tree):
    """Host only: (levels[256], n_gpus, access_cost[16], sibling_mask[16]) a tree stands for."""
    lv = np.zeros(256, np.uint8)
    ng = C.c_uint8(0)
    ac = np.zeros(16, np.uint8)
    sib = np.zeros(16, np.uint16)
    rc = _lib().gt_tree_tables(C.byref(tree), _p(lv, u8p), C.byref(ng), _p(ac, u8p), _p(sib, u16p))
    if rc != 0:
        raise GtError(rc, "gt_tree_tables")
    return lv, ng.value, ac, sib


class GtError(RuntimeError):
    def __init__(self, code, what, detail=""):
        self.code = code
        super().__init__("%s failed: %s (%d)%s" % (what, _lib().gt_strerror(code).decode(), code,
                                                   (": " + detail) if detail else ""))


_LIB = None


def lib_path():
    return os.path.join(HERE, "libgputopo.so")


def _lib():
    """Loads libgputopo.so (building it in-tree if the sources are newer).  No fallback."""
    global _LIB
    if _LIB is not None:
        return _LIB
    so = os.environ.get("GT_LIB")      # profiling aid: load another build of the same ABI
    if not so:
        so = lib_path()
        if _build.stale() and not os.environ.get("GT_NO_REBUILD"):   # GT_NO_REBUILD: use the built file as it is
            _build.build()
    L = C.CDLL(so)
    L.gt_strerror.restype = C.c_char_p
    L.gt_last_error.restype = C.c_char_p
    L.gt_last_error.argtypes = [C.c_void_p]
    L.gt_create.argtypes = [C.POINTER(GtConfig), C.POINTER(C.c_void_p)]
    L.gt_destroy.argtypes = [C.c_void_p]
    L.gt_destroy.restype = None
    L.gt_set_cost_lut.argtypes = [C.c_void_p, u8p]
    L.gt_lut_preset.argtypes = [C.c_int, u8p]
    L.gt_lut_cmax.argtypes = [u8p]
    L.gt_lut_cmax.restype = C.c_uint32
    for f in ("gt_score_int", "gt_score_milli"):
        getattr(L, f).argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
    L.gt_load_nodes.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, u8p, u8p, u8p, u16p, u16p]
    L.gt_update_free.argtypes = [C.c_void_p, u32p, u16p, C.c_uint32]
    L.gt_tree_tables.argtypes = [C.POINTER(GtTree), u8p, C.POINTER(C.c_uint8), u8p, u16p]
    L.gt_load_trees.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(GtTree), u16p]
    L.gt_gaia_select.argtypes = [C.c_void_p, C.c_uint32, u32p]
    L.gt_comm_mode.argtypes = [C.c_void_p]
    L.gt_level_from_p2plink.argtypes = [C.c_int32]
    L.gt_levels_from_p2plink.argtypes = [C.POINTER(C.c_int32), C.c_uint32, u8p]
    L.gt_wire_parse_extender_args.argtypes = [C.c_char_p, C.c_size_t, u32p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t,
                                              C.POINTER(C.c_int32)]
    L.gt_wire_host_priorities.argtypes = [C.c_char_p, C.c_uint32, u8p, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.gt_wire_parse_binding_args.argtypes = [C.c_char_p, C.c_size_t] + [C.c_char_p, C.c_size_t] * 4
    L.gt_wire_binding_result.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
    L.gt_set_node_names.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
    L.gt_ext_sort_json.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.gt_ext_bind_json.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_int64, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    L.gt_ext_pod_deleted.argtypes = [C.c_void_p, C.c_char_p]
    L.gt_score_all.argtypes = [C.c_void_p, C.c_void_p]
    L.gt_score_dirty.argtypes = [C.c_void_p]
    L.gt_best_node.argtypes = [C.c_void_p, u8p, C.c_uint32, u64p]
    L.gt_allocate.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, u16p, u16p]
    L.gt_commit.argtypes = [C.c_void_p, u32p, u16p, C.c_uint32]
    L.gt_release.argtypes = [C.c_void_p, u32p, u16p, C.c_uint32]
    L.gt_score_all_async.argtypes = [C.c_void_p]
    L.gt_sync.argtypes = [C.c_void_p]
    L.gt_stream.argtypes = [C.c_void_p]
    L.gt_stream.restype = C.c_void_p
    L.gt_device_best.argtypes = [C.c_void_p]
    L.gt_device_best.restype = C.c_void_p
    L.gt_num_nodes.argtypes = [C.c_void_p]
    L.gt_num_nodes.restype = C.c_uint32
    L.gt_launch_count.argtypes = [C.c_void_p]
    L.gt_launch_count.restype = C.c_uint64
    L.gt_count_subsets.argtypes = [C.c_void_p, u64p, u64p]
    L.gt_time_kernel.argtypes = [C.c_void_p, C.c_char_p, C.c_uint32, C.POINTER(C.c_float)]
    L.gt_comm_unique_id.argtypes = [u8p]
    if hasattr(L, "gt_host_free"):      # (an older build of the ABI loaded through GT_LIB for an A/B run may lack these)
        L.gt_time_pass_rotated.argtypes = [C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32, C.POINTER(C.c_float)]
        L.gt_best_node_local.argtypes = [C.c_void_p, u64p]
        L.gt_comm_share.argtypes = [C.c_void_p, C.c_void_p]
        L.gt_host_free.argtypes = [C.c_void_p]
        L.gt_host_free.restype = C.POINTER(C.c_uint16)
        L.gt_host_free_changed.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.gt_host_scores.argtypes = [C.c_void_p]
        L.gt_host_scores.restype = C.POINTER(C.c_uint8)
    if hasattr(L, "gt_score_all_async_many"):
        L.gt_score_all_async_many.argtypes = [C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32]
    if hasattr(L, "gt_score_all_async_many_timed"):
        L.gt_score_all_async_many_timed.argtypes = [C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
    L.gt_comm_init.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int]
    L.gt_synth_cluster.argtypes = [C.c_int, C.c_uint64, C.c_uint32, C.c_uint32, u8p, u8p, u16p, u8p, u8p, u16p]
    L.gt_schedule.argtypes = [C.c_void_p, u8p, u32p, C.c_uint32, u32p, u16p, u64p]
    L.gt_synth_events.argtypes = [C.c_int, C.c_uint64, C.c_uint32, u8p, u32p]
    L.gt_parse_topo_matrix.argtypes = [C.c_char_p, C.c_size_t, u8p, u32p]
    L.gt_annotation_encode.argtypes = [u8p, C.c_uint32, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
    L.gt_annotation_decode.argtypes = [C.c_char_p, C.c_size_t, u8p, u32p]
    L.gt_group_encode.argtypes = [C.c_uint16, C.c_char_p, C.c_size_t]
    L.gt_group_decode.argtypes = [C.c_char_p, u16p]
    L.gt_bind_annotations.argtypes = [C.c_uint16, C.c_int64, C.c_int, C.c_char_p, C.c_size_t]
    L.gt_ext_sort.argtypes = [C.c_void_p, C.c_uint32, u8p, u64p]
    L.gt_ext_bind.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int64, u16p, C.c_char_p, C.c_size_t]
    L.gt_plugin_allocate.argtypes = [C.c_char_p, C.c_int64, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    L.gt_score_nodes_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p,
                                        C.c_char_p, C.c_size_t]
    _LIB = L
    return L


def _p(a, t):
    return a.ctypes.data_as(t) if a is not None else None


def lut_preset(preset=LUT_DEFAULT):
    out = np.zeros(16, np.uint8)
    rc = _lib().gt_lut_preset(preset, _p(out, u8p))
    if rc:
        raise GtError(rc, "gt_lut_preset")
    return out


def lut_cmax(lut):
    """The normaliser of the score formula: largest cost over the link kinds of the table (codes 1..SYS)."""
    return int(_lib().gt_lut_cmax(_p(np.ascontiguousarray(lut, np.uint8), u8p)))


def score_int(s, length, cmax):
    return int(_lib().gt_score_int(s, length, cmax))


def score_milli(s, length, cmax):
    return int(_lib().gt_score_milli(s, length, cmax))


def synth_cluster(config, seed, first, count, gaia=False):
    """Synthetic cluster of SURVEY.md section 8d: (levels[count,256], n_gpus, free, cls[, access, sibling])."""
    lv = np.zeros((count, 256), np.uint8)
    n = np.zeros(count, np.uint8)
    fr = np.zeros(count, np.uint16)
    cl = np.zeros(count, np.uint8)
    ac = np.zeros((count, 16), np.uint8) if gaia else None
    sib = np.zeros((count, 16), np.uint16) if gaia else None
    rc = _lib().gt_synth_cluster(config, seed, first, count, _p(lv, u8p), _p(n, u8p), _p(fr, u16p), _p(cl, u8p),
                                 _p(ac, u8p), _p(sib, u16p))
    if rc:
        raise GtError(rc, "gt_synth_cluster")
    return (lv, n, fr, cl, ac, sib) if gaia else (lv, n, fr, cl)


def score_nodes_static(d_cost, d_n_gpus, d_free_mask, n_nodes, d_best, stream=None):
    """Allocation-free K0 + K1 on caller-owned DEVICE buffers (raw pointers); see gt_score_nodes_static
    in include/gputopo.h for who needs this and its limits.  Enqueues only: synchronise `stream` before
    reading d_best."""
    buf = C.create_string_buffer(256)
    rc = _lib().gt_score_nodes_static(d_cost, d_n_gpus, d_free_mask, n_nodes, d_best, stream, buf, len(buf))
    if rc:
        raise GtError(rc, "gt_score_nodes_static", buf.value.decode())


def synth_events(config, seed, n_events):
    k = np.zeros(n_events, np.uint8)
    arg = np.zeros(n_events, np.uint32)
    rc = _lib().gt_synth_events(config, seed, n_events, _p(k, u8p), _p(arg, u32p))
    if rc:
        raise GtError(rc, "gt_synth_events")
    return k, arg


def comm_unique_id():
    out = np.zeros(128, np.uint8)
    rc = _lib().gt_comm_unique_id(_p(out, u8p))
    if rc:
        raise GtError(rc, "gt_comm_unique_id")
    return out


class Allocator:
    """One context = one GPU's shard of the cluster (nodes [first_node, first_node + max_nodes))."""

    def __init__(self, max_nodes, device=0, first_node=0, k1_policy=K1_FIRST_UNUSED, stream=None, flags=0):
        self._L = _lib()
        cfg = GtConfig(C.sizeof(GtConfig), device, max_nodes, first_node, k1_policy, flags, stream)
        h = C.c_void_p()
        rc = self._L.gt_create(C.byref(cfg), C.byref(h))
        if rc:
            raise GtError(rc, "gt_create")
        self._h = h
        self.max_nodes = max_nodes
        self.first_node = first_node

    def _ck(self, rc, what):
        if rc:
            raise GtError(rc, what, self._L.gt_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.gt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def num_nodes(self):
        return int(self._L.gt_num_nodes(self._h))

    @property
    def launch_count(self):
        return int(self._L.gt_launch_count(self._h))

    @property
    def stream(self):
        return self._L.gt_stream(self._h)

    @property
    def device_best_ptr(self):
        return self._L.gt_device_best(self._h)

    def set_cost_lut(self, lut):
        lut = np.ascontiguousarray(lut, np.uint8)
        assert lut.size == 16
        self._ck(self._L.gt_set_cost_lut(self._h, _p(lut, u8p)), "gt_set_cost_lut")

    def load_nodes(self, levels, n_gpus, free, first=0, access_cost=None, sibling_mask=None):
        levels = np.ascontiguousarray(levels, np.uint8).reshape(-1, 256)
        n_gpus = np.ascontiguousarray(n_gpus, np.uint8)
        free = np.ascontiguousarray(free, np.uint16)
        count = len(n_gpus)
        assert levels.shape[0] == count and free.shape[0] == count
        ac = np.ascontiguousarray(access_cost, np.uint8) if access_cost is not None else None
        sib = np.ascontiguousarray(sibling_mask, np.uint16) if sibling_mask is not None else None
        self._ck(self._L.gt_load_nodes(self._h, first, count, _p(levels, u8p), _p(n_gpus, u8p), _p(ac, u8p),
                                       _p(sib, u16p), _p(free, u16p)), "gt_load_nodes")

    def load_trees(self, trees, free, first=0):
        """Nodes described by Gaia trees (list of GtTree): tables derived by the library, tree kept for gaia_select."""
        arr = (GtTree * len(trees))(*trees)
        free = np.ascontiguousarray(free, np.uint16)
        assert free.shape[0] == len(trees)
        self._ck(self._L.gt_load_trees(self._h, first, len(trees), arr, _p(free, u16p)), "gt_load_trees")

    def gaia_select(self, m):
        """Gaia Alg. 1 on the device for request size m: per node cost<<16|mask (m = 1 Singular, m >= 2 Link)."""
        out = np.empty(self.num_nodes, np.uint32)
        self._ck(self._L.gt_gaia_select(self._h, m, _p(out, u32p)), "gt_gaia_select")
        return out

    def comm_mode(self):
        """0 = no communicator, 1 = NCCL all-reduce, 2 = own kernel over peer memory"""
        return int(self._L.gt_comm_mode(self._h))

    def update_free(self, masks, node_ids=None):
        masks = np.ascontiguousarray(masks, np.uint16)
        ids = np.ascontiguousarray(node_ids, np.uint32) if node_ids is not None else None
        self._ck(self._L.gt_update_free(self._h, _p(ids, u32p), _p(masks, u16p), len(masks)), "gt_update_free")

    def score_all(self, out=None, fetch=True):
        """Prioritize for every node and every k.  Returns best[n_nodes,16] (u32 cost<<16|mask)."""
        n = self.num_nodes
        if fetch and out is None:
            out = np.empty((n, 16), np.uint32)
        ptr = out.ctypes.data if (fetch and out is not None) else None
        self._ck(self._L.gt_score_all(self._h, ptr), "gt_score_all")
        return out if fetch else None

    def score_all_into(self, host_ptr):
        self._ck(self._L.gt_score_all(self._h, host_ptr), "gt_score_all")

    def ext_sort_into(self, gpu_count, scores, place=None):
        """`sort` verb into a caller buffer (numpy u8[n_nodes], ideally pinned); place: optional u64[16]."""
        self._ck(self._L.gt_ext_sort(self._h, gpu_count, _p(scores, u8p), _p(place, u64p)), "gt_ext_sort")

    def host_free(self):
        """The context's pinned mirror of the free masks as a numpy view (max_nodes u16): edit in place, then
        host_free_changed()."""
        ptr = self._L.gt_host_free(self._h)
        if not ptr:
            raise GtError(-3, "gt_host_free", self._L.gt_last_error(self._h).decode())
        return np.ctypeslib.as_array(ptr, shape=(self.max_nodes,))

    def host_free_changed(self, first=0, count=None):
        self._ck(self._L.gt_host_free_changed(self._h, first, self.num_nodes if count is None else count),
                 "gt_host_free_changed")

    def host_scores(self):
        """The mapped pinned buffer the `sort` verb's scores land in (n_nodes u8), valid until the next ext_sort."""
        return np.ctypeslib.as_array(self._L.gt_host_scores(self._h), shape=(self.num_nodes,))

    def ext_sort_inplace(self, gpu_count, place=None):
        """`sort` verb without the copy out: scores are read from host_scores()."""
        self._ck(self._L.gt_ext_sort(self._h, gpu_count, None, _p(place, u64p)), "gt_ext_sort")

    def score_dirty(self):
        self._ck(self._L.gt_score_dirty(self._h), "gt_score_dirty")

    def score_all_async(self):
        self._ck(self._L.gt_score_all_async(self._h), "gt_score_all_async")

    def sync(self):
        self._ck(self._L.gt_sync(self._h), "gt_sync")

    def best_node(self, ks):
        ks = np.ascontiguousarray(ks, np.uint8)
        out = np.zeros(len(ks), np.uint64)
        self._ck(self._L.gt_best_node(self._h, _p(ks, u8p), len(ks), _p(out, u64p)), "gt_best_node")
        return out

    def allocate(self, node, k):
        m = C.c_uint16(0)
        c = C.c_uint16(0)
        self._ck(self._L.gt_allocate(self._h, node, k, C.byref(m), C.byref(c)), "gt_allocate")
        return int(m.value), int(c.value)

    def commit(self, nodes, masks):
        nodes = np.ascontiguousarray(nodes, np.uint32)
        masks = np.ascontiguousarray(masks, np.uint16)
        self._ck(self._L.gt_commit(self._h, _p(nodes, u32p), _p(masks, u16p), len(nodes)), "gt_commit")

    def release(self, nodes, masks):
        nodes = np.ascontiguousarray(nodes, np.uint32)
        masks = np.ascontiguousarray(masks, np.uint16)
        self._ck(self._L.gt_release(self._h, _p(nodes, u32p), _p(masks, u16p), len(nodes)), "gt_release")

    def schedule(self, ev_k, ev_arg):
        """Sequential sort -> bind loop over an event trace.  Returns (placed, node[], mask[])."""
        ev_k = np.ascontiguousarray(ev_k, np.uint8)
        ev_arg = np.ascontiguousarray(ev_arg, np.uint32)
        n = len(ev_k)
        on = np.zeros(n, np.uint32)
        om = np.zeros(n, np.uint16)
        placed = C.c_uint64(0)
        self._ck(self._L.gt_schedule(self._h, _p(ev_k, u8p), _p(ev_arg, u32p), n, _p(on, u32p), _p(om, u16p),
                                     C.byref(placed)), "gt_schedule")
        return int(placed.value), on, om

    def count_subsets(self):
        s = C.c_uint64(0)
        a = C.c_uint64(0)
        self._ck(self._L.gt_count_subsets(self._h, C.byref(s), C.byref(a)), "gt_count_subsets")
        return int(s.value), int(a.value)

    def time_kernel(self, which, iters=20):
        ms = C.c_float(0)
        self._ck(self._L.gt_time_kernel(self._h, which.encode(), iters, C.byref(ms)), "gt_time_kernel")
        return float(ms.value)

    def comm_share(self, owner):
        """Join the communicator of another Allocator of this process on the same device (it must be closed last)."""
        self._ck(self._L.gt_comm_share(self._h, owner._h), "gt_comm_share")
        self._comm_owner = owner      # keeps the lender alive

    def best_node_local(self):
        """This rank's own 16 placement keys (no exchange), recomputed from its best table."""
        out = np.zeros(16, np.uint64)
        self._ck(self._L.gt_best_node_local(self._h, _p(out, u64p)), "gt_best_node_local")
        return out

    def comm_init(self, unique_id, rank, world):
        uid = np.ascontiguousarray(unique_id, np.uint8)
        self._ck(self._L.gt_comm_init(self._h, _p(uid, u8p), rank, world), "gt_comm_init")


def score_all_async_many(allocators, passes):
    """`passes` scoring passes enqueued round-robin over the allocators in one call (no sync)."""
    arr = (C.c_void_p * len(allocators))(*[a._h.value for a in allocators])
    rc = _lib().gt_score_all_async_many(arr, len(allocators), passes)
    if rc:
        raise GtError(rc, "gt_score_all_async_many", "; ".join(_lib().gt_last_error(a._h).decode() for a in allocators[:1]))


def score_all_async_many_timed(allocators, passes, ev_start=None, ev_stop=None):
    """score_all_async_many between two CUDA events (raw cudaEvent_t handles, e.g. torch.cuda.Event.cuda_event of an event
    that has been recorded once), recorded by the library directly around its launches."""
    arr = (C.c_void_p * len(allocators))(*[a._h.value for a in allocators])
    rc = _lib().gt_score_all_async_many_timed(arr, len(allocators), passes, C.c_void_p(ev_start or 0), C.c_void_p(ev_stop or 0))
    if rc:
        raise GtError(rc, "gt_score_all_async_many_timed", "; ".join(_lib().gt_last_error(a._h).decode() for a in allocators[:1]))


def time_pass_rotated(allocators, iters=64):
    """(pass - K1, K1, pass) microseconds: passes back to back while the contexts take turns (cold L2 once their tables
    exceed it), K1 timed in place by its own global-timer stamps."""
    arr = (C.c_void_p * len(allocators))(*[a._h.value for a in allocators])
    out = (C.c_float * 3)()
    rc = _lib().gt_time_pass_rotated(arr, len(allocators), iters, out)
    if rc:
        raise GtError(rc, "gt_time_pass_rotated", _lib().gt_last_error(allocators[0]._h).decode())
    return tuple(float(x) for x in out)


def place_fields(key):
    key = int(key)
    return key >> 48, (key >> 16) & 0xFFFFFFFF, key & 0xFFFF


# ---------------------------------------------------------------- text formats around the path (host only)
def parse_topo_matrix(text):
    """`nvidia-smi topo -m` output -> (levels[256] u8, n_gpus).  design.md:27-29."""
    raw = text.encode() if isinstance(text, str) else text
    lv = np.zeros(256, np.uint8)
    n = C.c_uint32(0)
    rc = _lib().gt_parse_topo_matrix(raw, len(raw), _p(lv, u8p), C.byref(n))
    if rc:
        raise GtError(rc, "gt_parse_topo_matrix")
    return lv, int(n.value)


def annotation_encode(levels, n_gpus):
    """Level matrix -> node annotation text, one 'GPU_<ABBR>_<i>_<j>: <description>' line per pair (design.md:78-82)."""
    lv = np.ascontiguousarray(levels, np.uint8).reshape(256)
    buf = C.create_string_buffer(8192)
    ln = C.c_size_t(0)
    rc = _lib().gt_annotation_encode(_p(lv, u8p), n_gpus, buf, len(buf), C.byref(ln))
    if rc:
        raise GtError(rc, "gt_annotation_encode")
    return buf.value.decode()


def annotation_decode(text):
    raw = text.encode() if isinstance(text, str) else text
    lv = np.zeros(256, np.uint8)
    n = C.c_uint32(0)
    rc = _lib().gt_annotation_decode(raw, len(raw), _p(lv, u8p), C.byref(n))
    if rc:
        raise GtError(rc, "gt_annotation_decode")
    return lv, int(n.value)


def group_encode(mask):
    buf = C.create_string_buffer(64)
    rc = _lib().gt_group_encode(mask, buf, len(buf))
    if rc:
        raise GtError(rc, "gt_group_encode")
    return buf.value.decode()


def group_decode(text):
    m = C.c_uint16(0)
    rc = _lib().gt_group_decode(text.encode(), C.byref(m))
    if rc:
        raise GtError(rc, "gt_group_decode")
    return int(m.value)


def bind_annotations(mask, assume_time, assigned=False):
    buf = C.create_string_buffer(256)
    rc = _lib().gt_bind_annotations(mask, assume_time, 1 if assigned else 0, buf, len(buf))
    if rc:
        raise GtError(rc, "gt_bind_annotations")
    return dict(l.split(": ", 1) for l in buf.value.decode().splitlines())


# design.md:92-113: the scheduler policy that points kube-scheduler at the extender
EXTENDER_POLICY = {
    "kind": "Policy", "apiVersion": "v1",
    "extenders": [{
        "urlPrefix": "http://127.0.0.1:32743/gputopology-scheduler", "PrioritizeVerb": "sort", "bindVerb": "bind",
        "enableHttps": False, "nodeCacheCapable": True,
        "managedResources": [{"name": "aliyun.com/gpu", "ignoredByScheduler": False}], "ignorable": False,
    }],
}
RESOURCE_NAME = "aliyun.com/gpu"


def levels_from_p2plink(types):
    """n x n matrix of nvml.P2PLinkType values (the reference's gpuTopology map, design.md:57-70) -> levels[256]."""
    t = np.ascontiguousarray(types, np.int32)
    assert t.ndim == 2 and t.shape[0] == t.shape[1]
    out = np.zeros(256, np.uint8)
    rc = _lib().gt_levels_from_p2plink(t.ctypes.data_as(C.POINTER(C.c_int32)), t.shape[0], _p(out, u8p))
    if rc:
        raise GtError(rc, "gt_levels_from_p2plink")
    return out


def wire_parse_extender_args(body):
    """ExtenderArgs JSON (bytes or str) -> (gpu_count, pod uid, node names or None when the field is absent)."""
    if isinstance(body, str):
        body = body.encode()
    cnt, n = C.c_uint32(0), C.c_int32(0)
    uid = C.create_string_buffer(256)
    names = C.create_string_buffer(len(body) + 1)
    rc = _lib().gt_wire_parse_extender_args(body, len(body), C.byref(cnt), uid, len(uid), names, len(names), C.byref(n))
    if rc:
        raise GtError(rc, "gt_wire_parse_extender_args")
    lst = None if n.value < 0 else (names.value.decode().split("\n") if n.value else [])
    return cnt.value, uid.value.decode(), lst


def wire_host_priorities(names, scores):
    """HostPriorityList JSON text for node names and their scores."""
    scores = np.ascontiguousarray(scores, np.uint8)
    joined = "\n".join(names).encode()
    out = C.create_string_buffer(len(joined) * 6 + 32 * len(names) + 16)
    n = C.c_size_t(0)
    rc = _lib().gt_wire_host_priorities(joined, len(names), _p(scores, u8p), out, len(out), C.byref(n))
    if rc:
        raise GtError(rc, "gt_wire_host_priorities")
    return out.value.decode()


def wire_parse_binding_args(body):
    """ExtenderBindingArgs JSON -> dict(PodName, PodNamespace, PodUID, Node)."""
    if isinstance(body, str):
        body = body.encode()
    bufs = [C.create_string_buffer(len(body) + 1) for _ in range(4)]
    args = []
    for b in bufs:
        args += [b, len(b)]
    rc = _lib().gt_wire_parse_binding_args(body, len(body), *args)
    if rc:
        raise GtError(rc, "gt_wire_parse_binding_args")
    return dict(zip(("PodName", "PodNamespace", "PodUID", "Node"), (b.value.decode() for b in bufs)))


def wire_binding_result(error=""):
    out = C.create_string_buffer(len(error.encode()) * 6 + 32)
    rc = _lib().gt_wire_binding_result(error.encode(), out, len(out))
    if rc:
        raise GtError(rc, "gt_wire_binding_result")
    return out.value.decode()


class ExtenderHTTP:
    """What the reference's HTTP handlers do with a request body (design.md:92-121), on the C-ABI's JSON entry points:
    `handle("sort", body)` / `handle("bind", body)` return the reply body; no server is started here."""

    def __init__(self, allocator, node_names):
        self.alloc = allocator
        joined = "\n".join(node_names).encode()
        allocator._ck(allocator._L.gt_set_node_names(allocator._h, joined, len(joined)), "gt_set_node_names")
        self.annotations = {}      # pod uid -> annotations written by the last bind

    def handle(self, verb, body, now=0):
        if isinstance(body, str):
            body = body.encode()
        a = self.alloc
        if verb == "sort":
            out = C.create_string_buffer(64 * a.num_nodes + len(body) * 6 + 64)
            n = C.c_size_t(0)
            a._ck(a._L.gt_ext_sort_json(a._h, body, len(body), out, len(out), C.byref(n)), "gt_ext_sort_json")
            return out.raw[:n.value].decode()
        if verb == "bind":
            out = C.create_string_buffer(1024)
            ann = C.create_string_buffer(256)
            a._ck(a._L.gt_ext_bind_json(a._h, body, len(body), now, out, len(out), ann, len(ann)), "gt_ext_bind_json")
            if ann.value:
                uid = wire_parse_binding_args(body)["PodUID"]
                self.annotations[uid] = dict(l.split(": ", 1) for l in ann.value.decode().splitlines())
            return out.value.decode()
        raise ValueError("the extender has no %r verb (design.md:117: sort and bind only)" % verb)

    def pod_deleted(self, uid):
        a = self.alloc
        a._ck(a._L.gt_ext_pod_deleted(a._h, uid.encode()), "gt_ext_pod_deleted")


class Extender:
    """Mirror of the reference's scheduler extender for this path (design.md:88-121): `sort` answers with a
    HostPriorityList (score 0..10 per node name), `bind` picks the GPUs on the chosen node and returns the pod
    annotations of design.md:229-231.  There is no filter verb (design.md:117)."""

    def __init__(self, allocator, node_names):
        self.alloc = allocator
        self.names = list(node_names)
        self.index = {n: i for i, n in enumerate(self.names)}
        self.pods = {}     # pod uid -> (node index, mask): "the pods recorded on the device" design.md:234

    def sort(self, gpu_count, node_names=None):
        """ExtenderArgs{Pod, NodeNames} -> HostPriorityList [{Host, Score}] (nodeCacheCapable: names only)."""
        L = self.alloc._L
        scores = np.zeros(self.alloc.num_nodes, np.uint8)
        self.alloc._ck(L.gt_ext_sort(self.alloc._h, gpu_count, _p(scores, u8p), None), "gt_ext_sort")
        names = self.names if node_names is None else node_names
        return [{"Host": n, "Score": int(scores[self.index[n]])} for n in names]

    def bind(self, pod_name, pod_namespace, pod_uid, node, gpu_count, now):
        """ExtenderBindingArgs -> (ExtenderBindingResult, pod annotations)."""
        L = self.alloc._L
        mask = C.c_uint16(0)
        buf = C.create_string_buffer(256)
        if node not in self.index:
            return {"Error": "unknown node %s" % node}, {}
        rc = L.gt_ext_bind(self.alloc._h, self.index[node], gpu_count, now, C.byref(mask), buf, len(buf))
        if rc:
            return {"Error": "%s: %s" % (L.gt_strerror(rc).decode(), L.gt_last_error(self.alloc._h).decode())}, {}
        self.pods[pod_uid] = (self.index[node], int(mask.value))
        return {"Error": ""}, dict(l.split(": ", 1) for l in buf.value.decode().splitlines())

    def pod_deleted(self, pod_uid):
        node, mask = self.pods.pop(pod_uid)
        self.alloc.release([node], [mask])


class DevicePlugin:
    """Mirror of the reference's device plugin for this path: Allocate reads the device list bind wrote and
    returns it as NVIDIA_VISIBLE_DEVICES, flipping ALIYUN_COM_GPU_ASSIGNED to true (design.md:236-246)."""

    resource_name = RESOURCE_NAME      # ListAndWatch advertises this resource (design.md:84-86)

    @staticmethod
    def allocate(pod_annotations, now):
        text = "".join("%s: %s\n" % kv for kv in pod_annotations.items())
        env = C.create_string_buffer(128)
        ann = C.create_string_buffer(256)
        rc = _lib().gt_plugin_allocate(text.encode(), now, env, len(env), ann, len(ann))
        if rc:
            raise GtError(rc, "gt_plugin_allocate")
        k, v = env.value.decode().split("=", 1)
        return {"Envs": {k: v}}, dict(l.split(": ", 1) for l in ann.value.decode().splitlines())


def shard_bounds(total_nodes, rank, world):
    """Contiguous node range of a rank: [rank*N/G, (rank+1)*N/G)."""
    lo = (total_nodes * rank) // world
    hi = (total_nodes * (rank + 1)) // world
    return lo, hi


def merge_place(keys_per_shard):
    """Host-side merge of per-shard placement keys (one process driving several contexts, or ranks
    exchanging keys over a non-NCCL backend): element-wise minimum of the u64 keys."""
    out = np.asarray(keys_per_shard[0], np.uint64).copy()
    for k in keys_per_shard[1:]:
        out = np.minimum(out, np.asarray(k, np.uint64))
    return out


def allreduce_place_torch(keys, group=None):
    """Minimum of the placement keys over a torch.distributed group whose backend has no unsigned
    64-bit type (gloo): GT_NO_PLACE is mapped to INT64_MAX for the exchange."""
    import torch
    import torch.distributed as dist
    k = np.asarray(keys, np.uint64)
    no = k == np.uint64(NO_PLACE)
    t = torch.from_numpy(np.where(no, np.uint64(2 ** 63 - 1), k).astype(np.int64))
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    r = t.numpy().astype(np.uint64)
    return np.where(r == np.uint64(2 ** 63 - 1), np.uint64(NO_PLACE), r)
