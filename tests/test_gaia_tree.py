"""Gaia tree input format: host-side tables derived from a tree (CPU tests) and
Alg. 1 / 3 / 4 on the device through gt_load_trees + gt_gaia_select against the oracle (GPU tests)."""
import json
import os

import numpy as np
import pytest

import oracle_lib as ol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KA = json.load(open(os.path.join(ROOT, "tests", "golden", "known_answers.json")))
INF = 0xFFFF0000


def to_oracle(tree):
    return ol.Tree.from_buffer_copy(bytes(tree))


def spec_codes(spec):
    """oracle_lib tree spec ('PXB', ...) -> numeric level codes for the product's make_tree"""
    if isinstance(spec[0], str):
        return (ol.LV["SYS" if spec[0] == "SOC" else spec[0]],) + tuple(spec_codes(c) for c in spec[1:])
    return spec


def random_tree(gt, rng, n):
    """Random tree over GPUs 0..n-1: ids shuffled over the leaves, small access costs so that ties are common."""
    ids = list(rng.permutation(n))

    def build(group, depth):
        if len(group) == 1:
            return (int(group[0]), int(rng.randint(0, 4)))
        if depth >= 4 or (len(group) <= 4 and rng.rand() < 0.4):
            kids = [(int(g), int(rng.randint(0, 4))) for g in group]
        else:
            parts = rng.randint(2, min(4, len(group)) + 1)
            cuts = sorted(rng.choice(np.arange(1, len(group)), parts - 1, replace=False))
            kids = [build(p, depth + 1) for p in np.split(np.array(group), cuts)]
        return (int(rng.randint(1, 12)),) + tuple(kids)
    spec = build(ids, 0)
    if len(spec) == 2 and all(isinstance(x, int) for x in spec):   # a single GPU: give it a topology parent
        spec = (11, spec)
    return gt.make_tree(spec)


# ------------------------------------------------------------------ CPU: derived tables
@pytest.mark.parametrize("name,spec,n", [("F2", ol.F2_TREE, 8), ("F3", ol.F3_TREE, 4)])
def test_tree_tables_match_oracle_and_fixture(gt, oracle, name, spec, n):
    t = gt.make_tree(spec_codes(spec))
    lv, ng, ac, sib = gt.tree_tables(t)
    olv, oac, osib = oracle.tree_to_tables(ol.make_tree(spec), n)
    assert ng == n
    assert (lv == olv.reshape(-1)).all() and (ac == oac).all() and (sib == osib).all()
    cls_lv, cls_n, cls_ac, cls_sib = oracle.class_topology(ol.CLS["F2_PCIE" if name == "F2" else "F3_TREE4"])
    assert (lv == cls_lv.reshape(-1)).all() and (ac == cls_ac).all() and (sib == cls_sib).all()


def test_tree_tables_random_trees_match_oracle(gt, oracle):
    rng = np.random.RandomState(11)
    for _ in range(300):
        n = int(rng.randint(1, 17))
        t = random_tree(gt, rng, n)
        lv, ng, ac, sib = gt.tree_tables(t)
        olv, oac, osib = oracle.tree_to_tables(to_oracle(t), n)
        assert ng == n
        assert (lv == olv.reshape(-1)).all() and (ac == oac).all() and (sib == osib).all()
        m = lv.reshape(16, 16)
        assert (m == m.T).all() and (np.diag(m) == 0).all()


def test_malformed_trees_are_rejected(gt):
    good = gt.make_tree((11, (8, (0, 1), (1, 1)), (7, (2, 1), (3, 1))))
    gt.tree_tables(good)

    def broken(edit):
        t = gt.GtTree.from_buffer_copy(bytes(good))
        edit(t)
        with pytest.raises(gt.GtError) as e:
            gt.tree_tables(t)
        assert e.value.code == -1

    broken(lambda t: setattr(t, "n_nodes", 0))
    broken(lambda t: setattr(t, "n_nodes", 32))
    broken(lambda t: t.parent.__setitem__(1, 1))          # its own parent
    broken(lambda t: t.parent.__setitem__(4, 2))          # a leaf as a parent
    broken(lambda t: t.parent.__setitem__(1, 9))          # out of range
    broken(lambda t: t.gpu_id.__setitem__(3, 0))          # duplicate GPU id
    broken(lambda t: t.gpu_id.__setitem__(3, 16))         # GPU id out of range
    broken(lambda t: t.gpu_id.__setitem__(3, 5))          # hole in the ids
    broken(lambda t: t.access_cost.__setitem__(2, 256))
    broken(lambda t: t.level.__setitem__(1, 16))

    def cycle(t):                                          # 1 -> 4 -> 1 never reaches the root
        t.parent[1] = 4
        t.parent[4] = 1
    broken(cycle)


# ------------------------------------------------------------------ GPU: Alg. 1 / 3 / 4 on the device
def oracle_select(oracle, tree, lv, n, free, m, lut):
    mask = oracle.gaia_dispatch(to_oracle(tree), int(free) & ((1 << n) - 1), m)
    if mask == 0:
        return INF
    return (oracle.subset_cost(lv, lut, mask) << 16) | mask


@pytest.mark.gpu
@pytest.mark.parametrize("ka", ["KA2", "KA4b", "KA6", "KA3", "KA4a", "KA5"])
def test_gaia_known_answers_on_device(gt, oracle, ka):
    """The paper's worked examples (Tables I-IV, Fig. 5) answered by the device kernels from the tree input."""
    k = KA[ka]
    spec = ol.F2_TREE if k["topology"] == "F2_PCIE" else ol.F3_TREE
    tree = gt.make_tree(spec_codes(spec))
    n = 8 if k["topology"] == "F2_PCIE" else 4
    free = ((1 << n) - 1) & ~sum(1 << g for g in k["used"])
    with gt.Allocator(4, k1_policy=gt.K1_GAIA_SINGULAR) as a:
        a.load_trees([tree], np.array([free], np.uint16))
        got = int(a.gaia_select(k["m"])[0]) & 0xFFFF
    assert got in {sum(1 << g for g in e) for e in k["expect"]}


@pytest.mark.gpu
def test_gaia_select_matches_oracle_on_random_trees(gt, oracle):
    rng = np.random.RandomState(5)
    lut = oracle.lut(0)
    trees, frees, ns = [], [], []
    for i in range(600):
        n = int(rng.randint(1, 17)) if i % 3 else int(rng.choice([4, 8, 16]))
        trees.append(random_tree(gt, rng, n))
        ns.append(n)
        frees.append(int(rng.randint(0, 1 << 16)))       # bits above n must be ignored
    frees = np.array(frees, np.uint16)
    tables = [gt.tree_tables(t) for t in trees]
    with gt.Allocator(len(trees)) as a:
        a.load_trees(trees, frees)
        for m in range(1, 17):
            got = a.gaia_select(m)
            for i, t in enumerate(trees):
                if m == 1:
                    lv, n, ac, sib = tables[i]
                    g = oracle.gaia_singular(n, int(frees[i]) & ((1 << n) - 1), ac, sib)
                    want = INF if g < 0 else 1 << g
                else:
                    want = oracle_select(oracle, t, tables[i][0], ns[i], frees[i], m, lut)
                assert int(got[i]) == want, (i, m, hex(int(got[i])), hex(want))
        # the exhaustive path sees the same nodes: Link can never beat it, and its table is unchanged by the policy call
        best = a.score_all()
        for m in range(2, 17):
            got = a.gaia_select(m)
            feas = got != INF
            assert ((best[:, m - 1] != INF) == feas).all()
            assert ((got[feas] >> 16) >= (best[feas, m - 1] >> 16)).all()


@pytest.mark.gpu
def test_gaia_select_follows_free_mask_updates_and_needs_a_tree(gt, oracle):
    lut = oracle.lut(0)
    tree = gt.make_tree(spec_codes(ol.F2_TREE))
    lv = gt.tree_tables(tree)[0]
    with gt.Allocator(8) as a:
        a.load_trees([tree] * 5, np.full(5, 0xFF, np.uint16))
        a.score_all()
        a.commit([1, 3], [0x0C, 0x03])                    # dirty nodes, not yet rescored
        free = np.array([0xFF, 0xF3, 0xFF, 0xFC, 0xFF], np.uint16)
        for m in (2, 3, 5):
            got = a.gaia_select(m)
            for i in range(5):
                assert int(got[i]) == oracle_select(oracle, tree, lv, 8, free[i], m, lut)
        best = a.score_all()                              # the pending commit is still scored afterwards
        ref, _ = oracle.exhaustive_cluster(np.tile(lv, (5, 1)), np.full(5, 8, np.uint8), free, lut, 1)
        assert (best == ref).all()
    with gt.Allocator(4) as a:
        lvs, ng, fr, _ = oracle.synth_cluster(2, 1, 0, 4)
        a.load_nodes(lvs, ng, fr)
        with pytest.raises(gt.GtError) as e:
            a.gaia_select(2)
        assert e.value.code == -8
        assert (a.gaia_select(1) != INF).sum() == (fr & 0xFF != 0).sum()
        with pytest.raises(gt.GtError):
            a.gaia_select(17)
