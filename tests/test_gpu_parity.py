"""Parity of the CUDA path (through the C-ABI) against the CPU oracle: bit-exact, integer work.
Run on the B200 box with `-m gpu`."""
import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu


def all_masks_cluster(oracle, cls):
    lv, n, _, _ = oracle.class_topology(cls)
    N = 1 << n
    levels = np.tile(lv, (N, 1))
    n_gpus = np.full(N, n, np.uint8)
    free = np.arange(N, dtype=np.uint16)
    return levels, n_gpus, free


def check_cluster(gt, oracle, levels, n_gpus, free, lut=None, threads=8):
    lut = oracle.lut(0) if lut is None else lut
    want, cnt = oracle.exhaustive_cluster(levels, n_gpus, free, lut, threads)
    # the captured-graph variant of the pass must give the same table and placement keys
    with gt.Allocator(len(n_gpus), flags=gt.FLAG_GRAPH) as a:
        a.set_cost_lut(lut)
        a.load_nodes(levels, n_gpus, free)
        assert (a.score_all() == want).all()
        assert (a.best_node(np.arange(1, 17, dtype=np.uint8)) == oracle.cluster_argmin(want)).all()
    # rows copied out of the GPU-built (class, free mask) table (falls back to enumeration with many classes)
    with gt.Allocator(len(n_gpus), flags=gt.FLAG_MEMO_PASS) as a:
        a.set_cost_lut(lut)
        a.load_nodes(levels, n_gpus, free)
        assert (a.score_all() == want).all()
        assert (a.best_node(np.arange(1, 17, dtype=np.uint8)) == oracle.cluster_argmin(want)).all()
        a.update_free(np.zeros(len(n_gpus), np.uint16))
        assert (a.score_all() == ol.INFEASIBLE).all()
    with gt.Allocator(len(n_gpus)) as a:
        a.set_cost_lut(lut)
        a.load_nodes(levels, n_gpus, free)
        got = a.score_all()
        assert a.count_subsets()[0] == cnt
        bad = np.nonzero((got != want).any(axis=1))[0]
        assert len(bad) == 0, "first mismatch node %d free=%#x got=%s want=%s" % (
            bad[0], free[bad[0]], got[bad[0]], want[bad[0]])
        place = a.best_node(np.arange(1, 17, dtype=np.uint8))
        assert (place == oracle.cluster_argmin(want)).all()
    return got


@pytest.mark.parametrize("cls", ["DGX1", "F1_RING", "F2_PCIE", "F3_TREE4"])
def test_every_free_mask_small_topologies(gt, oracle, cls):
    """all 2^n free masks x all k for each fixture topology (3^8 = 6561 subsets)."""
    check_cluster(gt, oracle, *all_masks_cluster(oracle, ol.CLS[cls]))


@pytest.mark.parametrize("cls", ["NVS16", "DUAL8"])
def test_every_free_mask_16_gpu_topologies(gt, oracle, cls):
    """3^16 = 43 M subsets per topology, every free mask of a 16-GPU node."""
    check_cluster(gt, oracle, *all_masks_cluster(oracle, ol.CLS[cls]))


def test_random_matrices_and_luts(gt, oracle):
    """Arbitrary (even asymmetric) level matrices and random u8 cost tables, n from 1 to 16."""
    rng = np.random.RandomState(11)
    N = 6000
    levels = rng.randint(0, 16, (N, 256)).astype(np.uint8)
    n_gpus = rng.randint(0, 17, N).astype(np.uint8)
    free = rng.randint(0, 1 << 16, N).astype(np.uint16)
    n_gpus[:200] = 16
    free[:200] |= 0xFF00       # plenty of large free sets
    free[200:232] = 0xFFFF
    n_gpus[200:232] = 16
    lut = rng.randint(0, 256, 16).astype(np.uint8)
    check_cluster(gt, oracle, levels, n_gpus, free, lut)


def test_known_answers_through_the_abi(gt, oracle):
    """KA2, KA4b, KA6 (m >= 2) and the k = 1 'first unused device' convention, on the GPU."""
    rows = [("F2_PCIE", [], 2, 0b11), ("F3_TREE4", [], 2, 0b1100), ("F3_TREE4", [2], 2, 0b0011)]
    for cls, used, m, want in rows:
        lv, n, _, _ = oracle.class_topology(ol.CLS[cls])
        free = ((1 << n) - 1) & ~sum(1 << u for u in used)
        with gt.Allocator(4) as a:
            a.load_nodes(lv[None, :], [n], [free])
            a.score_all()
            mask, cost = a.allocate(0, m)
            assert mask == want
            assert a.allocate(0, 1)[0] == (free & -free)


@pytest.mark.parametrize("config,seed,N", [(2, 0x5EED0002, 1000), (3, 0x5EED0003, 3000), (33, 0x5EED0003, 3000),
                                          (4, 0x5EED0004, 20000)])
def test_synthetic_configs(gt, oracle, config, seed, N):
    lv, n, fr, _ = gt.synth_cluster(config, seed, 0, N)
    check_cluster(gt, oracle, lv, n, fr)


def test_edge_cases(gt, oracle):
    lv, n, _, _ = oracle.class_topology(ol.CLS["DGX1"])
    lut = oracle.lut(0)
    with gt.Allocator(8) as a:
        # a single node; an empty node; n_gpus = 0; free bits above n_gpus
        levels = np.tile(lv, (4, 1))
        a.load_nodes(levels, [8, 8, 0, 8], [0xFF, 0x00, 0xFFFF, 0xFF0F])
        got = a.score_all()
        want, _ = oracle.exhaustive_cluster(levels, np.array([8, 8, 0, 8], np.uint8),
                                            np.array([0xFF, 0, 0xFFFF, 0xFF0F], np.uint16), lut, 1)
        assert (got == want).all()
        assert (got[1] == ol.INFEASIBLE).all() and (got[2] == ol.INFEASIBLE).all()
        place = a.best_node([1, 8, 9])
        assert gt.place_fields(place[0]) == (0, 0, 1)
        assert gt.place_fields(place[1])[1:] == (0, 0xFF)
        assert int(place[2]) == gt.NO_PLACE
        with pytest.raises(gt.GtError):
            a.load_nodes(levels, [8, 8, 17, 8], [0, 0, 0, 0])
        with pytest.raises(gt.GtError):
            a.load_nodes(np.tile(lv, (9, 1)), [8] * 9, [0] * 9)


def test_idempotent_and_lut_change(gt, oracle):
    lv, n, fr, _ = gt.synth_cluster(4, 99, 0, 5000)
    with gt.Allocator(5000) as a:
        a.load_nodes(lv, n, fr)
        b1 = a.score_all()
        b2 = a.score_all()
        assert (b1 == b2).all()
        marks = oracle.lut(1)
        a.set_cost_lut(marks)
        b3 = a.score_all()
        want, _ = oracle.exhaustive_cluster(lv, n, fr, marks, 8)
        assert (b3 == want).all()


def test_commit_release_rescore_dirty(gt, oracle):
    """Host commit loop: commit / release mark nodes dirty, gt_score_dirty re-scores exactly those."""
    lut = oracle.lut(0)
    lv, n, fr, _ = gt.synth_cluster(2, 0x5EED0002, 0, 300)
    free = fr.copy()
    rng = np.random.RandomState(5)
    with gt.Allocator(300) as a:
        a.load_nodes(lv, n, free)
        a.score_all(fetch=False)
        live = []
        for step in range(60):
            ks = rng.randint(1, 9, 4).astype(np.uint8)
            for k in ks:
                place = a.best_node([k])[0]
                want_best, _ = oracle.exhaustive_cluster(lv, n, free, lut, 4)
                assert int(place) == int(oracle.cluster_argmin(want_best)[k - 1])
                if int(place) == gt.NO_PLACE:
                    continue
                cost, node, mask = gt.place_fields(place)
                assert a.allocate(node, int(k)) == (mask, cost)
                a.commit([node], [mask])
                free[node] &= ~mask & 0xFFFF
                live.append((node, mask))
                a.score_dirty()
            if live and step % 3 == 0:
                node, mask = live.pop(rng.randint(len(live)))
                a.release([node], [mask])
                free[node] |= mask
                a.score_dirty()
        got = a.score_all()
        want, _ = oracle.exhaustive_cluster(lv, n, free, lut, 4)
        assert (got == want).all()
        with pytest.raises(gt.GtError):
            a.commit([0], [0xFFFF])     # GPUs not free


def test_commit_batch_is_checked_cumulatively(gt, oracle):
    """One batch cannot hand the same GPU out twice, and a refused batch leaves no trace."""
    lv, n, fr, _ = gt.synth_cluster(2, 0x5EED0002, 0, 8)
    free = np.full(8, 0xFF, np.uint16)
    with gt.Allocator(8) as a:
        a.load_nodes(lv, n, free)
        before = a.score_all()
        with pytest.raises(gt.GtError) as e:
            a.commit([3, 5, 3], [0x03, 0x10, 0x06])      # GPU 1 of node 3 twice
        assert e.value.code == -8
        with pytest.raises(gt.GtError) as e:
            a.commit([2], [0x0100])                      # GPU 8 on an 8-GPU node
        assert e.value.code == -1
        with pytest.raises(gt.GtError):
            a.release([2], [0x0100])
        a.score_dirty()
        assert (a.score_all() == before).all()
        a.commit([3, 5, 3], [0x03, 0x10, 0x0C])          # disjoint: fine
        free[3] &= ~0x0F & 0xFFFF
        free[5] &= ~0x10 & 0xFFFF
        a.score_dirty()
        want, _ = oracle.exhaustive_cluster(lv, n, free, oracle.lut(0), 2)
        assert (a.score_all() == want).all()


def test_full_size_properties(gt, oracle):
    """BASELINE config 4 at its full 100 000 nodes: size-independent properties + a sampled oracle check."""
    N = 100000
    lv, n, fr, cl = gt.synth_cluster(4, 0x5EED0004, 0, N)
    with gt.Allocator(N) as a:
        a.load_nodes(lv, n, fr)
        best = a.score_all()
        mask = best & 0xFFFF
        cost = best >> 16
        pc = np.array([bin(int(x)).count("1") for x in (fr & ((1 << n.astype(np.uint32)) - 1))])
        k = np.arange(1, 17)[None, :]
        feas = k <= pc[:, None]
        assert ((best == ol.INFEASIBLE) == ~feas).all()
        # chosen GPUs are free, and exactly k of them
        assert ((mask & ~fr[:, None].astype(np.uint32))[feas] == 0).all()
        popc = np.zeros_like(mask)
        for b in range(16):
            popc += (mask >> b) & 1
        assert (popc[feas] == np.broadcast_to(k, best.shape)[feas]).all()
        # k = 1 costs nothing and takes the lowest free GPU; k = popcount takes everything
        has1 = pc >= 1
        assert (cost[has1, 0] == 0).all()
        lowest = (fr & (~fr + 1)).astype(np.uint32)
        assert (mask[has1, 0] == lowest[has1]).all()
        # sampled rows against the oracle
        idx = np.random.RandomState(1).choice(N, 4000, replace=False)
        want, _ = oracle.exhaustive_cluster(lv[idx], n[idx], fr[idx], oracle.lut(0), 8)
        assert (best[idx] == want).all()
        # the cluster argmin is a checksum over the whole table
        assert (a.best_node(np.arange(1, 17, dtype=np.uint8)) == oracle.cluster_argmin(best)).all()


def test_sharded_contexts_agree_with_one_context(gt, oracle):
    """Node sharding invariance on one GPU: 1, 2 and 4 contexts with global node
    offsets, keys merged with min, must name the same winners (global node index is inside the key)."""
    total = 40000
    lv, n, fr, _ = gt.synth_cluster(4, 0x5EED0004, 0, total)
    ks = np.arange(1, 17, dtype=np.uint8)
    want = oracle.cluster_argmin(oracle.exhaustive_cluster(lv, n, fr, oracle.lut(0), 8)[0])
    for world in (1, 2, 4):
        keys = []
        for r in range(world):
            lo, hi = gt.shard_bounds(total, r, world)
            with gt.Allocator(hi - lo, first_node=lo) as a:
                a.load_nodes(lv[lo:hi], n[lo:hi], fr[lo:hi])
                a.score_all(fetch=False)
                keys.append(a.best_node(ks))
        assert (gt.merge_place(keys) == want).all()


def test_schedule_config2_matches_oracle_replay(gt, oracle):
    """BASELINE config 2 at full size: 1 000 DGX-1 nodes, 10 000 pods with k uniform on 1..8, placed one at a
    time; every placement (node and GPU mask) equals the oracle's sequential loop."""
    lv, n, fr, _ = gt.synth_cluster(2, 0x5EED0002, 0, 1000)
    k, arg = gt.synth_events(2, 0x5EED0012, 10000)
    wp, wn, wm, wfree = oracle.replay(lv, n, fr, oracle.lut(0), k, arg)
    with gt.Allocator(1000) as a:
        a.load_nodes(lv, n, fr)
        placed, on, om = a.schedule(k, arg)
        assert placed == wp and (on == wn).all() and (om == wm).all()
        # the device tables follow the commit loop: a fresh scoring pass sees the final free masks
        got = a.score_all()
        want, _ = oracle.exhaustive_cluster(lv, n, wfree, oracle.lut(0), 8)
        assert (got == want).all()


def test_schedule_config5_replay_with_departures(gt, oracle):
    """Config 5 trace (arrivals and departures) on a mixed cluster, 200 000 events."""
    lv, n, fr, _ = gt.synth_cluster(4, 0x5EED0004, 0, 2000)
    k, arg = gt.synth_events(5, 0x5EED0005, 200000)
    wp, wn, wm, wfree = oracle.replay(lv, n, fr, oracle.lut(0), k, arg)
    with gt.Allocator(2000) as a:
        a.load_nodes(lv, n, fr)
        placed, on, om = a.schedule(k, arg)
        assert placed == wp and (on == wn).all() and (om == wm).all()
        assert placed > 50000


def test_schedule_per_event_gpu_path(gt, oracle):
    """Without the memo table (GT_FLAG_NO_MEMO) every event re-scores the touched node on the GPU."""
    lv, n, fr, _ = gt.synth_cluster(4, 77, 0, 300)
    k, arg = gt.synth_events(5, 123, 1500)
    wp, wn, wm, _ = oracle.replay(lv, n, fr, oracle.lut(0), k, arg)
    with gt.Allocator(300, flags=gt.FLAG_NO_MEMO) as a:
        a.load_nodes(lv, n, fr)
        placed, on, om = a.schedule(k, arg)
        assert placed == wp and (on == wn).all() and (om == wm).all()


def test_extender_sort_bind_allocate_flow(gt, oracle):
    """The reference's flow (imgs/gpu_topology_on_k8s.png steps 3-7) through the mirror: sort -> host choice ->
    bind -> Allocate, against the oracle's sequential loop, on real topologies parsed from text fixtures."""
    import os
    gold = os.path.join(os.path.dirname(__file__), "golden")
    tiles = [gt.parse_topo_matrix(open(os.path.join(gold, f + ".topo.txt")).read())
             for f in ("f1_ring", "dgx1", "f2_pcie", "f3_tree4")]
    N = 40
    lv = np.stack([tiles[i % 4][0] for i in range(N)])
    n = np.array([tiles[i % 4][1] for i in range(N)], np.uint8)
    free = np.array([(1 << int(x)) - 1 for x in n], np.uint16)
    names = ["node-%02d" % i for i in range(N)]
    lut = oracle.lut(0)
    cmax = oracle.cmax(lut)
    rng = np.random.RandomState(9)
    with gt.Allocator(N) as a:
        a.load_nodes(lv, n, free)
        ext = gt.Extender(a, names)
        for pod in range(120):
            m = int(rng.randint(1, 9))
            prio = ext.sort(m)
            want_best, _ = oracle.exhaustive_cluster(lv, n, free, lut, 2)
            for i, hp in enumerate(prio):
                b = int(want_best[i, m - 1])
                want_score = 0 if b == ol.INFEASIBLE else oracle.score_int(b >> 16, m * (m - 1) // 2, cmax)
                assert hp == {"Host": names[i], "Score": want_score}
            key = int(oracle.cluster_argmin(want_best)[m - 1])
            if key == gt.NO_PLACE:
                res, ann = ext.bind("p%d" % pod, "default", "uid-%d" % pod, names[0], m, 1000 + pod)
                if bin(int(free[0])).count("1") < m:
                    assert res["Error"]
                    continue
            cost, node, mask = gt.place_fields(key) if key != gt.NO_PLACE else (0, 0, 0)
            if key == gt.NO_PLACE:
                continue
            res, ann = ext.bind("p%d" % pod, "default", "uid-%d" % pod, names[node], m, 1000 + pod)
            assert res == {"Error": ""}
            assert ann == {"ALIYUN_COM_GPU_ASSIGNED": "false", "ALIYUN_COM_GPU_ASSUME_TIME": str(1000 + pod),
                           "ALIYUN_COM_GPU_GROUP": gt.group_encode(mask)}
            free[node] &= ~mask & 0xFFFF
            reply, upd = gt.DevicePlugin.allocate(ann, 2000 + pod)
            assert reply["Envs"]["NVIDIA_VISIBLE_DEVICES"] == gt.group_encode(mask)
            assert upd["ALIYUN_COM_GPU_ASSIGNED"] == "true"
            if pod % 5 == 4:      # a pod finishes
                uid = sorted(ext.pods)[0]
                nd, mk = ext.pods[uid]
                ext.pod_deleted(uid)
                free[nd] |= mk
        assert ext.bind("x", "default", "uid-x", "no-such-node", 1, 0)[0]["Error"]


def test_gaia_singular_policy_on_device(gt, oracle):
    """GT_K1_GAIA_SINGULAR: request size 1 follows Gaia Alg. 3 (KA3, KA4a, KA5 and a sweep over every free mask of
    the tree topologies); every other request size is unchanged."""
    for cls, n_expected in (("F2_PCIE", 8), ("F3_TREE4", 4), ("DGX1", 8)):
        lv, n, ac, sib = oracle.class_topology(ol.CLS[cls])
        N = 1 << n
        levels = np.tile(lv, (N, 1))
        free = np.arange(N, dtype=np.uint16)
        with gt.Allocator(N, k1_policy=gt.K1_GAIA_SINGULAR) as a:
            a.load_nodes(levels, np.full(N, n, np.uint8), free, access_cost=np.tile(ac, (N, 1)),
                         sibling_mask=np.tile(sib, (N, 1)))
            got = a.score_all()
            want, _ = oracle.exhaustive_cluster(levels, np.full(N, n, np.uint8), free, oracle.lut(0), 4)
            for f in range(1, N):
                want[f, 0] = 1 << oracle.gaia_singular(n, f, ac, sib)
            assert (got == want).all()
            place = a.best_node([1, 2])
            assert gt.place_fields(place[0]) == (0, 1, int(want[1, 0]) & 0xFFFF)
            assert int(place[1]) == int(oracle.cluster_argmin(want)[1])
            # dirty path keeps the policy
            a.commit([N - 1], [1])
            a.score_dirty()
            got2 = a.score_all()
            want[N - 1] = want[N - 2]
            assert (got2 == want).all()
    # the documented answers
    lv, n, ac, sib = oracle.class_topology(ol.CLS["F2_PCIE"])
    with gt.Allocator(1, k1_policy=gt.K1_GAIA_SINGULAR) as a:
        a.load_nodes(lv[None, :], [8], [0xFF & ~((1 << 4) | (1 << 6))], access_cost=ac[None, :], sibling_mask=sib[None, :])
        a.score_all()
        assert a.allocate(0, 1)[0] == 1 << 5            # KA3: GPU5
    lv, n, ac, sib = oracle.class_topology(ol.CLS["F3_TREE4"])
    with gt.Allocator(2, k1_policy=gt.K1_GAIA_SINGULAR) as a:
        a.load_nodes(np.tile(lv, (2, 1)), [4, 4], [0xF, 0xF & ~4], access_cost=np.tile(ac, (2, 1)),
                     sibling_mask=np.tile(sib, (2, 1)))
        a.score_all()
        assert a.allocate(0, 1)[0] in (1 << 2, 1 << 3)  # KA4a
        assert a.allocate(1, 1)[0] == 1 << 3            # KA5
        assert a.allocate(1, 2)[0] == 0b0011            # KA6
        # the commit loop keeps the policy too: a k = 1 pod on node 1 (cousin of GPU3 is used) takes GPU3 first
        placed, on, om = a.schedule(np.array([1], np.uint8), np.array([0], np.uint32))
        assert placed == 1 and int(on[0]) == 0 and int(om[0]) in (4, 8)


def test_farthest_policy_on_device(gt, oracle):
    """GT_K1_FARTHEST: request size 1 follows the prose of design.md:135-147 (the unused GPU with the largest distance
    sum to all GPUs of the node); every other request size is unchanged; dirty path, placement key and commit loop
    keep the policy."""
    lut = oracle.lut(0)
    for cls in ("F2_PCIE", "F3_TREE4", "DGX1", "F1_RING"):
        lv, n, _, _ = oracle.class_topology(ol.CLS[cls])
        N = 1 << n
        levels = np.tile(lv, (N, 1))
        free = np.arange(N, dtype=np.uint16)
        with gt.Allocator(N, k1_policy=gt.K1_FARTHEST) as a:
            a.load_nodes(levels, np.full(N, n, np.uint8), free)
            got = a.score_all()
            want, _ = oracle.exhaustive_cluster(levels, np.full(N, n, np.uint8), free, lut, 4)
            for f in range(1, N):
                want[f, 0] = 1 << oracle.k1_farthest(lv, n, f, lut)
            assert (got == want).all()
            place = a.best_node([1, 2])
            assert gt.place_fields(place[0]) == (0, 1, int(want[1, 0]) & 0xFFFF)
            assert int(place[1]) == int(oracle.cluster_argmin(want)[1])
            a.commit([N - 1], [1])
            a.score_dirty()
            want[N - 1] = want[N - 2]
            assert (a.score_all() == want).all()
    # F2 with everything free: GPU6 / GPU7 hang directly under the root, GPU6 is the farthest
    lv, n, _, _ = oracle.class_topology(ol.CLS["F2_PCIE"])
    with gt.Allocator(2, k1_policy=gt.K1_FARTHEST) as a:
        a.load_nodes(np.tile(lv, (2, 1)), [8, 8], [0xFF, 0xFF])
        a.score_all()
        assert a.allocate(0, 1)[0] == 1 << 6
        placed, on, om = a.schedule(np.array([1, 1], np.uint8), np.array([0, 0], np.uint32))
        assert placed == 2 and list(on) == [0, 0] and [int(m) for m in om] == [1 << 6, 1 << 7]
    # random matrices, cost tables and GPU counts
    rng = np.random.RandomState(77)
    N = 700
    levels = rng.randint(1, 16, (N, 256)).astype(np.uint8)
    ngs = rng.randint(1, 17, N).astype(np.uint8)
    free = rng.randint(0, 1 << 16, N).astype(np.uint16)
    rl = rng.randint(0, 256, 16).astype(np.uint8)
    with gt.Allocator(N, k1_policy=gt.K1_FARTHEST) as a:
        a.set_cost_lut(rl)
        a.load_nodes(levels, ngs, free)
        got = a.score_all()
        want, _ = oracle.exhaustive_cluster(levels, ngs, free, rl, 4)
        for i in range(N):
            g = oracle.k1_farthest(levels[i], int(ngs[i]), int(free[i]) & ((1 << int(ngs[i])) - 1), rl)
            if g >= 0:
                want[i, 0] = 1 << g
        assert (got == want).all()


def test_split_units_and_high_positions(gt, oracle):
    """Shapes the work decomposition treats specially: many nodes with >= 14 free GPUs (a node spans several
    warp units whose minima meet through atomics), 16-GPU nodes with <= 8 free GPUs at high indices, the maximum
    cost table (255 everywhere: cost sums up to 120*255), and awkward node counts."""
    rng = np.random.RandomState(21)
    N = 3000
    levels = rng.randint(1, 16, (N, 256)).astype(np.uint8)
    n_gpus = np.full(N, 16, np.uint8)
    free = np.zeros(N, np.uint16)
    for i in range(N):
        if i % 3 == 0:       # 14..16 free
            m = 0xFFFF
            for b in rng.choice(16, rng.randint(0, 3), replace=False):
                m &= ~(1 << b)
            free[i] = m
        elif i % 3 == 1:     # <= 8 free, all in the upper half
            free[i] = int(rng.randint(0, 256)) << 8
        else:
            free[i] = rng.randint(0, 1 << 16)
    check_cluster(gt, oracle, levels, n_gpus, free, np.full(16, 255, np.uint8))
    for n_nodes in (1, 31, 33, 127, 129, 257):
        lv, n, fr, _ = gt.synth_cluster(4, 1000 + n_nodes, 0, n_nodes)
        check_cluster(gt, oracle, lv, n, fr)


def test_partial_loads_and_capacity(gt, oracle):
    """Loading node ranges one after the other into a larger context equals loading them at once."""
    lv, n, fr, _ = gt.synth_cluster(4, 31337, 0, 5000)
    want, _ = oracle.exhaustive_cluster(lv, n, fr, oracle.lut(0), 8)
    with gt.Allocator(8000) as a:
        a.load_nodes(lv[:1234], n[:1234], fr[:1234], first=0)
        a.load_nodes(lv[1234:], n[1234:], fr[1234:], first=1234)
        assert a.num_nodes == 5000
        assert (a.score_all() == want).all()
        # reload a middle range with other free masks
        fr2 = fr.copy()
        fr2[2000:2100] = 0
        a.load_nodes(lv[2000:2100], n[2000:2100], fr2[2000:2100], first=2000)
        want2, _ = oracle.exhaustive_cluster(lv, n, fr2, oracle.lut(0), 8)
        assert (a.score_all() == want2).all()
        # whole-shard mask update
        a.update_free(fr)
        assert (a.score_all() == want).all()
        # scattered mask update through the dirty path
        ids = np.array([5, 4999, 2500], np.uint32)
        a.update_free(np.zeros(3, np.uint16), node_ids=ids)
        a.score_dirty()
        fr3 = fr.copy()
        fr3[ids] = 0
        want3, _ = oracle.exhaustive_cluster(lv, n, fr3, oracle.lut(0), 8)
        assert (a.score_all() == want3).all()
        assert (a.best_node(np.arange(1, 17, dtype=np.uint8)) == oracle.cluster_argmin(want3)).all()


def test_gpu_matches_committed_golden_sweep(gt, oracle):
    """The CUDA path against the committed vectors of tests/golden/sweep_small.npz (no oracle call involved)."""
    import os
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "sweep_small.npz"))
    for name in ("DGX1", "F1_RING", "F2_PCIE", "F3_TREE4"):
        lv, n, _, _ = oracle.class_topology(ol.CLS[name])
        N = 1 << n
        with gt.Allocator(N) as a:
            a.load_nodes(np.tile(lv, (N, 1)), np.full(N, n, np.uint8), np.arange(N, dtype=np.uint16))
            for preset in (0, 1):
                a.set_cost_lut(gt.lut_preset(preset))
                assert (a.score_all() == gold["%s_lut%d" % (name, preset)]).all()
    lv, n, fr, _ = gt.synth_cluster(2, 0x5EED0002, 0, 1000)
    assert (fr == gold["config2_free"]).all()
    with gt.Allocator(1000) as a:
        a.load_nodes(lv, n, fr)
        a.score_all(fetch=False)
        assert (a.best_node(np.arange(1, 17, dtype=np.uint8)) == gold["config2_place"]).all()


@pytest.mark.parametrize("config,N", [(4, 700), (4, 9000), (3, 5000), (33, 300)])
def test_sort_scores_left_by_the_pass(gt, oracle, config, N):
    """`sort` after a whole-shard mask update runs a full pass that leaves one score byte per node as it goes (nodes
    whose rows several work units build are scored afterwards from the table); same bytes as the formula on the oracle's
    table, through the copying call and through the zero-copy surface (pinned mirror in, mapped scores out)."""
    lut = oracle.lut(0)
    cmax = oracle.cmax(lut)
    lv, n, fr, _ = oracle.synth_cluster(config, 0x50A7 + N, 0, N)
    rng = np.random.RandomState(N)
    with gt.Allocator(N) as a:
        a.load_nodes(lv, n, fr)
        for it, m in enumerate((1, 2, 5, 8, 11, 16, 3)):
            cur = (fr & rng.randint(0, 1 << 16, N).astype(np.uint16)).astype(np.uint16) if it % 2 else fr
            want_best, _ = oracle.exhaustive_cluster(lv, n, cur, lut, 4)
            ln = m * (m - 1) // 2
            want = np.array([0 if b == ol.INFEASIBLE else oracle.score_int(int(b) >> 16, ln, cmax) for b in want_best[:, m - 1]],
                            np.uint8)
            place = np.zeros(16, np.uint64)
            if it % 3 == 2:
                scores = np.zeros(N, np.uint8)
                a.update_free(cur)
                a.ext_sort_into(m, scores, place)
            else:
                a.host_free()[:N] = cur
                a.host_free_changed()
                a.ext_sort_inplace(m, place)
                scores = a.host_scores().copy()
            assert (scores == want).all(), (config, N, m)
            assert (place == oracle.cluster_argmin(want_best)).all()
            assert (a.score_all() == want_best).all()
            # nothing changed: the next sort for another size comes from the table, not from a pass
            m2 = 1 + (m + 3) % 16
            a.ext_sort_inplace(m2)
            ln2 = m2 * (m2 - 1) // 2
            want2 = [0 if b == ol.INFEASIBLE else oracle.score_int(int(b) >> 16, ln2, cmax) for b in want_best[:, m2 - 1]]
            assert (a.host_scores() == np.array(want2, np.uint8)).all()


def test_pipeline_form_of_k1_matches_the_oracle():
    """k_score_pipe (GT_PIPE=1, an opt-in experiment: DESIGN.md section 3) must stay bit-exact while it is kept."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "import _pkg, oracle_lib\n"
        "gt = _pkg.load_package(); o = oracle_lib.load()\n"
        "for cfg, n in ((4, 30000), (3, 6000), (33, 5000)):\n"
        "    lv, ng, fr, _ = o.synth_cluster(cfg, 77 + n, 0, n)\n"
        "    with gt.Allocator(n) as a:\n"
        "        a.load_nodes(lv, ng, fr)\n"
        "        got = a.score_all(); keys = a.best_node(np.arange(1, 17, dtype=np.uint8))\n"
        "    want, _ = o.exhaustive_cluster(lv, ng, fr, o.lut(0), 8)\n"
        "    assert (got == want).all() and (keys == o.cluster_argmin(want)).all(), cfg\n"
        "print('pipe ok')\n"
    ) % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, GT_PIPE="1")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "pipe ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


def test_scoring_loop_in_one_call_and_local_keys(gt, oracle):
    """gt_score_all_async_many over several contexts = the same passes one by one; gt_best_node_local (K2 over the
    table) agrees with the keys K1 folded; the in-place K1 timing runs and leaves the contexts usable."""
    import torch
    stream = torch.cuda.Stream()
    ctxs, wants = [], []
    for r in range(3):
        lv, n, fr, _ = oracle.synth_cluster(4, 0xA11 + r, 0, 7000)
        a = gt.Allocator(7000, stream=stream.cuda_stream)
        a.load_nodes(lv, n, fr)
        ctxs.append(a)
        wants.append(oracle.exhaustive_cluster(lv, n, fr, oracle.lut(0), 4)[0])
    gt.score_all_async_many(ctxs, 7)
    ks = np.arange(1, 17, dtype=np.uint8)
    for a, want in zip(ctxs, wants):
        keys = a.best_node(ks)
        assert (keys == oracle.cluster_argmin(want)).all()
        assert (a.best_node_local() == keys).all()
        assert (a.score_all() == want).all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    e1.record(stream)
    gt.score_all_async_many_timed(ctxs, 6, e0.cuda_event, e1.cuda_event)     # the library records the caller's events
    torch.cuda.synchronize()
    assert 0.005 < e0.elapsed_time(e1) < 50
    rest, k1, whole = gt.time_pass_rotated(ctxs, 12)
    assert 0 < k1 < whole < 1000 and rest > 0
    for a, want in zip(ctxs, wants):
        assert (a.score_all() == want).all()
        a.close()


def test_masks_change_between_passes_that_are_not_waited_for(gt, oracle):
    """K0 of a pass runs ahead of the previous pass's K1 (other set of bin lists, other control block).  Every kind of
    mask update in between -- in place in the pinned mirror, uploaded, single nodes through commit -- must leave exactly
    the table of the LAST masks, and passes of several contexts interleaved on one stream must not disturb each other."""
    import torch
    N = 30000
    ks = np.arange(1, 17, dtype=np.uint8)
    lv, n, fr, _ = oracle.synth_cluster(4, 0xBEE5, 0, N)
    rng = np.random.default_rng(77)
    full = ((1 << n.astype(np.uint32)) - 1).astype(np.uint16)
    draws = [(rng.integers(0, 1 << 16, N).astype(np.uint16) & full) for _ in range(7)]
    stream = torch.cuda.Stream()
    with gt.Allocator(N, stream=stream.cuda_stream) as a, gt.Allocator(N, stream=stream.cuda_stream) as b:
        a.load_nodes(lv, n, fr)
        b.load_nodes(lv, n, draws[6])
        a.score_all_async()
        for i in range(6):                        # no synchronisation between these passes
            if i % 3 == 0:
                a.host_free()[:N] = draws[i]      # read in place by the next K0
                a.host_free_changed()
            elif i % 3 == 1:
                a.update_free(draws[i])           # whole-shard update through the caller's buffer
            else:
                a.host_free()[:N] = draws[i]
                a.host_free_changed()
                a.score_all_async()
                a.score_all_async()               # the same masks again: K0 fully ahead of the K1 in front
            b.score_all_async()                   # another cluster's pass in between, same stream
            a.score_all_async()
        want = oracle.exhaustive_cluster(lv, n, draws[5], oracle.lut(0), 8)[0]
        got = a.score_all()
        bad = np.nonzero((got != want).any(axis=1))[0]
        assert len(bad) == 0, "first mismatch node %d" % bad[0]
        assert (a.best_node(ks) == oracle.cluster_argmin(want)).all()
        want_b = oracle.exhaustive_cluster(lv, n, draws[6], oracle.lut(0), 8)[0]
        assert (b.best_node(ks) == oracle.cluster_argmin(want_b)).all()
        assert (b.score_all() == want_b).all()
        # single-node changes, a dirty pass, then full passes again
        free_now = draws[5].copy()
        nodes = np.nonzero(free_now)[0][:50].astype(np.uint32)
        masks = np.array([int(free_now[i]) & -int(free_now[i]) for i in nodes], np.uint16)     # lowest free GPU of each
        a.commit(nodes, masks)
        free_now[nodes] &= ~masks
        a.score_dirty()
        a.score_all_async()
        a.score_all_async()
        want2 = oracle.exhaustive_cluster(lv, n, free_now, oracle.lut(0), 8)[0]
        assert (a.score_all() == want2).all()
        assert (a.best_node(ks) == oracle.cluster_argmin(want2)).all()
