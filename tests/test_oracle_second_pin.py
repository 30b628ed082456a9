"""The C oracle against a second restatement that shares no code with it (tests/py_bruteforce.py, itertools over the
device groups, written from design.md:205-217 and the frozen tie-break).  The reference ships no allocator, so this
does not pin parity against a reference implementation ("parity unpinned" stands); it guards the oracle against its
own bugs beyond the six known answers."""
import numpy as np

import oracle_lib as ol
import py_bruteforce as bf


def _compare(oracle, lv, n, fr, lut):
    got, _ = oracle.exhaustive_cluster(lv, n, fr, lut, 4)
    rows = []
    for i in range(len(n)):
        want = bf.best_rows(lv[i], n[i], fr[i], lut)
        assert [int(x) for x in got[i]] == want, (i, int(n[i]), hex(int(fr[i])))
        rows.append(want)
    return got, rows


def test_config4_nodes_match_the_python_restatement(oracle):
    """1 500 nodes of the benchmark's heterogeneous cluster (16-GPU nodes with more than 11 free GPUs are capped
    to keep the pure-Python side in seconds; three of them run uncapped)."""
    lv, n, fr, _ = oracle.synth_cluster(4, 0x5EED0004, 0, 1500)
    fr = fr.copy()
    big = [i for i in range(len(n)) if bin(int(fr[i]) & ((1 << int(n[i])) - 1)).count("1") > 11]
    for i in big[3:]:
        fr[i] &= 0x5B7D          # 11 bits
    got, rows = _compare(oracle, lv, n, fr, oracle.lut(0))
    # the cluster choice too (lowest cost, then lowest node index), with a shard offset
    want = bf.best_node_keys(rows, first_node=70000)
    assert [int(x) for x in oracle.cluster_argmin(got, first=70000)] == want


def test_random_asymmetric_matrices_and_cost_tables_match(oracle):
    """600 nodes with random level matrices (not symmetric: only the upper triangle may count), random GPU counts,
    random free masks incl. bits above the GPU count, and random cost tables incl. 0 and 255."""
    rng = np.random.RandomState(20261107)
    N = 600
    lv = rng.randint(0, 16, (N, 256)).astype(np.uint8)
    n = rng.randint(1, 17, N).astype(np.uint8)
    fr = rng.randint(0, 1 << 16, N).astype(np.uint16)
    for i in range(N):                       # keep the Python side fast: at most 10 unused GPUs
        while bin(int(fr[i]) & ((1 << int(n[i])) - 1)).count("1") > 10:
            fr[i] &= fr[i] - 1
    for lut in (oracle.lut(1), rng.randint(0, 256, 16).astype(np.uint8),
                np.array([0, 255, 0, 255, 7, 7, 7, 1, 2, 3, 250, 251, 0, 0, 9, 9], np.uint8)):
        _compare(oracle, lv, n, fr, lut)


def test_empty_and_full_nodes(oracle):
    lv, nn, _, _ = oracle.class_topology(ol.CLS["DUAL8"])
    lv = np.tile(lv, (3, 1))
    n = np.array([nn, nn, 0], np.uint8)
    fr = np.array([0, 0xFFFF, 0xFFFF], np.uint16)
    _compare(oracle, lv, n, fr, oracle.lut(0))
