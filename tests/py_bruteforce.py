"""A second, independent restatement of the exhaustive contract -- TEST INFRASTRUCTURE ONLY.

Written from the reference text alone and sharing no code with oracle/gt_oracle.c:
  * a device group's communication cost is the sum of the link marks of its pairs -- `sum(topology)` over
    `len(topology)` = C(k,2) pairs, design.md:205-217 (worked example design.md:213-217);
  * the pair (g1, g2) with g1 < g2 is looked up as getTopology(g1, g2), design.md:169 -- the upper-triangle entry;
  * only unused devices take part, design.md:140, 164-168;
  * the lowest cost wins (`< min`, design.md:169); among equal costs the numerically lowest device mask
    (frozen in DESIGN.md section 1; the reference leaves ties open, design.md:188-190);
  * a single device costs nothing, design.md:19.
Plain itertools, no numpy in the inner loop, no bit tricks shared with the oracle's sub-mask walk."""
from itertools import combinations

INFEASIBLE = 0xFFFF0000


def best_rows(levels, n_gpus, free_mask, lut):
    """levels: 256 level codes (16x16 row-major); returns 16 ints, entry k-1 = cost << 16 | mask or INFEASIBLE."""
    lv = [int(x) for x in levels]
    weight = [int(x) for x in lut]
    n = min(int(n_gpus), 16)
    unused = [g for g in range(n) if (int(free_mask) >> g) & 1]
    out = [INFEASIBLE] * 16
    for k in range(1, len(unused) + 1):
        winner = None
        for group in combinations(unused, k):
            cost = 0
            for g1, g2 in combinations(group, 2):          # g1 < g2: combinations keeps the order of `unused`
                cost += weight[lv[g1 * 16 + g2] & 15]
            mask = 0
            for g in group:
                mask |= 1 << g
            cand = (cost, mask)
            if winner is None or cand < winner:
                winner = cand
        out[k - 1] = (winner[0] << 16) | winner[1]
    return out


def best_node_keys(rows, first_node=0):
    """Cluster choice per request size: lowest cost, then lowest node index;
    key = cost << 48 | node << 16 | mask, or 2^64-1 when no node can serve the size."""
    keys = [0xFFFFFFFFFFFFFFFF] * 16
    for i, row in enumerate(rows):
        for k, b in enumerate(row):
            if b == INFEASIBLE:
                continue
            key = ((b >> 16) << 48) | ((first_node + i) << 16) | (b & 0xFFFF)
            if key < keys[k]:
                keys[k] = key
    return keys
