/*
 * gt_oracle.c -- CPU oracle (plain C).  TEST INFRASTRUCTURE ONLY, see gt_oracle.h.
 * PARITY UNPINNED vs. a reference implementation (the reference ships none); pinned by KA1-KA6.
 */
#include "gt_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ link table, score */

/* design.md:33-44 lists the link kinds; design.md:196-203 gives the "mark" column. */
void gto_lut_preset(int preset, uint8_t lut[16])
{
    int i;
    for (i = 0; i < 16; i++) lut[i] = (uint8_t)i; /* default: cost == level code */
    if (preset == 1) {
        memset(lut, 0, 16);
        lut[GTO_LV_SYS] = 1; lut[GTO_LV_NODE] = 2; lut[GTO_LV_PHB] = 3;
        lut[GTO_LV_PXB] = 4; lut[GTO_LV_PIX] = 5; lut[GTO_LV_PSB] = 6;
        lut[GTO_LV_NV1] = 7; lut[GTO_LV_NV2] = 8; lut[GTO_LV_NV3] = 9; lut[GTO_LV_NV4] = 10;
        lut[GTO_LV_NVS] = 11;
    }
}

uint32_t gto_lut_cmax(const uint8_t lut[16])
{
    uint32_t m = 0; int i;
    /* the marks of the table of design.md:196-203 (+ NVSwitch): codes 1..SYS; user-defined codes 12..15 stay out */
    for (i = 1; i <= GTO_LV_SYS; i++) if (lut[i] > m) m = lut[i];
    return m;
}

/* design.md:208 with the precedence the worked example at design.md:213-217 fixes:
 * 10 * [1 - sum / (cmax * len)] */
int gto_score_milli(uint32_t sum, uint32_t len, uint32_t cmax)
{
    uint64_t den = (uint64_t)cmax * len;
    if (den == 0) return 10000;            /* design.md:19: one GPU ignores topology */
    if (sum >= den) return 0;
    return (int)((10000ull * (den - sum)) / den);
}

int gto_score_int(uint32_t sum, uint32_t len, uint32_t cmax)
{
    uint64_t den = (uint64_t)cmax * len;
    if (den == 0) return 10;
    if (sum >= den) return 0;
    return (int)((10ull * (den - sum)) / den);
}

/* ------------------------------------------------------------------ exhaustive (the contract) */

uint32_t gto_subset_cost(const uint8_t levels[256], const uint8_t lut[16], uint16_t mask)
{
    /* ids = the chosen device list; sum over its C(k,2) pairs (design.md:208 sum(topology), len(topology) terms) */
    int ids[GTO_MAX_GPUS], k = 0, a, b; uint32_t sum = 0, m = mask;
    while (m) { ids[k++] = __builtin_ctz(m); m &= m - 1; }
    for (a = 0; a < k; a++)
        for (b = a + 1; b < k; b++)
            sum += lut[levels[ids[a] * 16 + ids[b]] & 15];   /* getTopology(g1,g2), design.md:169 */
    return sum;
}

uint64_t gto_exhaustive_node(const uint8_t levels[256], int n_gpus, uint16_t free_mask,
                             const uint8_t lut[16], uint32_t out_best[16])
{
    uint32_t f, s; uint64_t scored = 0; int k;
    for (k = 0; k < 16; k++) out_best[k] = GTO_INFEASIBLE;
    if (n_gpus > GTO_MAX_GPUS) n_gpus = GTO_MAX_GPUS;
    if (n_gpus <= 0) return 0;
    f = free_mask & ((1u << n_gpus) - 1u);   /* "if gpu not used", design.md:140,164-168 */
    if (!f) return 0;
    /* every non-empty sub-mask of f in increasing numeric order: a strict '<' then keeps the
     * numerically lowest mask among equal costs */
    s = 0;
    for (;;) {
        uint32_t cost, key;
        s = ((s | ~f) + 1u) & f;
        if (!s) break;
        k = __builtin_popcount(s);
        cost = gto_subset_cost(levels, lut, (uint16_t)s);
        key = (cost << 16) | s;
        if (key < out_best[k - 1]) out_best[k - 1] = key;   /* "< min", design.md:169 */
        scored++;
    }
    return scored;
}

typedef struct {
    const uint8_t *levels; const uint8_t *n_gpus; const uint16_t *free_mask; const uint8_t *lut;
    uint32_t *out; uint32_t lo, hi; uint64_t scored;
} gto_job;

static void *gto_worker(void *p)
{
    gto_job *j = (gto_job *)p; uint32_t i;
    for (i = j->lo; i < j->hi; i++)
        j->scored += gto_exhaustive_node(j->levels + (size_t)i * 256, j->n_gpus[i], j->free_mask[i],
                                         j->lut, j->out + (size_t)i * 16);
    return 0;
}

uint64_t gto_exhaustive_cluster(const uint8_t *levels, const uint8_t *n_gpus, const uint16_t *free_mask,
                                uint32_t n_nodes, const uint8_t lut[16], uint32_t *out_best, int threads)
{
    gto_job *jobs; pthread_t *tid; int t; uint64_t total = 0;
    if (threads < 1) threads = 1;
    if ((uint32_t)threads > n_nodes) threads = n_nodes ? (int)n_nodes : 1;
    jobs = (gto_job *)calloc((size_t)threads, sizeof *jobs);
    tid = (pthread_t *)calloc((size_t)threads, sizeof *tid);
    for (t = 0; t < threads; t++) {
        jobs[t].levels = levels; jobs[t].n_gpus = n_gpus; jobs[t].free_mask = free_mask; jobs[t].lut = lut;
        jobs[t].out = out_best;
        jobs[t].lo = (uint32_t)(((uint64_t)n_nodes * t) / threads);
        jobs[t].hi = (uint32_t)(((uint64_t)n_nodes * (t + 1)) / threads);
        if (t) pthread_create(&tid[t], 0, gto_worker, &jobs[t]);
    }
    gto_worker(&jobs[0]);
    for (t = 1; t < threads; t++) pthread_join(tid[t], 0);
    for (t = 0; t < threads; t++) total += jobs[t].scored;
    free(jobs); free(tid);
    return total;
}

void gto_cluster_argmin(const uint32_t *best, uint32_t n_nodes, uint32_t first_node, uint64_t out_place[16])
{
    uint32_t i; int k;
    for (k = 0; k < 16; k++) out_place[k] = ~0ull;
    for (i = 0; i < n_nodes; i++)
        for (k = 0; k < 16; k++) {
            uint32_t b = best[(size_t)i * 16 + k]; uint64_t key;
            if (b == GTO_INFEASIBLE) continue;
            key = ((uint64_t)(b >> 16) << 48) | ((uint64_t)(first_node + i) << 16) | (b & 0xFFFFu);
            if (key < out_place[k]) out_place[k] = key;
        }
}

/* ------------------------------------------------------------------ design.md greedy, literal */

uint16_t gto_greedy_node(const uint8_t levels[256], int n_gpus, uint16_t free_mask,
                         const uint8_t lut[16], int m, uint32_t *out_cost)
{
    uint32_t f = free_mask & ((1u << n_gpus) - 1u), ids = 0; int g1, g2, cnt;
    if (out_cost) *out_cost = 0;
    if (m < 1 || __builtin_popcount(f) < m) return 0;
    if (m == 1) {                                   /* design.md:154-160: first unused device */
        for (g1 = 0; g1 < n_gpus; g1++) if ((f >> g1) & 1) return (uint16_t)(1u << g1);
        return 0;
    }
    {                                               /* design.md:162-174: min pair, strict '<' */
        uint32_t min = 0xFFFFFFFFu; int b1 = -1, b2 = -1;
        for (g1 = 0; g1 < n_gpus; g1++) {
            if (!((f >> g1) & 1)) continue;         /* "if gpu1 is used continue" */
            for (g2 = 0; g2 < n_gpus; g2++) {
                uint32_t d;
                if (g2 == g1) continue;             /* no self entry, design.md:19 */
                if (!((f >> g2) & 1)) continue;     /* "if gpu 2 is used continue" */
                d = lut[levels[(g1 < g2 ? g1 : g2) * 16 + (g1 < g2 ? g2 : g1)] & 15];
                if (d < min) { min = d; b1 = g1; b2 = g2; }
            }
        }
        ids = (1u << b1) | (1u << b2);
    }
    cnt = 2;
    while (cnt < m) {                               /* design.md:180-184 */
        uint32_t bests = 0xFFFFFFFFu; int bg = -1;
        for (g1 = 0; g1 < n_gpus; g1++) {
            uint32_t s = 0;
            if (!((f >> g1) & 1) || ((ids >> g1) & 1)) continue;
            for (g2 = 0; g2 < n_gpus; g2++)
                if ((ids >> g2) & 1)
                    s += lut[levels[(g1 < g2 ? g1 : g2) * 16 + (g1 < g2 ? g2 : g1)] & 15];
            if (s < bests) { bests = s; bg = g1; }
        }
        ids |= 1u << bg; cnt++;
    }
    if (out_cost) *out_cost = gto_subset_cost(levels, lut, (uint16_t)ids);
    return (uint16_t)ids;
}

uint64_t gto_greedy_cluster(const uint8_t *levels, const uint8_t *n_gpus, const uint16_t *free_mask,
                            uint32_t n_nodes, const uint8_t lut[16], uint32_t *out_best)
{
    uint32_t i; uint64_t decisions = 0;
    for (i = 0; i < n_nodes; i++) {
        int n = n_gpus[i] > GTO_MAX_GPUS ? GTO_MAX_GPUS : n_gpus[i], m, p;
        uint32_t f = n ? (free_mask[i] & ((1u << n) - 1u)) : 0;
        for (m = 0; m < 16; m++) out_best[(size_t)i * 16 + m] = GTO_INFEASIBLE;
        p = __builtin_popcount(f);
        for (m = 1; m <= p; m++) {
            uint32_t cost; uint16_t mask = gto_greedy_node(levels + (size_t)i * 256, n, (uint16_t)f, lut, m, &cost);
            out_best[(size_t)i * 16 + m - 1] = (cost << 16) | mask;
            decisions++;
        }
    }
    return decisions;
}

/* ------------------------------------------------------------------ Gaia Alg. 3 / 4 / 1 */

/* design.md:135-147 */
int gto_k1_farthest(const uint8_t levels[256], int n_gpus, uint16_t free_mask, const uint8_t lut[16])
{
    int i, j, best = -1; uint32_t best_sum = 0;
    if (n_gpus > GTO_MAX_GPUS) n_gpus = GTO_MAX_GPUS;
    for (i = 0; i < n_gpus; i++) {
        uint32_t sum = 0;
        if (!((free_mask >> i) & 1)) continue;                    /* "if gpu not used" */
        for (j = 0; j < n_gpus; j++)                              /* distance to all GPUs, upper-triangle entry */
            if (j != i) sum += lut[levels[(i < j ? i : j) * 16 + (i < j ? j : i)] & 15];
        if (best < 0 || sum > best_sum) { best = i; best_sum = sum; }   /* the longest path sum; first on ties */
    }
    return best;
}

/* gpu_scheduler_algorithm_1.png lines 1-10 */
int gto_gaia_singular(int n_gpus, uint16_t free_mask, const uint8_t access_cost[16],
                      const uint16_t sibling_mask[16])
{
    uint32_t f = free_mask & ((1u << n_gpus) - 1u); int i, best = -1;
    for (i = 0; i < n_gpus; i++) {          /* line 2: Resources is 1 and Cousin.Resources is 0 */
        if (!((f >> i) & 1)) continue;
        if (sibling_mask[i] == 0 || (sibling_mask[i] & f) != 0) continue;
        if (best < 0 || access_cost[i] < access_cost[best]) best = i;   /* line 6 */
    }
    if (best >= 0) return best;
    for (i = 0; i < n_gpus; i++) {          /* line 9: lowest AccessCost with Resources 1 */
        if (!((f >> i) & 1)) continue;
        if (best < 0 || access_cost[i] < access_cost[best]) best = i;
    }
    return best;
}

static int tree_free_leaves(const gto_tree *t, int node, uint32_t f, int *leaves, int cap)
{   /* free leaves under `node`, sorted by (access cost, gpu id) */
    int i, cnt = 0;
    for (i = 0; i < t->n_nodes; i++) {
        int a = i, j;
        if (t->gpu_id[i] < 0 || !((f >> t->gpu_id[i]) & 1)) continue;
        while (a != node && a != 0) a = t->parent[a];
        if (a != node) continue;
        if (cnt < cap) leaves[cnt++] = i;
        for (j = cnt - 1; j > 0; j--) {
            int x = leaves[j], y = leaves[j - 1];
            if (t->access_cost[x] < t->access_cost[y] ||
                (t->access_cost[x] == t->access_cost[y] && t->gpu_id[x] < t->gpu_id[y])) {
                leaves[j] = y; leaves[j - 1] = x;
            } else break;
        }
    }
    return cnt;
}

static int tree_subtree_size(const gto_tree *t, int node)
{
    int i, cnt = 0;
    for (i = 0; i < t->n_nodes; i++) {
        int a = i;
        if (t->gpu_id[i] < 0) continue;
        while (a != node && a != 0) a = t->parent[a];
        if (a == node) cnt++;
    }
    return cnt;
}

/* gpu_scheduler_algorithm_2.jpg.  Lines 1-8: from every free leaf climb while Resources < m; the
 * node where the climb stops (possibly the root) joins Sc.  Line 10 "m x Children.AccessCost" is
 * undefined for a candidate whose children are topology nodes (they carry 0); the reading used
 * here: the access cost of the most expensive of the m cheapest free leaves beneath it -- equal to
 * the child cost when the children are leaves (the only case the paper's examples exercise).
 * Equal TotalCost -> the candidate spanning fewer GPUs, then the lower first gpu id.
 * Line 13: S = the m cheapest free leaves of BestNode. */
uint16_t gto_gaia_link(const gto_tree *t, uint16_t free_mask, int m)
{
    uint32_t f = free_mask; int i, best = -1, best_cost = 0, best_size = 0, leaves[GTO_MAX_GPUS];
    uint16_t out = 0;
    if (m < 2) return 0;
    for (i = 0; i < t->n_nodes; i++) {
        int a, cnt, cost, size;
        if (t->gpu_id[i] < 0 || !((f >> t->gpu_id[i]) & 1)) continue;
        a = i;
        for (;;) {
            cnt = tree_free_leaves(t, a, f, leaves, GTO_MAX_GPUS);
            if (cnt >= m || a == 0) break;
            a = t->parent[a];
        }
        if (cnt < m) continue;
        cost = m * t->access_cost[leaves[m - 1]];
        size = tree_subtree_size(t, a);
        if (best < 0 || cost < best_cost || (cost == best_cost && size < best_size)) {
            best = a; best_cost = cost; best_size = size;
        }
    }
    if (best < 0) return 0;
    tree_free_leaves(t, best, f, leaves, GTO_MAX_GPUS);
    for (i = 0; i < m; i++) out |= (uint16_t)(1u << t->gpu_id[leaves[i]]);
    return out;
}

/* gpu_scheduler_algorithm.jpg (Alg. 1): m == 1 -> Singular, m > 1 -> Link; Fragment skipped (gaia.md:36) */
uint16_t gto_gaia_dispatch(const gto_tree *t, uint16_t free_mask, int m)
{
    if (m == 1) {
        uint8_t lv[256], ac[16]; uint16_t sib[16]; int n = 0, i, g;
        for (i = 0; i < t->n_nodes; i++) if (t->gpu_id[i] >= n) n = t->gpu_id[i] + 1;
        gto_tree_to_tables(t, n, lv, ac, sib);
        g = gto_gaia_singular(n, free_mask, ac, sib);
        return g < 0 ? 0 : (uint16_t)(1u << g);
    }
    return gto_gaia_link(t, free_mask, m);
}

void gto_tree_to_tables(const gto_tree *t, int n_gpus, uint8_t levels[256], uint8_t access_cost[16],
                        uint16_t sibling_mask[16])
{
    int leaf_of[GTO_MAX_GPUS], i, j;
    memset(levels, 0, 256); memset(access_cost, 0, 16); memset(sibling_mask, 0, 32);
    for (i = 0; i < GTO_MAX_GPUS; i++) leaf_of[i] = -1;
    for (i = 0; i < t->n_nodes; i++) if (t->gpu_id[i] >= 0) leaf_of[t->gpu_id[i]] = i;
    for (i = 0; i < n_gpus; i++) {
        if (leaf_of[i] < 0) continue;
        access_cost[i] = (uint8_t)t->access_cost[leaf_of[i]];
        for (j = 0; j < n_gpus; j++) {
            int a, b, da = 0, db = 0, x;
            if (j == i || leaf_of[j] < 0) continue;
            if (t->parent[leaf_of[i]] == t->parent[leaf_of[j]]) sibling_mask[i] |= (uint16_t)(1u << j);
            /* lowest common ancestor: its label is the pair's link level */
            a = leaf_of[i]; b = leaf_of[j];
            for (x = a; x != 0; x = t->parent[x]) da++;
            for (x = b; x != 0; x = t->parent[x]) db++;
            while (da > db) { a = t->parent[a]; da--; }
            while (db > da) { b = t->parent[b]; db--; }
            while (a != b) { a = t->parent[a]; b = t->parent[b]; }
            levels[i * 16 + j] = (uint8_t)t->level[a];
        }
    }
}

/* ------------------------------------------------------------------ synthetic clusters */

uint64_t gto_splitmix64_at(uint64_t seed, uint64_t index)
{
    uint64_t z = seed + (index + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

static void set_pair(uint8_t lv[256], int i, int j, int code)
{
    lv[i * 16 + j] = (uint8_t)code; lv[j * 16 + i] = (uint8_t)code;
}

static void fill_all(uint8_t lv[256], int n, int code)
{
    int i, j;
    for (i = 0; i < n; i++) for (j = 0; j < n; j++) if (i != j) lv[i * 16 + j] = (uint8_t)code;
}

void gto_class_topology(int cls, uint8_t levels[256], uint8_t *n_gpus, uint8_t access_cost[16],
                        uint16_t sibling_mask[16])
{
    int i, j;
    memset(levels, 0, 256); memset(access_cost, 0, 16); memset(sibling_mask, 0, 32);
    switch (cls) {
    case GTO_CLASS_DGX1: {   /* DGX-1V hybrid cube-mesh (synthetic fixture) */
        static const int nv1[][2] = {{0,1},{0,2},{1,3},{2,6},{3,7},{4,5},{4,6},{5,7}};
        static const int nv2[][2] = {{0,3},{0,4},{1,2},{1,5},{2,3},{4,7},{5,6},{6,7}};
        *n_gpus = 8; fill_all(levels, 8, GTO_LV_SYS);
        for (i = 0; i < 8; i++) set_pair(levels, nv1[i][0], nv1[i][1], GTO_LV_NV1);
        for (i = 0; i < 8; i++) set_pair(levels, nv2[i][0], nv2[i][1], GTO_LV_NV2);
        for (i = 0; i < 8; i++) access_cost[i] = 1;
        break; }
    case GTO_CLASS_F1_RING: { /* imgs/gpu_topology_on_machine.png: NV3 ring 0-1-3-2-7-6-4-5-0, rest PHB */
        static const int ring[8] = {0, 1, 3, 2, 7, 6, 4, 5};
        *n_gpus = 8; fill_all(levels, 8, GTO_LV_PHB);
        for (i = 0; i < 8; i++) set_pair(levels, ring[i], ring[(i + 1) % 8], GTO_LV_NV3);
        for (i = 0; i < 8; i++) access_cost[i] = 1;
        break; }
    case GTO_CLASS_F2_PCIE: { /* gpu_topology_tree.png (Gaia Fig. 4) */
        *n_gpus = 8; fill_all(levels, 8, GTO_LV_SYS);
        for (i = 0; i < 6; i++) for (j = 0; j < 6; j++) if (i != j) levels[i * 16 + j] = GTO_LV_PHB;
        for (i = 0; i < 4; i++) for (j = 0; j < 4; j++) if (i != j) levels[i * 16 + j] = GTO_LV_PXB;
        set_pair(levels, 0, 1, GTO_LV_PIX); set_pair(levels, 4, 5, GTO_LV_PXB); set_pair(levels, 6, 7, GTO_LV_PXB);
        for (i = 0; i < 8; i++) { access_cost[i] = (uint8_t)(i / 2 + 1); sibling_mask[i] = (uint16_t)(1u << (i ^ 1)); }
        break; }
    case GTO_CLASS_F3_TREE4: { /* Gaia Fig. 7: SOC{ PXB{0,1} cost 2, PIX{2,3} cost 1 } */
        *n_gpus = 4; fill_all(levels, 4, GTO_LV_SYS);
        set_pair(levels, 0, 1, GTO_LV_PXB); set_pair(levels, 2, 3, GTO_LV_PIX);
        access_cost[0] = access_cost[1] = 2; access_cost[2] = access_cost[3] = 1;
        for (i = 0; i < 4; i++) sibling_mask[i] = (uint16_t)(1u << (i ^ 1));
        break; }
    case GTO_CLASS_NVS16:     /* config 3: every pair at the NVSwitch level */
        *n_gpus = 16; fill_all(levels, 16, GTO_LV_NVS);
        for (i = 0; i < 16; i++) access_cost[i] = 1;
        break;
    case GTO_CLASS_DUAL8:     /* config 3 non-degenerate variant: two 8-GPU boards, SYS across */
    default:
        *n_gpus = 16; fill_all(levels, 16, GTO_LV_SYS);
        for (i = 0; i < 16; i++) for (j = 0; j < 16; j++)
            if (i != j && (i / 8) == (j / 8)) levels[i * 16 + j] = GTO_LV_NVS;
        for (i = 0; i < 16; i++) access_cost[i] = (uint8_t)(1 + i / 8);
        break;
    }
}

#define GTO_DRAWS_PER_NODE 32u
#define GTO_P_USED_Q32 1503238553u /* floor(0.35 * 2^32) */

void gto_synth_cluster(int config, uint64_t seed, uint32_t first, uint32_t count, uint8_t *levels,
                       uint8_t *n_gpus, uint16_t *free_mask, uint8_t *class_out)
{
    uint32_t i;
    for (i = 0; i < count; i++) {
        uint64_t base = (uint64_t)(first + i) * GTO_DRAWS_PER_NODE;
        int cls, g; uint8_t n, ac[16]; uint16_t sib[16], f = 0;
        switch (config) {
        case 1: case 2: cls = GTO_CLASS_DGX1; break;
        case 3: cls = GTO_CLASS_NVS16; break;
        case 33: cls = GTO_CLASS_DUAL8; break;
        default: {  /* config 4 mix: 40% DGX-1, 20% F1, 20% F2, 10% F3, 10% NVS16 */
            uint32_t u = (uint32_t)(gto_splitmix64_at(seed, base) % 100u);
            cls = u < 40 ? GTO_CLASS_DGX1 : u < 60 ? GTO_CLASS_F1_RING : u < 80 ? GTO_CLASS_F2_PCIE
                : u < 90 ? GTO_CLASS_F3_TREE4 : GTO_CLASS_NVS16;
            break; }
        }
        gto_class_topology(cls, levels + (size_t)i * 256, &n, ac, sib);
        if (config == 1) {
            f = 0xFF;
        } else if (config == 3 || config == 33) {   /* popcount uniform on 4..14, partial Fisher-Yates */
            int perm[16], c, t;
            for (t = 0; t < 16; t++) perm[t] = t;
            c = 4 + (int)(gto_splitmix64_at(seed, base + 1) % 11u);
            for (t = 0; t < c; t++) {
                int j = t + (int)(gto_splitmix64_at(seed, base + 2 + (uint64_t)t) % (uint64_t)(16 - t));
                int tmp = perm[t]; perm[t] = perm[j]; perm[j] = tmp;
                f |= (uint16_t)(1u << perm[t]);
            }
        } else {
            for (g = 0; g < n; g++)  /* each GPU used independently with p = 0.35 */
                if ((uint32_t)(gto_splitmix64_at(seed, base + 1 + (uint64_t)g) >> 32) >= GTO_P_USED_Q32)
                    f |= (uint16_t)(1u << g);
        }
        n_gpus[i] = n; free_mask[i] = f;
        if (class_out) class_out[i] = (uint8_t)cls;
    }
}

/* ------------------------------------------------------------------ sequential commit loop */

void gto_synth_events(int config, uint64_t seed, uint32_t n_events, uint8_t *ev_k, uint32_t *ev_arg)
{
    uint32_t i;
    for (i = 0; i < n_events; i++) {
        uint64_t a = gto_splitmix64_at(seed, 2ull * i), b = gto_splitmix64_at(seed, 2ull * i + 1);
        if (config == 5 && (uint32_t)(a >> 32) >= 2362232012u /* floor(0.55*2^32) */) {
            ev_k[i] = 0; ev_arg[i] = (uint32_t)(b >> 32);      /* departure of live pod (arg % live) */
        } else {
            ev_k[i] = (uint8_t)(1 + b % 8u); ev_arg[i] = 0;    /* arrival, k uniform 1..8 */
        }
    }
}

uint64_t gto_replay(const uint8_t *levels, const uint8_t *n_gpus, uint16_t *free_mask, uint32_t n_nodes,
                    const uint8_t lut[16], const uint8_t *ev_k, const uint32_t *ev_arg, uint32_t n_events,
                    uint32_t *out_node, uint16_t *out_mask)
{
    /* design.md:118 Prioritize = best combination per node; kube-scheduler takes the best node
     * (lowest cost, then lowest index: SURVEY.md A10); design.md:119 Bind commits it.  best[] is
     * recomputed for the one node a commit or release touched -- same result as re-scoring all. */
    uint32_t *best = (uint32_t *)malloc((size_t)n_nodes * 16 * sizeof *best);
    uint32_t *live_node = (uint32_t *)malloc((size_t)(n_events + 1) * sizeof *live_node);
    uint16_t *live_mask = (uint16_t *)malloc((size_t)(n_events + 1) * sizeof *live_mask);
    uint32_t live = 0, e, i; uint64_t placed = 0;
    for (i = 0; i < n_nodes; i++)
        gto_exhaustive_node(levels + (size_t)i * 256, n_gpus[i], free_mask[i], lut, best + (size_t)i * 16);
    for (e = 0; e < n_events; e++) {
        out_node[e] = 0xFFFFFFFFu; out_mask[e] = 0;
        if (ev_k[e] == 0) {
            uint32_t idx, nd;
            if (!live) continue;
            idx = ev_arg[e] % live; nd = live_node[idx];
            free_mask[nd] |= live_mask[idx];
            out_node[e] = nd; out_mask[e] = live_mask[idx];
            live--; live_node[idx] = live_node[live]; live_mask[idx] = live_mask[live];
            gto_exhaustive_node(levels + (size_t)nd * 256, n_gpus[nd], free_mask[nd], lut, best + (size_t)nd * 16);
        } else {
            int k = ev_k[e]; uint64_t bk = ~0ull; uint32_t nd; uint16_t m;
            for (i = 0; i < n_nodes; i++) {
                uint32_t b = best[(size_t)i * 16 + (k - 1)]; uint64_t key;
                if (b == GTO_INFEASIBLE) continue;
                key = ((uint64_t)(b >> 16) << 48) | ((uint64_t)i << 16) | (b & 0xFFFFu);
                if (key < bk) bk = key;
            }
            if (bk == ~0ull) continue;             /* unplaceable: counted and dropped */
            nd = (uint32_t)((bk >> 16) & 0xFFFFFFFFu); m = (uint16_t)(bk & 0xFFFFu);
            free_mask[nd] &= (uint16_t)~m;
            out_node[e] = nd; out_mask[e] = m;
            live_node[live] = nd; live_mask[live] = m; live++; placed++;
            gto_exhaustive_node(levels + (size_t)nd * 256, n_gpus[nd], free_mask[nd], lut, best + (size_t)nd * 16);
        }
    }
    free(best); free(live_node); free(live_mask);
    return placed;
}
