/*
 * gt_oracle.h -- CPU oracle for the topology-aware GPU allocation path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build, load or
 * call it, and there only as the checker or as the timed CPU baseline.
 *
 * PARITY UNPINNED with respect to a reference implementation: hellolijj/gpu-topology-on-k8s ships
 * a design document and no allocator source, so there is nothing to compile
 * under oracle/_ref.  This file is a plain-C restatement of the algorithms the reference documents
 * describe, pinned by the six known answers those documents contain (SURVEY.md section 4, KA1-KA6;
 * tests/test_oracle_known_answers.py) and by the greedy-vs-exhaustive divergence counts an independent script
 * produced for that survey (18 / 86 costlier, 28 / 30 equal-cost-different choices of 769 on F1 / DGX-1).
 *
 * Citations: design.md:N = /root/reference/design.md line N; "Alg. 1/3/4" = the Gaia paper
 * algorithms reproduced in /root/reference/reference/gaia_gpu_topology/gpu_scheduler_algorithm*.{jpg,png}.
 */
#ifndef GT_ORACLE_H
#define GT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GTO_MAX_GPUS 16
#define GTO_INFEASIBLE 0xFFFF0000u /* mask 0, cost 0xFFFF: fewer than k free GPUs */

/* Link-level codes (design.md:33-44 table; NVSwitch added below NV4).  Lower code = closer. */
enum {
    GTO_LV_X = 0, GTO_LV_NVS = 1, GTO_LV_NV4 = 2, GTO_LV_NV3 = 3, GTO_LV_NV2 = 4, GTO_LV_NV1 = 5,
    GTO_LV_PSB = 6, GTO_LV_PIX = 7, GTO_LV_PXB = 8, GTO_LV_PHB = 9, GTO_LV_NODE = 10, GTO_LV_SYS = 11
};

/* cost LUT presets: 0 = default (cost == level code, lower is closer),
 * 1 = the literal mark table of design.md:196-203 (SYS=1 .. PSB=6, NV1..NV4 = 7..10 as in the NVML enum) */
void gto_lut_preset(int preset, uint8_t lut[16]);

/* design.md:205-217: 10 - 10*sum/(cmax*len).  Exact rational carried as integers.
 * gto_score_int  = floor(10*(cmax*len-sum)/(cmax*len)); len==0 (single GPU, design.md:19) -> 10.
 * gto_score_milli = floor(10000*(cmax*len-sum)/(cmax*len)); len==0 -> 10000. */
int gto_score_int(uint32_t sum, uint32_t len, uint32_t cmax);
int gto_score_milli(uint32_t sum, uint32_t len, uint32_t cmax);
uint32_t gto_lut_cmax(const uint8_t lut[16]);

/* Pair-cost sum of one subset: sum over i<j in mask of lut[levels[i*16+j]]  (design.md:208 "sum(topology)") */
uint32_t gto_subset_cost(const uint8_t levels[256], const uint8_t lut[16], uint16_t mask);

/* THE CONTRACT (north_star): every k-subset of the free set, cost = pair-cost sum, lower wins
 * (design.md:169), ties -> numerically lowest mask.  out_best[k-1] = cost<<16 | mask for k = 1..16,
 * GTO_INFEASIBLE when fewer than k GPUs are free. Returns the number of subsets scored. */
uint64_t gto_exhaustive_node(const uint8_t levels[256], int n_gpus, uint16_t free_mask,
                             const uint8_t lut[16], uint32_t out_best[16]);

/* Same over a cluster, nodes split over `threads` host threads (pthreads). Returns subsets scored. */
uint64_t gto_exhaustive_cluster(const uint8_t *levels, const uint8_t *n_gpus, const uint16_t *free_mask,
                                uint32_t n_nodes, const uint8_t lut[16], uint32_t *out_best, int threads);

/* Cluster argmin per k: key = cost<<48 | node<<16 | mask, lowest wins;
 * 0xFFFF...  when no node is feasible.  node index = first_node + local index. */
void gto_cluster_argmin(const uint32_t *best, uint32_t n_nodes, uint32_t first_node, uint64_t out_place[16]);

/* design.md:153-186, literal: m==1 first unused device; m>=2 min-pair seed with strict '<' in scan
 * order, then repeatedly add the unused GPU with minimal distance sum to the chosen ids.
 * Returns the chosen mask (0 when fewer than m are free); *out_cost = pair-cost sum. */
uint16_t gto_greedy_node(const uint8_t levels[256], int n_gpus, uint16_t free_mask,
                         const uint8_t lut[16], int m, uint32_t *out_cost);

/* The reference design's own heuristic over a cluster: for every node and every m = 1..popcount(free) the greedy
 * choice (timed as the "what the reference would have run" CPU baseline).  out_best as gto_exhaustive_cluster.
 * Returns the number of (node, m) decisions made. */
uint64_t gto_greedy_cluster(const uint8_t *levels, const uint8_t *n_gpus, const uint16_t *free_mask,
                            uint32_t n_nodes, const uint8_t lut[16], uint32_t *out_best);

/* design.md:135-147, the prose variant for request size 1: the unused GPU "far away from the GPU group" -- the one
 * whose distance sum to ALL GPUs of the node is largest; the CPU-affinity tie-break of design.md:144-145 needs an input
 * this path does not have, ties go to the lowest id.  Returns gpu id or -1. */
int gto_k1_farthest(const uint8_t levels[256], int n_gpus, uint16_t free_mask, const uint8_t lut[16]);

/* Gaia Alg. 3 Singular (m == 1): candidates = free GPU whose siblings ("cousin") are all used
 * (and that has at least one sibling); lowest access cost, then lowest id; if none, the free GPU
 * with lowest access cost.  Returns gpu id or -1. */
int gto_gaia_singular(int n_gpus, uint16_t free_mask, const uint8_t access_cost[16],
                      const uint16_t sibling_mask[16]);

/* Gaia resource-access cost tree (paper p.255, gpu_topology_tree.png) */
typedef struct {
    int n_nodes;            /* tree nodes, index 0 = root */
    int parent[2 * GTO_MAX_GPUS];
    int gpu_id[2 * GTO_MAX_GPUS];     /* >= 0 for leaves, -1 for topology nodes */
    int access_cost[2 * GTO_MAX_GPUS];/* leaves only; topology nodes carry 0 */
    int level[2 * GTO_MAX_GPUS];      /* GTO_LV_* label of a topology node (mode of its children) */
} gto_tree;

/* Gaia Alg. 4 Link (m > 1) on a tree; returns chosen mask or 0.  See gt_oracle.c for the reading
 * of "Children.AccessCost" used (the paper leaves it open). */
uint16_t gto_gaia_link(const gto_tree *t, uint16_t free_mask, int m);
/* Gaia Alg. 1 dispatch: m==1 -> Singular, m>1 -> Link (Fragment, 0<m<1, is skipped: gaia.md:36) */
uint16_t gto_gaia_dispatch(const gto_tree *t, uint16_t free_mask, int m);
/* derive level matrix / access_cost / sibling_mask from a tree */
void gto_tree_to_tables(const gto_tree *t, int n_gpus, uint8_t levels[256], uint8_t access_cost[16],
                        uint16_t sibling_mask[16]);

/* ---- synthetic clusters, splitmix64, random access per node ---- */
uint64_t gto_splitmix64_at(uint64_t seed, uint64_t index); /* index-th output (0-based) of the stream */

enum { GTO_CLASS_DGX1 = 0, GTO_CLASS_F1_RING = 1, GTO_CLASS_F2_PCIE = 2, GTO_CLASS_F3_TREE4 = 3,
       GTO_CLASS_NVS16 = 4, GTO_CLASS_DUAL8 = 5, GTO_NUM_CLASSES = 6 };
/* fixture topology of a class: level matrix, gpu count, Gaia access cost + sibling masks */
void gto_class_topology(int cls, uint8_t levels[256], uint8_t *n_gpus, uint8_t access_cost[16],
                        uint16_t sibling_mask[16]);

/* config: 1 single DGX-1 all free; 2 DGX-1 cluster p_used=0.35; 3 NVS16 popcount 4..14;
 * 33 = config 3 non-degenerate variant (DUAL8 boards); 4 heterogeneous mix p_used=0.35.
 * Fills nodes [first, first+count). class_out may be NULL. */
void gto_synth_cluster(int config, uint64_t seed, uint32_t first, uint32_t count, uint8_t *levels,
                       uint8_t *n_gpus, uint16_t *free_mask, uint8_t *class_out);

/* ---- sequential commit loop (design.md:117-119 Prioritize then Bind, one pod at a time) ---- */
/* events: k > 0 arrival requesting k GPUs; k == 0 departure of live pod number `arg` (index into the
 * live list, swap-remove order).  out_node/out_mask per event (0xFFFFFFFF / 0 when unplaceable or departure).
 * free_mask is updated in place.  Returns number of placements. */
uint64_t gto_replay(const uint8_t *levels, const uint8_t *n_gpus, uint16_t *free_mask, uint32_t n_nodes,
                    const uint8_t lut[16], const uint8_t *ev_k, const uint32_t *ev_arg, uint32_t n_events,
                    uint32_t *out_node, uint16_t *out_mask);
/* event generators: config 2 (10k arrivals, k uniform 1..8) and config 5 (arrival p=0.55 else departure) */
void gto_synth_events(int config, uint64_t seed, uint32_t n_events, uint8_t *ev_k, uint32_t *ev_arg);

#ifdef __cplusplus
}
#endif
#endif
