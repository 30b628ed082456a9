/*
 * gputopo.h -- C-ABI of libgputopo.so: the topology-aware GPU allocation hot path of
 * hellolijj/gpu-topology-on-k8s (design.md section 2.2 "scheduler extender scoring") on B200.
 *
 * The reference is a design document with Go snippets and no allocator source; "replaces" below
 * cites the place in the reference documents that specifies the behaviour each entry point
 * implements (design.md:N = /root/reference/design.md line N; Alg. n = Gaia paper algorithms in
 * /root/reference/reference/gaia_gpu_topology/).  A Go extender binds these through cgo
 * (INTEGRATION.md).
 *
 * Conventions: every buffer is caller-owned; inputs are copied before the call returns and no
 * caller pointer is retained (cgo pointer rules).  All functions return 0 on success or a negative
 * gt_status; nothing throws or aborts across this boundary.  "Fewer than k free GPUs" is not an
 * error: the entry is GT_INFEASIBLE (mask 0, cost 0xFFFF).  A context is used by one caller at a
 * time (not thread-safe), matching kube-scheduler's one-pod-at-a-time loop.  There is no CPU
 * fallback: without a CUDA device gt_create fails with GT_ERR_NO_DEVICE.
 */
#ifndef GPUTOPO_H
#define GPUTOPO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GT_MAX_GPUS 16
#define GT_INFEASIBLE 0xFFFF0000u

/* best[node][k-1] entry: cost<<16 | gpu mask */
#define GT_BEST_COST(b) ((uint32_t)(b) >> 16)
#define GT_BEST_MASK(b) ((uint32_t)(b) & 0xFFFFu)
/* placement key: cost<<48 | global node index<<16 | gpu mask; GT_NO_PLACE when no node can host the request */
#define GT_NO_PLACE 0xFFFFFFFFFFFFFFFFull
#define GT_PLACE_COST(p) ((uint32_t)((uint64_t)(p) >> 48))
#define GT_PLACE_NODE(p) ((uint32_t)(((uint64_t)(p) >> 16) & 0xFFFFFFFFu))
#define GT_PLACE_MASK(p) ((uint32_t)((uint64_t)(p) & 0xFFFFu))

typedef enum gt_status {
    GT_OK = 0,
    GT_ERR_INVALID = -1,     /* bad argument */
    GT_ERR_NO_DEVICE = -2,   /* no usable CUDA device / driver */
    GT_ERR_CUDA = -3,        /* a CUDA call failed; see gt_last_error */
    GT_ERR_NOMEM = -4,
    GT_ERR_RANGE = -5,       /* node index / count beyond the context's capacity */
    GT_ERR_PARSE = -6,       /* malformed topology text / annotation */
    GT_ERR_NCCL = -7,
    GT_ERR_STATE = -8        /* call order (e.g. allocate before score) */
} gt_status;

/* Link-level codes.  replaces: the P2PLinkType table design.md:33-44 (+ NVSwitch, absent there). Lower = closer. */
enum {
    GT_LV_X = 0, GT_LV_NVS = 1, GT_LV_NV4 = 2, GT_LV_NV3 = 3, GT_LV_NV2 = 4, GT_LV_NV1 = 5,
    GT_LV_PSB = 6, GT_LV_PIX = 7, GT_LV_PXB = 8, GT_LV_PHB = 9, GT_LV_NODE = 10, GT_LV_SYS = 11
};
enum { GT_LUT_DEFAULT = 0, GT_LUT_DESIGN_MARKS = 1 };
/* request-size-1 policy (the reference gives three): the pseudo-code of design.md:154-160 (first unused
 * device; what the exhaustive path yields by itself), Gaia Alg. 3 Singular, and the prose of design.md:135-147 (the unused
 * GPU farthest from the GPU group: largest distance sum to all GPUs of the node; lowest id on ties -- the CPU-affinity
 * tie-break of design.md:144-145 needs an input this path does not have) */
enum { GT_K1_FIRST_UNUSED = 0, GT_K1_GAIA_SINGULAR = 1, GT_K1_FARTHEST = 2 };
enum { GT_FLAG_GRAPH = 2 /* replay K0 -> K1 as a captured CUDA graph instead of two plain launches (measured slower) */,
       GT_FLAG_NO_MEMO = 4 /* gt_schedule re-scores on the GPU per event instead of using the memo table */,
       GT_FLAG_MEMO_PASS = 8 /* full passes copy rows out of the GPU-built (topology class, free mask) table instead of
                                enumerating again; falls back to enumeration when there are too many classes */ };

typedef struct gt_ctx gt_ctx;

typedef struct gt_config {
    uint32_t struct_size;  /* sizeof(gt_config) */
    int32_t device;        /* CUDA device ordinal */
    uint32_t max_nodes;    /* capacity of this context (its shard of the cluster) */
    uint32_t first_node;   /* global index of local node 0: placement keys carry first_node + local */
    uint32_t k1_policy;    /* GT_K1_* */
    uint32_t flags;        /* GT_FLAG_* */
    void *stream;          /* optional cudaStream_t to launch on; NULL = the library creates one */
} gt_config;

/* ---- life cycle ---- */
int gt_create(const gt_config *cfg, gt_ctx **out);
void gt_destroy(gt_ctx *ctx);
const char *gt_strerror(int status);
const char *gt_last_error(const gt_ctx *ctx);

/* ---- link table / score (host only) ---- */
/* replaces: bandwidth-weight column design.md:33-47 (left "TODO") and mark table design.md:196-203 */
int gt_lut_preset(int preset, uint8_t out_lut[16]);
int gt_set_cost_lut(gt_ctx *ctx, const uint8_t lut[16]);
/* replaces: score formula design.md:205-217.  floor(10*(cmax*len-sum)/(cmax*len)), len = C(k,2) */
int gt_score_int(uint32_t sum, uint32_t len, uint32_t cmax);
/* replaces: the "6" of design.md:208 -- the largest mark of the table: max of the cost table over the link kinds
 * the table and the parsers know (codes 1..GT_LV_SYS).  User-defined codes 12..15 do not stretch the scale; a
 * combination whose cost exceeds cmax*len scores 0.  This is what gt_ext_sort normalises by. */
uint32_t gt_lut_cmax(const uint8_t lut[16]);
int gt_score_milli(uint32_t sum, uint32_t len, uint32_t cmax);

/* ---- topology tables ---- */
/* replaces: type gpuTopology map[uint]map[uint]gpuTopologyType design.md:64-70 filled from the node
 * annotations design.md:78-82; free = complement of the GPUs in ALIYUN_COM_GPU_GROUP of the pods on
 * the node (design.md:231).  levels: count x 256 bytes (16x16 row-major GT_LV_* codes, entry [i][j]
 * with i<j is the one read); n_gpus: count bytes; access_cost (count x 16) and sibling_mask
 * (count x 16 u16) are optional (NULL) and only used by GT_K1_GAIA_SINGULAR; free_mask: count u16. */
int gt_load_nodes(gt_ctx *ctx, uint32_t first, uint32_t count, const uint8_t *levels, const uint8_t *n_gpus,
                  const uint8_t *access_cost, const uint16_t *sibling_mask, const uint16_t *free_mask);
/* node_ids == NULL: masks are for nodes [0, count) */
int gt_update_free(gt_ctx *ctx, const uint32_t *node_ids, const uint16_t *masks, uint32_t count);

/* ---- the hot path ---- */
/* replaces: PrioritizeVerb "sort" design.md:99,118 -- for every node the best GPU combination for
 * every request size k = 1..16 (every k-subset of the free set, pair-cost sum design.md:208, lowest
 * wins design.md:169, ties -> numerically lowest mask).  out_best: n_nodes x 16 u32 or NULL to leave
 * the table on the device. */
int gt_score_all(gt_ctx *ctx, uint32_t *out_best);
/* same, for the nodes touched by gt_update_free / gt_commit / gt_release since the last scoring */
int gt_score_dirty(gt_ctx *ctx);
/* replaces: kube-scheduler's host choice after "sort" (design.md:115-119): for each request size
 * k[i] the placement key of the best node (lowest cost, then lowest node index); with a
 * communicator attached the minimum over all ranks. */
int gt_best_node(gt_ctx *ctx, const uint8_t *k, uint32_t nreq, uint64_t *out_place);
/* replaces: bindVerb "bind" GPU choice design.md:119 and what the device plugin's Allocate hands
 * out design.md:239 -- the best combination of k GPUs on one node (local index). */
int gt_allocate(gt_ctx *ctx, uint32_t node, uint32_t k, uint16_t *out_mask, uint16_t *out_cost);
/* replaces: the state update of bind design.md:223-234 (pod recorded on the device) / pod deletion.
 * Host commit loop: clears (commit) or sets (release) the bits and marks the node dirty. */
int gt_commit(gt_ctx *ctx, const uint32_t *nodes, const uint16_t *masks, uint32_t count);
int gt_release(gt_ctx *ctx, const uint32_t *nodes, const uint16_t *masks, uint32_t count);

/* replaces: the scheduling cycle itself -- kube-scheduler handing pods one at a time to "sort" then
 * "bind" (design.md:115-119), each pod seeing the previous pod's effect.  Events: ev_k[i] > 0 = a pod
 * asking for ev_k[i] GPUs (placed on the best node, lowest cost then lowest index, and committed);
 * ev_k[i] == 0 = the live pod number (ev_arg[i] % live pods) leaves and its GPUs are released.
 * out_node / out_mask per event (0xFFFFFFFF / 0 when unplaceable or nothing to release).
 * The GPU scores every (topology class, free mask) once (3^n subsets per class) into a memo table and
 * the host commit loop looks rows up; with too many distinct topologies for that table (> 1 GiB), or
 * GT_FLAG_NO_MEMO, the touched node is re-scored on the GPU after every event.  No CPU scoring.
 * On a sharded cluster (gt_comm_init) the call is collective: every rank passes the same events and gets the same
 * outputs (global node indices).  The table is built in `world` slices, one per GPU, and gathered; the host loop --
 * sequential by definition -- then runs replicated on every rank over the gathered node states (8 bytes per node),
 * each rank applying the placements that land on its own shard to its tables.  No exchange per event. */
int gt_schedule(gt_ctx *ctx, const uint8_t *ev_k, const uint32_t *ev_arg, uint32_t n_events, uint32_t *out_node,
                uint16_t *out_mask, uint64_t *out_placed);

/* ---- Gaia resource-access cost tree as an input format ---- */
/* replaces: the tree T of the Gaia paper p. 255 (reference/gaia_gpu_topology/gpu_topology_tree.png, gaia.md:16-22):
 * topology nodes labelled with a link level, GPU leaves with an id and an access cost.  Index 0 is the root;
 * parent[0] is ignored.  gpu_id >= 0 marks a leaf; level[] is read for topology nodes (GT_LV_* code of the link
 * its children share), access_cost[] for leaves (0..255). */
#define GT_MAX_TREE_NODES 32
typedef struct gt_tree {
    int32_t n_nodes; /* 1 .. 31 (at most 2*16-1) */
    int32_t parent[GT_MAX_TREE_NODES];
    int32_t gpu_id[GT_MAX_TREE_NODES];
    int32_t access_cost[GT_MAX_TREE_NODES];
    int32_t level[GT_MAX_TREE_NODES];
} gt_tree;
/* host only: the tables a tree stands for -- pair level = label of the lowest common ancestor, sibling set =
 * leaves under the same parent, access cost per GPU.  GT_ERR_INVALID for a malformed tree (cycle, duplicate or
 * out-of-range GPU id, leaf used as a parent). */
int gt_tree_tables(const gt_tree *tree, uint8_t out_levels[256], uint8_t *out_n_gpus, uint8_t out_access_cost[16],
                   uint16_t out_sibling_mask[16]);
/* gt_load_nodes for nodes described by trees: derives the tables above, loads them, and keeps the tree shape on
 * the device for gt_gaia_select. */
int gt_load_trees(gt_ctx *ctx, uint32_t first, uint32_t count, const gt_tree *trees, const uint16_t *free_mask);
/* replaces: Gaia Alg. 1 dispatch (gpu_scheduler_algorithm.jpg; Fragment, 0 < m < 1, is skipped: gaia.md:36) on the
 * device, one answer per node for request size m: m == 1 -> Alg. 3 Singular (gpu_scheduler_algorithm_1.png),
 * m >= 2 -> Alg. 4 Link (gpu_scheduler_algorithm_2.jpg: from every free leaf climb until m leaves are free beneath,
 * lowest m x access cost wins, then the smaller subtree; hand out its m cheapest free leaves).
 * out_best[node] = pair-cost sum of the chosen set << 16 | mask (same word as gt_score_all's table), 0xFFFF0000
 * when the node cannot serve m (for m >= 2 also a node that was loaded without a tree).  GT_ERR_STATE for m >= 2
 * when no tree was ever loaded, GT_ERR_INVALID for m outside 1..16. */
int gt_gaia_select(gt_ctx *ctx, uint32_t m, uint32_t *out_best);

/* ---- the extender / device-plugin surface around the path (text formats are host only) ---- */
/* replaces: topology discovery output design.md:27-29 -- the `nvidia-smi topo -m` table
 * (imgs/gpu_topology_on_machine.png) -> 16x16 level codes.  NV# with # > 4 maps to GT_LV_NVS. */
int gt_parse_topo_matrix(const char *text, size_t len, uint8_t out_levels[256], uint32_t *out_n);
/* replaces: filling `type gpuTopology map[uint]map[uint]gpuTopologyType` (gpuTopologyType = nvml.P2PLinkType,
 * design.md:57-70) -- the numeric enum of the NVML Go bindings the design links to (design.md:50-52; the bindings are
 * not in the reference and no version is pinned): 0 unknown, then the rows of the table of design.md:33-44 in order,
 * 1 cross CPU .. 6 same board, 7..10 one to four NVLink links; 11 and 12 (five, six links) -> GT_LV_NVS; unknown is
 * taken as the farthest level (SYS).  types: n x n row-major, the diagonal is ignored.  Returns the GT_LV_* code,
 * -1 / GT_ERR_INVALID for a value outside the enum. */
int gt_level_from_p2plink(int32_t p2p_link_type);
int gt_levels_from_p2plink(const int32_t *types, uint32_t n, uint8_t out_levels[256]);
/* replaces: node annotations "GPU_<ABBR>_<i>_<j>: <description>" design.md:78-82 (one line per pair i<j) */
int gt_annotation_encode(const uint8_t levels[256], uint32_t n_gpus, char *out, size_t cap, size_t *out_len);
int gt_annotation_decode(const char *text, size_t len, uint8_t out_levels[256], uint32_t *out_n);
/* replaces: ALIYUN_COM_GPU_GROUP value "0,1,2,3" design.md:231 */
int gt_group_encode(uint16_t mask, char *out, size_t cap);
int gt_group_decode(const char *text, uint16_t *out_mask);
/* replaces: the pod annotations bind writes design.md:229-231 (assigned = 0) */
int gt_bind_annotations(uint16_t mask, int64_t assume_time, int assigned, char *out, size_t cap);
/* replaces: PrioritizeVerb "sort" reply design.md:99,118: score 0..10 per node (HostPriorityList) for a pod
 * asking for m GPUs; re-scores whatever changed first.  out_scores: n_nodes bytes.  out_place: optional,
 * 16 placement keys (as gt_best_node for k = 1..16) fetched in the same synchronisation. */
int gt_ext_sort(gt_ctx *ctx, uint32_t m, uint8_t *out_scores, uint64_t *out_place);
/* The same cycle without host copies, for an extender that keeps the informer's view in pinned memory:
 * gt_host_free returns the context's own pinned mirror of the free masks (max_nodes u16; waits for an upload, or a pass
 * whose binning kernel reads the mirror in place, that is still in flight: call it again before every round of edits,
 * do not keep the pointer across scoring calls that were not waited for) -- write masks in place, then gt_host_free_changed(first, count) says which nodes changed (the whole
 * shard: one bulk upload and a full pass by the next scoring call).  gt_ext_sort with out_scores == NULL leaves the
 * reply in the mapped pinned buffer gt_host_scores returns (n_nodes bytes, valid until the next gt_ext_sort). */
uint16_t *gt_host_free(gt_ctx *ctx);
int gt_host_free_changed(gt_ctx *ctx, uint32_t first, uint32_t count);
const uint8_t *gt_host_scores(gt_ctx *ctx);
/* replaces: bindVerb "bind" design.md:100,119,223-234: choose the GPUs on `node`, commit, format annotations */
int gt_ext_bind(gt_ctx *ctx, uint32_t node, uint32_t m, int64_t assume_time, uint16_t *out_mask,
                char *out_annotations, size_t cap);
/* replaces: device plugin Allocate design.md:236-246: GROUP annotation -> "NVIDIA_VISIBLE_DEVICES=0,1"
 * and the updated ASSIGNED / ASSUME_TIME annotations */
int gt_plugin_allocate(const char *pod_annotations, int64_t now, char *out_env, size_t env_cap,
                       char *out_annotations, size_t ann_cap);

/* ---- the extender's HTTP bodies ---- */
/* The request and reply types are upstream Kubernetes scheduler-extender types the reference links to
 * (design.md:121) and does not contain: ExtenderArgs{Pod, Nodes, NodeNames} -> HostPriorityList [{Host, Score}] for
 * `sort`, ExtenderBindingArgs{PodName, PodNamespace, PodUID, Node} -> ExtenderBindingResult{Error} for `bind`.
 * Field names are matched without regard to case, as Go's encoding/json does.  With these a Go handler is
 * `body -> C.gt_ext_sort_json -> w.Write`. */
#define GT_RESOURCE_NAME "aliyun.com/gpu" /* managedResources name, design.md:102-108 */
/* host only.  out_gpu_count: sum over the pod's containers of resources.limits[GT_RESOURCE_NAME] (requests where a
 * container states no limit); a fractional or negative quantity is GT_ERR_PARSE (Gaia's Fragment case, 0 < m < 1,
 * is not served: gaia.md:36).  out_uid: pod metadata.uid.  out_names: the NodeNames joined by '\n';
 * out_n_names = -1 when the field is absent or null (= every node).  Buffers may be NULL; too small: GT_ERR_RANGE. */
int gt_wire_parse_extender_args(const char *json, size_t len, uint32_t *out_gpu_count, char *out_uid, size_t uid_cap,
                                char *out_names, size_t names_cap, int32_t *out_n_names);
/* host only.  names: n node names separated by '\n'; scores: n bytes -> [{"Host":"a","Score":7},...] */
int gt_wire_host_priorities(const char *names, uint32_t n, const uint8_t *scores, char *out, size_t cap,
                            size_t *out_len);
int gt_wire_parse_binding_args(const char *json, size_t len, char *out_pod_name, size_t name_cap,
                               char *out_pod_namespace, size_t ns_cap, char *out_pod_uid, size_t uid_cap,
                               char *out_node, size_t node_cap);
int gt_wire_binding_result(const char *error, char *out, size_t cap);
/* node names in node-index order, separated by '\n' (what the informer knows the nodes as): needed by the two
 * handlers below.  count must equal the number of loaded nodes when they are called. */
int gt_set_node_names(gt_ctx *ctx, const char *names, size_t len);
/* replaces: the whole `sort` handler (design.md:99,117-118): ExtenderArgs JSON in, HostPriorityList JSON out
 * (scores for the NodeNames of the request in their order, or for every node).  Remembers the pod's GPU count by
 * its uid for the bind that follows.  Unknown node names score 0. */
int gt_ext_sort_json(gt_ctx *ctx, const char *request, size_t len, char *out, size_t cap, size_t *out_len);
/* replaces: the whole `bind` handler (design.md:100,119,223-234): ExtenderBindingArgs JSON in; picks the GPUs for the
 * pod seen by the last gt_ext_sort_json with that uid, commits them, records pod -> (node, GPUs) and returns the
 * ExtenderBindingResult JSON (Error set for an unknown pod / node or when the node cannot serve the request; the
 * return code is GT_OK whenever a reply body was produced) plus the pod annotations of design.md:229-231. */
int gt_ext_bind_json(gt_ctx *ctx, const char *request, size_t len, int64_t assume_time, char *out, size_t cap,
                     char *out_annotations, size_t ann_cap);
/* pod deletion seen by the informer: releases the GPUs gt_ext_bind_json recorded for the uid (GT_ERR_STATE: unknown) */
int gt_ext_pod_deleted(gt_ctx *ctx, const char *pod_uid);

/* ---- asynchronous / measurement surface ---- */
/* enqueue scoring + cluster argmin, no copies, no sync.  Passes enqueued back to back overlap at their seams: the
 * binning kernel of a pass reads the masks and fills its bin lists while the scoring kernel of the pass before it is
 * still finishing (separate lists and control blocks; results of a pass are complete before the next one touches the
 * table).  Mask updates through the API are ordered with that automatically. */
int gt_score_all_async(gt_ctx *ctx);
/* `passes` scoring passes enqueued round-robin over n_ctx contexts (ctxs[i % n_ctx] for pass i), no copies, no sync:
 * a scoring loop whose per-pass host cost is the two kernel launches, not a call through the binding */
int gt_score_all_async_many(gt_ctx **ctxs, uint32_t n_ctx, uint32_t passes);
/* the same loop between two CUDA events of the caller (cudaEvent_t handles, either may be NULL), recorded on the first
 * context's stream directly in front of the first launch and behind the last one */
int gt_score_all_async_many_timed(gt_ctx **ctxs, uint32_t n_ctx, uint32_t passes, void *ev_start, void *ev_stop);
int gt_sync(gt_ctx *ctx);
void *gt_stream(gt_ctx *ctx);            /* the cudaStream_t work is enqueued on */
const uint32_t *gt_device_best(gt_ctx *ctx);   /* device pointer of best[n_nodes][16] */
uint32_t gt_num_nodes(const gt_ctx *ctx);
uint64_t gt_launch_count(const gt_ctx *ctx);   /* kernels this context has launched */
/* candidate subsets one full scoring pass covers: sum over nodes of 2^popcount(free)-1; optionally
 * the algorithmic pair-cost additions sum C(f,2)*2^(f-2) */
int gt_count_subsets(gt_ctx *ctx, uint64_t *out_subsets, uint64_t *out_pair_adds);
/* time `iters` back-to-back launches of one kernel ("step", "step_nograph", "bin", "score", "argmin", "int32_peak") with
 * CUDA events on the context's stream; out_ms = average per launch */
int gt_time_kernel(gt_ctx *ctx, const char *which, uint32_t iters, float *out_ms);
/* full passes back to back while n_ctx contexts on one device and stream take turns, so that every launch finds its
 * tables in DRAM once the replicas exceed the L2; K1 stamps the global timer itself (first CTA past the dependency wait,
 * last CTA done), nothing sits between the launches: out_us = {pass - K1, K1 in place, pass} */
int gt_time_pass_rotated(gt_ctx **ctxs, uint32_t n_ctx, uint32_t iters, float out_us[3]);
/* this rank's own 16 placement keys recomputed from its best table, no exchange: the min over ranks of these is what
 * gt_best_node must return on every rank (checked by bench.py and the multi-GPU tests) */
int gt_best_node_local(gt_ctx *ctx, uint64_t out_place[16]);

/* ---- scoring without any allocation ---- */
/* No counterpart in the reference.
 * Some hosts of this library cannot allocate device memory at the moment they need a scoring pass:
 * an extender embedded in a process whose memory pool already owns the whole GPU (cudaMalloc there
 * fails or fragments the pool), or a latency-critical scheduling loop that cannot afford the
 * device-wide synchronisation cudaMalloc/cudaFree imply.  For them this one call runs K0 + K1
 * (bin by popcount, then score every k-subset) on buffers the CALLER already owns, and keeps the
 * kernels' scratch (bin counters and bin lists, ~408 KiB) in a fixed array inside the library image:
 * every GPU that loads libgputopo.so holds that array until the process exits.
 *   d_cost      n_nodes x 256 u8 pair costs on the device (16x16 per node, symmetric with a zero
 *               diagonal; with the default cost table a cost equals the GT_LV_* level code, so the
 *               level tiles of a symmetric topology matrix can be passed as they are)
 *   d_n_gpus    n_nodes u8, d_free_mask n_nodes u16, d_best n_nodes x 16 u32 (output, as gt_score_all)
 *   stream      cudaStream_t to enqueue on (NULL = the legacy default stream); the call returns
 *               once the work is enqueued -- synchronise the stream before reading d_best
 * Limits and rules:
 *   - n_nodes <= GT_STATIC_MAX_NODES, otherwise GT_ERR_RANGE (and a message in err_buf, if given);
 *   - the device that owns the buffers must be current; the first call on each device looks up the
 *     array's device address (one-time set-up) and should therefore happen before any stream capture;
 *   - that first call is also where the CUDA runtime loads this library's device code onto the GPU
 *     (lazy module loading): the driver reserves the array and the kernel images then, and that step
 *     CAN need device memory.  Make the first call per device at start-up, before a pool takes the
 *     whole GPU (a call with n_nodes = 1 on dummy buffers will do), or run with
 *     CUDA_MODULE_LOADING=EAGER so that the load happens when the library is opened.  Every later
 *     call on that device reserves nothing (the test checks the free-memory figure across calls);
 *   - all four buffers must be device (or managed) memory of the current device: host pointers are
 *     refused with GT_ERR_INVALID before anything is launched;
 *   - all calls on a device share the one array: the counters are cleared at the start of every call,
 *     the rest is overwritten, but two calls that overlap on the GPU (different streams without an
 *     ordering between them, or several host threads) race -- keep them stream-ordered;
 *   - nothing here calls cudaMalloc, cudaFree or any allocating library routine. */
#define GT_STATIC_MAX_NODES 4096u
int gt_score_nodes_static(const uint8_t *d_cost, const uint8_t *d_n_gpus, const uint16_t *d_free_mask,
                          uint32_t n_nodes, uint32_t *d_best, void *stream, char *err_buf, size_t err_len);

/* ---- multi-GPU: nodes are sharded, one context per GPU; the only exchange is a min over 16 keys ---- */
/* One process per GPU; rank 0 makes the id, every rank calls gt_comm_init with it.  From then on the 16 placement keys
 * of every pass are min-combined over the ranks: gt_best_node and gt_ext_sort(.., out_place) return the cluster's
 * winners (replaces nothing in the reference -- kube-scheduler picks the host there, design.md:115-119).  All ranks
 * must issue the same sequence of passes and key requests on a context.  A peer that does not show up within 2 s
 * turns the call that wanted the keys into GT_ERR_STATE; nothing on the GPU waits without a time limit. */
int gt_comm_unique_id(uint8_t out_id[128]);
int gt_comm_init(gt_ctx *ctx, const uint8_t id[128], int rank, int world);
/* further contexts of one process on the same device (other clusters, replicas) join the NCCL communicator of `owner`
 * instead of creating their own (they get an inbox of their own for the peer-memory exchange, so their passes stay
 * independent); collective: all ranks share the same way, in the same order; `owner` must be destroyed last */
int gt_comm_share(gt_ctx *ctx, gt_ctx *owner);
/* how the placement keys travel between ranks: 0 = no communicator, 1 = NCCL all-reduce on a side stream, 2 = peer
 * memory over NVLink (ranks are separate processes that could map each other's inbox through CUDA IPC): the work
 * rides on one warp of the scoring kernel of the following passes and on the call that asks for the keys */
int gt_comm_mode(const gt_ctx *ctx);

/* ---- synthetic clusters for tests and benchmarks (splitmix64) ---- */
int gt_synth_cluster(int config, uint64_t seed, uint32_t first, uint32_t count, uint8_t *levels,
                     uint8_t *n_gpus, uint16_t *free_mask, uint8_t *class_out, uint8_t *access_cost,
                     uint16_t *sibling_mask);
/* event traces of config 2 (arrivals only, k uniform 1..8) and config 5 (arrival p = 0.55, else departure) */
int gt_synth_events(int config, uint64_t seed, uint32_t n_events, uint8_t *ev_k, uint32_t *ev_arg);

#ifdef __cplusplus
}
#endif
#endif
