// api.cu -- the C-ABI of libgputopo.so (include/gputopo.h): context, tables in HBM, launches.
// No CPU fallback lives here: scoring only ever happens in the kernels of kernels.cuh.
#include "../../include/gputopo.h"
#include "kernels.cuh"
#include "wire.h"

#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

namespace {

struct NcclId { char internal[128]; };
struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(NcclId *) = nullptr;
    int (*CommInitRank)(void **, int, NcclId, int) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
constexpr int kNcclUint64 = 5;  // ncclUint64 in nccl.h
constexpr int kNcclUint8 = 1;
constexpr int kNcclMin = 3;     // ncclMin

NcclApi *nccl_api()
{
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names) {
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) return nullptr;
    api.GetUniqueId = (int (*)(NcclId *))dlsym(api.handle, "ncclGetUniqueId");
    api.CommInitRank = (int (*)(void **, int, NcclId, int))dlsym(api.handle, "ncclCommInitRank");
    api.AllReduce = (int (*)(const void *, void *, size_t, int, int, void *, cudaStream_t))dlsym(api.handle, "ncclAllReduce");
    api.AllGather = (int (*)(const void *, void *, size_t, int, void *, cudaStream_t))dlsym(api.handle, "ncclAllGather");
    api.CommDestroy = (int (*)(void *))dlsym(api.handle, "ncclCommDestroy");
    api.GetErrorString = (const char *(*)(int))dlsym(api.handle, "ncclGetErrorString");
    if (!api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.CommDestroy) {
        api.handle = nullptr;
        return nullptr;
    }
    return &api;
}

}  // namespace

struct gt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    uint32_t cap = 0, n_nodes = 0, first_node = 0, k1_policy = 0, flags = 0;
    int num_sms = 148;
    uint8_t lut[16];
    // HBM tables (SoA)
    uint8_t *d_levels = nullptr, *d_cost = nullptr, *d_ngpus = nullptr;
    uint8_t *d_access = nullptr;
    uint16_t *d_sibling = nullptr;
    gt::TreeRow *d_tree = nullptr;              // Gaia trees (gt_load_trees); allocated on first use
    uint32_t *d_gaia_out = nullptr;             // answers of gt_gaia_select
    uint16_t *d_free = nullptr;
    uint32_t *d_best = nullptr;
    uint32_t *d_bin_cnt = nullptr, *d_bin_list = nullptr;   // scratch of K0/K1 (control block, node lists)
    uint16_t *d_bin_free = nullptr;                         // scratch: free masks beside the node lists
    unsigned long long *d_place = nullptr;      // [8][16]: [1] scratch of gt_best_node_local, [2],[3],[6] the key buffers K1 folds into, [4],[5],[7] the cluster-wide keys of the three latest passes (peer-memory exchange)
    unsigned long long *place_cur = nullptr;    // where the latest placement keys are
    cudaGraphExec_t step_exec[3] = {nullptr, nullptr, nullptr};   // captured K0 + K1 of a full pass, per key buffer
    uint32_t step_nodes = 0;                    // n_nodes the graphs were captured for
    bool graph_failed = false;
    uint32_t bin_parity = 0;                    // placement-key buffer the next pass writes
    // multi-GPU: the 128-byte min-exchange runs on its own stream so it overlaps the next pass
    cudaStream_t comm_stream = nullptr;         // shared by the contexts of this process on the same device (see comm_stream_for)
    bool own_comm_stream = false;
    cudaEvent_t ev_pass = nullptr, ev_comm[2] = {nullptr, nullptr};
    bool comm_used[2] = {false, false};
    int comm_last = -1;                         // parity whose exchange is the latest one in flight
    // peer-memory exchange (k_exchange): every rank's inbox mapped through CUDA IPC; NCCL stays the fallback
    bool p2p = false;
    // extender handlers on JSON bodies: node names, pods between sort and bind, pods bound
    std::vector<std::string> node_names;
    std::unordered_map<std::string, uint32_t> node_index;
    std::unordered_map<std::string, uint32_t> pod_request;                       // uid -> GPUs asked for (from sort)
    std::deque<std::string> pod_request_order;                                   // oldest first: the table is bounded
    std::unordered_map<std::string, std::pair<uint32_t, uint16_t>> pod_bound;    // uid -> (node, GPUs)
    unsigned long long *d_inbox = nullptr;      // own inbox: keys[world][2][16] + flags[world][2]
    unsigned long long *peer_inbox[gt::kMaxPeers] = {};
    unsigned long long *d_xflags = nullptr;     // [3]: number of a pass whose exchange gave up waiting for a peer (0 = none)
    // peer-memory exchange (kernels.cuh exchange_step): passes whose keys travel, how far posting and reducing got
    uint64_t x_seq = 0, x_posted = 0, x_reduced = 0;
    bool comm_borrowed = false;                 // the NCCL communicator and the communication stream belong to another context
    uint32_t *d_ids = nullptr;
    uint16_t *d_masks = nullptr;
    uint32_t *d_peak_out = nullptr;
    uint8_t *d_scores = nullptr;                // one score byte per node, left by a pass that ran for a `sort` call
    uint32_t want_score_m = 0;                  // request size the pass being enqueued shall leave scores for (0 = none)
    uint32_t scores_m = 0;                      // request size d_scores is valid for (0 = stale)
    uint8_t *h_scores = nullptr;                // mapped pinned: the score kernels write the reply straight to the host
    uint32_t *d_scores_map = nullptr;           // device view of h_scores
    unsigned long long *h_place_map = nullptr, *d_place_map = nullptr;   // same for the 16 placement keys
#ifdef GT_TRACE
    unsigned long long *d_trace = nullptr;
#endif
    // pinned host mirrors
    uint16_t *h_free = nullptr;
    uint16_t *d_free_map = nullptr;             // device view of h_free (mapped): K0 reads a whole-shard update in place
    unsigned long long *ktime_next = nullptr;   // measurement: where the next K1 leaves its time stamps
    bool masks_in_place = false;                // the K0 being enqueued reads the masks from the mirror, not from d_free
    bool k0_inputs_pending = true;              // a write to the device masks / GPU counts / id list has been queued since the last K0
    uint32_t ctl_idx = 0;                       // which of the three control blocks the next K0 / K1 pair uses (zero between passes)
    uint32_t list_half = 0;                     // which of the two sets of bin lists the next K0 / K1 pair uses
    bool k0_may_run_ahead = false;              // the last K1 launched rotated blocks and lists: the next K0 touches nothing it reads
    bool capturing = false;                     // kernels being recorded into a graph are replayed later: no per-call promises
    uint8_t *h_ngpus = nullptr;
    uint32_t *h_ids = nullptr;
    uint16_t *h_masks = nullptr;
    unsigned long long *h_place = nullptr;
    uint32_t *h_row = nullptr;
    std::vector<uint8_t> dirty_flag;
    std::vector<uint32_t> dirty;
    bool free_uploaded = true;    // device free[] equals the mirror except for `dirty`
    cudaEvent_t ev_h2d = nullptr; // last copy that reads the pinned mirrors (they may not change before it ran)
    bool h2d_pending = false;
    bool mirror_read_in_flight = false;         // a K0 that reads the mirror in place has been queued and the stream not waited for since
    bool best_valid = false, place_valid = false;
    uint64_t launches = 0;
    void *comm = nullptr;
    int rank = 0, world = 1;
    // commit loop (gt_schedule): topology classes seen by gt_load_nodes and the memo table the GPU fills
    std::unordered_map<std::string, uint32_t> class_of_tile;   // 256 level bytes + n_gpus -> class id
    std::vector<std::string> class_tile;
    std::vector<uint32_t> node_class;
    std::vector<uint32_t> memo;            // [sum over classes of 2^n][16] best rows, filled by K0+K1
    std::vector<size_t> memo_off;          // first row of each class
    uint32_t memo_classes = 0;             // classes covered by `memo`
    uint64_t lut_version = 0, memo_lut_version = ~0ull;
    // GT_FLAG_MEMO_PASS: the same table on the device, plus each node's class and each class's first row
    uint32_t *d_memo = nullptr, *d_row0 = nullptr, *d_node_class = nullptr;
    uint32_t dmemo_classes = 0;
    uint64_t dmemo_lut_version = ~0ull;
    bool classes_uploaded = false;
    bool memo_pass_failed = false;
    char err[256];
};

namespace {

int fail(gt_ctx *c, int code, const char *fmt, const char *detail = "")
{
    if (c) snprintf(c->err, sizeof c->err, fmt, detail);
    return code;
}

#define GT_CUDA(c, call)                                                          \
    do {                                                                          \
        cudaError_t e__ = (call);                                                 \
        if (e__ != cudaSuccess) {                                                 \
            if (c) snprintf((c)->err, sizeof(c)->err, "%.160s: %s", #call, cudaGetErrorString(e__)); \
            return e__ == cudaErrorMemoryAllocation ? GT_ERR_NOMEM : GT_ERR_CUDA; \
        }                                                                         \
    } while (0)

uint4 lut_as_uint4(const uint8_t lut[16])
{
    uint4 v;
    memcpy(&v, lut, 16);
    return v;
}

int apply_lut(gt_ctx *c, uint32_t first, uint32_t count)
{
    if (!count) return GT_OK;
    const size_t n16 = (size_t)count * 16;   // tile rows: 16 per node, 16 bytes each
    const int grid = (int)std::min<size_t>((n16 + 255) / 256, (size_t)c->num_sms * 8);
    gt::k_apply_lut<<<grid, 256, 0, c->stream>>>(c->d_levels + (size_t)first * 256, c->d_cost + (size_t)first * 256,
                                                 n16, lut_as_uint4(c->lut));
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    return GT_OK;
}

__global__ void k_scatter_free(const uint32_t *ids, const uint16_t *masks, uint32_t n, uint16_t *free_mask)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) free_mask[ids[i]] = masks[i];
}

// the pinned mirrors are about to be modified: a copy that still reads them must have run
int mirror_writable(gt_ctx *c)
{
    if (c->h2d_pending) {
        GT_CUDA(c, cudaEventSynchronize(c->ev_h2d));
        c->h2d_pending = false;
    }
    // a K0 reading the mirror in place leaves no event behind: one between K0 and K1 would make K1 an ordinary launch
    // instead of a programmatic dependent.  Callers that wait for their passes (the `sort` cycle) never get here with
    // the flag set (stream_waited_for); one that edits the mirror while a pass is in flight waits for that pass.
    if (c->mirror_read_in_flight) {
        GT_CUDA(c, cudaStreamSynchronize(c->stream));
        c->mirror_read_in_flight = false;
    }
    return GT_OK;
}

void stream_waited_for(gt_ctx *c) { c->mirror_read_in_flight = false; }

// bring the device free[] up to date with the host mirror
int flush_free(gt_ctx *c, bool all)
{
    if (!c->n_nodes) return GT_OK;
    if (all || !c->free_uploaded) {
        c->k0_inputs_pending = true;
        GT_CUDA(c, cudaMemcpyAsync(c->d_free, c->h_free, (size_t)c->n_nodes * 2, cudaMemcpyHostToDevice, c->stream));
        GT_CUDA(c, cudaEventRecord(c->ev_h2d, c->stream));
        c->h2d_pending = true;
        c->free_uploaded = true;
        return GT_OK;
    }
    const uint32_t n = (uint32_t)c->dirty.size();
    if (!n) return GT_OK;
    int rcw = mirror_writable(c);   // the id / mask staging buffers are reused
    if (rcw != GT_OK) return rcw;
    for (uint32_t i = 0; i < n; i++) {
        c->h_ids[i] = c->dirty[i];
        c->h_masks[i] = c->h_free[c->dirty[i]];
    }
    c->k0_inputs_pending = true;
    GT_CUDA(c, cudaMemcpyAsync(c->d_ids, c->h_ids, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    GT_CUDA(c, cudaMemcpyAsync(c->d_masks, c->h_masks, (size_t)n * 2, cudaMemcpyHostToDevice, c->stream));
    k_scatter_free<<<(n + 255) / 256, 256, 0, c->stream>>>(c->d_ids, c->d_masks, n, c->d_free);
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    GT_CUDA(c, cudaEventRecord(c->ev_h2d, c->stream));
    c->h2d_pending = true;
    return GT_OK;
}

void clear_dirty(gt_ctx *c)
{
    for (uint32_t id : c->dirty) c->dirty_flag[id] = 0;
    c->dirty.clear();
}

void mark_dirty(gt_ctx *c, uint32_t node)
{
    if (!c->dirty_flag[node]) {
        c->dirty_flag[node] = 1;
        c->dirty.push_back(node);
    }
    c->place_valid = false;
}

// key buffer K1 folds into: rows [2], [3] and [6] of d_place (two alternate without peers, three rotate with the
// peer-memory exchange, where the keys of a pass are posted a pass later than they are produced)
unsigned long long *bin_place(gt_ctx *c, uint32_t slot) { return c->d_place + (slot == 2 ? 96 : 32 + 16 * slot); }
uint32_t *cur_ctl(gt_ctx *c) { return c->d_bin_cnt + c->ctl_idx * gt::kCtlWords; }
uint32_t *cur_list(gt_ctx *c) { return c->d_bin_list + (size_t)c->list_half * gt::kBins * c->cap; }
uint16_t *cur_flist(gt_ctx *c) { return c->d_bin_free + (size_t)c->list_half * gt::kBins * c->cap; }

gt::ScoreParams score_params(gt_ctx *c, unsigned long long *fold_into, bool keep_bins)
{
    gt::ScoreParams p = {};
    p.cost = c->d_cost; p.best = c->d_best; p.ctl = cur_ctl(c); p.bin_list = cur_list(c);
    p.self_clean = 1u; p.ctl_next = nullptr; p.trigger_early = 0u;
    p.bin_free = cur_flist(c); p.place = fold_into; p.cap = c->cap; p.first_node = c->first_node;
    p.fold_place = fold_into ? 1u : 0u; p.keep_bins = keep_bins ? 1u : 0u; p.zero = 0;
    p.xw = gt::XWork{};
    p.scores = nullptr; p.score_k = 0; p.score_den = 0;
    if (c->want_score_m && fold_into) {   // full pass on behalf of a `sort` call: one score byte per node on the way
        const uint32_t m = c->want_score_m;
        p.scores = c->d_scores; p.score_k = m; p.score_den = gt_lut_cmax(c->lut) * (m * (m - 1) / 2);
    }
#ifdef GT_TRACE
    p.trace = c->d_trace;
#endif
    return p;
}

// K0.  The control block it counts into is zero (three take turns, see launch_score); preset_place: K1 will fold.
int launch_bin(gt_ctx *c, const uint32_t *d_ids, uint32_t n, unsigned long long *preset_place, const gt::XWork *xw = nullptr)
{
    // programmatic dependent of the kernel in front of it (the previous pass's K1 in a scoring loop): the launch
    // latency between two passes is spent while that kernel still runs; K0 waits inside for its completion
    cudaLaunchConfig_t cfg = {};
    // multi-GPU: CTA 0 is an extra one that carries a step of the key exchange (k_bin)
    const gt::XWork xwork = xw ? *xw : gt::XWork{};
    cfg.gridDim = dim3((n + gt::kCta - 1) / gt::kCta + (xwork.world ? 1u : 0u));
    cfg.blockDim = dim3(gt::kCta);
    cfg.stream = c->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static const bool no_pdl = getenv("GT_NO_PDL") != nullptr || getenv("GT_NO_PDL_K0") != nullptr;   // profiling aid
    cfg.numAttrs = no_pdl ? 0 : 1;
    const bool in_place = c->masks_in_place && !d_ids;
    // whole-shard K0 whose inputs nothing in the queue is still writing: it may fetch and count the masks before its
    // wait for the kernel in front of it, i.e. behind the tail of the previous pass's K1 (GT_NO_EARLY_K0: profiling aid)
    static const bool no_early = getenv("GT_NO_EARLY_K0") != nullptr;
    // ... and when the K1 in front of it works on the other set of bin lists and another control block (see
    // launch_score), K0 also reserves its ranges and fills its lists before the wait: 2 (GT_NO_K0_AHEAD: level 1 at most)
    static const bool no_ahead = getenv("GT_NO_K0_AHEAD") != nullptr;
    uint32_t early = !no_pdl && !no_early && !d_ids && !c->k0_inputs_pending && !c->capturing ? 1u : 0u;
    if (early && c->k0_may_run_ahead && !no_ahead) early = 2u;
    GT_CUDA(c, cudaLaunchKernelEx(&cfg, gt::k_bin, (const uint8_t *)c->d_ngpus,
                                  (const uint16_t *)(in_place ? c->d_free_map : c->d_free), d_ids, n,
                                  c->cap, cur_ctl(c), cur_list(c), cur_flist(c), c->d_best, preset_place,
                                  (uint8_t *)((c->want_score_m && preset_place && !d_ids) ? c->d_scores : nullptr),
                                  (uint16_t *)(in_place ? c->d_free : nullptr), early, xwork));
    c->k0_inputs_pending = false;    // this K0, or its wait, is behind every write queued so far
    if (in_place) {       // the mirror may not change before this K0 has read it (mirror_writable)
        static const bool eager_event = getenv("GT_MIRROR_EVENT") != nullptr;     // A/B aid: the event of earlier builds
        if (eager_event) {
            GT_CUDA(c, cudaEventRecord(c->ev_h2d, c->stream));
            c->h2d_pending = true;
        } else {
            c->mirror_read_in_flight = true;
        }
        c->free_uploaded = true;
        c->masks_in_place = false;
    }
    c->launches++;
    return GT_OK;
}

constexpr uint32_t kPipeMinNodes = 4096;

// k_score_pipe needs more dynamic shared memory than the default limit: raised once per device
int pipe_ready(gt_ctx *c)
{
    static std::mutex mu;
    static bool done[64] = {};
    std::lock_guard<std::mutex> lock(mu);
    if (c->device >= 0 && c->device < 64 && done[c->device]) return GT_OK;
    GT_CUDA(c, cudaFuncSetAttribute(gt::k_score_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(gt::pipe::Shared)));
    if (c->device >= 0 && c->device < 64) done[c->device] = true;
    return GT_OK;
}

int launch_score(gt_ctx *c, uint32_t n, unsigned long long *fold, bool keep_bins = false, const gt::XWork *xw = nullptr)
{
    // persistent warps: every resident CTA slot, fewer when there is little work
    const uint32_t grid = std::min<uint32_t>(n / 8 + gt::kBins, (uint32_t)c->num_sms * gt::kScoreMinCtas);
    // programmatic dependent launch: K1's CTAs are scheduled while K0 drains and wait inside the kernel
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(gt::kScoreCta);
    cfg.stream = c->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    static const bool no_pdl = getenv("GT_NO_PDL") != nullptr;   // profiling aid
    cfg.numAttrs = no_pdl ? 0 : 1;
    gt::ScoreParams prm = score_params(c, fold, keep_bins);
    prm.ktime = c->ktime_next;
    if (xw) prm.xw = *xw;     // multi-GPU: one warp of this K1 does the exchange work of the two passes before it
    // GT_PIPE=1: the producer / consumer form of K1 for full passes -- same results (it passes every parity test), but
    // its one producer warp per CTA cannot describe units as fast as the consumers finish them, so it measures 2-3x
    // slower than k_score (profiles/r2_ab_pipe_*.txt) and stays an experiment
    static const bool use_pipe = getenv("GT_PIPE") != nullptr;
    if (fold && use_pipe && n >= kPipeMinNodes) {
        int rcp = pipe_ready(c);
        if (rcp != GT_OK) return rcp;
        cfg.gridDim = dim3(std::min<uint32_t>(n / 8 + gt::kBins, (uint32_t)c->num_sms * gt::pipe::kCtasPerSm));
        cfg.blockDim = dim3(gt::pipe::kCta);
        cfg.dynamicSmemBytes = sizeof(gt::pipe::Shared);
        GT_CUDA(c, cudaLaunchKernelEx(&cfg, gt::k_score_pipe, prm));
        c->k0_may_run_ahead = false;
        c->launches++;
        c->place_cur = fold;
        return GT_OK;
    }
    // the two control blocks take turns (see k_score) unless this launch has to leave its own block usable again:
    // timing runs that keep the bins or want the end-of-kernel stamp, and kernels recorded into a graph
    static const bool no_flip = getenv("GT_NO_CTL_FLIP") != nullptr;     // profiling aid
    const bool flip = !no_flip && !keep_bins && !prm.ktime && !c->capturing && !(c->flags & GT_FLAG_GRAPH);
    if (flip) {
        // three blocks in rotation: this pass's, the next pass's (zeroed by the K1 before this one, so the next K0 may
        // count into it while this K1 still runs) and the one the pass before this used, which this K1 zeroes
        prm.self_clean = 0u;
        prm.ctl_next = c->d_bin_cnt + ((c->ctl_idx + 2u) % 3u) * gt::kCtlWords;
        static const bool no_ahead = getenv("GT_NO_K0_AHEAD") != nullptr;
        prm.trigger_early = no_ahead ? 0u : 1u;   // dependents may become resident as this grid's CTAs leave
    }
    GT_CUDA(c, cudaLaunchKernelEx(&cfg, gt::k_score, prm));
    if (flip) {
        c->ctl_idx = (c->ctl_idx + 1u) % 3u;
        c->list_half ^= 1u;
    }
    c->k0_may_run_ahead = flip;
    c->launches++;
    if (fold) c->place_cur = fold;
    return GT_OK;
}

// request-size-1 policy of the Gaia paper, applied after K1 (no-op for the default policy)
int launch_singular(gt_ctx *c, const uint32_t *d_ids, uint32_t n, unsigned long long *place)
{
    if (c->k1_policy == GT_K1_FIRST_UNUSED || !n) return GT_OK;
    gt::k_singular<<<(n + gt::kCta - 1) / gt::kCta, gt::kCta, 0, c->stream>>>(
        c->d_ngpus, c->d_free, c->d_access, c->d_sibling, c->d_cost, c->k1_policy, d_ids, n, c->d_best, place,
        c->first_node);
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    return GT_OK;
}

// Multi-GPU without peer memory (ranks that cannot map each other's memory, more than 8 ranks, GT_NO_P2P): min of the
// 16 placement keys over all ranks as an NCCL all-reduce on the communication stream, ordered behind the pass by an
// event; consumers wait for ev_comm[parity].
int exchange_min(gt_ctx *c, unsigned long long *place, uint32_t parity)
{
    if (!c->comm || c->p2p) return GT_OK;
    GT_CUDA(c, cudaEventRecord(c->ev_pass, c->stream));
    GT_CUDA(c, cudaStreamWaitEvent(c->comm_stream, c->ev_pass, 0));
    NcclApi *api = nccl_api();
    int e = api->AllReduce(place, place, 16, kNcclUint64, kNcclMin, c->comm, c->comm_stream);
    if (e != 0) return fail(c, GT_ERR_NCCL, "ncclAllReduce: %s", api->GetErrorString ? api->GetErrorString(e) : "?");
    GT_CUDA(c, cudaEventRecord(c->ev_comm[parity], c->comm_stream));
    c->comm_used[parity] = true;
    c->comm_last = (int)parity;
    return GT_OK;
}

// ---- Multi-GPU over peer memory: the exchange is piggybacked on the passes (kernels.cuh, exchange_step).
// Key buffers of pass number s: this rank's own keys (what K1 folds into) and the cluster-wide result.
unsigned long long *x_local(gt_ctx *c, uint64_t s) { return bin_place(c, (uint32_t)(s % 3u)); }
unsigned long long *x_global(gt_ctx *c, uint64_t s) { const uint32_t q = (uint32_t)(s % 3u); return c->d_place + (q == 2 ? 112 : 64 + 16 * q); }
bool lazy_exchange(const gt_ctx *c) { return c->comm && c->p2p; }

gt::XWork x_blank(gt_ctx *c)
{
    gt::XWork w = {};
    for (int q = 0; q < c->world; q++) w.inbox[q] = c->peer_inbox[q];
    w.err_flag = c->d_xflags + 3;
    w.rank = (uint32_t)c->rank;
    w.world = 0;          // set once there is work
    return w;
}

// Protocol state of this rank: its keys of passes 1..x_posted are in every inbox, the cluster-wide keys of passes
// 1..x_reduced are in x_global; x_reduced <= x_posted <= x_reduced + 1.  One step (exchange_step) is, in this order,
// reduce(x_reduced + 1) if that pass is posted, post(x_posted + 1), reduce of the pass just posted -- each only as far
// as the caller's bounds allow (a pass can be posted once its K1 is complete).
gt::XWork x_next(gt_ctx *c, uint64_t max_post, uint64_t max_reduce)
{
    gt::XWork w = x_blank(c);
    if (c->x_reduced < c->x_posted && c->x_reduced + 1 <= max_reduce) {
        const uint64_t s = ++c->x_reduced;
        w.seq_reduce1 = s; w.out1 = x_global(c, s);
        w.world = (uint32_t)c->world;
    }
    if (c->x_posted == c->x_reduced && c->x_posted + 1 <= max_post) {
        const uint64_t s = ++c->x_posted;
        w.seq_post = s; w.post_keys = x_local(c, s);
        w.world = (uint32_t)c->world;
        if (s <= max_reduce) { w.seq_reduce2 = s; w.out2 = x_global(c, s); c->x_reduced = s; }
    }
    return w;
}

int launch_exchange_step(gt_ctx *c, const gt::XWork &w, cudaStream_t on = nullptr)
{
    if (!w.world) return GT_OK;
    gt::k_exchange_step<<<1, 32, 0, on ? on : c->stream>>>(w);
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    return GT_OK;
}

// one-warp kernels until the protocol is at (max_post, max_reduce) or as far as it can get
int x_catch_up(gt_ctx *c, uint64_t max_post, uint64_t max_reduce, cudaStream_t on = nullptr)
{
    for (;;) {
        gt::XWork w = x_next(c, max_post, max_reduce);
        if (!w.world) return GT_OK;
        int rc = launch_exchange_step(c, w, on);
        if (rc != GT_OK) return rc;
    }
}

// somebody wants the cluster-wide keys of the latest pass: whatever is still open, in one-warp kernels.
// keys_host / on: the last step also hands the keys to the caller's mapped buffer; they run on another stream (gt_ext_sort).
int x_consume(gt_ctx *c, unsigned long long *keys_host = nullptr, cudaStream_t on = nullptr)
{
    const uint64_t L = c->x_seq;
    if (!L) return GT_OK;
    while (c->x_reduced < L) {
        gt::XWork w = x_next(c, L, L);
        if (!w.world) break;
        if (c->x_reduced >= L) w.keys_host = keys_host;
        int rc = launch_exchange_step(c, w, on);
        if (rc != GT_OK) return rc;
    }
    return GT_OK;
}

// where the placement keys a caller may see are: the cluster-wide ones with peers, else what the pass left
const unsigned long long *cluster_keys(gt_ctx *c)
{
    return lazy_exchange(c) && c->x_seq ? x_global(c, c->x_seq) : c->place_cur;
}

// an exchange kernel gave up waiting for a peer (kernels.cuh, kExchangeTimeoutNs): the placement keys of this context
// are this rank's own, not the cluster's
int exchange_failed(gt_ctx *c)
{
    return fail(c, GT_ERR_STATE, "the min-exchange of the placement keys timed out waiting for a peer: a rank is gone, or the "
                                 "ranks issued the passes of their contexts in different orders%s");
}

// NCCL fallback: before a pass overwrites place[parity] the all-reduce that last used that buffer must be finished
int wait_place_free(gt_ctx *c, uint32_t parity)
{
    if (c->comm && !c->p2p && c->comm_used[parity]) GT_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_comm[parity], 0));
    return GT_OK;
}

// full pass as one graph launch: K0 -> K1, captured once per node count and place parity
int launch_step_graph(gt_ctx *c, uint32_t parity)
{
    if (c->step_nodes != c->n_nodes) {
        for (auto &g : c->step_exec)
            if (g) { cudaGraphExecDestroy(g); g = nullptr; }
    }
    if (!c->step_exec[parity]) {
        cudaGraph_t graph = nullptr;
        const uint64_t before = c->launches;
        GT_CUDA(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
        c->capturing = true;
        const uint32_t want = c->want_score_m;
        c->want_score_m = 0;            // the captured kernels are replayed for every later pass: no per-call parameters
        int rc = launch_bin(c, nullptr, c->n_nodes, bin_place(c, parity));
        if (rc == GT_OK) rc = launch_score(c, c->n_nodes, bin_place(c, parity));
        if (rc == GT_OK) rc = launch_singular(c, nullptr, c->n_nodes, bin_place(c, parity));
        cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
        c->capturing = false;
        c->want_score_m = want;
        c->launches = before;
        if (rc != GT_OK || e != cudaSuccess || !graph) {
            cudaGetLastError();
            if (graph) cudaGraphDestroy(graph);
            return GT_ERR_CUDA;
        }
        e = cudaGraphInstantiate(&c->step_exec[parity], graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) {
            cudaGetLastError();
            c->step_exec[parity] = nullptr;
            return GT_ERR_CUDA;
        }
        c->step_nodes = c->n_nodes;
    }
    GT_CUDA(c, cudaGraphLaunch(c->step_exec[parity], c->stream));
    c->launches += c->k1_policy != GT_K1_FIRST_UNUSED ? 3 : 2;
    c->place_cur = bin_place(c, parity);
    return GT_OK;
}

int launch_argmin(gt_ctx *c, unsigned long long *place)
{
    GT_CUDA(c, cudaMemsetAsync(place, 0xFF, 16 * sizeof(unsigned long long), c->stream));
    const size_t quarters = (size_t)c->n_nodes * 4;      // 16-byte pieces; 8 per thread
    const uint32_t grid = (uint32_t)std::min<size_t>((quarters + gt::kCta * 8 - 1) / (gt::kCta * 8), (size_t)c->num_sms * 4);
    gt::k_argmin<<<std::max(grid, 1u), gt::kCta, 0, c->stream>>>(c->d_best, c->n_nodes, c->first_node, place);
    c->place_cur = place;
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    return GT_OK;
}

extern "C" {
int build_memo(gt_ctx *c);   // defined with gt_schedule (inside the extern "C" block: same linkage here)
}

// GT_FLAG_MEMO_PASS: one row copy per node out of the GPU-built (class, free mask) table
int launch_memo_pass(gt_ctx *c, unsigned long long *place)
{
    int rc = build_memo(c);            // K0 + K1 over every (class, mask); GT_ERR_RANGE when there are too many classes
    if (rc != GT_OK) return rc;
    const uint32_t ncls = (uint32_t)c->class_tile.size();
    if (c->dmemo_classes != ncls || c->dmemo_lut_version != c->lut_version) {
        cudaFree(c->d_memo); cudaFree(c->d_row0);
        c->d_memo = c->d_row0 = nullptr;
        std::vector<uint32_t> row0(ncls);
        for (uint32_t k = 0; k < ncls; k++) row0[k] = (uint32_t)c->memo_off[k];
        GT_CUDA(c, cudaMalloc(&c->d_memo, c->memo.size() * 4));
        GT_CUDA(c, cudaMalloc(&c->d_row0, (size_t)ncls * 4));
        GT_CUDA(c, cudaMemcpy(c->d_memo, c->memo.data(), c->memo.size() * 4, cudaMemcpyHostToDevice));
        GT_CUDA(c, cudaMemcpy(c->d_row0, row0.data(), (size_t)ncls * 4, cudaMemcpyHostToDevice));
        c->dmemo_classes = ncls;
        c->dmemo_lut_version = c->lut_version;
    }
    if (!c->classes_uploaded) {
        if (!c->d_node_class) GT_CUDA(c, cudaMalloc(&c->d_node_class, (size_t)c->cap * 4));
        GT_CUDA(c, cudaMemcpy(c->d_node_class, c->node_class.data(), (size_t)c->n_nodes * 4, cudaMemcpyHostToDevice));
        c->classes_uploaded = true;
    }
    GT_CUDA(c, cudaMemsetAsync(place, 0xFF, 16 * sizeof(unsigned long long), c->stream));
    const size_t quarters = (size_t)c->n_nodes * 4;
    const uint32_t grid = (uint32_t)std::min<size_t>((quarters + gt::kCta * 4 - 1) / (gt::kCta * 4), (size_t)c->num_sms * 8);
    gt::k_memo_pass<<<std::max(grid, 1u), gt::kCta, 0, c->stream>>>(c->d_memo, c->d_row0, c->d_node_class, c->d_ngpus,
                                                                   c->d_free, c->n_nodes, c->first_node, c->d_best, place);
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    c->place_cur = place;
    return GT_OK;
}

// Full pass (d_ids == nullptr): K0 + K1 with the cluster argmin folded into K1, linked by programmatic
// dependent launch (two plain launches; as a captured graph only with GT_FLAG_GRAPH).
// Listed (dirty) nodes: K0 + K1 over the list, then K2 over the whole table.  n == 0: K2 only.
int enqueue_scoring(gt_ctx *c, const uint32_t *d_ids, uint32_t n)
{
    int rc;
    const bool memo_pass = !d_ids && n && (c->flags & GT_FLAG_MEMO_PASS) && !c->memo_pass_failed;
    const bool graph = !d_ids && n && (c->flags & GT_FLAG_GRAPH) && !c->graph_failed;
    // with peers over peer memory the pass has a number, its key buffers go by that number, and the exchange work of
    // the two passes before it rides on its K1 (or, on the paths without a K1 of their own, in a one-warp kernel)
    static const bool probe_no_x = getenv("GT_PROBE_NO_EXCHANGE") != nullptr;   // timing probe: passes of a rank with peers, nothing exchanged
    const bool lazy = lazy_exchange(c) && !probe_no_x;
    const uint64_t seq = lazy ? c->x_seq + 1 : 0;
    const uint32_t par = lazy ? (uint32_t)(seq % 3u) : c->bin_parity;
    if ((rc = wait_place_free(c, par)) != GT_OK) return rc;
    // Plain K0 -> K1 passes: one warp of K1 carries the step, reduce(seq-2) and post(seq-1).  GT_X_ON_K0: an extra CTA
    // of K0 instead, before K0's dependency wait, i.e. beside the last units of the previous K1 -- what it may post then
    // is the pass before the previous one, reduce(seq-3) and post(seq-2).  Parity-clean, but the step's NVLink round
    // trips (2-3 us) outlast the tail of that K1 and K0 ends late: 1 us per pass slower (profiles/r2_mp_probe_v4.txt).
    // Either way the key buffer this pass folds into last held pass seq-3, which must be posted before K0 presets it.
    static const bool x_on_k1 = getenv("GT_X_ON_K0") == nullptr;
    if (lazy && seq > 4 && (rc = x_catch_up(c, seq - 3, seq - 4)) != GT_OK) return rc;
    gt::XWork step = gt::XWork{};
    c->scores_m = 0;
    bool step_taken = false;
    if (!d_ids && n) {
        bool done = false;
        if (memo_pass) {
            rc = launch_memo_pass(c, bin_place(c, par));
            if (rc == GT_ERR_RANGE) c->memo_pass_failed = true;   // too many topology classes: enumerate as usual
            else if (rc != GT_OK) return rc;
            else done = true;
        }
        if (!done && graph) {   // measured slower than two plain launches: opt-in
            done = launch_step_graph(c, par) == GT_OK;
            if (!done) c->graph_failed = true;   // capture unavailable (e.g. the stream is being captured): plain launches
        }
        if (!done) {
            if (lazy) step = x_on_k1 ? x_next(c, seq - 1, seq >= 2 ? seq - 2 : 0) : x_next(c, seq >= 2 ? seq - 2 : 0, seq >= 3 ? seq - 3 : 0);
            if ((rc = launch_bin(c, nullptr, n, bin_place(c, par), !x_on_k1 && step.world ? &step : nullptr)) != GT_OK) return rc;
            if ((rc = launch_score(c, n, bin_place(c, par), false, x_on_k1 && step.world ? &step : nullptr)) != GT_OK) return rc;
            step_taken = true;
            if ((rc = launch_singular(c, nullptr, n, bin_place(c, par))) != GT_OK) return rc;
            c->scores_m = c->want_score_m;
        }
    } else {
        if (n) {
            if ((rc = launch_bin(c, d_ids, n, nullptr)) != GT_OK) return rc;
            if ((rc = launch_score(c, n, nullptr)) != GT_OK) return rc;
            if ((rc = launch_singular(c, d_ids, n, nullptr)) != GT_OK) return rc;
        }
        if ((rc = launch_argmin(c, bin_place(c, par))) != GT_OK) return rc;
    }
    // paths without a step of their own: one-warp kernels behind the pass, reduce(seq-2) and post(seq-1)
    if (lazy && !step_taken && (rc = x_catch_up(c, seq - 1, seq >= 2 ? seq - 2 : 0)) != GT_OK) return rc;
    if (lazy) c->x_seq = seq;
    else if (!probe_no_x && (rc = exchange_min(c, c->place_cur, par)) != GT_OK) return rc;
    c->bin_parity ^= 1u;
    c->best_valid = true;
    c->place_valid = true;
    return GT_OK;
}

}  // namespace

extern "C" {

const char *gt_strerror(int status)
{
    switch (status) {
    case GT_OK: return "ok";
    case GT_ERR_INVALID: return "invalid argument";
    case GT_ERR_NO_DEVICE: return "no usable CUDA device";
    case GT_ERR_CUDA: return "CUDA error";
    case GT_ERR_NOMEM: return "out of memory";
    case GT_ERR_RANGE: return "node index out of range";
    case GT_ERR_PARSE: return "parse error";
    case GT_ERR_NCCL: return "NCCL error";
    case GT_ERR_STATE: return "call out of order";
    default: return "unknown status";
    }
}

const char *gt_last_error(const gt_ctx *ctx) { return ctx ? ctx->err : ""; }

int gt_create(const gt_config *cfg, gt_ctx **out)
{
    if (!cfg || !out || cfg->struct_size < sizeof(gt_config) || cfg->max_nodes == 0) return GT_ERR_INVALID;
    if (cfg->k1_policy > GT_K1_FARTHEST) return GT_ERR_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
        cudaGetLastError();
        return GT_ERR_NO_DEVICE;
    }
    if (cfg->device < 0 || cfg->device >= ndev) return GT_ERR_INVALID;
    gt_ctx *c = new (std::nothrow) gt_ctx();
    if (!c) return GT_ERR_NOMEM;
    c->err[0] = 0;
    c->device = cfg->device;
    c->cap = cfg->max_nodes;
    c->first_node = cfg->first_node;
    c->k1_policy = cfg->k1_policy;
    c->flags = cfg->flags;
    gt_lut_preset(GT_LUT_DEFAULT, c->lut);
    int rc = [&]() -> int {
        GT_CUDA(c, cudaSetDevice(c->device));
        cudaDeviceProp prop;
        GT_CUDA(c, cudaGetDeviceProperties(&prop, c->device));
        c->num_sms = prop.multiProcessorCount;
        if (cfg->stream) {
            c->stream = (cudaStream_t)cfg->stream;
        } else {
            GT_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
            c->own_stream = true;
        }
        const size_t N = c->cap;
        GT_CUDA(c, cudaMalloc(&c->d_levels, N * 256));
        GT_CUDA(c, cudaMalloc(&c->d_cost, N * 256));
        GT_CUDA(c, cudaMalloc(&c->d_ngpus, N));
        GT_CUDA(c, cudaMalloc(&c->d_access, N * 16));
        GT_CUDA(c, cudaMalloc(&c->d_sibling, N * 32));
        GT_CUDA(c, cudaMalloc(&c->d_free, N * 2));
        GT_CUDA(c, cudaMalloc(&c->d_best, N * 64));
        GT_CUDA(c, cudaMalloc(&c->d_bin_cnt, 3 * gt::kCtlWords * sizeof(uint32_t)));
        GT_CUDA(c, cudaMemset(c->d_bin_cnt, 0, 3 * gt::kCtlWords * sizeof(uint32_t)));
        GT_CUDA(c, cudaMalloc(&c->d_bin_free, 2 * (size_t)gt::kBins * N * sizeof(uint16_t)));   // two sets: K0 of the next pass
        GT_CUDA(c, cudaMalloc(&c->d_bin_list, 2 * (size_t)gt::kBins * N * sizeof(uint32_t)));   // fills one while K1 reads the other
        GT_CUDA(c, cudaMalloc(&c->d_place, 128 * sizeof(unsigned long long)));
        GT_CUDA(c, cudaMemset(c->d_place, 0xFF, 128 * sizeof(unsigned long long)));
        c->place_cur = c->d_place + 32;
        GT_CUDA(c, cudaMalloc(&c->d_ids, N * 4));
        GT_CUDA(c, cudaMalloc(&c->d_masks, N * 2));
        GT_CUDA(c, cudaMalloc(&c->d_peak_out, 64));
        GT_CUDA(c, cudaMalloc(&c->d_xflags, 4 * sizeof(unsigned long long)));
        GT_CUDA(c, cudaMemset(c->d_xflags, 0, 4 * sizeof(unsigned long long)));
        GT_CUDA(c, cudaMalloc(&c->d_scores, (N + 15) / 16 * 16));
        GT_CUDA(c, cudaHostAlloc(&c->h_scores, (N + 15) / 16 * 16, cudaHostAllocMapped));
        GT_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_scores_map, c->h_scores, 0));
        GT_CUDA(c, cudaHostAlloc(&c->h_place_map, 17 * sizeof(unsigned long long), cudaHostAllocMapped));
        c->h_place_map[16] = 0;
        GT_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_place_map, c->h_place_map, 0));
#ifdef GT_TRACE
        GT_CUDA(c, cudaMalloc(&c->d_trace, (size_t)c->num_sms * 8 * 4 * gt::kTraceSlots * 8));
        GT_CUDA(c, cudaMemset(c->d_trace, 0, (size_t)c->num_sms * 8 * 4 * gt::kTraceSlots * 8));
#endif
        GT_CUDA(c, cudaMemset(c->d_ngpus, 0, N));
        GT_CUDA(c, cudaMemset(c->d_free, 0, N * 2));
        GT_CUDA(c, cudaMemset(c->d_levels, 0, N * 256));
        GT_CUDA(c, cudaMemset(c->d_cost, 0, N * 256));
        GT_CUDA(c, cudaMemset(c->d_access, 0, N * 16));
        GT_CUDA(c, cudaMemset(c->d_sibling, 0, N * 32));
        GT_CUDA(c, cudaHostAlloc(&c->h_free, N * 2, cudaHostAllocMapped));
        GT_CUDA(c, cudaHostGetDevicePointer((void **)&c->d_free_map, c->h_free, 0));
        GT_CUDA(c, cudaHostAlloc(&c->h_ngpus, N, cudaHostAllocDefault));
        GT_CUDA(c, cudaHostAlloc(&c->h_ids, N * 4, cudaHostAllocDefault));
        GT_CUDA(c, cudaHostAlloc(&c->h_masks, N * 2, cudaHostAllocDefault));
        GT_CUDA(c, cudaHostAlloc(&c->h_place, 17 * sizeof(unsigned long long), cudaHostAllocDefault));
        c->h_place[16] = 0;
        GT_CUDA(c, cudaHostAlloc(&c->h_row, 64, cudaHostAllocDefault));
        GT_CUDA(c, cudaEventCreateWithFlags(&c->ev_h2d, cudaEventDisableTiming));
        memset(c->h_free, 0, N * 2);
        memset(c->h_ngpus, 0, N);
        c->dirty_flag.assign(N, 0);
        c->node_class.assign(N, 0);
        return GT_OK;
    }();
    if (rc != GT_OK) {
        fprintf(stderr, "gt_create: %s\n", c->err);
        gt_destroy(c);
        return rc;
    }
    *out = c;
    return GT_OK;
}

void gt_destroy(gt_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm && !c->comm_borrowed) {
        NcclApi *api = nccl_api();
        if (api) api->CommDestroy(c->comm);
    }
    cudaFree(c->d_levels); cudaFree(c->d_cost); cudaFree(c->d_ngpus); cudaFree(c->d_access); cudaFree(c->d_sibling);
    cudaFree(c->d_free); cudaFree(c->d_best); cudaFree(c->d_bin_cnt); cudaFree(c->d_bin_list); cudaFree(c->d_bin_free); cudaFree(c->d_memo); cudaFree(c->d_row0); cudaFree(c->d_node_class); cudaFree(c->d_place);
    cudaFree(c->d_tree); cudaFree(c->d_gaia_out);
    for (int q = 0; q < gt::kMaxPeers; q++)
        if (c->peer_inbox[q] && c->peer_inbox[q] != c->d_inbox) cudaIpcCloseMemHandle(c->peer_inbox[q]);
    cudaFree(c->d_inbox); cudaFree(c->d_xflags);
    cudaFree(c->d_ids); cudaFree(c->d_masks); cudaFree(c->d_peak_out); cudaFree(c->d_scores); cudaFreeHost(c->h_scores); cudaFreeHost(c->h_place_map);
    cudaFreeHost(c->h_free); cudaFreeHost(c->h_ngpus); cudaFreeHost(c->h_ids); cudaFreeHost(c->h_masks);
    cudaFreeHost(c->h_place); cudaFreeHost(c->h_row);
    for (auto &g : c->step_exec)
        if (g) cudaGraphExecDestroy(g);
    if (c->comm_stream) {
        cudaStreamSynchronize(c->comm_stream);
        if (c->own_comm_stream) cudaStreamDestroy(c->comm_stream);
    }
    if (c->ev_h2d) cudaEventDestroy(c->ev_h2d);
    if (c->ev_pass) cudaEventDestroy(c->ev_pass);
    for (auto &e : c->ev_comm)
        if (e) cudaEventDestroy(e);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int gt_set_cost_lut(gt_ctx *c, const uint8_t lut[16])
{
    if (!c || !lut) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    memcpy(c->lut, lut, 16);
    c->lut_version++;
    c->best_valid = c->place_valid = false;
    int rc = apply_lut(c, 0, c->n_nodes);
    if (rc != GT_OK) return rc;
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    return GT_OK;
}

int gt_load_nodes(gt_ctx *c, uint32_t first, uint32_t count, const uint8_t *levels, const uint8_t *n_gpus,
                  const uint8_t *access_cost, const uint16_t *sibling_mask, const uint16_t *free_mask)
{
    if (!c || !levels || !n_gpus || !free_mask) return GT_ERR_INVALID;
    if ((uint64_t)first + count > c->cap) return fail(c, GT_ERR_RANGE, "gt_load_nodes: first+count exceeds max_nodes%s");
    if (!count) return GT_OK;
    for (uint32_t i = 0; i < count; i++)
        if (n_gpus[i] > GT_MAX_GPUS) return fail(c, GT_ERR_INVALID, "gt_load_nodes: n_gpus > 16%s");
    GT_CUDA(c, cudaSetDevice(c->device));
    {
        int rcw = mirror_writable(c);
        if (rcw != GT_OK) return rcw;
    }
    GT_CUDA(c, cudaMemcpyAsync(c->d_levels + (size_t)first * 256, levels, (size_t)count * 256, cudaMemcpyHostToDevice, c->stream));
    memcpy(c->h_ngpus + first, n_gpus, count);
    memcpy(c->h_free + first, free_mask, (size_t)count * 2);
    {   // topology classes: nodes with the same level matrix and GPU count share one memo table
        const bool gaia = c->k1_policy == GT_K1_GAIA_SINGULAR;
        std::string key(gaia ? 257 + 16 + 32 : 257, '\0');
        for (uint32_t i = 0; i < count; i++) {
            memcpy(&key[0], levels + (size_t)i * 256, 256);
            key[256] = (char)n_gpus[i];
            if (gaia) {   // the request-1 policy reads per-GPU access costs and sibling sets: part of the class
                if (access_cost) memcpy(&key[257], access_cost + (size_t)i * 16, 16);
                if (sibling_mask) memcpy(&key[273], sibling_mask + (size_t)i * 16, 32);
            }
            auto it = c->class_of_tile.find(key);
            if (it == c->class_of_tile.end()) {
                it = c->class_of_tile.emplace(key, (uint32_t)c->class_tile.size()).first;
                c->class_tile.push_back(key);
            }
            c->node_class[first + i] = it->second;
        }
        c->classes_uploaded = false;
    }
    c->k0_inputs_pending = true;
    GT_CUDA(c, cudaMemcpyAsync(c->d_ngpus + first, c->h_ngpus + first, count, cudaMemcpyHostToDevice, c->stream));
    GT_CUDA(c, cudaMemcpyAsync(c->d_free + first, c->h_free + first, (size_t)count * 2, cudaMemcpyHostToDevice, c->stream));
    if (access_cost)
        GT_CUDA(c, cudaMemcpyAsync(c->d_access + (size_t)first * 16, access_cost, (size_t)count * 16, cudaMemcpyHostToDevice, c->stream));
    if (sibling_mask)
        GT_CUDA(c, cudaMemcpyAsync(c->d_sibling + (size_t)first * 16, sibling_mask, (size_t)count * 32, cudaMemcpyHostToDevice, c->stream));
    c->n_nodes = std::max(c->n_nodes, first + count);
    int rc = apply_lut(c, first, count);
    if (rc != GT_OK) return rc;
    GT_CUDA(c, cudaStreamSynchronize(c->stream));   // caller buffers are free to change after return
    c->best_valid = c->place_valid = false;
    return GT_OK;
}

int gt_load_trees(gt_ctx *c, uint32_t first, uint32_t count, const gt_tree *trees, const uint16_t *free_mask)
{
    if (!c || !trees || !free_mask) return GT_ERR_INVALID;
    if ((uint64_t)first + count > c->cap) return fail(c, GT_ERR_RANGE, "gt_load_trees: first+count exceeds max_nodes%s");
    if (!count) return GT_OK;
    std::vector<uint8_t> levels((size_t)count * 256), ng(count), access((size_t)count * 16);
    std::vector<uint16_t> sibling((size_t)count * 16);
    std::vector<gt::TreeRow> rows(count);
    for (uint32_t i = 0; i < count; i++) {
        if (i && memcmp(trees + i, trees + i - 1, sizeof(gt_tree)) == 0) {   // fleets repeat a few machine types
            memcpy(&levels[(size_t)i * 256], &levels[(size_t)(i - 1) * 256], 256);
            ng[i] = ng[i - 1];
            memcpy(&access[(size_t)i * 16], &access[(size_t)(i - 1) * 16], 16);
            memcpy(&sibling[(size_t)i * 16], &sibling[(size_t)(i - 1) * 16], 32);
            rows[i] = rows[i - 1];
            continue;
        }
        if (gt::tree_derive(trees + i, &levels[(size_t)i * 256], &ng[i], &access[(size_t)i * 16], &sibling[(size_t)i * 16],
                            &rows[i]) != GT_OK)
            return fail(c, GT_ERR_INVALID, "gt_load_trees: malformed tree%s");
    }
    GT_CUDA(c, cudaSetDevice(c->device));
    if (!c->d_tree) {
        GT_CUDA(c, cudaMalloc(&c->d_tree, (size_t)c->cap * sizeof(gt::TreeRow)));
        GT_CUDA(c, cudaMemsetAsync(c->d_tree, 0, (size_t)c->cap * sizeof(gt::TreeRow), c->stream));
    }
    GT_CUDA(c, cudaMemcpyAsync(c->d_tree + first, rows.data(), (size_t)count * sizeof(gt::TreeRow), cudaMemcpyHostToDevice, c->stream));
    return gt_load_nodes(c, first, count, levels.data(), ng.data(), access.data(), sibling.data(), free_mask);   // syncs the stream
}

int gt_gaia_select(gt_ctx *c, uint32_t m, uint32_t *out_best)
{
    if (!c || !out_best || m < 1 || m > 16) return GT_ERR_INVALID;
    if (m >= 2 && !c->d_tree) return fail(c, GT_ERR_STATE, "gt_gaia_select: no tree loaded (gt_load_trees)%s");
    if (!c->n_nodes) return GT_OK;
    GT_CUDA(c, cudaSetDevice(c->device));
    int rc = flush_free(c, !c->free_uploaded);   // device masks = host mirror; the dirty list stays for the next scoring
    if (rc != GT_OK) return rc;
    if (!c->d_gaia_out) GT_CUDA(c, cudaMalloc(&c->d_gaia_out, (size_t)c->cap * 4));
    gt::k_gaia_select<<<(c->n_nodes + gt::kCta - 1) / gt::kCta, gt::kCta, 0, c->stream>>>(
        c->d_tree, c->d_ngpus, c->d_free, c->d_access, c->d_sibling, c->d_cost, c->n_nodes, m, c->d_gaia_out);
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    GT_CUDA(c, cudaMemcpyAsync(out_best, c->d_gaia_out, (size_t)c->n_nodes * 4, cudaMemcpyDeviceToHost, c->stream));
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    return GT_OK;
}

int gt_update_free(gt_ctx *c, const uint32_t *node_ids, const uint16_t *masks, uint32_t count)
{
    if (!c || !masks) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    int rcw = mirror_writable(c);
    if (rcw != GT_OK) return rcw;
    if (!node_ids) {
        if (count > c->n_nodes) return fail(c, GT_ERR_RANGE, "gt_update_free: count exceeds loaded nodes%s");
        memcpy(c->h_free, masks, (size_t)count * 2);
        if (count == c->n_nodes) {       // whole shard: one bulk copy, every node rescored by the next gt_score_all
            c->free_uploaded = false;
            clear_dirty(c);
            c->best_valid = c->place_valid = false;
            return GT_OK;
        }
        for (uint32_t i = 0; i < count; i++) mark_dirty(c, i);
        return GT_OK;
    }
    for (uint32_t i = 0; i < count; i++)
        if (node_ids[i] >= c->n_nodes) return fail(c, GT_ERR_RANGE, "gt_update_free: node id out of range%s");
    for (uint32_t i = 0; i < count; i++) {
        c->h_free[node_ids[i]] = masks[i];
        mark_dirty(c, node_ids[i]);
    }
    return GT_OK;
}

int gt_commit(gt_ctx *c, const uint32_t *nodes, const uint16_t *masks, uint32_t count)
{
    if (!c || !nodes || !masks) return GT_ERR_INVALID;
    for (uint32_t i = 0; i < count; i++) {
        if (nodes[i] >= c->n_nodes) return fail(c, GT_ERR_RANGE, "gt_commit: node id out of range%s");
        if (masks[i] & ~((1u << c->h_ngpus[nodes[i]]) - 1u)) return fail(c, GT_ERR_INVALID, "gt_commit: GPU index beyond the node's GPU count%s");
    }
    int rcw = mirror_writable(c);
    if (rcw != GT_OK) return rcw;
    // entries are checked against the masks as the earlier entries of this batch left them: the same GPU cannot be
    // handed out twice by one call; on a conflict the batch is undone
    for (uint32_t i = 0; i < count; i++) {
        if ((c->h_free[nodes[i]] & masks[i]) != masks[i]) {
            for (uint32_t j = 0; j < i; j++) c->h_free[nodes[j]] |= masks[j];
            return fail(c, GT_ERR_STATE, "gt_commit: GPU already in use%s");
        }
        c->h_free[nodes[i]] &= (uint16_t)~masks[i];
    }
    for (uint32_t i = 0; i < count; i++) mark_dirty(c, nodes[i]);
    return GT_OK;
}

int gt_release(gt_ctx *c, const uint32_t *nodes, const uint16_t *masks, uint32_t count)
{
    if (!c || !nodes || !masks) return GT_ERR_INVALID;
    for (uint32_t i = 0; i < count; i++) {
        if (nodes[i] >= c->n_nodes) return fail(c, GT_ERR_RANGE, "gt_release: node id out of range%s");
        if (masks[i] & ~((1u << c->h_ngpus[nodes[i]]) - 1u)) return fail(c, GT_ERR_INVALID, "gt_release: GPU index beyond the node's GPU count%s");
    }
    int rcw = mirror_writable(c);
    if (rcw != GT_OK) return rcw;
    for (uint32_t i = 0; i < count; i++) {
        c->h_free[nodes[i]] |= masks[i];
        mark_dirty(c, nodes[i]);
    }
    return GT_OK;
}

int gt_score_all_async(gt_ctx *c)
{
    if (!c) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    // a whole-shard mask update needs no upload of its own: K0 reads the pinned mirror in place and refreshes the
    // device copy as it bins (plain K0 -> K1 passes only; GT_NO_ZEROCOPY: always upload first, profiling aid)
    static const bool no_zero_copy = getenv("GT_NO_ZEROCOPY") != nullptr;
    const bool plain = c->n_nodes && !(c->flags & (GT_FLAG_GRAPH | GT_FLAG_MEMO_PASS));
    if (plain && !no_zero_copy && (!c->free_uploaded || !c->dirty.empty())) {
        c->masks_in_place = true;
    } else {
        int rc = flush_free(c, !c->free_uploaded || !c->dirty.empty());
        if (rc != GT_OK) return rc;
    }
    clear_dirty(c);
    int rc = enqueue_scoring(c, nullptr, c->n_nodes);
    c->masks_in_place = false;
    return rc;
}

// a scoring loop without a trip through the caller's language per pass: passes enqueued round-robin over the contexts
int gt_score_all_async_many(gt_ctx **ctxs, uint32_t n_ctx, uint32_t passes)
{
    if (!ctxs || !n_ctx) return GT_ERR_INVALID;
    for (uint32_t i = 0; i < passes; i++) {
        int rc = gt_score_all_async(ctxs[i % n_ctx]);
        if (rc != GT_OK) return rc;
    }
    return GT_OK;
}

// the same loop between two CUDA events of the caller (cudaEvent_t, may be NULL), recorded on the first context's stream
// right before the first launch and right after the last: a timed region without the caller's own call overhead
// between the start event and the first kernel
int gt_score_all_async_many_timed(gt_ctx **ctxs, uint32_t n_ctx, uint32_t passes, void *ev_start, void *ev_stop)
{
    if (!ctxs || !n_ctx || !ctxs[0]) return GT_ERR_INVALID;
    gt_ctx *c0 = ctxs[0];
    GT_CUDA(c0, cudaSetDevice(c0->device));
    if (ev_start) GT_CUDA(c0, cudaEventRecord(static_cast<cudaEvent_t>(ev_start), c0->stream));
    int rc = gt_score_all_async_many(ctxs, n_ctx, passes);
    if (rc != GT_OK) return rc;
    if (ev_stop) GT_CUDA(c0, cudaEventRecord(static_cast<cudaEvent_t>(ev_stop), c0->stream));
    return GT_OK;
}

int gt_sync(gt_ctx *c)
{
    if (!c) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    if (c->comm_stream) GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    return GT_OK;
}

int gt_score_all(gt_ctx *c, uint32_t *out_best)
{
    int rc = gt_score_all_async(c);
    if (rc != GT_OK) return rc;
    if (out_best && c->n_nodes)
        GT_CUDA(c, cudaMemcpyAsync(out_best, c->d_best, (size_t)c->n_nodes * 64, cudaMemcpyDeviceToHost, c->stream));
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    return GT_OK;
}

int gt_score_dirty(gt_ctx *c)
{
    if (!c) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    if (!c->best_valid || !c->free_uploaded) return gt_score_all(c, nullptr);
    const uint32_t n = (uint32_t)c->dirty.size();
    int rc = flush_free(c, false);           // fills d_ids with the dirty list
    if (rc != GT_OK) return rc;
    clear_dirty(c);
    rc = enqueue_scoring(c, c->d_ids, n);
    if (rc != GT_OK) return rc;
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    return GT_OK;
}

int gt_best_node(gt_ctx *c, const uint8_t *k, uint32_t nreq, uint64_t *out_place)
{
    if (!c || (nreq && (!k || !out_place))) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    if (!c->best_valid) return fail(c, GT_ERR_STATE, "gt_best_node: score first%s");
    if (!c->dirty.empty() || !c->free_uploaded) return fail(c, GT_ERR_STATE, "gt_best_node: free masks changed, rescore first%s");
    if (!c->place_valid) {
        int rc = enqueue_scoring(c, nullptr, 0);
        if (rc != GT_OK) return rc;
    }
    if (lazy_exchange(c)) {
        int rc = x_consume(c);
        if (rc != GT_OK) return rc;
    } else if (c->comm && c->comm_last >= 0) {
        GT_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_comm[c->comm_last], 0));
    }
    GT_CUDA(c, cudaMemcpyAsync(c->h_place, cluster_keys(c), 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    if (c->comm && c->p2p)
        GT_CUDA(c, cudaMemcpyAsync(c->h_place + 16, c->d_xflags + 3, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    if (c->comm && c->p2p && c->h_place[16] != 0) return exchange_failed(c);
    for (uint32_t i = 0; i < nreq; i++)
        out_place[i] = (k[i] >= 1 && k[i] <= 16) ? (uint64_t)c->h_place[k[i] - 1] : GT_NO_PLACE;
    return GT_OK;
}

int gt_allocate(gt_ctx *c, uint32_t node, uint32_t k, uint16_t *out_mask, uint16_t *out_cost)
{
    if (!c || !out_mask || !out_cost || k < 1 || k > 16) return GT_ERR_INVALID;
    if (node >= c->n_nodes) return fail(c, GT_ERR_RANGE, "gt_allocate: node id out of range%s");
    if (!c->best_valid) return fail(c, GT_ERR_STATE, "gt_allocate: score first%s");
    if (c->dirty_flag[node] || !c->free_uploaded) return fail(c, GT_ERR_STATE, "gt_allocate: node changed, rescore first%s");
    GT_CUDA(c, cudaSetDevice(c->device));
    GT_CUDA(c, cudaMemcpyAsync(c->h_row, c->d_best + (size_t)node * 16, 64, cudaMemcpyDeviceToHost, c->stream));
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    *out_mask = (uint16_t)GT_BEST_MASK(c->h_row[k - 1]);
    *out_cost = (uint16_t)GT_BEST_COST(c->h_row[k - 1]);
    return GT_OK;
}

void *gt_stream(gt_ctx *c) { return c ? (void *)c->stream : nullptr; }
const uint32_t *gt_device_best(gt_ctx *c) { return c ? c->d_best : nullptr; }
uint32_t gt_num_nodes(const gt_ctx *c) { return c ? c->n_nodes : 0; }
uint64_t gt_launch_count(const gt_ctx *c) { return c ? c->launches : 0; }

int gt_count_subsets(gt_ctx *c, uint64_t *out_subsets, uint64_t *out_pair_adds)
{
    if (!c) return GT_ERR_INVALID;
    uint64_t s = 0, a = 0;
    for (uint32_t i = 0; i < c->n_nodes; i++) {
        const uint32_t ng = c->h_ngpus[i];
        const uint32_t f = c->h_free[i] & ((1u << ng) - 1u);
        const uint32_t p = (uint32_t)__builtin_popcount(f);
        s += (1ull << p) - 1;
        if (p >= 2) a += (uint64_t)(p * (p - 1) / 2) << (p - 2);
    }
    if (out_subsets) *out_subsets = s;
    if (out_pair_adds) *out_pair_adds = a;
    return GT_OK;
}

int gt_time_kernel(gt_ctx *c, const char *which, uint32_t iters, float *out_ms)
{
    if (!c || !which || !out_ms || !iters) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    // a local measurement: no collective may be issued from here (other ranks are not in this call)
    struct CommOff {
        gt_ctx *c; void *saved;
        explicit CommOff(gt_ctx *x) : c(x), saved(x->comm) { x->comm = nullptr; }
        ~CommOff() { c->comm = saved; c->place_valid = false; }
    } comm_off(c);
    if (comm_off.saved && c->p2p && c->x_reduced < c->x_seq)    // these passes reuse the key buffers of the last two passes
        return fail(c, GT_ERR_STATE, "gt_time_kernel: the keys of the last pass were not exchanged yet: gt_best_node on every rank first%s");
    if (c->comm_stream) GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));   // no exchange may still be reading the key buffers
    int rc = flush_free(c, !c->free_uploaded || !c->dirty.empty());
    if (rc != GT_OK) return rc;
    clear_dirty(c);
    cudaEvent_t e0, e1;
    GT_CUDA(c, cudaEventCreate(&e0));
    GT_CUDA(c, cudaEventCreate(&e1));
    const std::string w(which);
    const uint32_t peak_iters = 2048;
    auto one = [&]() -> int {
        if (w == "bin") {   // K0 alone: its counters have to be cleared by hand between launches
            GT_CUDA(c, cudaMemsetAsync(cur_ctl(c), 0, gt::kCtlWords * sizeof(uint32_t), c->stream));
            return launch_bin(c, nullptr, c->n_nodes, bin_place(c, 0));
        }
        if (w == "score") return launch_score(c, c->n_nodes, bin_place(c, 0), true);
        if (w == "argmin") return launch_argmin(c, bin_place(c, c->bin_parity));
        if (w == "step") return enqueue_scoring(c, nullptr, c->n_nodes);
        if (w == "step_nograph") {
            int r2 = launch_bin(c, nullptr, c->n_nodes, bin_place(c, 0));
            return r2 != GT_OK ? r2 : launch_score(c, c->n_nodes, bin_place(c, 0));
        }
        if (w == "noop") { gt::k_noop<<<1, 32, 0, c->stream>>>(c->d_peak_out); return GT_OK; }
        if (w == "noop_grid") { gt::k_noop<<<c->num_sms * 4, 128, 0, c->stream>>>(c->d_peak_out); return GT_OK; }
        if (w == "noop_smem") { gt::k_noop_smem<<<c->num_sms * 4, 128, 0, c->stream>>>(c->d_peak_out); return GT_OK; }
        if (w == "int32_peak") {
            gt::k_int32_peak<<<c->num_sms * 8, gt::kCta, 0, c->stream>>>(c->d_peak_out, peak_iters, 12345u);
            c->launches++;
            return GT_OK;
        }
        return GT_ERR_INVALID;
    };
    if (w == "score" || w == "argmin") {   // bins / best table must exist
        if ((rc = launch_bin(c, nullptr, c->n_nodes, bin_place(c, 0))) != GT_OK) return rc;
        if ((rc = launch_score(c, c->n_nodes, bin_place(c, 0), true)) != GT_OK) return rc;
    }
    if ((rc = one()) != GT_OK) return rc;   // warm-up
    GT_CUDA(c, cudaEventRecord(e0, c->stream));
    for (uint32_t i = 0; i < iters; i++)
        if ((rc = one()) != GT_OK) return rc;
    GT_CUDA(c, cudaEventRecord(e1, c->stream));
    GT_CUDA(c, cudaEventSynchronize(e1));
    float ms = 0;
    GT_CUDA(c, cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    // leave the pipeline consistent: control block zero, best table and placement keys current
    GT_CUDA(c, cudaMemsetAsync(cur_ctl(c), 0, gt::kCtlWords * sizeof(uint32_t), c->stream));
    if ((rc = enqueue_scoring(c, nullptr, c->n_nodes)) != GT_OK) return rc;
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    *out_ms = ms / (float)iters;
    return GT_OK;
}

// Full passes back to back (programmatic launches intact, as in a scoring loop) while the contexts take turns: with
// more replicas than fit the L2 every launch finds its tables in DRAM.  K1 stamps the global timer when its first CTA
// gets past the dependency wait and when its last CTA finishes: out_us = {pass - K1 (K0 and the two hand-overs), K1
// in place, pass}, averages over `iters` passes.
int gt_time_pass_rotated(gt_ctx **ctxs, uint32_t n_ctx, uint32_t iters, float out_us[3])
{
    if (!ctxs || !n_ctx || !iters || !out_us) return GT_ERR_INVALID;
    gt_ctx *c0 = ctxs[0];
    if (!c0) return GT_ERR_INVALID;
    for (uint32_t i = 0; i < n_ctx; i++)
        if (!ctxs[i] || ctxs[i]->device != c0->device || ctxs[i]->stream != c0->stream || !ctxs[i]->n_nodes)
            return fail(c0, GT_ERR_INVALID, "gt_time_pass_rotated: contexts must share device and stream%s");
    GT_CUDA(c0, cudaSetDevice(c0->device));
    iters = std::min<uint32_t>(iters, 4096);
    int rc = GT_OK;
    for (uint32_t i = 0; i < n_ctx && rc == GT_OK; i++) {
        gt_ctx *c = ctxs[i];
        if (lazy_exchange(c) && c->x_reduced < c->x_seq)
            return fail(c0, GT_ERR_STATE, "gt_time_pass_rotated: keys of the last pass not exchanged yet: gt_best_node on every rank first%s");
        if (c->comm_stream) GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));
        rc = flush_free(c, !c->free_uploaded || !c->dirty.empty());
        clear_dirty(c);
    }
    if (rc != GT_OK) return rc;
    struct Buf { unsigned long long *p = nullptr; ~Buf() { cudaFree(p); } } stamps;
    GT_CUDA(c0, cudaMalloc(&stamps.p, (size_t)iters * 2 * sizeof(unsigned long long)));
    GT_CUDA(c0, cudaMemsetAsync(stamps.p, 0, (size_t)iters * 2 * sizeof(unsigned long long), c0->stream));
    for (uint32_t it = 0; it < iters + n_ctx && rc == GT_OK; it++) {   // first round over the contexts: warm-up
        gt_ctx *c = ctxs[it % n_ctx];
        void *saved = c->comm;
        c->comm = nullptr;   // a local measurement: other ranks are not in this call
        rc = launch_bin(c, nullptr, c->n_nodes, bin_place(c, 0));
        c->ktime_next = it >= n_ctx ? stamps.p + 2 * (size_t)(it - n_ctx) : nullptr;
        if (rc == GT_OK) rc = launch_score(c, c->n_nodes, bin_place(c, 0));
        c->ktime_next = nullptr;
        c->comm = saved;
        c->place_valid = false;
    }
    GT_CUDA(c0, cudaStreamSynchronize(c0->stream));
    if (rc != GT_OK) return rc;
    std::vector<unsigned long long> h((size_t)iters * 2);
    GT_CUDA(c0, cudaMemcpy(h.data(), stamps.p, h.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    double k1 = 0;
    for (uint32_t it = 0; it < iters; it++) k1 += (double)(h[2 * it + 1] - h[2 * it]);
    const double pass = iters > 1 ? (double)(h[2 * (iters - 1) + 1] - h[1]) / (iters - 1) : (double)(h[1] - h[0]);
    out_us[1] = (float)(k1 / iters * 1e-3);
    out_us[2] = (float)(pass * 1e-3);
    out_us[0] = out_us[2] - out_us[1];
    return GT_OK;
}

// This rank's own 16 placement keys, recomputed from its best table (K2) without any exchange: what a checker
// min-combines over ranks to verify the exchanged keys.
int gt_best_node_local(gt_ctx *c, uint64_t out_place[16])
{
    if (!c || !out_place) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    if (!c->best_valid || !c->dirty.empty() || !c->free_uploaded) return fail(c, GT_ERR_STATE, "gt_best_node_local: score first%s");
    unsigned long long *scratch = c->d_place + 16;     // [1]: not used by any pass
    GT_CUDA(c, cudaMemsetAsync(scratch, 0xFF, 16 * sizeof(unsigned long long), c->stream));
    const size_t quarters = (size_t)c->n_nodes * 4;
    const uint32_t grid = (uint32_t)std::min<size_t>((quarters + gt::kCta * 8 - 1) / (gt::kCta * 8), (size_t)c->num_sms * 4);
    if (c->n_nodes) {
        gt::k_argmin<<<std::max(grid, 1u), gt::kCta, 0, c->stream>>>(c->d_best, c->n_nodes, c->first_node, scratch);
        c->launches++;
        GT_CUDA(c, cudaGetLastError());
    }
    GT_CUDA(c, cudaMemcpyAsync(c->h_place, scratch, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    memcpy(out_place, c->h_place, 16 * sizeof(uint64_t));
    return GT_OK;
}

}  // extern "C"

namespace {

// One communication stream per device and process: the exchanges of several contexts on a device (bench.py rotates
// eight cluster replicas) run one after the other in the order their passes were issued.  With a stream each, the
// exchange kernels of all of them sat on the SMs at once, each waiting for its own K1, and every pass paid for it
// (K1's programmatic launch is slower while other grids are resident).  The price is an ordering rule NCCL users know:
// every rank issues the passes of its contexts on a device in the same order.  GT_PRIVATE_COMM_STREAM: one per context.
int comm_stream_for(gt_ctx *c)
{
    static std::mutex mu;
    static cudaStream_t shared[64] = {};
    static const bool priv = getenv("GT_PRIVATE_COMM_STREAM") != nullptr;
    if (priv || c->device < 0 || c->device >= 64) {
        GT_CUDA(c, cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
        c->own_comm_stream = true;
        return GT_OK;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!shared[c->device]) GT_CUDA(c, cudaStreamCreateWithFlags(&shared[c->device], cudaStreamNonBlocking));
    c->comm_stream = shared[c->device];
    c->own_comm_stream = false;
    return GT_OK;
}

// Peer-memory exchange: every rank allocates an inbox, the CUDA IPC handles travel once through ncclAllGather and
// each rank maps the others' inboxes.  All ranks agree (min all-reduce of a success word) whether to use it;
// otherwise -- ranks without peer access, more than kMaxPeers ranks, GT_NO_P2P set -- the exchange stays the NCCL
// all-reduce.  Ranks must be separate processes (an IPC handle cannot be opened by the process that made it).
int setup_peer_exchange(gt_ctx *c)
{
    NcclApi *api = nccl_api();
    const int world = c->world;
    if (world < 2) return GT_OK;
    unsigned long long ok = (world <= gt::kMaxPeers && api->AllGather && getenv("GT_NO_P2P") == nullptr) ? 1 : 0;
    const size_t inbox_words = (size_t)world * 32 + (size_t)world * 2;
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof mine);
    struct Scratch {       // set-up buffers: freed on every way out
        unsigned char *handles = nullptr;
        unsigned long long *ok = nullptr;
        ~Scratch() { cudaFree(handles); cudaFree(ok); }
    } tmp;
    unsigned char *&d_handles = tmp.handles;
    unsigned long long *&d_ok = tmp.ok;
    std::vector<cudaIpcMemHandle_t> all((size_t)world);
    GT_CUDA(c, cudaMalloc(&d_handles, (size_t)(world + 1) * sizeof(cudaIpcMemHandle_t)));
    GT_CUDA(c, cudaMalloc(&d_ok, sizeof(unsigned long long)));
    if (ok) {
        if (cudaMalloc(&c->d_inbox, inbox_words * 8) != cudaSuccess || cudaMemset(c->d_inbox, 0, inbox_words * 8) != cudaSuccess ||
            cudaIpcGetMemHandle(&mine, c->d_inbox) != cudaSuccess) {
            cudaGetLastError();
            ok = 0;
        }
    }
    // handles to everybody (a zero handle from a rank that could not make one)
    GT_CUDA(c, cudaMemcpy(d_handles + (size_t)world * sizeof mine, &mine, sizeof mine, cudaMemcpyHostToDevice));
    int e = api->AllGather ? api->AllGather(d_handles + (size_t)world * sizeof mine, d_handles, sizeof mine, kNcclUint8, c->comm, c->comm_stream) : 0;
    if (e != 0) return fail(c, GT_ERR_NCCL, "ncclAllGather: %s", api->GetErrorString ? api->GetErrorString(e) : "?");
    GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    GT_CUDA(c, cudaMemcpy(all.data(), d_handles, (size_t)world * sizeof mine, cudaMemcpyDeviceToHost));
    if (ok) {
        for (int q = 0; q < world && ok; q++) {
            if (q == c->rank) { c->peer_inbox[q] = c->d_inbox; continue; }
            void *ptr = nullptr;
            if (cudaIpcOpenMemHandle(&ptr, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                ok = 0;
            } else {
                c->peer_inbox[q] = (unsigned long long *)ptr;
            }
        }
    }
    // unanimous or not at all
    GT_CUDA(c, cudaMemcpy(d_ok, &ok, sizeof ok, cudaMemcpyHostToDevice));
    e = api->AllReduce(d_ok, d_ok, 1, kNcclUint64, kNcclMin, c->comm, c->comm_stream);
    if (e != 0) return fail(c, GT_ERR_NCCL, "ncclAllReduce: %s", api->GetErrorString ? api->GetErrorString(e) : "?");
    GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    GT_CUDA(c, cudaMemcpy(&ok, d_ok, sizeof ok, cudaMemcpyDeviceToHost));
    c->p2p = ok != 0;
    return GT_OK;
}

}  // namespace

extern "C" {

int gt_comm_mode(const gt_ctx *c) { return !c || !c->comm ? 0 : (c->p2p ? 2 : 1); }

int gt_comm_unique_id(uint8_t out_id[128])
{
    NcclApi *api = nccl_api();
    if (!api || !out_id) return GT_ERR_NCCL;
    NcclId id;
    if (api->GetUniqueId(&id) != 0) return GT_ERR_NCCL;
    memcpy(out_id, id.internal, 128);
    return GT_OK;
}

int gt_comm_init(gt_ctx *c, const uint8_t id[128], int rank, int world)
{
    if (!c || !id || world < 1 || rank < 0 || rank >= world) return GT_ERR_INVALID;
    if (c->comm) return fail(c, GT_ERR_STATE, "gt_comm_init: the context already has a communicator%s");
    NcclApi *api = nccl_api();
    if (!api) return fail(c, GT_ERR_NCCL, "libnccl.so.2 not loadable%s");
    GT_CUDA(c, cudaSetDevice(c->device));
    NcclId nid;
    memcpy(nid.internal, id, 128);
    int e = api->CommInitRank(&c->comm, world, nid, rank);
    if (e != 0) {
        c->comm = nullptr;
        return fail(c, GT_ERR_NCCL, "ncclCommInitRank: %s", api->GetErrorString ? api->GetErrorString(e) : "?");
    }
    c->rank = rank;
    c->world = world;
    int rcs = comm_stream_for(c);
    if (rcs != GT_OK) return rcs;
    GT_CUDA(c, cudaEventCreateWithFlags(&c->ev_pass, cudaEventDisableTiming));
    for (auto &ev : c->ev_comm) GT_CUDA(c, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    // one exchange now, so connections are set up before any timed or latency-critical pass
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    e = api->AllReduce(c->d_place, c->d_place, 16, kNcclUint64, kNcclMin, c->comm, c->comm_stream);
    if (e != 0) return fail(c, GT_ERR_NCCL, "ncclAllReduce (warm-up): %s", api->GetErrorString ? api->GetErrorString(e) : "?");
    GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    c->place_valid = false;
    return setup_peer_exchange(c);
}

// A second context of the same process and device joins the NCCL communicator (and the communication stream) of
// `owner` instead of creating its own -- communicator set-up is the slow part; its inbox for the peer-memory exchange
// is its own, so its passes and exchanges are independent of the owner's.  Collective: every rank shares the same
// way, in the same order.  `owner` must outlive `c`.
int gt_comm_share(gt_ctx *c, gt_ctx *owner)
{
    if (!c || !owner || c == owner) return GT_ERR_INVALID;
    if (c->comm) return fail(c, GT_ERR_STATE, "gt_comm_share: the context already has a communicator%s");
    if (!owner->comm) return fail(c, GT_ERR_STATE, "gt_comm_share: the lending context has no communicator%s");
    if (owner->device != c->device) return fail(c, GT_ERR_INVALID, "gt_comm_share: contexts on different devices%s");
    GT_CUDA(c, cudaSetDevice(c->device));
    GT_CUDA(c, cudaStreamSynchronize(owner->comm_stream));
    c->comm = owner->comm;
    c->comm_borrowed = true;
    c->rank = owner->rank;
    c->world = owner->world;
    c->comm_stream = owner->comm_stream;
    c->own_comm_stream = false;
    GT_CUDA(c, cudaEventCreateWithFlags(&c->ev_pass, cudaEventDisableTiming));
    for (auto &ev : c->ev_comm) GT_CUDA(c, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    c->place_valid = false;
    return setup_peer_exchange(c);
}

// ---------------------------------------------------------------------------------------------
// gt_schedule: the sequential sort -> bind loop, one event at a time.
namespace {

constexpr size_t kMemoMaxBytes = (size_t)1 << 30;

// GPU fills best rows for (topology class, free mask) pairs: rows [row_lo, row_hi) of the row space in which class k
// owns rows off[k] .. off[k] + 2^n_k - 1 (row = off[k] + free mask).  K0 + K1 over a pseudo-cluster with one node per row.
int memo_rows(gt_ctx *c, const std::vector<std::string> &classes, const std::vector<size_t> &off, size_t row_lo, size_t row_hi,
              uint32_t *out_rows)
{
    if (row_hi <= row_lo) return GT_OK;
    const size_t rows = row_hi - row_lo;
    gt_config cfg = {};
    cfg.struct_size = sizeof cfg;
    cfg.device = c->device;
    cfg.max_nodes = (uint32_t)rows;
    cfg.k1_policy = c->k1_policy;
    const bool gaia = c->k1_policy == GT_K1_GAIA_SINGULAR;
    std::vector<uint8_t> ac;
    std::vector<uint16_t> sb;
    gt_ctx *m = nullptr;
    int rc = gt_create(&cfg, &m);
    if (rc != GT_OK) return fail(c, rc, "gt_schedule: memo context: %s", gt_strerror(rc));
    rc = gt_set_cost_lut(m, c->lut);
    std::vector<uint8_t> lv, ng;
    std::vector<uint16_t> fr;
    const size_t batch = (size_t)1 << 16;
    for (uint32_t k = 0; k < classes.size() && rc == GT_OK; k++) {
        const uint32_t n = (uint8_t)classes[k][256];
        const size_t c_lo = std::max(off[k], row_lo), c_hi = std::min(off[k] + ((size_t)1 << n), row_hi);
        for (size_t b0 = c_lo; b0 < c_hi && rc == GT_OK; b0 += batch) {
            const size_t bn = std::min(batch, c_hi - b0);
            lv.resize(bn * 256); ng.assign(bn, (uint8_t)n); fr.resize(bn);
            for (size_t i = 0; i < bn; i++) {
                memcpy(&lv[i * 256], classes[k].data(), 256);
                fr[i] = (uint16_t)(b0 + i - off[k]);
            }
            if (gaia) {
                ac.resize(bn * 16); sb.resize(bn * 16);
                for (size_t i = 0; i < bn; i++) {
                    memcpy(&ac[i * 16], classes[k].data() + 257, 16);
                    memcpy(&sb[i * 16], classes[k].data() + 273, 32);
                }
            }
            rc = gt_load_nodes(m, (uint32_t)(b0 - row_lo), (uint32_t)bn, lv.data(), ng.data(),
                               gaia ? ac.data() : nullptr, gaia ? sb.data() : nullptr, fr.data());
        }
    }
    if (rc == GT_OK) rc = gt_score_all(m, out_rows);
    if (rc != GT_OK) fail(c, rc, "gt_schedule: memo build: %s", gt_last_error(m));
    c->launches += gt_launch_count(m);
    gt_destroy(m);
    return rc;
}

int build_memo(gt_ctx *c)
{
    const uint32_t ncls = (uint32_t)c->class_tile.size();
    if (c->memo_classes == ncls && c->memo_lut_version == c->lut_version) return GT_OK;
    size_t rows = 0;
    c->memo_off.assign(ncls, 0);
    for (uint32_t k = 0; k < ncls; k++) {
        c->memo_off[k] = rows;
        rows += (size_t)1 << (uint8_t)c->class_tile[k][256];
    }
    if (rows * 64 > kMemoMaxBytes || rows > 0x7FFFFFFFull) return GT_ERR_RANGE;
    c->memo.resize(rows * 16);
    int rc = memo_rows(c, c->class_tile, c->memo_off, 0, rows, c->memo.data());
    if (rc != GT_OK) return rc;
    c->memo_classes = ncls;
    c->memo_lut_version = c->lut_version;
    return GT_OK;
}

// per-k tournament tree over the nodes' placement keys
struct MinTree {
    uint32_t leaves = 1;
    std::vector<unsigned long long> t[16];
    void init(uint32_t n)
    {
        leaves = 1;
        while (leaves < n) leaves <<= 1;
        for (auto &v : t) v.assign((size_t)2 * leaves, ~0ull);
    }
    void build()
    {
        for (auto &v : t)
            for (uint32_t i = leaves - 1; i >= 1; i--) v[i] = std::min(v[2 * i], v[2 * i + 1]);
    }
    void set(int k, uint32_t node, unsigned long long key)
    {
        auto &v = t[k];
        uint32_t i = leaves + node;
        v[i] = key;
        for (i >>= 1; i >= 1; i >>= 1) {
            const unsigned long long m = std::min(v[2 * i], v[2 * i + 1]);
            if (v[i] == m) break;
            v[i] = m;
        }
    }
};

inline unsigned long long place_key(uint32_t b, uint32_t global_node)
{
    return b == GT_INFEASIBLE ? ~0ull
                              : ((unsigned long long)(b >> 16) << 48) | ((unsigned long long)global_node << 16) | (b & 0xFFFFu);
}

}  // namespace

// ---- gt_schedule on a sharded cluster (one rank per GPU).  The commit loop is sequential by definition: every pod
// sees the previous pod's effect on whichever shard it landed.  What shards is the GPU work -- the (class, free mask)
// table is built in `world` slices, one per GPU, and gathered -- while the host loop itself runs replicated: every
// rank gathers the cluster's node states once (class, GPU count, free mask: 8 bytes per node), replays the same
// events over the whole cluster with the same table, arrives at the same placements, and applies the ones that hit
// its own shard to its tables.  No per-event exchange; outputs are identical on every rank.
namespace {

// all-gather of `bytes` per rank through the communicator (NCCL, staged through a scratch device buffer)
int allgather_host(gt_ctx *c, const void *mine, size_t bytes, std::vector<uint8_t> &all)
{
    NcclApi *api = nccl_api();
    if (!api || !api->AllGather) return fail(c, GT_ERR_NCCL, "gt_schedule: ncclAllGather unavailable%s");
    struct Buf { uint8_t *p = nullptr; ~Buf() { cudaFree(p); } } d;
    const size_t w = (size_t)c->world;
    GT_CUDA(c, cudaMalloc(&d.p, bytes * (w + 1)));
    GT_CUDA(c, cudaMemcpy(d.p + bytes * w, mine, bytes, cudaMemcpyHostToDevice));
    int e = api->AllGather(d.p + bytes * w, d.p, bytes, kNcclUint8, c->comm, c->comm_stream);
    if (e != 0) return fail(c, GT_ERR_NCCL, "ncclAllGather: %s", api->GetErrorString ? api->GetErrorString(e) : "?");
    GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    all.resize(bytes * w);
    GT_CUDA(c, cudaMemcpy(all.data(), d.p, bytes * w, cudaMemcpyDeviceToHost));
    return GT_OK;
}

struct NodeState { uint32_t cls; uint16_t free; uint8_t ngpus, pad; };

int schedule_sharded(gt_ctx *c, const uint8_t *ev_k, const uint32_t *ev_arg, uint32_t n_events, uint32_t *out_node,
                     uint16_t *out_mask, uint64_t *out_placed)
{
    if (c->flags & GT_FLAG_NO_MEMO) return fail(c, GT_ERR_STATE, "gt_schedule: GT_FLAG_NO_MEMO is not available on a sharded context%s");
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    if (c->comm_stream) GT_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    int rc;
    if ((rc = mirror_writable(c)) != GT_OK) return rc;
    const uint32_t W = (uint32_t)c->world;
    std::vector<uint8_t> all;
    // shard sizes, first nodes, class counts
    const uint32_t hdr[4] = {c->n_nodes, c->first_node, (uint32_t)c->class_tile.size(), 0};
    if ((rc = allgather_host(c, hdr, sizeof hdr, all)) != GT_OK) return rc;
    std::vector<uint32_t> n_of(W), first_of(W), ncls_of(W), base_of(W);
    uint32_t max_n = 1, max_cls = 1, N = 0;
    for (uint32_t r = 0; r < W; r++) {
        const uint32_t *h = reinterpret_cast<const uint32_t *>(all.data()) + 4 * r;
        n_of[r] = h[0]; first_of[r] = h[1]; ncls_of[r] = h[2];
        base_of[r] = N; N += h[0];
        max_n = std::max(max_n, h[0]); max_cls = std::max(max_cls, h[2]);
    }
    // one class list for the whole cluster: every rank's classes in rank order, duplicates folded
    const size_t key_len = c->k1_policy == GT_K1_GAIA_SINGULAR ? 257 + 16 + 32 : 257;
    std::vector<uint8_t> mine((size_t)max_cls * key_len, 0);
    for (size_t k = 0; k < c->class_tile.size(); k++) memcpy(&mine[k * key_len], c->class_tile[k].data(), key_len);
    if ((rc = allgather_host(c, mine.data(), mine.size(), all)) != GT_OK) return rc;
    std::unordered_map<std::string, uint32_t> gid_of;
    std::vector<std::string> gclass;
    std::vector<uint32_t> my_map(c->class_tile.size());
    for (uint32_t r = 0; r < W; r++)
        for (uint32_t k = 0; k < ncls_of[r]; k++) {
            std::string key(reinterpret_cast<const char *>(&all[((size_t)r * max_cls + k) * key_len]), key_len);
            auto it = gid_of.find(key);
            if (it == gid_of.end()) {
                it = gid_of.emplace(key, (uint32_t)gclass.size()).first;
                gclass.push_back(key);
            }
            if (r == (uint32_t)c->rank) my_map[k] = it->second;
        }
    // the table: this rank scores its slice of the rows, everybody gets all of them
    std::vector<size_t> off(gclass.size());
    size_t rows = 0;
    for (size_t k = 0; k < gclass.size(); k++) { off[k] = rows; rows += (size_t)1 << (uint8_t)gclass[k][256]; }
    if (rows * 64 > kMemoMaxBytes) return fail(c, GT_ERR_RANGE, "gt_schedule: too many distinct topologies for the table%s");
    const size_t per = (rows + W - 1) / W, lo = std::min(rows, per * (size_t)c->rank), hi = std::min(rows, lo + per);
    std::vector<uint32_t> slice(per * 16, GT_INFEASIBLE);
    if ((rc = memo_rows(c, gclass, off, lo, hi, slice.data())) != GT_OK) return rc;
    if ((rc = allgather_host(c, slice.data(), per * 64, all)) != GT_OK) return rc;
    std::vector<uint32_t> memo(rows * 16);
    memcpy(memo.data(), all.data(), rows * 64);       // slices are contiguous in rank order; the tail of the last is padding
    // node states
    std::vector<NodeState> st(max_n);
    for (uint32_t i = 0; i < c->n_nodes; i++) st[i] = NodeState{my_map[c->node_class[i]], c->h_free[i], c->h_ngpus[i], 0};
    if ((rc = allgather_host(c, st.data(), (size_t)max_n * sizeof(NodeState), all)) != GT_OK) return rc;
    std::vector<NodeState> node(N);
    std::vector<uint32_t> gnode(N);
    for (uint32_t r = 0; r < W; r++)
        for (uint32_t i = 0; i < n_of[r]; i++) {
            node[base_of[r] + i] = reinterpret_cast<const NodeState *>(all.data())[(size_t)r * max_n + i];
            gnode[base_of[r] + i] = first_of[r] + i;
        }
    auto pos_of = [&](uint32_t g) -> uint32_t {
        for (uint32_t r = 0; r < W; r++)
            if (g >= first_of[r] && g - first_of[r] < n_of[r]) return base_of[r] + (g - first_of[r]);
        return 0xFFFFFFFFu;
    };
    const uint32_t my_lo = base_of[c->rank], my_hi = my_lo + c->n_nodes;
    // the replicated loop
    MinTree tree;
    tree.init(std::max(N, 1u));
    auto row_of = [&](uint32_t pos) -> const uint32_t * {
        const NodeState &s2 = node[pos];
        return &memo[(off[s2.cls] + (s2.free & ((1u << s2.ngpus) - 1u))) * 16];
    };
    for (uint32_t pos = 0; pos < N; pos++) {
        const uint32_t *row = row_of(pos);
        for (int k = 0; k < 16; k++) tree.t[k][tree.leaves + pos] = place_key(row[k], gnode[pos]);
    }
    tree.build();
    auto refresh = [&](uint32_t pos) {
        const uint32_t *row = row_of(pos);
        for (int k = 0; k < 16; k++) tree.set(k, pos, place_key(row[k], gnode[pos]));
        if (pos >= my_lo && pos < my_hi) {           // lands on this rank's shard: its tables follow
            c->h_free[pos - my_lo] = node[pos].free;
            mark_dirty(c, pos - my_lo);
        }
    };
    uint64_t placed = 0;
    std::vector<uint32_t> live_pos;
    std::vector<uint16_t> live_mask;
    for (uint32_t e = 0; e < n_events; e++) {
        out_node[e] = 0xFFFFFFFFu; out_mask[e] = 0;
        if (ev_k[e] == 0) {
            if (live_pos.empty()) continue;
            const size_t idx = ev_arg[e] % live_pos.size();
            const uint32_t pos = live_pos[idx]; const uint16_t m = live_mask[idx];
            node[pos].free |= m;
            out_node[e] = gnode[pos]; out_mask[e] = m;
            live_pos[idx] = live_pos.back(); live_pos.pop_back();
            live_mask[idx] = live_mask.back(); live_mask.pop_back();
            refresh(pos);
        } else if (ev_k[e] <= 16) {
            const unsigned long long key = tree.t[ev_k[e] - 1][1];
            if (key == ~0ull) continue;
            const uint32_t pos = pos_of(GT_PLACE_NODE(key)); const uint16_t m = (uint16_t)GT_PLACE_MASK(key);
            if (pos == 0xFFFFFFFFu) return fail(c, GT_ERR_STATE, "gt_schedule: overlapping or inconsistent shard ranges%s");
            node[pos].free &= (uint16_t)~m;
            out_node[e] = gnode[pos]; out_mask[e] = m;
            live_pos.push_back(pos); live_mask.push_back(m);
            placed++;
            refresh(pos);
        }
    }
    if (out_placed) *out_placed = placed;
    return GT_OK;
}

}  // namespace

int gt_schedule(gt_ctx *c, const uint8_t *ev_k, const uint32_t *ev_arg, uint32_t n_events, uint32_t *out_node,
                uint16_t *out_mask, uint64_t *out_placed)
{
    if (!c || (n_events && (!ev_k || !ev_arg || !out_node || !out_mask))) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    if (c->comm) return schedule_sharded(c, ev_k, ev_arg, n_events, out_node, out_mask, out_placed);
    uint64_t placed = 0;
    std::vector<uint32_t> live_node;
    std::vector<uint16_t> live_mask;
    int rc = (c->flags & GT_FLAG_NO_MEMO) ? GT_ERR_RANGE : build_memo(c);
    if (rc != GT_OK && rc != GT_ERR_RANGE) return rc;
    if (rc == GT_ERR_RANGE) {
        // too many distinct topologies for a memo table: every event re-scores the touched node on the GPU
        if ((rc = gt_score_dirty(c)) != GT_OK) return rc;
        for (uint32_t e = 0; e < n_events; e++) {
            out_node[e] = 0xFFFFFFFFu; out_mask[e] = 0;
            if (ev_k[e] == 0) {
                if (live_node.empty()) continue;
                const size_t idx = ev_arg[e] % live_node.size();
                const uint32_t nd = live_node[idx]; const uint16_t m = live_mask[idx];
                if ((rc = gt_release(c, &nd, &m, 1)) != GT_OK) return rc;
                out_node[e] = c->first_node + nd; out_mask[e] = m;
                live_node[idx] = live_node.back(); live_node.pop_back();
                live_mask[idx] = live_mask.back(); live_mask.pop_back();
            } else {
                uint64_t key;
                const uint8_t k = ev_k[e];
                if ((rc = gt_best_node(c, &k, 1, &key)) != GT_OK) return rc;
                if (key == GT_NO_PLACE) continue;
                const uint32_t nd = GT_PLACE_NODE(key) - c->first_node; const uint16_t m = (uint16_t)GT_PLACE_MASK(key);
                if ((rc = gt_commit(c, &nd, &m, 1)) != GT_OK) return rc;
                out_node[e] = c->first_node + nd; out_mask[e] = m;
                live_node.push_back(nd); live_mask.push_back(m);
                placed++;
            }
            if ((rc = gt_score_dirty(c)) != GT_OK) return rc;
        }
        if (out_placed) *out_placed = placed;
        return GT_OK;
    }
    // memo path: the GPU has scored every (class, free mask); the host loop only looks rows up
    if ((rc = mirror_writable(c)) != GT_OK) return rc;
    const uint32_t N = c->n_nodes;
    MinTree tree;
    tree.init(std::max(N, 1u));
    auto row_of = [&](uint32_t nd) -> const uint32_t * {
        const uint32_t cls = c->node_class[nd];
        const uint32_t n = c->h_ngpus[nd];
        const uint32_t f = c->h_free[nd] & ((1u << n) - 1u);
        return &c->memo[(c->memo_off[cls] + f) * 16];
    };
    for (uint32_t nd = 0; nd < N; nd++) {
        const uint32_t *row = row_of(nd);
        for (int k = 0; k < 16; k++) tree.t[k][tree.leaves + nd] = place_key(row[k], c->first_node + nd);
    }
    tree.build();
    auto refresh = [&](uint32_t nd) {
        const uint32_t *row = row_of(nd);
        for (int k = 0; k < 16; k++) tree.set(k, nd, place_key(row[k], c->first_node + nd));
        mark_dirty(c, nd);
    };
    for (uint32_t e = 0; e < n_events; e++) {
        out_node[e] = 0xFFFFFFFFu; out_mask[e] = 0;
        if (ev_k[e] == 0) {
            if (live_node.empty()) continue;
            const size_t idx = ev_arg[e] % live_node.size();
            const uint32_t nd = live_node[idx]; const uint16_t m = live_mask[idx];
            c->h_free[nd] |= m;
            out_node[e] = c->first_node + nd; out_mask[e] = m;
            live_node[idx] = live_node.back(); live_node.pop_back();
            live_mask[idx] = live_mask.back(); live_mask.pop_back();
            refresh(nd);
        } else if (ev_k[e] <= 16) {
            const unsigned long long key = tree.t[ev_k[e] - 1][1];
            if (key == ~0ull) continue;
            const uint32_t nd = GT_PLACE_NODE(key) - c->first_node; const uint16_t m = (uint16_t)GT_PLACE_MASK(key);
            c->h_free[nd] &= (uint16_t)~m;
            out_node[e] = c->first_node + nd; out_mask[e] = m;
            live_node.push_back(nd); live_mask.push_back(m);
            placed++;
            refresh(nd);
        }
    }
    if (out_placed) *out_placed = placed;
    return GT_OK;
}

// "sort" verb (design.md:99,118): one score 0..10 per node for a pod asking for m GPUs
int gt_ext_sort(gt_ctx *c, uint32_t m, uint8_t *out_scores, uint64_t *out_place)
{
    if (!c || m < 1 || m > 16) return GT_ERR_INVALID;
    GT_CUDA(c, cudaSetDevice(c->device));
    int rc = GT_OK;
    if (!c->best_valid || !c->free_uploaded) {
        c->want_score_m = m;             // the pass leaves one score byte per node for this request size as it goes
        rc = gt_score_all_async(c);
        c->want_score_m = 0;
    } else if (!c->dirty.empty()) {
        rc = gt_score_dirty(c);
    }
    if (rc != GT_OK) return rc;
    if (!c->n_nodes) return GT_OK;
    const uint32_t cmax = gt_lut_cmax(c->lut);
    bool keys_by_step = false;
    if (out_place && lazy_exchange(c) && c->x_seq && c->x_reduced < c->x_seq) {
        // the exchange of the pass that just went out runs on the communication stream, beside the score write-back, and
        // hands the keys to the reply buffer itself
        GT_CUDA(c, cudaEventRecord(c->ev_pass, c->stream));
        GT_CUDA(c, cudaStreamWaitEvent(c->comm_stream, c->ev_pass, 0));
        if ((rc = x_consume(c, c->d_place_map, c->comm_stream)) != GT_OK) return rc;
        GT_CUDA(c, cudaEventRecord(c->ev_comm[0], c->comm_stream));
        keys_by_step = true;
    } else if (out_place && c->comm && c->comm_last >= 0) {
        GT_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_comm[c->comm_last], 0));
    }
    const unsigned long long *keys = out_place ? cluster_keys(c) : c->place_cur;
    // the kernels store scores and keys into mapped pinned memory: no device-to-host copy behind them; launched as
    // programmatic dependents of the pass, so their launch latency is spent while K1 still runs
    // (block sizes of 32 / 64 threads instead of 256, and finishing the bytes on the device for a copy-engine transfer
    // instead of storing into mapped memory, were measured: no change / 4.5 us slower per cycle, profiles/r2_scores_dma_ab.txt)
    const uint32_t scores_cta = (uint32_t)gt::kCta;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(scores_cta);
    cfg.stream = c->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    unsigned long long *okeys = (out_place && !keys_by_step) ? c->d_place_map : nullptr;
    const unsigned long long *xerr = c->d_xflags + 3;
    uint32_t *out_dev = c->d_scores_map;
    if (c->scores_m == m) {
        const uint32_t threads = (c->n_nodes + 15) / 16;
        cfg.gridDim = dim3((std::max(threads, 32u) + scores_cta - 1) / scores_cta);
        GT_CUDA(c, cudaLaunchKernelEx(&cfg, gt::k_scores_out, (const uint8_t *)c->d_scores, (const uint32_t *)c->d_best, c->n_nodes, m,
                                      cmax * (m * (m - 1) / 2), reinterpret_cast<uint4 *>(out_dev), keys, okeys, xerr));
    } else {
        const uint32_t threads = (c->n_nodes + 3) / 4;
        cfg.gridDim = dim3((std::max(threads, 32u) + scores_cta - 1) / scores_cta);
        GT_CUDA(c, cudaLaunchKernelEx(&cfg, gt::k_sort_scores, (const uint32_t *)c->d_best, c->n_nodes, m, cmax, out_dev,
                                      keys, okeys, xerr));
    }
    c->launches++;
    GT_CUDA(c, cudaGetLastError());
    if (keys_by_step) GT_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_comm[0], 0));   // later passes stay behind the exchange
    GT_CUDA(c, cudaStreamSynchronize(c->stream));
    stream_waited_for(c);
    if (out_scores) memcpy(out_scores, c->h_scores, c->n_nodes);   // NULL: the caller reads gt_host_scores() in place
    if (out_place) {
        if (c->comm && c->p2p && c->h_place_map[16] != 0) return exchange_failed(c);
        memcpy(out_place, c->h_place_map, 16 * sizeof(uint64_t));
    }
    return GT_OK;
}

// Zero-copy surface for callers that keep their cluster state in pinned memory: the context's own pinned mirror of the
// free masks (edit in place, then say what changed) and the mapped buffer the `sort` verb's scores land in.
uint16_t *gt_host_free(gt_ctx *c)
{
    if (!c) return nullptr;
    if (mirror_writable(c) != GT_OK) return nullptr;   // an upload that still reads the mirror has to finish first
    return c->h_free;
}

int gt_host_free_changed(gt_ctx *c, uint32_t first, uint32_t count)
{
    if (!c) return GT_ERR_INVALID;
    if ((uint64_t)first + count > c->n_nodes) return fail(c, GT_ERR_RANGE, "gt_host_free_changed: range exceeds loaded nodes%s");
    if (first == 0 && count == c->n_nodes) {       // whole shard: one bulk copy, every node rescored by the next pass
        c->free_uploaded = false;
        clear_dirty(c);
        c->best_valid = c->place_valid = false;
        return GT_OK;
    }
    for (uint32_t i = 0; i < count; i++) mark_dirty(c, first + i);
    return GT_OK;
}

const uint8_t *gt_host_scores(gt_ctx *c) { return c ? c->h_scores : nullptr; }

// "bind" verb (design.md:100,119,223-234): GPUs for the pod on the chosen node, commit, pod annotations
int gt_ext_bind(gt_ctx *c, uint32_t node, uint32_t m, int64_t assume_time, uint16_t *out_mask, char *out_annotations,
                size_t cap)
{
    if (!c || !out_mask) return GT_ERR_INVALID;
    if (node >= c->n_nodes) return fail(c, GT_ERR_RANGE, "gt_ext_bind: node id out of range%s");
    int rc = GT_OK;
    if (!c->best_valid || !c->free_uploaded) rc = gt_score_all(c, nullptr);
    else if (c->dirty_flag[node]) rc = gt_score_dirty(c);
    if (rc != GT_OK) return rc;
    uint16_t mask = 0, cost = 0;
    if ((rc = gt_allocate(c, node, m, &mask, &cost)) != GT_OK) return rc;
    if (mask == 0) return fail(c, GT_ERR_STATE, "gt_ext_bind: fewer free GPUs than requested%s");
    if ((rc = gt_commit(c, &node, &mask, 1)) != GT_OK) return rc;
    *out_mask = mask;
    if (out_annotations && cap) rc = gt_bind_annotations(mask, assume_time, 0, out_annotations, cap);
    return rc;
}

int gt_set_node_names(gt_ctx *c, const char *names, size_t len)
{
    if (!c || (!names && len)) return GT_ERR_INVALID;
    std::vector<std::string> v;
    size_t s0 = 0;
    for (size_t i = 0; i <= len; i++)
        if (i == len || names[i] == '\n') {
            if (i > s0 || i < len) v.emplace_back(names + s0, i - s0);
            s0 = i + 1;
        }
    if (v.size() > c->cap) return fail(c, GT_ERR_RANGE, "gt_set_node_names: more names than max_nodes%s");
    c->node_index.clear();
    for (uint32_t i = 0; i < v.size(); i++)
        if (v[i].empty() || !c->node_index.emplace(v[i], i).second)
            return fail(c, GT_ERR_INVALID, "gt_set_node_names: empty or duplicate name '%s'", v[i].c_str());
    c->node_names.swap(v);
    return GT_OK;
}

constexpr size_t kMaxPendingPods = 65536;   // pods seen by sort and not (yet) bound

int gt_ext_sort_json(gt_ctx *c, const char *request, size_t len, char *out, size_t cap, size_t *out_len)
{
    if (!c || !request || !out) return GT_ERR_INVALID;
    if (c->node_names.size() != c->n_nodes) return fail(c, GT_ERR_STATE, "gt_ext_sort_json: gt_set_node_names first%s");
    gt::ExtenderArgsWire a;
    int rc = gt::parse_extender_args(request, len, &a);
    if (rc != GT_OK) return fail(c, rc, "gt_ext_sort_json: malformed ExtenderArgs%s");
    if (a.gpu_count < 1 || a.gpu_count > 16)
        return fail(c, GT_ERR_INVALID, "gt_ext_sort_json: the pod asks for no or more than 16 " GT_RESOURCE_NAME "%s");
    std::vector<uint8_t> scores(c->n_nodes ? c->n_nodes : 1);
    if ((rc = gt_ext_sort(c, a.gpu_count, scores.data(), nullptr)) != GT_OK) return rc;
    if (!a.uid.empty()) {
        // pods that were scored and then bound elsewhere, found unschedulable or re-submitted under a new uid never
        // reach bind or pod_deleted: the table keeps the kMaxPendingPods most recent uids
        if (c->pod_request.emplace(a.uid, a.gpu_count).second) {
            c->pod_request_order.push_back(a.uid);
            while (c->pod_request_order.size() > kMaxPendingPods) {
                c->pod_request.erase(c->pod_request_order.front());   // no-op when bind already took it
                c->pod_request_order.pop_front();
            }
        } else {
            c->pod_request[a.uid] = a.gpu_count;
        }
    }
    std::string s = "[";
    auto item = [&](const std::string &name, unsigned score, bool first) {
        if (!first) s.push_back(',');
        s += "{\"Host\":";
        gt::json_escape(name, &s);
        s += ",\"Score\":" + std::to_string(score) + "}";
    };
    if (a.has_names) {
        for (size_t i = 0; i < a.names.size(); i++) {
            auto it = c->node_index.find(a.names[i]);
            item(a.names[i], it == c->node_index.end() ? 0u : scores[it->second], i == 0);
        }
    } else {
        for (uint32_t i = 0; i < c->n_nodes; i++) item(c->node_names[i], scores[i], i == 0);
    }
    s.push_back(']');
    if (out_len) *out_len = s.size();
    if (s.size() + 1 > cap) return fail(c, GT_ERR_RANGE, "gt_ext_sort_json: reply buffer too small%s");
    memcpy(out, s.c_str(), s.size() + 1);
    return GT_OK;
}

int gt_ext_bind_json(gt_ctx *c, const char *request, size_t len, int64_t assume_time, char *out, size_t cap,
                     char *out_annotations, size_t ann_cap)
{
    if (!c || !request || !out) return GT_ERR_INVALID;
    if (out_annotations && ann_cap) out_annotations[0] = 0;
    gt::BindingArgsWire b;
    int rc = gt::parse_binding_args(request, len, &b);
    if (rc != GT_OK) return fail(c, rc, "gt_ext_bind_json: malformed ExtenderBindingArgs%s");
    std::string err;
    auto pod = c->pod_request.find(b.pod_uid);
    auto node = c->node_index.find(b.node);
    if (pod == c->pod_request.end()) err = "pod " + b.pod_uid + " was not seen by sort";
    else if (node == c->node_index.end() || node->second >= c->n_nodes) err = "unknown node " + b.node;
    else if (c->pod_bound.count(b.pod_uid)) err = "pod " + b.pod_uid + " is already bound";
    else {
        uint16_t mask = 0;
        rc = gt_ext_bind(c, node->second, pod->second, assume_time, &mask, out_annotations, ann_cap);
        if (rc == GT_OK) {
            c->pod_bound[b.pod_uid] = std::make_pair(node->second, mask);
            c->pod_request.erase(pod);
        } else {
            err = std::string(gt_strerror(rc)) + ": " + c->err;
        }
    }
    return gt_wire_binding_result(err.c_str(), out, cap);
}

int gt_ext_pod_deleted(gt_ctx *c, const char *pod_uid)
{
    if (!c || !pod_uid) return GT_ERR_INVALID;
    c->pod_request.erase(pod_uid);
    auto it = c->pod_bound.find(pod_uid);
    if (it == c->pod_bound.end()) return fail(c, GT_ERR_STATE, "gt_ext_pod_deleted: no GPUs recorded for pod %s", pod_uid);
    const uint32_t node = it->second.first;
    const uint16_t mask = it->second.second;
    c->pod_bound.erase(it);
    return gt_release(c, &node, &mask, 1);
}

int gt_synth_events(int config, uint64_t seed, uint32_t n_events, uint8_t *ev_k, uint32_t *ev_arg)
{
    if (!ev_k || !ev_arg || (config != 2 && config != 5)) return GT_ERR_INVALID;
    auto sm = [&](uint64_t index) {
        uint64_t z = seed + (index + 1) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    };
    for (uint32_t i = 0; i < n_events; i++) {
        const uint64_t a = sm(2ull * i), b = sm(2ull * i + 1);
        if (config == 5 && (uint32_t)(a >> 32) >= 2362232012u) {   // floor(0.55 * 2^32): arrival below, departure above
            ev_k[i] = 0; ev_arg[i] = (uint32_t)(b >> 32);
        } else {
            ev_k[i] = (uint8_t)(1 + b % 8u); ev_arg[i] = 0;
        }
    }
    return GT_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// gt_score_nodes_static: K0 + K1 with their scratch in a fixed device array (see gputopo.h).
// Layout: 64 words of control block (bin counters, K1 ticket; padded to 256 B), then 17 bin lists of
// GT_STATIC_MAX_NODES node ids (u32) each, then the same lists' free masks (u16).
static __device__ __align__(256) uint32_t g_static_scratch[64 + gt::kBins * GT_STATIC_MAX_NODES + gt::kBins * GT_STATIC_MAX_NODES / 2];

namespace {
constexpr int kStaticMaxDevices = 64;
constexpr uint32_t kStaticCntWords = 64;
// g_static_scratch names a host-side stand-in here, not device memory: the address a kernel may use
// is looked up once per device with cudaGetSymbolAddress and remembered.
uint32_t *g_static_dev_ptr[kStaticMaxDevices];
int g_static_sms[kStaticMaxDevices];
std::mutex g_static_mu;

int static_fail(int code, char *buf, size_t len, const char *msg, const char *detail = "")
{
    if (buf && len) snprintf(buf, len, "%s%s", msg, detail);
    return code;
}
}  // namespace

extern "C" int gt_score_nodes_static(const uint8_t *d_cost, const uint8_t *d_n_gpus, const uint16_t *d_free_mask,
                                     uint32_t n_nodes, uint32_t *d_best, void *stream, char *err_buf, size_t err_len)
{
    if (err_buf && err_len) err_buf[0] = 0;
#ifdef GT_TRACE
    return static_fail(GT_ERR_STATE, err_buf, err_len, "trace build: gt_score_nodes_static is disabled");
#endif
    if (n_nodes > GT_STATIC_MAX_NODES) {
        char d[96];
        snprintf(d, sizeof d, "%u nodes requested, the built-in scratch holds at most %u", n_nodes, GT_STATIC_MAX_NODES);
        return static_fail(GT_ERR_RANGE, err_buf, err_len, "gt_score_nodes_static: ", d);
    }
    if (n_nodes == 0) return GT_OK;
    if (!d_cost || !d_n_gpus || !d_free_mask || !d_best)
        return static_fail(GT_ERR_INVALID, err_buf, err_len, "gt_score_nodes_static: null buffer");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return static_fail(GT_ERR_NO_DEVICE, err_buf, err_len, "gt_score_nodes_static: ", cudaGetErrorString(e));
    }
    if (dev < 0 || dev >= kStaticMaxDevices)
        return static_fail(GT_ERR_INVALID, err_buf, err_len, "gt_score_nodes_static: device ordinal not supported");
    uint32_t *scratch;
    int sms;
    {
        std::lock_guard<std::mutex> lock(g_static_mu);
        if (!g_static_dev_ptr[dev]) {   // one-time set-up for this device (it is current)
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_static_scratch);
            if (e != cudaSuccess || !p) {
                cudaGetLastError();
                return static_fail(GT_ERR_CUDA, err_buf, err_len, "cudaGetSymbolAddress(g_static_scratch): ",
                                   cudaGetErrorString(e));
            }
            int n_sm = 0;
            e = cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
            if (e != cudaSuccess || n_sm <= 0) {
                cudaGetLastError();
                return static_fail(GT_ERR_CUDA, err_buf, err_len, "cudaDeviceGetAttribute: ", cudaGetErrorString(e));
            }
            g_static_sms[dev] = n_sm;
            g_static_dev_ptr[dev] = static_cast<uint32_t *>(p);
        }
        scratch = g_static_dev_ptr[dev];
        sms = g_static_sms[dev];
    }
    // a host pointer in any of the four would fault inside the kernels (or, where the GPU can reach host memory,
    // quietly use the wrong bytes): refuse it here
    const struct { const void *p; const char *name; } bufs[4] = {
        {d_cost, "d_cost"}, {d_n_gpus, "d_n_gpus"}, {d_free_mask, "d_free_mask"}, {d_best, "d_best"}};
    for (const auto &b : bufs) {
        cudaPointerAttributes attr;
        e = cudaPointerGetAttributes(&attr, b.p);
        if (e != cudaSuccess || (attr.type != cudaMemoryTypeDevice && attr.type != cudaMemoryTypeManaged) ||
            (attr.type == cudaMemoryTypeDevice && attr.device != dev)) {
            cudaGetLastError();
            char d[96];
            snprintf(d, sizeof d, "%s is not device memory of the current device", b.name);
            return static_fail(GT_ERR_INVALID, err_buf, err_len, "gt_score_nodes_static: ", d);
        }
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    uint32_t *cnt = scratch, *list = scratch + kStaticCntWords;
    // the array keeps the previous call's contents: the counters must start at zero, the lists are
    // overwritten up to the new counts
    e = cudaMemsetAsync(cnt, 0, kStaticCntWords * sizeof(uint32_t), s);
    if (e != cudaSuccess) return static_fail(GT_ERR_CUDA, err_buf, err_len, "cudaMemsetAsync: ", cudaGetErrorString(e));
    uint16_t *flist = reinterpret_cast<uint16_t *>(list + (size_t)gt::kBins * GT_STATIC_MAX_NODES);
    gt::k_bin<<<(n_nodes + gt::kCta - 1) / gt::kCta, gt::kCta, 0, s>>>(d_n_gpus, d_free_mask, nullptr, n_nodes,
                                                                     GT_STATIC_MAX_NODES, cnt, list, flist, d_best,
                                                                     nullptr, nullptr, nullptr, 0u, gt::XWork{});
    gt::ScoreParams prm = {};
    prm.cost = d_cost; prm.best = d_best; prm.ctl = cnt; prm.bin_list = list; prm.bin_free = flist;
    prm.place = nullptr; prm.cap = GT_STATIC_MAX_NODES; prm.first_node = 0; prm.fold_place = 0; prm.keep_bins = 0; prm.zero = 0;
    prm.scores = nullptr; prm.score_k = 0; prm.score_den = 0;
    const uint32_t grid = std::min<uint32_t>(n_nodes / 8 + gt::kBins, (uint32_t)sms * gt::kScoreMinCtas);
    gt::k_score<<<grid, gt::kScoreCta, 0, s>>>(prm);
    e = cudaGetLastError();
    if (e != cudaSuccess) return static_fail(GT_ERR_CUDA, err_buf, err_len, "kernel launch: ", cudaGetErrorString(e));
    return GT_OK;
}

#ifdef GT_TRACE
// debug builds only: kernel-level timeline of the last passes (ring of 64); resets the ring
extern "C" int gt_debug_ktrace(gt_ctx *c, unsigned long long *out /* 64 x 8 */, uint32_t *out_pass)
{
    if (!c || !out || !out_pass) return GT_ERR_INVALID;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaMemcpyFromSymbol(out, gt::g_ktrace, sizeof(unsigned long long) * 64 * 8);
    cudaMemcpyFromSymbol(out_pass, gt::g_kpass, 4);
    unsigned long long init[64][8];
    for (auto &r : init) { r[0] = ~0ull; r[1] = 0; r[2] = ~0ull; r[3] = ~0ull; r[4] = 0; r[5] = ~0ull; r[6] = r[7] = 0; }
    cudaMemcpyToSymbol(gt::g_ktrace, init, sizeof init);
    return GT_OK;
}

// debug builds only: the per-warp event trace of the last k_score launch
extern "C" int gt_debug_trace(gt_ctx *c, unsigned long long *out, uint32_t n_words)
{
    if (!c || !out) return GT_ERR_INVALID;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    const size_t have = (size_t)c->num_sms * 8 * 4 * gt::kTraceSlots;
    cudaMemcpy(out, c->d_trace, std::min<size_t>(have, n_words) * 8, cudaMemcpyDeviceToHost);
    cudaMemset(c->d_trace, 0, have * 8);
    return GT_OK;
}
#endif
