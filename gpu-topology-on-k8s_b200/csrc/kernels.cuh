// kernels.cuh -- sm_100a kernels of the topology-aware GPU allocation path.
//
//   k_apply_lut : level codes -> u8 pair costs (load time; design.md:33-47 weight column as a runtime LUT)
//   k_bin       : K0, bins nodes by popcount(free) so that every enumeration shape below is static
//   k_score     : K1, scores every k-subset of every node's free set and keeps, per k, the lowest
//                 (cost, mask) -- design.md:162-186 replaced by the exhaustive optimum (north_star)
//   k_argmin    : K2, best node per request size k over the best table
//
// Enumeration.  The free GPUs of a node are compressed to positions 0..p-1 (order preserving, so the
// numeric order of masks is preserved).  Subsets of positions 0..7 are walked by a fully unrolled
// depth-first tree: a tree node is a subset, its children add one smaller position, and
//     key(S + {b}) = key(S) + R_b(S) + bit_b          (one IADD3; key = cost<<16 | mask)
//     R_i(S + {b}) = R_i(S) + c[i][b]   for i < b     (pair costs pre-shifted by 16)
// so a subset costs ~1 add for the key, ~1 add of R upkeep and 1 min, instead of C(k,2) additions.
// popcount(S) is the recursion depth, i.e. static, so the per-k minima live in registers.
// Nodes with p <= 8 free GPUs take one thread each; nodes with p > 8 spread the 2^(p-8) subsets H of
// positions 8..p-1 over lanes, each lane running the 8-position tree seeded with
// key0 = cost(H)<<16 | mask(H) and R_b = cross cost of position b against H.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gaia_tree.h"

namespace gt {

constexpr int kMaxG = 16;
constexpr int kBins = 17;            // popcount 0..16
constexpr int kCta = 256;
constexpr uint32_t kInf = 0xFFFF0000u;
#ifndef GT_SCORE_CTA
#define GT_SCORE_CTA 128    // (32 / 64: no change; 320 / 640 through dynamic shared memory: slower, profiles/r2_score_cta_ab.txt)
#endif
constexpr int kScoreCta = GT_SCORE_CTA;   // K1: 4 independent warps per CTA
#ifndef GT_K1_TRIGGER
#define GT_K1_TRIGGER 2   // where K1 lets the next kernel's CTAs be scheduled: 0 = at exit, 1 = after its own wait, 2 = after its work loop
#endif
#ifndef GT_SCORE_MIN_CTAS
#define GT_SCORE_MIN_CTAS 5
#endif
constexpr int kScoreMinCtas = GT_SCORE_MIN_CTAS;   // resident CTAs per SM the register budget is set for
constexpr int kTilesPerWarp = 16;    // nodes per warp unit at p = 9 (2 lanes per node)
constexpr int kTileStride = 272;     // bytes per staged 16x16 cost tile (256 + 16: shifts banks between tiles)
#ifndef GT_SMALL_SMEM
#define GT_SMALL_SMEM 1   // one-lane-per-node units read their pair costs from tiles the warp staged in shared memory
                          // (0: byte gathers from global memory; A/B in profiles/r2_cold_ab_*.txt)
#endif
constexpr int kStageBytes = (GT_SMALL_SMEM ? 32 : kTilesPerWarp) * kTileStride;   // staging buffer of a warp: 16 full tiles (32 for small units)
constexpr int kSplitP = 14;          // nodes with p >= kSplitP span several warp units

// Multi-GPU: one step of the min-exchange of the 16 placement keys over peer memory (see exchange_step below).
constexpr int kMaxPeers = 8;
struct XWork {
    unsigned long long *inbox[kMaxPeers];   // inbox base of every rank (own address for the own rank); world == 0: no work
    const unsigned long long *post_keys;    // this rank's 16 keys of pass `seq_post`
    unsigned long long *out1, *out2;        // where the cluster-wide keys of passes seq_reduce1 / seq_reduce2 go
    unsigned long long *keys_host;          // optional: mapped host memory, receives the 16 keys of seq_reduce2 + err_flag's word
    unsigned long long *err_flag;           // pass number of a reduction that gave up waiting for a peer
    unsigned long long seq_reduce1, seq_post, seq_reduce2;   // 0 = nothing to do
    uint32_t rank, world;
};

struct ScoreParams {
    const uint8_t *cost;      // [N][256] u8 pair cost, row-major 16x16
    uint32_t *best;           // [N][16]
    uint32_t *ctl;            // control block, see kCtl*
    const uint32_t *bin_list; // [17][cap] node ids by p
    const uint16_t *bin_free; // [17][cap] their free masks (already limited to n_gpus bits)
    unsigned long long *place;// [16] cluster argmin, folded into when fold_place != 0
    uint32_t cap, first_node;
    uint32_t fold_place;      // full pass: K1 folds its rows into place (K0 preset it)
    uint32_t keep_bins;       // timing runs: leave the bin counts for another launch
    uint32_t self_clean;      // 1: the last CTA to finish zeroes this control block (it costs every CTA an atomic round trip
                              //    before it may exit); 0: the block stays as it is and ctl_next is the next pass's
    uint32_t *ctl_next;       // the control block the pass before this one used: CTA 0 zeroes it for the pass after next (null: none)
    uint32_t trigger_early;   // 1: dependents may be scheduled once every CTA is past its wait (they become resident as
                              //    CTAs leave: a K0 that runs ahead, see k_bin); 0: once every CTA has finished its units
    uint32_t zero;            // always 0: keeps the ticket address from looking warp-uniform to ptxas (see k_score)
    unsigned long long *ktime;      // measurement: [0] globaltimer when CTA 0 got past the dependency wait, [1] when the last CTA finished (null = off)
    XWork xw;                       // multi-GPU: exchange work of the two passes before this one, done by one warp of K1
    uint8_t *scores;                // `sort` verb inside the pass: HostPriority score 0..10 of every node for request size
    uint32_t score_k, score_den;    // score_k (1..16; 0 = off), den = cmax * C(score_k, 2); see score_byte()
#ifdef GT_TRACE
    unsigned long long *trace;   // [warps][kTraceSlots] (globaltimer ns << 8 | event id), debug builds only
#endif
};
#ifdef GT_TRACE
constexpr int kTraceSlots = 64;
__device__ __forceinline__ unsigned long long gt_now()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// pass timeline: per pass (ring of 64) [K0 first entry, K0 last exit, K1 first entry, K1 after wait (first), K1 last exit]
__device__ unsigned long long g_ktrace[64][8];
__device__ unsigned int g_kpass;
#define GT_KT_MIN(slot, t) atomicMin(&g_ktrace[g_kpass & 63u][slot], (t))
#define GT_KT_MAX(slot, t) atomicMax(&g_ktrace[g_kpass & 63u][slot], (t))
#define GT_TR(id)                                                                         \
    do {                                                                                  \
        if (lane == 0 && tr_n < kTraceSlots) tr[tr_n++] = (gt_now() << 8) | (unsigned)(id); \
    } while (0)
#else
#define GT_TR(id) do { } while (0)
#endif
// scratch control block (u32 words): [0..16] bin counts, [17] K1 unit ticket, [18] K1 finished CTAs (self_clean launches).
// Three of them take turns (api.cu, launch_score): the one a pass uses was zeroed by the K1 two passes earlier.
constexpr int kCtlWords = 32;
constexpr int kCtlTicket = 17, kCtlDone = 18;
// K1 hands out work units bin by bin, costliest unit type first (measured per-unit times, profiles/): the
// 32-node units of p = 7, 8 and the multi-node units of p = 9 .. 12 are the long ones, tiny bins come last.
// Kept as immediates (5 bits per entry), not in constant memory: a cold constant-cache miss in front of every
// warp's first unit costs more than the whole decode.
__host__ __device__ constexpr int bin_order_host(int i)
{
#ifndef GT_BIN_ORDER
#define GT_BIN_ORDER 6
#endif
    // A/B'd on the final hand-over structure, us per pass at 100 000 nodes (profiles/r2_bin_order_ab_v2.txt, _v3.txt):
    // 0: 14.97  2: 14.81  3: 14.72  4: 14.67  5: 14.68  6: 14.62  7: 14.69  8: 14.75   (1 lost by 1.3-1.9 us earlier in the round)
#if GT_BIN_ORDER == 0     // round 1
    constexpr int order[kBins] = {9, 10, 8, 13, 11, 12, 7, 15, 14, 16, 6, 5, 4, 3, 2, 1, 0};
#elif GT_BIN_ORDER == 1   // many-lanes-per-node units strictly before one-lane-per-node units
    constexpr int order[kBins] = {12, 11, 13, 16, 15, 14, 9, 10, 8, 7, 6, 5, 4, 3, 2, 1, 0};
#elif GT_BIN_ORDER == 2   // longest unit first by the per-unit times of profiles/r2_k_score_warp_timeline_cold_v2.txt
    constexpr int order[kBins] = {7, 12, 8, 11, 16, 9, 13, 10, 14, 5, 6, 4, 15, 3, 2, 1, 0};
#elif GT_BIN_ORDER == 3   // the round-1 order with the 128-subset one-lane-per-node units in front
    constexpr int order[kBins] = {7, 9, 10, 8, 13, 11, 12, 15, 14, 16, 6, 5, 4, 3, 2, 1, 0};
#elif GT_BIN_ORDER == 4
    constexpr int order[kBins] = {7, 12, 8, 11, 9, 10, 13, 15, 14, 16, 6, 5, 4, 3, 2, 1, 0};
#elif GT_BIN_ORDER == 5
    constexpr int order[kBins] = {7, 8, 9, 10, 13, 11, 12, 16, 14, 15, 6, 5, 4, 3, 2, 1, 0};
#elif GT_BIN_ORDER == 6   // default: as 3, the split units by descending unit time, p = 5 before p = 6
    constexpr int order[kBins] = {7, 9, 10, 8, 13, 11, 12, 16, 15, 14, 5, 6, 4, 3, 2, 1, 0};
#elif GT_BIN_ORDER == 7
    constexpr int order[kBins] = {7, 12, 9, 10, 8, 13, 11, 15, 14, 16, 6, 5, 4, 3, 2, 1, 0};
#else
    constexpr int order[kBins] = {7, 9, 10, 8, 13, 11, 12, 15, 14, 16, 6, 5, 4, 0, 1, 2, 3};
#endif
    return order[i];
}
__host__ __device__ constexpr unsigned long long bin_order_pack(int from, int to)
{
    unsigned long long v = 0;
    for (int i = from; i < to; i++) v |= (unsigned long long)bin_order_host(i) << (5 * (i - from));
    return v;
}
__device__ __forceinline__ int bin_order(int i)
{
    constexpr unsigned long long lo = bin_order_pack(0, 12), hi = bin_order_pack(12, kBins);
    return (int)(((i < 12) ? (lo >> (5 * i)) : (hi >> (5 * (i - 12)))) & 31ull);
}

// ---------------------------------------------------------------- multi-GPU: min-exchange of the 16 placement keys
constexpr unsigned long long kExchangeTimeoutNs = 2000000000ull;   // a peer that does not show up within 2 s is not coming
// spin until *flag >= want; false when the time is up (the waits below must never hang the GPU: ranks that issue their
// passes in different orders, or a rank that died, end in an error on the host instead)
__device__ __forceinline__ bool wait_flag_bounded(const unsigned long long *flag, unsigned long long want, unsigned ns)
{
    unsigned long long t0 = 0;
    for (uint32_t spins = 0;; spins++) {
        if (*reinterpret_cast<const volatile unsigned long long *>(flag) >= want) return true;
        __nanosleep(ns);
        if ((spins & 255u) == 255u) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (!t0) t0 = t;
            else if (t - t0 > kExchangeTimeoutNs) return false;
        }
    }
}

// system-scope fence of the message-passing pattern below (flags written and read as volatile = relaxed.sys accesses):
// acquire-release is all it needs; __threadfence_system() is the sequentially consistent one
__device__ __forceinline__ void fence_sys()
{
#ifdef GT_X_FENCE_SC
    __threadfence_system();
#else
    asm volatile("fence.acq_rel.sys;" ::: "memory");
#endif
}

// The exchange, one warp, no kernel of its own in a loop of passes.  Every rank owns an inbox that all peers have
// mapped (CUDA IPC): keys[world][2][16] then flags[world][2], slot parity = pass number & 1.
//   post(s):   store this rank's 16 keys of pass s into slot [rank][s & 1] of every rank's inbox, fence, then raise
//              the flag beside them to s (the flag is how a reader knows the keys are there);
//   reduce(s): wait until the flags of all ranks in the own inbox show s, take the minimum of the world x 16 keys.
// A pass's keys are final when its K1 has finished, so K1 of pass s carries reduce(s-2) and post(s-1) -- in that
// order -- in one warp of its first CTA: the NVLink round trips disappear behind 10 us of scoring, nothing extra is
// launched and nothing waits on the scoring stream.  Whoever wants the keys of the latest pass L runs the same step as
// a one-warp kernel: reduce(L-1), post(L), reduce(L) (k_exchange_step).  Slot reuse is safe without credits: a rank
// overwrites slot s & 1 with post(s) only after its own reduce(s-1), i.e. after it saw every peer's post(s-1), which
// each peer issued after its own reduce(s-2) -- the last reader of the slot's previous content.  The same chain keeps
// the ranks within a pass of each other.  Waits are bounded; a rank that gives up reports it through err_flag.
__device__ __forceinline__ void exchange_step(const XWork &w, uint32_t lane)
{
    unsigned long long *mine = w.inbox[w.rank];
    const size_t flags = (size_t)w.world * 32;
    auto reduce = [&](unsigned long long seq, unsigned long long *out) {
        if (!seq) return;
        const uint32_t par = (uint32_t)(seq & 1u);
        bool ok = true;
        if (lane < w.world) ok = wait_flag_bounded(mine + flags + lane * 2 + par, seq, 32);
        if (!__all_sync(0xFFFFFFFFu, ok)) {
            if (lane == 0) *reinterpret_cast<volatile unsigned long long *>(w.err_flag) = seq;
            return;
        }
        fence_sys();      // acquire: the keys are read after the flags were seen
        if (lane < 16) {
            unsigned long long m = ~0ull;
            for (uint32_t r = 0; r < w.world; r++) {
                const unsigned long long v = *reinterpret_cast<volatile unsigned long long *>(mine + (size_t)(r * 2 + par) * 16 + lane);
                m = v < m ? v : m;
            }
            out[lane] = m;
        }
    };
    reduce(w.seq_reduce1, w.out1);
    if (w.seq_post) {
        const uint32_t slot = w.rank * 2 + (uint32_t)(w.seq_post & 1u);
        if (lane < 16) {
            const unsigned long long key = w.post_keys[lane];
            for (uint32_t q = 0; q < w.world; q++)
                *reinterpret_cast<volatile unsigned long long *>(w.inbox[q] + (size_t)slot * 16 + lane) = key;
        }
        fence_sys();      // release: keys before flags, on every peer
        __syncwarp();
        if (lane < w.world) *reinterpret_cast<volatile unsigned long long *>(w.inbox[lane] + flags + slot) = w.seq_post;
    }
    reduce(w.seq_reduce2, w.out2);
    __syncwarp();
    if (w.keys_host) {       // the caller's copy, straight from here: the score write-back runs beside this kernel
        const unsigned long long *last = w.seq_reduce2 ? w.out2 : w.out1;     // the step's last reduction
        if (lane < 16) w.keys_host[lane] = last[lane];
        if (lane == 16) w.keys_host[16] = *reinterpret_cast<volatile unsigned long long *>(w.err_flag);
    }
}

__global__ void __launch_bounds__(32) k_exchange_step(XWork w)
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    exchange_step(w, threadIdx.x);
}

// HostPriority score of a best-table entry (design.md:205-217): floor(10 * (den - cost) / den), den = cmax * C(m,2);
// a single GPU (den = 0) scores 10, an infeasible node 0.  kScorePending marks nodes whose rows several work units
// build together (p >= 14): k_scores_out computes their score once the pass is over.
constexpr uint32_t kScorePending = 0xFFu;
__device__ __forceinline__ uint32_t score_byte(uint32_t b, uint32_t den)
{
    if (b == kInf) return 0u;
    const uint32_t cost = b >> 16;
    return den == 0u ? 10u : (cost >= den ? 0u : (10u * (den - cost)) / den);
}

// ---------------------------------------------------------------- load time
// cost[i][j] = cost[j][i] = lut[levels[min(i,j)][max(i,j)]], cost[i][i] = 0: the upper triangle of the level
// matrix is the one that counts (gputopo.h), and the scoring kernels rely on symmetric tiles with a
// zero diagonal (row sums over a GPU mask then equal pair sums).  One thread per tile row.
__global__ void k_apply_lut(const uint8_t *__restrict__ levels, uint8_t *__restrict__ cost, size_t n_rows,
                            uint4 lut4)
{
    __shared__ uint8_t s_lut[16];
    if (threadIdx.x < 4) reinterpret_cast<uint32_t *>(s_lut)[threadIdx.x] = (&lut4.x)[threadIdx.x];
    __syncthreads();
    size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; r < n_rows; r += stride) {
        const uint8_t *tile = levels + (r >> 4) * 256;
        const uint32_t i = (uint32_t)(r & 15);
        uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
        for (uint32_t j = 0; j < 16; j++) {
            const uint32_t lo = min(i, j), hi = max(i, j);
            const uint32_t v = (i == j) ? 0u : (uint32_t)s_lut[tile[lo * 16 + hi] & 15];
            w[j >> 2] |= v << ((j & 3) * 8);
        }
        reinterpret_cast<uint4 *>(cost)[r] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---------------------------------------------------------------- K0: bin nodes by popcount(free)
// ids == nullptr: nodes [0, n); else the listed (dirty) nodes.  ctl[0..16] must be zero on entry.
// place != nullptr: preset the cluster argmin that K1 folds into.
__global__ void __launch_bounds__(kCta) k_bin(const uint8_t *__restrict__ n_gpus,
                                              const uint16_t *__restrict__ free_mask,
                                              const uint32_t *__restrict__ ids, uint32_t n, uint32_t cap,
                                              uint32_t *__restrict__ ctl, uint32_t *__restrict__ bin_list,
                                              uint16_t *__restrict__ bin_free, uint32_t *__restrict__ best,
                                              unsigned long long *__restrict__ place, uint8_t *__restrict__ scores,
                                              uint16_t *__restrict__ free_copy, uint32_t early, XWork xw)
{
    __shared__ uint32_t s_cnt[kBins], s_base[kBins];
    const uint32_t tid = threadIdx.x;
    // Multi-GPU: the grid has one CTA more than the nodes need, and that one (CTA 0, the first to become resident)
    // carries a step of the key exchange in one warp -- before the dependency wait when K0 runs ahead, i.e. beside the
    // last units of the previous pass's K1.  What it posts are the keys of the pass before that one (complete: the K1
    // in front waited for its K0, which waited for that pass's K1).
    const uint32_t has_x = xw.world != 0 ? 1u : 0u;
    if (has_x && blockIdx.x == 0) {
        if (!early) asm volatile("griddepcontrol.wait;" ::: "memory");
        if (tid < 32) exchange_step(xw, tid);
        if (early) asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;");
        return;
    }
    const uint32_t blk = blockIdx.x - has_x;
#ifdef GT_TRACE
    if (tid == 0) GT_KT_MIN(0, gt_now());
#endif
    if (tid < kBins) s_cnt[tid] = 0;
    // K0 is itself launched as a programmatic dependent of whatever kernel precedes it (normally the previous
    // pass's K1): its CTAs become resident as that grid drains and wait until it has completed and flushed.
    // early != 0 (the host sets it when nothing queued in front of this launch writes the masks or the id list):
    // the masks are fetched and counted BEFORE that wait, behind the tail of the previous kernel; everything that
    // kernel may still read or write (control block, bin lists, best rows, scores, the device copy of the masks)
    // is touched after it.
    if (!early) asm volatile("griddepcontrol.wait;" ::: "memory");
#ifdef GT_TRACE
    if (tid == 0) GT_KT_MIN(5, gt_now());
#endif
    // (fetching the CTA's 256 masks as 32 16-byte loads by one warp -- one 512-byte PCIe read instead of eight 64-byte
    // ones when the mirror is read in place -- was measured: 0.3 us slower per resident pass for the extra barrier,
    // no gain end to end: profiles/r2_k0_wide_loads_ab.txt)
    __syncthreads();
    const uint32_t i = blk * blockDim.x + tid;
    int p = -1;
    uint32_t slot = 0, node = 0, f = 0, raw = 0;
    if (i < n) {
        node = ids ? ids[i] : i;
        uint32_t ng = n_gpus[node];
        ng = ng > (uint32_t)kMaxG ? (uint32_t)kMaxG : ng;
        raw = free_mask[node];
        f = raw & ((1u << ng) - 1u);
        p = __popc(f);
        slot = atomicAdd(&s_cnt[p], 1u);
    }
    if (early == 1) asm volatile("griddepcontrol.wait;" ::: "memory");
    // early == 2: the kernel in front is a K1 that reads the other set of bin lists and another control block (this
    // one was zeroed a pass ago): ranges are reserved and lists filled before the wait as well, so that a K0 whose CTAs
    // became resident while that K1 was still finishing its last units has only the presets below left to do
    __syncthreads();
    if (tid < kBins && s_cnt[tid]) s_base[tid] = atomicAdd(&ctl[tid], s_cnt[tid]);
    __syncthreads();
    if (p >= 0) {
        const size_t at = (size_t)p * cap + s_base[p] + slot;
        bin_list[at] = node;
        bin_free[at] = (uint16_t)f;
    }
    if (early == 2) asm volatile("griddepcontrol.wait;" ::: "memory");
#ifndef GT_K0_TRIGGER_AFTER_WAIT
#define GT_K0_TRIGGER_AFTER_WAIT 1
#endif
    // every CTA of this grid is resident and has done its share when it gets here: K1's CTAs may be scheduled while the
    // presets below are written (they wait for this grid to complete before they read anything)
    if (GT_K0_TRIGGER_AFTER_WAIT && early == 2) asm volatile("griddepcontrol.launch_dependents;");
    // what the previous pass's kernels may still have been reading or writing
    if (place && blk == 0 && tid < 16) place[tid] = ~0ull;
    if (p >= 0) {
        // free_mask may be the caller's pinned host mirror read in place (whole-shard update): the device copy that
        // later passes over single nodes read is refreshed on the way, no separate upload
        if (free_copy) free_copy[node] = (uint16_t)raw;
        if (p >= kSplitP) {   // several warp units of K1 meet in this row through atomicMin
            uint4 *o = reinterpret_cast<uint4 *>(best + (size_t)node * 16);
            const uint4 inf = make_uint4(kInf, kInf, kInf, kInf);
            o[0] = inf; o[1] = inf; o[2] = inf; o[3] = inf;
            if (scores) scores[node] = (uint8_t)kScorePending;
        }
    }
    // programmatic dependent launch: from here on K1's CTAs may be scheduled while this grid drains; K1 waits
    // (griddepcontrol.wait) for the whole grid before it reads the bins.  Triggering at the top of K0 instead
    // was measured 3-6 us slower per pass: the waiting K1 CTAs crowd K0 off the SMs.  (A K0 of one 128-thread,
    // 32-register CTA per SM, small enough to sit resident beside K1's CTAs and start the moment K1 ends, was
    // measured too: 20.2 instead of 17.8 us per pass, profiles/r2_cold_ab_k0_persist_trigger.txt.)
#ifdef GT_TRACE
    if (tid == 0) GT_KT_MAX(1, gt_now());
#endif
    asm volatile("griddepcontrol.launch_dependents;");
}

// 16 bytes global -> shared without a register stop-over (LDGSTS).  Staging loops issue all their copies first and
// wait once: one DRAM round trip per unit, however many tiles it has.
__device__ __forceinline__ void copy16_async(void *dst, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void copies_landed()
{
    asm volatile("cp.async.wait_all;" ::: "memory");
}

// ---------------------------------------------------------------- the unrolled subset tree
template <int P, int B, int D>
struct Children;

// visit S + {b}: S has D elements, all > b
template <int P, int b, int D>
__device__ __forceinline__ void visit_child(uint32_t key, const uint32_t (&R)[P], const uint32_t (&c)[P * P],
                                            const uint32_t (&bit)[P], uint32_t (&acc)[P + 1])
{
    const uint32_t nkey = key + R[b] + bit[b];
    acc[D + 1] = min(acc[D + 1], nkey);
    if constexpr (b > 0) {
        uint32_t nR[P];
#pragma unroll
        for (int i = 0; i < P; i++) nR[i] = (i < b) ? R[i] + c[i * P + b] : 0u;
        Children<P, b, D + 1>::run(nkey, nR, c, bit, acc);
    }
}

template <int P, int B, int D>
struct Children {   // children of S add a position in [0, B)
    static __device__ __forceinline__ void run(uint32_t key, const uint32_t (&R)[P], const uint32_t (&c)[P * P],
                                               const uint32_t (&bit)[P], uint32_t (&acc)[P + 1])
    {
        if constexpr (B > 0) {
            visit_child<P, B - 1, D>(key, R, c, bit, acc);
            Children<P, B - 1, D>::run(key, R, c, bit, acc);
        }
    }
};

// ---------------------------------------------------------------- K1 small: one lane per node, p == P <= 8
// tile: the node's cost tile -- in global memory (SMEM = false: the pair costs are gathered with independent
// byte loads, one round trip; staging the tiles through shared memory first was measured slower, it adds a
// second dependent stage in front of every unit) or staged in shared memory (SMEM = true, unused).
template <int P, bool SMEM>
__device__ __forceinline__ void score_small(const uint8_t *tile, uint32_t f, uint32_t (&out)[16])
{
    if constexpr (P > 0) {
        uint32_t pos[P], bit[P];
#pragma unroll
        for (int a = 0; a < P; a++) {
            const int g = __ffs(f) - 1;
            f &= f - 1;
            pos[a] = (uint32_t)g;
            bit[a] = 1u << g;
        }
        uint32_t c[P * P];
#pragma unroll
        for (int a = 0; a < P; a++)
#pragma unroll
            for (int b = 0; b < P; b++) {
                if (a < b) {
                    const uint8_t v = SMEM ? tile[pos[a] * 16 + pos[b]] : __ldg(tile + pos[a] * 16 + pos[b]);
                    c[a * P + b] = (uint32_t)v << 16;
                } else {
                    c[a * P + b] = 0u;
                }
            }
        uint32_t acc[P + 1], R[P];
#pragma unroll
        for (int k = 0; k <= P; k++) acc[k] = kInf;
#pragma unroll
        for (int a = 0; a < P; a++) R[a] = 0u;
        Children<P, P, 0>::run(0u, R, c, bit, acc);
#pragma unroll
        for (int k = 1; k <= P; k++) out[k - 1] = acc[k];
    }
}

template <bool SMEM>
__device__ __forceinline__ void score_small_any(int p, const uint8_t *tile, uint32_t f, uint32_t (&out)[16])
{
    switch (p) {
    case 1: score_small<1, SMEM>(tile, f, out); break;
    case 2: score_small<2, SMEM>(tile, f, out); break;
    case 3: score_small<3, SMEM>(tile, f, out); break;
    case 4: score_small<4, SMEM>(tile, f, out); break;
    case 5: score_small<5, SMEM>(tile, f, out); break;
    case 6: score_small<6, SMEM>(tile, f, out); break;
    case 7: score_small<7, SMEM>(tile, f, out); break;
    case 8: score_small<8, SMEM>(tile, f, out); break;
    default: break;
    }
}

// entry k-1 of a row held in registers (one-lane-per-node units have at most 8 feasible sizes)
__device__ __forceinline__ uint32_t pick_row(const uint32_t (&out)[16], uint32_t k)
{
    uint32_t v = kInf;
#pragma unroll
    for (int q = 0; q < 8; q++)
        if (k == (uint32_t)(q + 1)) v = out[q];
    return v;
}

// The cluster argmin is carried in registers: lane k (< 16) of every warp holds the warp's best
// (value = cost<<16|mask, node) for request size k+1; it reaches shared / global memory once, at warp exit.
struct WarpBest {
    uint32_t v, node;
    __device__ __forceinline__ void take(uint32_t nv, uint32_t nn)
    {
        if ((nv >> 16) < (v >> 16) || ((nv >> 16) == (v >> 16) && nn < node)) { v = nv; node = nn; }
    }
};

// One lane per node: the lowest (cost, node) of the warp's candidates for request sizes K0+1 .. K0+4 goes to
// lanes K0 .. K0+3.  Straight-line on purpose: the four reductions of each stage are independent, so their
// latencies overlap (a branchy per-k version measured ~250 ns per k).  Rows beyond p hold kInf and fold to a no-op.
#ifndef GT_FOLD_SELECT
#define GT_FOLD_SELECT 1
#endif
template <int K0>
__device__ __forceinline__ void fold_small_warp4(WarpBest &wb, const uint32_t (&out)[16], uint32_t node, uint32_t lane)
{
    uint32_t cmin[4], nmin[4], bv[4];
    bool tie[4];
#pragma unroll
    for (int q = 0; q < 4; q++) cmin[q] = __reduce_min_sync(0xFFFFFFFFu, out[K0 + q] & 0xFFFF0000u);
#pragma unroll
    for (int q = 0; q < 4; q++) {
        tie[q] = (out[K0 + q] & 0xFFFF0000u) == cmin[q];
        nmin[q] = __reduce_min_sync(0xFFFFFFFFu, tie[q] ? node : 0xFFFFFFFFu);
    }
#pragma unroll
    for (int q = 0; q < 4; q++) {
        const uint32_t owner = __ffs(__ballot_sync(0xFFFFFFFFu, tie[q] && node == nmin[q])) - 1;
        bv[q] = __shfl_sync(0xFFFFFFFFu, out[K0 + q], (int)(owner & 31u));
    }
#if GT_FOLD_SELECT
    // lane K0 + q takes candidate q: pick the lane's candidate first, compare once (four predicated compare chains
    // were 5 % of K1's instructions)
    const uint32_t q = lane - (uint32_t)K0;
    uint32_t v = bv[0], n = nmin[0], cm = cmin[0];
#pragma unroll
    for (int j = 1; j < 4; j++)
        if (q == (uint32_t)j) { v = bv[j]; n = nmin[j]; cm = cmin[j]; }
    if (q < 4u && cm != kInf) wb.take(v, n);
#else
#pragma unroll
    for (int q = 0; q < 4; q++)
        if (lane == (uint32_t)(K0 + q) && cmin[q] != kInf) wb.take(bv[q], nmin[q]);
#endif
}

// ---------------------------------------------------------------- K1 big: p > 8, 2^(p-8) lanes per node
// 4 bits -> 4 bytes of 0xFF
__device__ __forceinline__ uint32_t bytemask4(uint32_t nib)
{
    return ((nib * 0x00204081u) & 0x01010101u) * 0xFFu;
}
// sum of the bytes of a tile row selected by a byte mask
__device__ __forceinline__ uint32_t row_sum(const uint4 row, const uint32_t (&m)[4])
{
    uint32_t s = __dp4a(row.x & m[0], 0x01010101u, 0u);
    s = __dp4a(row.y & m[1], 0x01010101u, s);
    s = __dp4a(row.z & m[2], 0x01010101u, s);
    return __dp4a(row.w & m[3], 0x01010101u, s);
}

// one lane of a big node: subsets H (given by h over the free GPUs beyond the 8 lowest) x the 8-position tree.
// raw: the node's 16x16 cost tile in shared memory, symmetric, zero diagonal, 16-byte aligned.
#ifndef GT_BIG_TICKET_MID
#define GT_BIG_TICKET_MID 4     // the tree's top-level children b = 7 .. this are walked before `mid()` runs, the rest (2^b - 1
#endif                          // subsets) after it: K1 asks for its next unit there, in front of the last 15 subsets and the
                                // shared-memory minima, 0.5-0.8 us before the unit's tree phase ends.  us per pass: 8 (behind the
                                // tree): 14.32, 6: 14.17, 5: 14.16, 4: 14.12, 3: 14.20, 2: 14.18, 1: 14.15 (profiles/
                                // r2_big_ticket_mid_ab*.txt); the same inside the P = 7, 8 trees of one-lane-per-node units
                                // instead of in front of them: +0.1 (r2_small_ticket_mid_ab_v2.txt)
template <class Mid>
__device__ __forceinline__ void score_big_lane(const uint8_t *raw, uint32_t f, int hb, uint32_t h, uint32_t *best_row, Mid mid)
{
    uint32_t pos[8], bit[8], c[64], R[8], acc[9];
#pragma unroll
    for (int a = 0; a < 8; a++) {
        const int g = __ffs(f) - 1;
        f &= f - 1;
        pos[a] = (uint32_t)g;
        bit[a] = 1u << g;
    }
    // H = the free GPUs beyond the 8 lowest whose bit in h is set
    uint32_t hmask = 0;
    for (int j = 0; j < hb; j++) {
        const uint32_t low = f & (0u - f);
        f ^= low;
        if ((h >> j) & 1u) hmask |= low;
    }
    uint32_t m[4];
#pragma unroll
    for (int q = 0; q < 4; q++) m[q] = bytemask4((hmask >> (4 * q)) & 15u);
#pragma unroll
    for (int a = 0; a < 8; a++)
#pragma unroll
        for (int b = 0; b < 8; b++)
            c[a * 8 + b] = (a < b) ? ((uint32_t)raw[pos[a] * 16 + pos[b]] << 16) : 0u;
    // cross cost of tree position b against H, and the cost inside H (each pair counted twice)
#pragma unroll
    for (int b = 0; b < 8; b++)
        R[b] = row_sum(*reinterpret_cast<const uint4 *>(raw + pos[b] * 16), m) << 16;
    uint32_t ch2 = 0;
    for (uint32_t hm = hmask; hm; hm &= hm - 1)
        ch2 += row_sum(*reinterpret_cast<const uint4 *>(raw + (__ffs(hm) - 1) * 16), m);
    const int nsel = __popc(h);
    const uint32_t key0 = ((ch2 >> 1) << 16) | hmask;
#pragma unroll
    for (int k = 0; k <= 8; k++) acc[k] = kInf;
    if (h) acc[0] = key0;
#if GT_BIG_TICKET_MID >= 8
    Children<8, 8, 0>::run(key0, R, c, bit, acc);
    mid();
#else
    visit_child<8, 7, 0>(key0, R, c, bit, acc);
    if (GT_BIG_TICKET_MID < 7) visit_child<8, 6, 0>(key0, R, c, bit, acc);
    if (GT_BIG_TICKET_MID < 6) visit_child<8, 5, 0>(key0, R, c, bit, acc);
    if (GT_BIG_TICKET_MID < 5) visit_child<8, 4, 0>(key0, R, c, bit, acc);
    mid();
    Children<8, (GT_BIG_TICKET_MID < 4 ? 4 : GT_BIG_TICKET_MID), 0>::run(key0, R, c, bit, acc);
#endif
#pragma unroll
    for (int j = 0; j <= 8; j++)
        if (acc[j] != kInf) atomicMin(best_row + nsel + j, acc[j]);
}

// Warp-private staging: the 16x16 cost tiles of the unit's nodes (coalesced 16-byte loads) and the
// per-k minima of each node.  No CTA barrier in K1's work loop.
struct WarpScratch {
    __align__(16) uint8_t raw[kStageBytes];
    uint32_t best[kTilesPerWarp][kBins];
};

#ifndef GT_LATE_TICKET
#define GT_LATE_TICKET 1   // ask for the next unit after the tree instead of at the start of the current unit: a ticket taken
                           // at the start reserves a unit that then waits behind a long one while other warps run dry
                           // (0.5 us per pass at 100 000 nodes, profiles/r2_cold_ab_late_ticket.txt)
#endif
// next unit number from the global ticket.  ptxas turns an atomic add on a provably warp-uniform address (even from
// inline PTX, even atom.inc) into a warp-aggregated atomic whose leader broadcast (SHFL) waits for the reply at once:
// the address therefore carries a run-time zero times the lane index.
// Inside `if (lane == 0)` the compiler knows the lane and folds the product away again (cuobjdump: VOTEU / FLO / POPC /
// @P0 ATOMG / SHFL.IDX right behind it: 0.5 us of waiting per unit).  So the whole warp executes the instruction under a
// predicate and no branch tells the compiler which lane it is; lanes other than 0 get 0 back.
__device__ __forceinline__ uint32_t take_ticket(const ScoreParams &prm, uint32_t lane)
{
    uint32_t next = 0;
    asm volatile("{\n\t.reg .pred p;\n\tsetp.eq.u32 p, %2, 0;\n\t@p atom.global.add.u32 %0, [%1], 1;\n\t}"
                 : "+r"(next) : "l"(prm.ctl + kCtlTicket + prm.zero * lane), "r"(lane) : "memory");
    return next;
}

// hb = p - 8 bits of H.  hb <= 5: a warp holds 32 >> hb nodes.  hb > 5: a node spans 2^(hb-5) warp units
// whose minima meet in global memory (rows preset to kInf by k_bin).
__device__ __forceinline__ void score_big_warp(const ScoreParams &prm, WarpScratch &ws, WarpBest &wb,
                                               int p, uint32_t lu, uint32_t cnt, uint32_t lane, uint32_t &next
#ifdef GT_TRACE
                                               , unsigned long long *tr, int &tr_n
#endif
)
{
    const int hb = p - 8;
    const uint32_t *list = prm.bin_list + (size_t)p * prm.cap;
    const uint16_t *flist = prm.bin_free + (size_t)p * prm.cap;
    uint32_t first, live, slot, h;
    if (hb <= 5) {
        const uint32_t slots = 32u >> hb;
        first = lu * slots;
        live = min(slots, cnt - first);
        slot = lane >> hb;
        h = lane & ((1u << hb) - 1u);
    } else {
        first = lu >> (hb - 5);
        live = 1;
        slot = 0;
        h = ((lu & ((1u << (hb - 5)) - 1u)) << 5) | lane;
    }
    const bool active = slot < live;
    const uint32_t node = __ldg(list + first + (active ? slot : 0u));
    const uint32_t f = __ldg(flist + first + (active ? slot : 0u));
    if (f == 0x12345u) GT_TR(99);
    GT_TR(10);   // list entries arrived
    // stage tiles: 16 lanes x 16 B per tile, two tiles per copy instruction; all copies are issued before any is
    // waited for (a load -> store loop paid one memory round trip per iteration, eight of them for a p = 9 unit)
#pragma unroll
    for (int it = 0; it < kTilesPerWarp / 2; it++) {
        const uint32_t t = 2u * (uint32_t)it + (lane >> 4);
        const uint32_t nd = __shfl_sync(0xFFFFFFFFu, node, (int)((t << (hb <= 5 ? hb : 0)) & 31u));   // slot t's first lane
        if (t < live)
            copy16_async(ws.raw + t * kTileStride + (lane & 15u) * 16, prm.cost + (size_t)nd * 256 + (lane & 15u) * 16);
    }
    for (uint32_t s2 = 0; s2 < live; s2++)
        if (lane < (uint32_t)kBins) ws.best[s2][lane] = kInf;
    copies_landed();
    __syncwarp();
    GT_TR(11);   // tiles staged
#ifndef GT_BIG_TICKET_BEFORE_TREE
#define GT_BIG_TICKET_BEFORE_TREE 0   // 1 was measured: 1 us per pass slower (a unit reserved for 3 us while other warps run dry)
#endif
    if (GT_LATE_TICKET && GT_BIG_TICKET_BEFORE_TREE) next = take_ticket(prm, lane);
    // (the ticket instruction is a predicated one executed by whoever is here; lane 0 always is: slot 0 is live)
    if (active)
        score_big_lane(ws.raw + slot * kTileStride, f, hb, h, &ws.best[slot][0], [&]() {
            if (GT_LATE_TICKET && !GT_BIG_TICKET_BEFORE_TREE && GT_BIG_TICKET_MID < 8) next = take_ticket(prm, lane);
        });
    __syncwarp();
    if (GT_LATE_TICKET && !GT_BIG_TICKET_BEFORE_TREE && GT_BIG_TICKET_MID >= 8) next = take_ticket(prm, lane);   // (lane 0 only, see there) the reply arrives during the write-out
    GT_TR(12);   // set-up + tree + atomics done
    if (hb <= 5) {
        // half-warps write one row per pass (k = (lane & 15) + 1) and keep the best candidate per k
        WarpBest hb2;
        hb2.v = kInf; hb2.node = 0xFFFFFFFFu;
        // (node ids re-read from the bin list, an L1 hit: taking them from the registers of the slot's lanes with a
        //  shuffle measured 0.1 us per pass slower, profiles/r2_writeout_ab.txt)
        for (uint32_t s2 = lane >> 4; s2 < live; s2 += 2) {
            const uint32_t nd = __ldg(list + first + s2);
            const uint32_t v = ws.best[s2][(lane & 15u) + 1u];
            prm.best[(size_t)nd * 16 + (lane & 15u)] = v;
            if (prm.score_k == (lane & 15u) + 1u) prm.scores[nd] = (uint8_t)score_byte(v, prm.score_den);
            hb2.take(v, nd);
        }
        __syncwarp();
        const uint32_t ov = __shfl_xor_sync(0xFFFFFFFFu, hb2.v, 16), on = __shfl_xor_sync(0xFFFFFFFFu, hb2.node, 16);
        hb2.take(ov, on);
        if (lane < 16) wb.take(hb2.v, hb2.node);
    } else if (lane < 16) {
        const uint32_t v = ws.best[0][lane + 1];
        if (v != kInf) {
            atomicMin(prm.best + (size_t)node * 16 + lane, v);
            wb.take(v, node);   // a partial minimum of a node is a real subset of it: folding it is exact
        }
    }
    __syncwarp();
}

// Persistent warps pull work units (costliest unit types first) from a global ticket; the ticket for the next unit is
// asked for late in the current one (take_ticket, GT_SMALL_TICKET_AT), its reply is read at the end of the unit.
//   bin p <= 8 : 32 nodes per unit (one lane each)
//   bin 9..13  : 32 >> (p-8) nodes per unit
//   bin >= 14  : 2^(p-13) units per node
__global__ void __launch_bounds__(kScoreCta, kScoreMinCtas) k_score(ScoreParams prm)
{
    __shared__ WarpScratch s_ws[kScoreCta / 32];
    __shared__ unsigned long long s_place[16];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
#ifdef GT_TRACE
    unsigned long long *tr = prm.trace + (size_t)(blockIdx.x * (kScoreCta / 32) + warp) * kTraceSlots;
    int tr_n = 0;
#endif
    GT_TR(1);
#ifdef GT_TRACE
    const unsigned long long kt_entry = gt_now();
#endif
    if (tid < 16) s_place[tid] = ~0ull;
    asm volatile("griddepcontrol.wait;" ::: "memory");   // K0's bins are complete and visible (no-op without PDL)
    if (prm.ktime && blockIdx.x == 0 && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(prm.ktime[0]));
    // two control blocks take turns: the one the previous pass used is free from here on (its K1 has completed) and
    // is zeroed now for the next pass's K0, so that no CTA of this pass has to find out whether it is the last one
    if (prm.ctl_next && blockIdx.x == 0 && tid < (uint32_t)kCtlWords) prm.ctl_next[tid] = 0u;
    GT_TR(2);
    if (GT_K1_TRIGGER == 1 || prm.trigger_early) asm volatile("griddepcontrol.launch_dependents;");
#ifdef GT_TRACE
    if (tid == 0) { GT_KT_MIN(2, kt_entry); GT_KT_MIN(3, gt_now()); }
#endif
    // units per bin in hand-out order and their inclusive scan, in every warp for itself (17 lanes, registers only):
    // nobody waits at a barrier for one warp's round trip to the control block
    uint32_t my_cnt = 0, my_end = 0;
    {
        uint32_t units = 0;
        if (lane < (uint32_t)kBins) {
            const int p = bin_order((int)lane);
            my_cnt = prm.ctl[p];
            if (p >= kSplitP) units = my_cnt << (p - 13);
            else if (p > 8) units = (my_cnt + (32u >> (p - 8)) - 1) >> (13 - p);
            else units = (my_cnt + 31u) >> 5;
#ifdef GT_BINMASK   // profiling builds only: score just the bins in the mask (results are incomplete)
            if (!((GT_BINMASK >> p) & 1u)) units = 0;
#endif
        }
#ifdef GT_TRACE
        if (__any_sync(0xFFFFFFFFu, units == 0xFFFFFFF1u)) GT_TR(99);
        GT_TR(6);    // control block arrived
#endif
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, units, d);
            if (lane >= (uint32_t)d) units += o;
        }
        my_end = lane < (uint32_t)kBins ? units : 0xFFFFFFFFu;
    }
    GT_TR(7);    // scan done
    __syncthreads();      // s_place is set
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, my_end, kBins - 1);
    WarpScratch &ws = s_ws[warp];
    GT_TR(3);
    // first unit = this warp's global index; later ones come from the ticket (take_ticket).
    // Multi-GPU: the last warp of CTA 0 carries the exchange step of the two passes before this one (see exchange_step).
    // It gets no unit of its own to start with -- the step costs it 2-3 us of NVLink round trips, and a unit waiting
    // behind that would end the kernel late -- and joins the others through the ticket afterwards.
    const uint32_t gw = blockIdx.x * (kScoreCta / 32) + warp, xw_warp = kScoreCta / 32 - 1;
    const bool has_x = prm.xw.world != 0;
    const uint32_t n_warps = gridDim.x * (kScoreCta / 32) - (has_x ? 1u : 0u);   // warps that start on a unit of their own
    // (dealing the first units round-robin over the CTAs, or handing the longest units out first, measured slower
    // both with warm and with cold L2: profiles/r2_cold_ab_orders_staging.txt)
    uint32_t u = has_x && gw > xw_warp ? gw - 1 : gw;
    if (has_x && gw == xw_warp) {
        exchange_step(prm.xw, lane);
        uint32_t t = 0;
        t = take_ticket(prm, lane);
        u = __shfl_sync(0xFFFFFFFFu, t, 0) + n_warps;
    }
    WarpBest wb;
    wb.v = kInf; wb.node = 0xFFFFFFFFu;
#if GT_MIX_FROM
    // hand-out positions [0, GT_MIX_FROM) (long, issue-bound units) alternate with positions [GT_MIX_FROM, GT_MIX_TO]
    // (short, latency-bound units) until one of the two runs out; what is left follows in order
    const uint32_t mix_a = __shfl_sync(0xFFFFFFFFu, my_end, GT_MIX_FROM - 1);
    const uint32_t mix_b = __shfl_sync(0xFFFFFFFFu, my_end, GT_MIX_TO) - mix_a;
    const uint32_t mix_m = min(mix_a, mix_b);
#endif
    while (u < total) {
        uint32_t next = 0;
        if (!GT_LATE_TICKET) next = take_ticket(prm, lane);
#if GT_MIX_FROM
        if (u < 2u * mix_m) u = (u & 1u) ? mix_a + (u >> 1) : (u >> 1);
        else if (mix_a > mix_b && u - 2u * mix_b < mix_a - mix_b) u -= mix_b;
#endif
        // `next` is not touched again before the end of this unit: the reply has a whole unit to arrive
        // which bin: lane l keeps the cumulative unit count after hand-out position l and that bin's node count; one
        // ballot instead of a walk through shared memory (a chain of dependent loads in front of every unit: 0.35 us per
        // pass, profiles/r2_ballot_lookup_ab.txt)
        const int i = __popc(__ballot_sync(0xFFFFFFFFu, lane < (uint32_t)kBins && u >= my_end));
        const int p = bin_order(i);
        const uint32_t lu = u - (i ? __shfl_sync(0xFFFFFFFFu, my_end, i - 1) : 0u);
        const uint32_t cnt = __shfl_sync(0xFFFFFFFFu, my_cnt, i);
        GT_TR(32 + p);   // unit start
        if (p > 8) {
#ifdef GT_TRACE
            score_big_warp(prm, ws, wb, p, lu, cnt, lane, next, tr, tr_n);
#else
            score_big_warp(prm, ws, wb, p, lu, cnt, lane, next);
#endif
        } else {
            const uint32_t idx = lu * 32u + lane;
            const bool ok = idx < cnt;
            uint32_t out[16];
#pragma unroll
            for (int k = 0; k < 16; k++) out[k] = kInf;
            uint32_t node = 0xFFFFFFFFu, f = 0;
            if (ok) {
                const size_t at = (size_t)p * prm.cap + idx;
                node = __ldg(prm.bin_list + at);
                f = __ldg(prm.bin_free + at);
            }
#if GT_SMALL_SMEM
            // The unit's tiles travel as whole 16-byte rows, two nodes per load instruction (a half-warp per node,
            // one lane per row), and only the rows of free GPUs: a handful of wide loads that touch one or two lines
            // per node, instead of up to 28 byte loads per lane that touch 32 lines each -- those kept the SM's
            // load pipe busy for a microsecond per unit.  Lanes then pick their pair costs out of shared memory.
            {
                const uint32_t row = lane & 15u;
                __syncwarp();
                // (every lane copying the rows of its own node instead -- p copy instructions, no shuffles, 32 lines
                // each -- measured 0.2 us slower per pass: profiles/r2_cold_ab_staging_occupancy_scorecta.txt)
                if (!__any_sync(0xFFFFFFFFu, f >> 8)) {      // rows 0..7 only: four nodes per copy instruction
#pragma unroll
                    for (int t = 0; t < 8; t++) {
                        const int s2 = 4 * t + (int)(lane >> 3);
                        const uint32_t nd = __shfl_sync(0xFFFFFFFFu, node, s2);
                        const uint32_t ff = __shfl_sync(0xFFFFFFFFu, f, s2);
                        if ((ff >> (lane & 7u)) & 1u)
                            copy16_async(ws.raw + s2 * kTileStride + (lane & 7u) * 16, prm.cost + (size_t)nd * 256 + (lane & 7u) * 16);
                    }
                } else {
#pragma unroll
                    for (int t = 0; t < 16; t++) {       // asynchronous copies, all in flight together: one round trip
                        const int s2 = 2 * t + (int)(lane >> 4);
                        const uint32_t nd = __shfl_sync(0xFFFFFFFFu, node, s2);
                        const uint32_t ff = __shfl_sync(0xFFFFFFFFu, f, s2);     // 0 for slots beyond the bin's end
                        if ((ff >> row) & 1u)
                            copy16_async(ws.raw + s2 * kTileStride + row * 16, prm.cost + (size_t)nd * 256 + row * 16);
                    }
                }
                copies_landed();
                __syncwarp();
            }
#ifndef GT_SMALL_TICKET_AT
#define GT_SMALL_TICKET_AT 9    // one-lane-per-node units of bins below this p ask for the next unit in front of their tree, once
#endif                          // the tiles are staged (us per pass: 0: 14.54, 5: 14.55, 7: 14.40, 8: 14.42, 9: 14.32;
                                // many-lanes-per-node units the same way: 15.36; everybody at the start of the unit: 15.9;
                                // profiles/r2_small_ticket_ab*.txt, r2_ticket_timing_ab_v3.txt)
            if (GT_LATE_TICKET && p < GT_SMALL_TICKET_AT) next = take_ticket(prm, lane);
            if (ok) {
                score_small_any<true>(p, ws.raw + lane * kTileStride, f, out);
#else
            if (ok) {
                score_small_any<false>(p, prm.cost + (size_t)node * 256, f, out);
#endif
                GT_TR(20);   // small: gathers + tree done
                uint4 *o = reinterpret_cast<uint4 *>(prm.best + (size_t)node * 16);
                o[0] = make_uint4(out[0], out[1], out[2], out[3]);
                o[1] = make_uint4(out[4], out[5], out[6], out[7]);
                o[2] = make_uint4(kInf, kInf, kInf, kInf);
                o[3] = make_uint4(kInf, kInf, kInf, kInf);
                if (prm.score_k) prm.scores[node] = (uint8_t)score_byte(pick_row(out, prm.score_k), prm.score_den);
            }
#if GT_SMALL_SMEM
            if (GT_LATE_TICKET && p >= GT_SMALL_TICKET_AT) next = take_ticket(prm, lane);   // (lane 0 only) the reply arrives during the fold
#else
            if (GT_LATE_TICKET) next = take_ticket(prm, lane);
#endif
            if (prm.fold_place) {
                fold_small_warp4<0>(wb, out, node, lane);
                if (p > 4) fold_small_warp4<4>(wb, out, node, lane);
            }
        }
        GT_TR(13);   // written out / folded
        u = __shfl_sync(0xFFFFFFFFu, next, 0) + n_warps;
        GT_TR(14);   // next ticket known
    }
    if (prm.fold_place && lane < 16 && wb.v != kInf) {
        const unsigned long long key = ((unsigned long long)(wb.v >> 16) << 48) |
                                       ((unsigned long long)(prm.first_node + wb.node) << 16) | (wb.v & 0xFFFFu);
        atomicMin(&s_place[lane], key);
    }
    GT_TR(4);
    if (GT_K1_TRIGGER == 2 && !prm.trigger_early) asm volatile("griddepcontrol.launch_dependents;");
    __syncthreads();
    GT_TR(5);
    if (prm.fold_place && tid < 16 && s_place[tid] != ~0ull) atomicMin(prm.place + tid, s_place[tid]);
    // self_clean: the last CTA to finish leaves the control block zeroed for the next pass
    if (tid == 0 && prm.self_clean) {
#ifdef GT_TRACE
        GT_KT_MAX(4, gt_now());
        __threadfence();
#endif
        const uint32_t done = atomicAdd(prm.ctl + kCtlDone, 1u);
        if (done == gridDim.x - 1) {
#ifdef GT_TRACE
            g_kpass++;
#endif
            if (!prm.keep_bins)
                for (int q = 0; q < kBins; q++) prm.ctl[q] = 0u;
            prm.ctl[kCtlTicket] = 0u;
            prm.ctl[kCtlDone] = 0u;
            if (prm.ktime) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(prm.ktime[1]));
        }
    }
}

// ---------------------------------------------------------------- K1 as a producer / consumer pipeline (full passes)
// Same enumeration, same units, same results as k_score; what changes is who waits for memory.  In k_score every warp
// walks the chain ticket -> bin-list entry -> cost tiles (DRAM) -> set-up -> tree -> write-out by itself, twice per
// pass on average, and the SM's issue slots stand idle through the first three links.  Here one CTA of 20 warps per SM
// splits the roles: warp 0 (the producer) walks this SM's share of the units ahead of everybody else, reads their
// bin-list entries and streams each node's tile rows into a ring of shared-memory slots with asynchronous 16-byte
// copies (cp.async: global -> shared without a register stop-over, their completion counted by the slot's mbarrier);
// the other 19 warps (consumers) take filled slots in order, run the unit out of shared memory, write its rows and hand
// the slot back.  Units are dealt round-robin over the SMs in cost order (unit u belongs to CTA u mod gridDim.x): every SM
// gets the same number of every unit type +-1, no ticket atomics.  Slot = 4608 bytes of tile rows (16 full tiles of a
// many-lanes-per-node unit, or 32 nodes' rows 0..highest free GPU of a one-lane-per-node unit, 16 bytes of padding
// between nodes against bank conflicts; a unit whose rows do not fit travels as two half units) + the unit's descriptor.
namespace pipe {
#ifndef GT_PIPE_WARPS
#define GT_PIPE_WARPS 20
#endif
#ifndef GT_PIPE_CTAS
#define GT_PIPE_CTAS 1
#endif
#ifndef GT_PIPE_SLOTS
#define GT_PIPE_SLOTS 32
#endif
constexpr int kWarps = GT_PIPE_WARPS;      // 1 producer + 19 consumers: what 96 registers per thread leave room for
constexpr int kCtasPerSm = GT_PIPE_CTAS;
constexpr int kCta = kWarps * 32;
constexpr int kSlots = GT_PIPE_SLOTS;
constexpr int kSlotRaw = 4608;
constexpr uint32_t kPoison = 0xFFFFFFFFu;  // descriptor of the slots that tell the consumers to stop
constexpr int kBatch = 8;                  // units whose bin-list entries the producer has in flight together

struct Slot {
    __align__(128) uint8_t raw[kSlotRaw];
    uint32_t node[32];     // per slot lane: node (local index), 0xFFFFFFFF = none
    uint32_t f[32];        // its free mask
    uint16_t off[32];      // where its rows start in raw
    uint32_t p, lu, live, pad;
};
struct Shared {
    Slot slot[kSlots];
    uint32_t best[kWarps][kTilesPerWarp][kBins];   // per consumer warp: per-k minima of the unit's nodes
    unsigned long long full[kSlots], empty[kSlots];
    unsigned long long place[16];
    uint32_t cnt[kBins], end[kBins];
    uint32_t next, n_units;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try(unsigned long long *bar, uint32_t parity)
{
    uint32_t ok;
    asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// a wait that cannot hang the GPU: a pipeline bug ends the kernel with an error after a second instead
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t parity)
{
    if (mbar_try(bar, parity)) return;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (uint32_t spins = 0;; spins++) {
        if (mbar_try(bar, parity)) return;
        if ((spins & 1023u) == 1023u) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > 1000000000ull) __trap();
        }
    }
}
// the arrival that fires once all of this thread's earlier asynchronous copies have landed.  (One bulk copy per node --
// cp.async.bulk, 64-256 bytes each -- was measured first: the copy engine delivered about 2 bytes per ns per SM on
// such small scattered pieces, 3x slower than the kernel it was meant to feed: profiles/r2_ab_pipe_bulkcopy_v1.txt.)
__device__ __forceinline__ void mbar_arrive_after_copies(unsigned long long *bar)
{
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// where unit u of the pass lives: bin (hand-out order index i, advanced from the previous unit), its nodes
struct UnitAt { uint32_t p, lu, first, live; };
__device__ __forceinline__ UnitAt locate(const Shared &sh, uint32_t u, int &i)
{
    while (u >= sh.end[i]) i++;
    UnitAt a;
    a.p = (uint32_t)bin_order(i);
    a.lu = u - (i ? sh.end[i - 1] : 0u);
    const uint32_t cnt = sh.cnt[a.p];
    if (a.p > 8) {
        const uint32_t hb = a.p - 8;
        if (hb <= 5) {
            const uint32_t slots = 32u >> hb;
            a.first = a.lu * slots;
            a.live = min(slots, cnt - a.first);
        } else {
            a.first = a.lu >> (hb - 5);
            a.live = 1;
        }
    } else {
        a.first = a.lu * 32u;
        a.live = min(32u, cnt - a.first);
    }
    return a;
}
}  // namespace pipe

__global__ void __launch_bounds__(pipe::kCta, pipe::kCtasPerSm) k_score_pipe(ScoreParams prm)
{
    using namespace pipe;
    extern __shared__ __align__(128) uint8_t pipe_smem[];
    Shared &sh = *reinterpret_cast<Shared *>(pipe_smem);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (tid == 0) {
        // full: the 32 producer lanes' copies + the descriptor's release by lane 0; empty: the consumer's lane 0
        for (int q = 0; q < kSlots; q++) { mbar_init(&sh.full[q], 33); mbar_init(&sh.empty[q], 1); }
        sh.next = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 16) sh.place[tid] = ~0ull;
    asm volatile("griddepcontrol.wait;" ::: "memory");   // K0's bins are complete and visible
    if (prm.ktime && blockIdx.x == 0 && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(prm.ktime[0]));
    if (warp == 0) {
        uint32_t units = 0, cnt = 0;
        int p = 0;
        if (lane < (uint32_t)kBins) {
            p = bin_order((int)lane);
            cnt = prm.ctl[p];
            sh.cnt[p] = cnt;
            if (p >= kSplitP) units = cnt << (p - 13);
            else if (p > 8) units = (cnt + (32u >> (p - 8)) - 1) >> (13 - p);
            else units = (cnt + 31u) >> 5;
#ifdef GT_BINMASK   // profiling builds only: score just the bins in the mask (results are incomplete)
            if (!((GT_BINMASK >> p) & 1u)) units = 0;
#endif
        }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, units, d);
            if (lane >= (uint32_t)d) units += o;
        }
        if (lane < (uint32_t)kBins) sh.end[lane] = units;
        const uint32_t total = __shfl_sync(0xFFFFFFFFu, units, kBins - 1);
        if (lane == 0) sh.n_units = total > blockIdx.x ? (total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;
    }
    __syncthreads();
    WarpBest wb;
    wb.v = kInf; wb.node = 0xFFFFFFFFu;

    if (warp == 0) {
        // ------------------------------------------------ producer
        const uint32_t n_units = sh.n_units, G = gridDim.x;
        uint32_t q = 0;                 // slots filled so far (a unit takes one slot, sometimes two)
        int i_load = 0, i_use = 0;
        // lane t describes node t of the unit (node, free mask, bytes of tile rows wanted, where they go in the slot);
        // the copies themselves are spread over the lanes 16 bytes (one tile row) at a time: `per` lanes per node
        auto fill = [&](uint32_t p, uint32_t lu, uint32_t live, uint32_t node, uint32_t f, uint32_t bytes, uint32_t off) {
            const uint32_t s = q % kSlots;
            if (q >= (uint32_t)kSlots) mbar_wait(&sh.empty[s], ((q / kSlots) - 1u) & 1u);
            Slot &sl = sh.slot[s];
            sl.node[lane] = node; sl.f[lane] = f; sl.off[lane] = (uint16_t)off;
            if (lane == 0) { sl.p = p; sl.lu = lu; sl.live = live; }
            const uint32_t most = __reduce_max_sync(0xFFFFFFFFu, bytes);
            // 8 lanes (rows) per node, 4 nodes per copy instruction; straight-line so that the shuffles overlap
            const uint32_t row = lane & 7u, sub = lane >> 3;
            if (most) {
#pragma unroll
                for (int it = 0; it < 8; it++) {
                    const uint32_t j = (uint32_t)it * 4u + sub;
                    const uint32_t nj = __shfl_sync(0xFFFFFFFFu, node, j), bj = __shfl_sync(0xFFFFFFFFu, bytes, j);
                    const uint32_t oj = __shfl_sync(0xFFFFFFFFu, off, j);
                    if (row * 16u < bj) copy16_async(sl.raw + oj + row * 16u, prm.cost + (size_t)nj * 256 + row * 16u);
                }
            }
            if (most > 128u) {      // rows 8..15 of the nodes that have them
#pragma unroll
                for (int it = 0; it < 8; it++) {
                    const uint32_t j = (uint32_t)it * 4u + sub;
                    const uint32_t nj = __shfl_sync(0xFFFFFFFFu, node, j), bj = __shfl_sync(0xFFFFFFFFu, bytes, j);
                    const uint32_t oj = __shfl_sync(0xFFFFFFFFu, off, j);
                    if ((row + 8u) * 16u < bj)
                        copy16_async(sl.raw + oj + (row + 8u) * 16u, prm.cost + (size_t)nj * 256 + (row + 8u) * 16u);
                }
            }
            mbar_arrive_after_copies(&sh.full[s]);
            __syncwarp();
            if (lane == 0) mbar_arrive(&sh.full[s]);      // releases the descriptor written above
            q++;
        };
        for (uint32_t j0 = 0; j0 < n_units; j0 += kBatch) {
            uint32_t nd[kBatch], fr[kBatch];
#pragma unroll
            for (int b = 0; b < kBatch; b++) {       // the batch's bin-list entries: one round trip for all of them
                nd[b] = 0xFFFFFFFFu; fr[b] = 0;
                if (j0 + b < n_units) {
                    const UnitAt a = locate(sh, blockIdx.x + (j0 + b) * G, i_load);
                    if (lane < a.live) {
                        const size_t at = (size_t)a.p * prm.cap + a.first + lane;
                        nd[b] = __ldg(prm.bin_list + at);
                        fr[b] = __ldg(prm.bin_free + at);
                    }
                }
            }
#pragma unroll
            for (int b = 0; b < kBatch; b++) {
                if (j0 + b >= n_units) break;
                const UnitAt a = locate(sh, blockIdx.x + (j0 + b) * G, i_use);
                const uint32_t node = nd[b], f = fr[b];
                if (a.p > 8) {           // whole tiles, 272 bytes apart
                    fill(a.p, a.lu, a.live, node, f, lane < a.live ? 256u : 0u, lane * (uint32_t)kTileStride);
                } else {                 // rows 0 .. highest free GPU, 16 bytes of padding behind every node
                    const uint32_t bytes = f ? 16u * (32u - (uint32_t)__clz(f)) : 0u;
                    uint32_t incl = bytes + 16u;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
                        if (lane >= (uint32_t)d) incl += o;
                    }
                    const uint32_t all = __shfl_sync(0xFFFFFFFFu, incl, 31);
                    if (all <= (uint32_t)kSlotRaw) {
                        fill(a.p, a.lu, a.live, node, f, bytes, incl - bytes - 16u);
                    } else {             // many 16-GPU nodes among them: two half units
                        const uint32_t half0 = __shfl_sync(0xFFFFFFFFu, incl, 15);
                        const uint32_t n2 = __shfl_down_sync(0xFFFFFFFFu, node, 16), f2 = __shfl_down_sync(0xFFFFFFFFu, f, 16);
                        const uint32_t b2 = __shfl_down_sync(0xFFFFFFFFu, bytes, 16), i2 = __shfl_down_sync(0xFFFFFFFFu, incl, 16);
                        const bool lo = lane < 16;
                        fill(a.p, a.lu, min(a.live, 16u), lo ? node : 0xFFFFFFFFu, lo ? f : 0u, lo ? bytes : 0u,
                             lo ? incl - bytes - 16u : 0u);
                        fill(a.p, a.lu, a.live > 16u ? a.live - 16u : 0u, lo ? n2 : 0xFFFFFFFFu, lo ? f2 : 0u, lo ? b2 : 0u,
                             lo ? i2 - b2 - 16u - half0 : 0u);
                    }
                }
            }
        }
        for (int w = 1; w < kWarps; w++) fill(kPoison, 0, 0, 0xFFFFFFFFu, 0, 0, 0);   // one stop sign per consumer
    } else {
        // ------------------------------------------------ consumers
        uint32_t (*my_best)[kBins] = sh.best[warp];
        if (prm.xw.world && blockIdx.x == 0 && warp == 1) exchange_step(prm.xw, lane);   // multi-GPU, see there
        for (;;) {
            uint32_t q = 0;
            if (lane == 0) q = atomicAdd(&sh.next, 1u);
            q = __shfl_sync(0xFFFFFFFFu, q, 0);
            const uint32_t s = q % kSlots;
            mbar_wait(&sh.full[s], (q / kSlots) & 1u);
            Slot &sl = sh.slot[s];
            const uint32_t p = sl.p;
            if (p == kPoison) {
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh.empty[s]);
                break;
            }
            const uint32_t live = sl.live;
            if (p > 8) {
                const int hb = (int)p - 8;
                uint32_t slot, h;
                if (hb <= 5) { slot = lane >> hb; h = lane & ((1u << hb) - 1u); }
                else { slot = 0; h = ((sl.lu & ((1u << (hb - 5)) - 1u)) << 5) | lane; }
                const bool active = slot < live;
                for (uint32_t s2 = 0; s2 < live; s2++)
                    if (lane < (uint32_t)kBins) my_best[s2][lane] = kInf;
                __syncwarp();
                if (active) score_big_lane(sl.raw + slot * kTileStride, sl.f[slot], hb, h, &my_best[slot][0], []() {});
                __syncwarp();
                if (hb <= 5) {
                    WarpBest hb2;
                    hb2.v = kInf; hb2.node = 0xFFFFFFFFu;
                    for (uint32_t s2 = lane >> 4; s2 < live; s2 += 2) {
                        const uint32_t nd = sl.node[s2];
                        const uint32_t v = my_best[s2][(lane & 15u) + 1u];
                        prm.best[(size_t)nd * 16 + (lane & 15u)] = v;
                        if (prm.score_k == (lane & 15u) + 1u) prm.scores[nd] = (uint8_t)score_byte(v, prm.score_den);
                        hb2.take(v, nd);
                    }
                    __syncwarp();
                    const uint32_t ov = __shfl_xor_sync(0xFFFFFFFFu, hb2.v, 16), on = __shfl_xor_sync(0xFFFFFFFFu, hb2.node, 16);
                    hb2.take(ov, on);
                    if (lane < 16) wb.take(hb2.v, hb2.node);
                } else if (lane < 16) {
                    const uint32_t v = my_best[0][lane + 1];
                    const uint32_t nd = sl.node[0];
                    if (v != kInf) {
                        atomicMin(prm.best + (size_t)nd * 16 + lane, v);
                        wb.take(v, nd);
                    }
                }
            } else {
                const bool ok = lane < live;
                uint32_t out[16];
#pragma unroll
                for (int k = 0; k < 16; k++) out[k] = kInf;
                const uint32_t node = sl.node[lane], f = sl.f[lane];
                if (ok) {
                    score_small_any<true>((int)p, sl.raw + sl.off[lane], f, out);
                    uint4 *o = reinterpret_cast<uint4 *>(prm.best + (size_t)node * 16);
                    o[0] = make_uint4(out[0], out[1], out[2], out[3]);
                    o[1] = make_uint4(out[4], out[5], out[6], out[7]);
                    o[2] = make_uint4(kInf, kInf, kInf, kInf);
                    o[3] = make_uint4(kInf, kInf, kInf, kInf);
                    if (prm.score_k) prm.scores[node] = (uint8_t)score_byte(pick_row(out, prm.score_k), prm.score_den);
                }
                if (prm.fold_place) {
                    fold_small_warp4<0>(wb, out, node, lane);
                    if (p > 4) fold_small_warp4<4>(wb, out, node, lane);
                }
            }
            __syncwarp();     // every lane is done with the slot's rows and descriptor
            if (lane == 0) mbar_arrive(&sh.empty[s]);
        }
        if (prm.fold_place && lane < 16 && wb.v != kInf) {
            const unsigned long long key = ((unsigned long long)(wb.v >> 16) << 48) |
                                           ((unsigned long long)(prm.first_node + wb.node) << 16) | (wb.v & 0xFFFFu);
            atomicMin(&sh.place[lane], key);
        }
    }
    asm volatile("griddepcontrol.launch_dependents;");
    __syncthreads();
    if (prm.fold_place && tid < 16 && sh.place[tid] != ~0ull) atomicMin(prm.place + tid, sh.place[tid]);
    if (tid == 0) {           // the last CTA to finish leaves the control block zeroed for the next pass
        const uint32_t done = atomicAdd(prm.ctl + kCtlDone, 1u);
        if (done == gridDim.x - 1) {
            if (!prm.keep_bins)
                for (int q = 0; q < kBins; q++) prm.ctl[q] = 0u;
            prm.ctl[kCtlTicket] = 0u;
            prm.ctl[kCtlDone] = 0u;
            if (prm.ktime) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(prm.ktime[1]));
        }
    }
}

// ---------------------------------------------------------------- multi-GPU: min-exchange of the 16 placement keys
// One CTA per GPU on the communication stream, all GPUs of one NVLink domain (peer memory mapped through CUDA IPC).
// Every rank stores its 16 keys into every rank's inbox, fences, raises its flag there with the exchange number, waits
// until all flags of its own inbox carry that number and reduces the world x 16 keys with min -- one store and one
// flag round trip over NVLink, no host involvement and nothing in front of the next pass on the scoring stream:
// the kernel can be enqueued before K1 has run and waits for K1's `seq_flag` itself.
// Inbox (u64 words): keys[world][2 parities][16], then flags[world][2].  Parities alternate with the exchange number, and a
// rank reuses a parity only after its previous exchange on the other parity completed, i.e. after every peer has
// entered that one -- by stream order the peers have then finished reading this parity.
// ---------------------------------------------------------------- K2: best node per k
// place[16] must hold ~0 (or a previous partial minimum) on entry.  Each thread reads whole 16-byte
// quarters of rows, so its four k are fixed: k = 4*(tid&3) + {1,2,3,4}.
__global__ void __launch_bounds__(kCta) k_argmin(const uint32_t *__restrict__ best, uint32_t n,
                                                 uint32_t first_node, unsigned long long *__restrict__ place)
{
    __shared__ unsigned long long s_min[16];
    const uint32_t tid = threadIdx.x;
    if (tid < 16) s_min[tid] = ~0ull;
    __syncthreads();
    unsigned long long m[4] = {~0ull, ~0ull, ~0ull, ~0ull};
    const size_t total = (size_t)n * 4;   // uint4 quarters
    const uint4 *b4 = reinterpret_cast<const uint4 *>(best);
    const size_t stride = (size_t)gridDim.x * kCta;
    size_t e = (size_t)blockIdx.x * kCta + tid;
    auto take = [&](const uint4 &v, size_t idx) {
        const unsigned long long node = ((unsigned long long)(first_node + (uint32_t)(idx >> 2))) << 16;
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const unsigned long long key = ((unsigned long long)(w[q] >> 16) << 48) | node | (w[q] & 0xFFFFu);
            if (w[q] != kInf && key < m[q]) m[q] = key;
        }
    };
    for (; e + 3 * stride < total; e += 4 * stride) {   // four independent 16-byte loads in flight
        const uint4 v0 = __ldg(b4 + e), v1 = __ldg(b4 + e + stride), v2 = __ldg(b4 + e + 2 * stride),
                    v3 = __ldg(b4 + e + 3 * stride);
        take(v0, e); take(v1, e + stride); take(v2, e + 2 * stride); take(v3, e + 3 * stride);
    }
    for (; e < total; e += stride) take(__ldg(b4 + e), e);
    // lanes with equal (lane & 3) hold the same four k
#pragma unroll
    for (int q = 0; q < 4; q++) {
#pragma unroll
        for (int sh = 16; sh >= 4; sh >>= 1) {
            const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, m[q], sh);
            m[q] = o < m[q] ? o : m[q];
        }
    }
    if ((tid & 31u) < 4u) {
#pragma unroll
        for (int q = 0; q < 4; q++)
            if (m[q] != ~0ull) atomicMin(&s_min[(tid & 3u) * 4 + q], m[q]);
    }
    __syncthreads();
    if (tid < 16 && s_min[tid] != ~0ull) atomicMin(&place[tid], s_min[tid]);
}

// ---------------------------------------------------------------- memo pass (GT_FLAG_MEMO_PASS)
// The best row of a node depends only on (topology class, free mask): with few classes the GPU scores every
// (class, mask) once (K0 + K1 over a pseudo-cluster) and a pass becomes one row copy per node plus the cluster
// argmin.  Same thread layout as k_argmin: a thread moves 16-byte quarters of rows, so its four k are fixed.
__global__ void __launch_bounds__(kCta) k_memo_pass(const uint32_t *__restrict__ memo, const uint32_t *__restrict__ row0,
                                                    const uint32_t *__restrict__ node_class,
                                                    const uint8_t *__restrict__ n_gpus,
                                                    const uint16_t *__restrict__ free_mask, uint32_t n,
                                                    uint32_t first_node, uint32_t *__restrict__ best,
                                                    unsigned long long *__restrict__ place)
{
    __shared__ unsigned long long s_min[16];
    const uint32_t tid = threadIdx.x;
    if (tid < 16) s_min[tid] = ~0ull;
    __syncthreads();
    unsigned long long m[4] = {~0ull, ~0ull, ~0ull, ~0ull};
    const size_t total = (size_t)n * 4;
    const size_t stride = (size_t)gridDim.x * kCta;
    for (size_t e = (size_t)blockIdx.x * kCta + tid; e < total; e += stride) {
        const uint32_t node = (uint32_t)(e >> 2), q = (uint32_t)(e & 3);
        uint32_t ng = n_gpus[node];
        ng = ng > (uint32_t)kMaxG ? (uint32_t)kMaxG : ng;
        const uint32_t f = (uint32_t)free_mask[node] & ((1u << ng) - 1u);
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(memo + ((size_t)row0[node_class[node]] + f) * 16) + q);
        reinterpret_cast<uint4 *>(best)[e] = v;
        const unsigned long long nd = ((unsigned long long)(first_node + node)) << 16;
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const unsigned long long key = ((unsigned long long)(w[j] >> 16) << 48) | nd | (w[j] & 0xFFFFu);
            if (w[j] != kInf && key < m[j]) m[j] = key;
        }
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
#pragma unroll
        for (int sh = 16; sh >= 4; sh >>= 1) {
            const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, m[j], sh);
            m[j] = o < m[j] ? o : m[j];
        }
    }
    if ((tid & 31u) < 4u) {
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (m[j] != ~0ull) atomicMin(&s_min[(tid & 3u) * 4 + j], m[j]);
    }
    __syncthreads();
    if (tid < 16 && s_min[tid] != ~0ull) atomicMin(&place[tid], s_min[tid]);
}

// ---------------------------------------------------------------- INT32 add peak (roofline denominator)
// 8 independent IADD chains per thread; ops = grid*block*iters*8*unroll
__global__ void __launch_bounds__(kCta) k_int32_peak(uint32_t *out, uint32_t iters, uint32_t seed)
{
    uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3u, a2 = a0 * 5u, a3 = a0 * 7u, a4 = a0 * 11u, a5 = a0 * 13u,
             a6 = a0 * 17u, a7 = a0 * 19u;
    const uint32_t d = seed | 1u;
    for (uint32_t i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a0) : "r"(d));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a1) : "r"(d));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a2) : "r"(d));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a3) : "r"(d));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a4) : "r"(d));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a5) : "r"(d));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a6) : "r"(d));
            asm volatile("add.u32 %0, %0, %1;" : "+r"(a7) : "r"(d));
        }
    }
    const uint32_t r = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
    if (r == 0x12345678u) out[0] = r;
}

// ---------------------------------------------------------------- request-size-1 policies beside "first unused"
// policy 1, Gaia Alg. 3 "Singular" (gpu_scheduler_algorithm_1.png): candidates = free GPUs that have siblings and whose
// siblings are all used; lowest access cost (then lowest id) wins; without candidates the free GPU with the lowest
// access cost.
// policy 2, the prose of design.md:135-147: the free GPU "far away from the GPU group" -- the largest pair-cost sum
// to all GPUs of the node (a row sum of the symmetric, zero-diagonal cost tile), lowest id on ties (the CPU-affinity
// tie-break of design.md:144-145 needs an input this path does not have).
// Overrides best[node][0] (cost stays 0) and, for the node that holds the k = 1 placement key, its mask.
__global__ void __launch_bounds__(kCta) k_singular(const uint8_t *__restrict__ n_gpus,
                                                   const uint16_t *__restrict__ free_mask,
                                                   const uint8_t *__restrict__ access,
                                                   const uint16_t *__restrict__ sibling,
                                                   const uint8_t *__restrict__ cost, uint32_t policy,
                                                   const uint32_t *__restrict__ ids, uint32_t n,
                                                   uint32_t *__restrict__ best, unsigned long long *place,
                                                   uint32_t first_node)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t node = ids ? ids[i] : i;
    uint32_t ng = n_gpus[node];
    ng = ng > (uint32_t)kMaxG ? (uint32_t)kMaxG : ng;
    const uint32_t f = (uint32_t)free_mask[node] & ((1u << ng) - 1u);
    if (!f) return;
    int pick = -1, pick_any = -1;
    if (policy == 2u) {
        uint32_t far = 0;
        for (uint32_t r = f; r; r &= r - 1) {
            const int g = __ffs(r) - 1;
            const uint4 row = __ldg(reinterpret_cast<const uint4 *>(cost + (size_t)node * 256 + g * 16));
            const uint32_t w[4] = {row.x, row.y, row.z, row.w};
            uint32_t sum = 0;
            for (uint32_t j = 0; j < ng; j++) sum += (w[j >> 2] >> (8 * (j & 3))) & 0xFFu;
            if (pick < 0 || sum > far) { far = sum; pick = g; }
        }
    } else {
        uint32_t c1 = 0xFFFFFFFFu, c_any = 0xFFFFFFFFu;
        for (uint32_t r = f; r; r &= r - 1) {
            const int g = __ffs(r) - 1;
            const uint32_t ac = access[(size_t)node * 16 + g], sib = sibling[(size_t)node * 16 + g];
            if (ac < c_any) { c_any = ac; pick_any = g; }
            if (sib != 0 && (sib & f) == 0 && ac < c1) { c1 = ac; pick = g; }
        }
    }
    const uint32_t v = 1u << (pick >= 0 ? pick : pick_any);
    best[(size_t)node * 16] = v;
    if (place) {
        const unsigned long long key = place[0];
        if (key != ~0ull && (uint32_t)((key >> 16) & 0xFFFFFFFFu) == first_node + node)
            place[0] = (key & ~0xFFFFull) | v;
    }
}

// ---------------------------------------------------------------- Gaia Alg. 1 dispatch over tree inputs
// m == 1: Alg. 3 Singular as in k_singular.  m >= 2: Alg. 4 Link (gpu_scheduler_algorithm_2.jpg): from every free
// leaf (tree-index order) climb while fewer than m leaves are free beneath; the node where the climb stops is a
// candidate with TotalCost = m x access cost of the m-th cheapest free leaf beneath it; lowest TotalCost, then the
// candidate spanning fewer GPUs, then the first one met wins; its m cheapest free leaves (access cost, then id) are
// handed out.  One thread per cluster node; the 128-byte tree row sits in the thread's local memory.
// out[node] = pair-cost sum of the chosen set << 16 | mask, kInf when the node cannot serve m.
__global__ void __launch_bounds__(kCta) k_gaia_select(const TreeRow *__restrict__ rows, const uint8_t *__restrict__ n_gpus,
                                                      const uint16_t *__restrict__ free_mask,
                                                      const uint8_t *__restrict__ access,
                                                      const uint16_t *__restrict__ sibling,
                                                      const uint8_t *__restrict__ cost, uint32_t n, uint32_t m,
                                                      uint32_t *__restrict__ out)
{
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= n) return;
    uint32_t ng = n_gpus[node];
    ng = ng > (uint32_t)kMaxG ? (uint32_t)kMaxG : ng;
    const uint32_t f = (uint32_t)free_mask[node] & ((1u << ng) - 1u);
    union { uint4 q; uint8_t b[16]; } ac;
    ac.q = __ldg(reinterpret_cast<const uint4 *>(access + (size_t)node * 16));
    uint32_t result = kInf;
    if (m == 1) {
        if (f) {
            int pick = -1, pick_any = -1;
            uint32_t c1 = 0xFFFFFFFFu, c_any = 0xFFFFFFFFu;
            for (uint32_t r = f; r; r &= r - 1) {
                const int g = __ffs(r) - 1;
                const uint32_t a = ac.b[g], sib = sibling[(size_t)node * 16 + g];
                if (a < c_any) { c_any = a; pick_any = g; }
                if (sib != 0 && (sib & f) == 0 && a < c1) { c1 = a; pick = g; }
            }
            result = 1u << (pick >= 0 ? pick : pick_any);
        }
    } else if ((uint32_t)__popc(f) >= m && rows != nullptr) {
        union { uint4 q[8]; TreeRow r; } t;
        const uint4 *src = reinterpret_cast<const uint4 *>(rows + node);
#pragma unroll
        for (int q = 0; q < 8; q++) t.q[q] = __ldg(src + q);
        const uint32_t nt = t.r.n_tree;
        int best = -1;
        uint32_t best_cost = 0, best_size = 0;
        for (uint32_t i = 0; i < nt; i++) {
            if (!(t.r.parent[i] & 0x80u) || !(t.r.sub[i] & f)) continue;   // free leaves only
            uint32_t a = i;
            while ((uint32_t)__popc(t.r.sub[a] & f) < m && a != 0) a = t.r.parent[a] & 0x7Fu;
            const uint32_t under = t.r.sub[a] & f;
            if ((uint32_t)__popc(under) < m) continue;
            uint32_t seen = 0, mth = 0;
            for (int j = 0; j < 16; j++) {
                const uint32_t g = t.r.order[j];
                if (g == 0xFFu) break;
                if ((under >> g) & 1u)
                    if (++seen == m) { mth = g; break; }
            }
            const uint32_t tc = m * ac.b[mth], size = (uint32_t)__popc(t.r.sub[a]);
            if (best < 0 || tc < best_cost || (tc == best_cost && size < best_size)) {
                best = (int)a; best_cost = tc; best_size = size;
            }
        }
        if (best >= 0) {
            const uint32_t under = t.r.sub[best] & f;
            uint32_t mask = 0, seen = 0;
            for (int j = 0; j < 16 && seen < m; j++) {
                const uint32_t g = t.r.order[j];
                if (g == 0xFFu) break;
                if ((under >> g) & 1u) { mask |= 1u << g; seen++; }
            }
            uint32_t sum = 0;
            const uint8_t *tile = cost + (size_t)node * 256;
            for (uint32_t r1 = mask; r1; r1 &= r1 - 1) {
                const int g = __ffs(r1) - 1;
                for (uint32_t r2 = r1 & (r1 - 1); r2; r2 &= r2 - 1) sum += tile[g * 16 + __ffs(r2) - 1];
            }
            result = (sum << 16) | mask;
        }
    }
    out[node] = result;
}

// ---------------------------------------------------------------- "sort" verb: HostPriority scores 0..10
// score = floor(10 * (cmax*len - cost) / (cmax*len)), len = C(m,2) (design.md:205-217); infeasible -> 0
// Four nodes per thread, one 4-byte store: `scores` (and `out_place`) may be mapped pinned host memory, so the
// reply goes straight over PCIe in 128-byte warp stores and needs no separate device-to-host copy.
__global__ void __launch_bounds__(kCta) k_sort_scores(const uint32_t *__restrict__ best, uint32_t n, uint32_t m,
                                                      uint32_t cmax, uint32_t *__restrict__ scores4,
                                                      const unsigned long long *__restrict__ place,
                                                      unsigned long long *__restrict__ out_place,
                                                      const unsigned long long *__restrict__ xerr)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    asm volatile("griddepcontrol.wait;" ::: "memory");   // launched while the pass still runs: wait for its rows
    if (out_place && t < 16) out_place[t] = place[t];
    if (out_place && t == 16) out_place[16] = *xerr;     // non-zero: a min-exchange gave up on a peer
    const uint32_t i0 = t * 4;
    if (i0 >= n) return;
    const uint32_t den = cmax * (m * (m - 1) / 2);
    uint32_t packed = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint32_t sc = 0;
        if (i0 + q < n) {
            const uint32_t b = best[(size_t)(i0 + q) * 16 + (m - 1)];
            if (b != kInf) {
                const uint32_t cost = b >> 16;
                sc = den == 0 ? 10u : (cost >= den ? 0u : (10u * (den - cost)) / den);
            }
        }
        packed |= sc << (8 * q);
    }
    scores4[t] = packed;
}

// The pass already left one score byte per node on the device (ScoreParams::scores): this moves them to the host
// reply buffer, 16 nodes per thread and one 16-byte store (`out16` may be mapped pinned host memory), filling in the
// nodes K1 could not score on the fly (kScorePending: rows built by several work units) from the best table.
__global__ void __launch_bounds__(kCta) k_scores_out(const uint8_t *__restrict__ scores, const uint32_t *__restrict__ best,
                                                     uint32_t n, uint32_t m, uint32_t den, uint4 *__restrict__ out16,
                                                     const unsigned long long *__restrict__ place,
                                                     unsigned long long *__restrict__ out_place,
                                                     const unsigned long long *__restrict__ xerr)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    asm volatile("griddepcontrol.wait;" ::: "memory");   // launched while the pass still runs: wait for its rows
    if (out_place && t < 16) out_place[t] = place[t];
    if (out_place && t == 16) out_place[16] = *xerr;     // non-zero: a min-exchange gave up on a peer
    const uint32_t i0 = t * 16;
    if (i0 >= n) return;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (i0 + 16 <= n) {
        v = __ldg(reinterpret_cast<const uint4 *>(scores) + t);
    } else {
        uint8_t *b = reinterpret_cast<uint8_t *>(&v);
        for (uint32_t q = 0; i0 + q < n; q++) b[q] = scores[i0 + q];
    }
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int q = 0; q < 4; q++) {
        if (w[q] & 0x80808080u) {     // scores are 0..10: a set top bit is the pending mark
#pragma unroll
            for (int r = 0; r < 4; r++)
                if (((w[q] >> (8 * r)) & 0xFFu) == kScorePending) {
                    const uint32_t node = i0 + q * 4 + r;
                    const uint32_t sc = node < n ? score_byte(best[(size_t)node * 16 + (m - 1)], den) : 0u;
                    w[q] = (w[q] & ~(0xFFu << (8 * r))) | (sc << (8 * r));
                }
        }
    }
    out16[t] = make_uint4(w[0], w[1], w[2], w[3]);
}

// launch-cadence probes for profiling (no work)
__global__ void k_noop(uint32_t *out) { if (out == nullptr && threadIdx.x == 9999) out[0] = 0; }
__global__ void __launch_bounds__(128, 4) k_noop_smem(uint32_t *out)
{
    __shared__ uint32_t s[10000];
    s[threadIdx.x] = threadIdx.x;
    __syncthreads();
    if (s[(threadIdx.x + 1) & 127] == 0xFFFFFFFFu) out[0] = 0;
}

}  // namespace gt
