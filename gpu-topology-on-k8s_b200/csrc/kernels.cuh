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

namespace gt {

constexpr int kMaxG = 16;
constexpr int kBins = 17;            // popcount 0..16
constexpr int kCta = 256;
constexpr uint32_t kInf = 0xFFFF0000u;
constexpr int kBigSlots = 128;       // nodes per CTA unit at p = 9 (2 lanes per node)
constexpr int kSlotStride = 264;     // bytes per compressed cost tile in shared memory (256 + pad)

struct ScoreParams {
    const uint8_t *cost;      // [N][256] u8 pair cost, row-major 16x16
    const uint8_t *n_gpus;    // [N]
    const uint16_t *free_mask;// [N]
    uint32_t *best;           // [N][16]
    const uint32_t *bin_cnt;  // [17]
    const uint32_t *bin_list; // [17][cap]
    uint32_t cap;
};

// ---------------------------------------------------------------- load time
__global__ void k_apply_lut(const uint8_t *__restrict__ levels, uint8_t *__restrict__ cost, size_t n_bytes16,
                            uint4 lut4)
{
    __shared__ uint8_t s_lut[16];
    if (threadIdx.x < 4) reinterpret_cast<uint32_t *>(s_lut)[threadIdx.x] = (&lut4.x)[threadIdx.x];
    __syncthreads();
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n_bytes16; i += stride) {
        uint4 v = reinterpret_cast<const uint4 *>(levels)[i];
        uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; q++) {
            uint32_t x = w[q];
            w[q] = (uint32_t)s_lut[x & 15] | ((uint32_t)s_lut[(x >> 8) & 15] << 8) |
                   ((uint32_t)s_lut[(x >> 16) & 15] << 16) | ((uint32_t)s_lut[(x >> 24) & 15] << 24);
        }
        reinterpret_cast<uint4 *>(cost)[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
}

// ---------------------------------------------------------------- K0: bin nodes by popcount(free)
// ids == nullptr: nodes [0, n); else the listed (dirty) nodes.
__global__ void __launch_bounds__(kCta) k_bin(const uint8_t *__restrict__ n_gpus,
                                              const uint16_t *__restrict__ free_mask,
                                              const uint32_t *__restrict__ ids, uint32_t n, uint32_t cap,
                                              uint32_t *__restrict__ bin_cnt, uint32_t *__restrict__ bin_list)
{
    __shared__ uint32_t s_cnt[kBins], s_base[kBins];
    const uint32_t tid = threadIdx.x;
    if (tid < kBins) s_cnt[tid] = 0;
    __syncthreads();
    const uint32_t i = blockIdx.x * blockDim.x + tid;
    int p = -1;
    uint32_t slot = 0, node = 0;
    if (i < n) {
        node = ids ? ids[i] : i;
        uint32_t ng = n_gpus[node];
        ng = ng > (uint32_t)kMaxG ? (uint32_t)kMaxG : ng;
        const uint32_t f = (uint32_t)free_mask[node] & ((1u << ng) - 1u);
        p = __popc(f);
        slot = atomicAdd(&s_cnt[p], 1u);
    }
    __syncthreads();
    if (tid < kBins && s_cnt[tid]) s_base[tid] = atomicAdd(&bin_cnt[tid], s_cnt[tid]);
    __syncthreads();
    if (p >= 0) bin_list[(size_t)p * cap + s_base[p] + slot] = node;
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

// ---------------------------------------------------------------- K1 small: one thread per node, p == P <= 8
template <int P>
__device__ __forceinline__ void score_small(const ScoreParams &prm, uint32_t node)
{
    uint32_t out[16];
#pragma unroll
    for (int k = 0; k < 16; k++) out[k] = kInf;
    if constexpr (P > 0) {
        uint32_t ng = prm.n_gpus[node];
        ng = ng > (uint32_t)kMaxG ? (uint32_t)kMaxG : ng;
        uint32_t f = (uint32_t)prm.free_mask[node] & ((1u << ng) - 1u);
        uint32_t pos[P], bit[P];
#pragma unroll
        for (int a = 0; a < P; a++) {
            const int g = __ffs(f) - 1;
            f &= f - 1;
            pos[a] = (uint32_t)g;
            bit[a] = 1u << g;
        }
        const uint8_t *cm = prm.cost + (size_t)node * 256;
        uint32_t c[P * P];
#pragma unroll
        for (int a = 0; a < P; a++)
#pragma unroll
            for (int b = 0; b < P; b++)
                c[a * P + b] = (a < b) ? ((uint32_t)__ldg(cm + pos[a] * 16 + pos[b]) << 16) : 0u;
        uint32_t acc[P + 1], R[P];
#pragma unroll
        for (int k = 0; k <= P; k++) acc[k] = kInf;
#pragma unroll
        for (int a = 0; a < P; a++) R[a] = 0u;
        Children<P, P, 0>::run(0u, R, c, bit, acc);
#pragma unroll
        for (int k = 1; k <= P; k++) out[k - 1] = acc[k];
    }
    uint4 *o = reinterpret_cast<uint4 *>(prm.best + (size_t)node * 16);
    o[0] = make_uint4(out[0], out[1], out[2], out[3]);
    o[1] = make_uint4(out[4], out[5], out[6], out[7]);
    o[2] = make_uint4(out[8], out[9], out[10], out[11]);
    o[3] = make_uint4(out[12], out[13], out[14], out[15]);
}

// ---------------------------------------------------------------- K1 big: p > 8, 2^(p-8) lanes per node
struct BigSmem {
    uint8_t cc[kBigSlots * kSlotStride];   // compressed pair costs cc[slot][a*16+b], symmetric
    uint16_t bit[kBigSlots][16];           // 1 << gpu index of compressed position a (0 beyond p)
    uint32_t best[kBigSlots][kBins];       // per k minima of the slot's node
    uint32_t node[kBigSlots];
};

__device__ __forceinline__ void score_big(const ScoreParams &prm, BigSmem &sm, int p, uint32_t unit,
                                          uint32_t cnt)
{
    const uint32_t tid = threadIdx.x;
    const int hb = p - 8;                       // bits of H
    const uint32_t slots = (uint32_t)kCta >> hb; // nodes in this unit
    const uint32_t first = unit * slots;
    const uint32_t live = min(slots, cnt - first);
    const uint32_t npairs = (uint32_t)(p * (p - 1) / 2);

    // positions of the free GPUs, per slot
    if (tid < live) {
        const uint32_t node = prm.bin_list[(size_t)p * prm.cap + first + tid];
        sm.node[tid] = node;
        uint32_t ng = prm.n_gpus[node];
        ng = ng > (uint32_t)kMaxG ? (uint32_t)kMaxG : ng;
        uint32_t f = (uint32_t)prm.free_mask[node] & ((1u << ng) - 1u);
#pragma unroll
        for (int a = 0; a < 16; a++) {
            const int g = f ? __ffs(f) - 1 : 0;
            f &= f - 1;
            sm.bit[tid][a] = (a < p) ? (uint16_t)(1u << g) : (uint16_t)0;
        }
    }
    for (uint32_t e = tid; e < live * kBins; e += kCta) sm.best[e / kBins][e % kBins] = kInf;
    __syncthreads();
    // compressed pair costs: pairs ordered (0,1),(0,2),(1,2),(0,3),... so the first C(p,2) cover positions < p
    for (uint32_t e = tid; e < live * npairs; e += kCta) {
        const uint32_t s = e / npairs, q = e - s * npairs;
        // b = largest with b(b-1)/2 <= q
        uint32_t b = (uint32_t)((1.0f + sqrtf(1.0f + 8.0f * (float)q)) * 0.5f);
        while (b * (b - 1) / 2 > q) b--;
        while ((b + 1) * b / 2 <= q) b++;
        const uint32_t a = q - b * (b - 1) / 2;
        const uint32_t ga = __ffs((uint32_t)sm.bit[s][a]) - 1, gb = __ffs((uint32_t)sm.bit[s][b]) - 1;
        const uint8_t v = __ldg(prm.cost + (size_t)sm.node[s] * 256 + ga * 16 + gb);
        sm.cc[s * kSlotStride + a * 16 + b] = v;
        sm.cc[s * kSlotStride + b * 16 + a] = v;
    }
    __syncthreads();

    const uint32_t slot = tid >> hb;
    const uint32_t h = tid & ((1u << hb) - 1u);
    if (slot < live) {
        const uint8_t *cc = sm.cc + slot * kSlotStride;
        uint32_t c[64], bit[8], R[8], acc[9];
#pragma unroll
        for (int a = 0; a < 8; a++) {
            bit[a] = sm.bit[slot][a];
#pragma unroll
            for (int b = 0; b < 8; b++) c[a * 8 + b] = (a < b) ? ((uint32_t)cc[a * 16 + b] << 16) : 0u;
        }
        uint32_t ch = 0, hmask = 0;
#pragma unroll
        for (int b = 0; b < 8; b++) R[b] = 0u;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if ((h >> j) & 1u) {
                hmask += sm.bit[slot][8 + j];
#pragma unroll
                for (int b = 0; b < 8; b++) R[b] += (uint32_t)cc[b * 16 + 8 + j];
#pragma unroll
                for (int i = 0; i < 8; i++)
                    if (i < j && ((h >> i) & 1u)) ch += (uint32_t)cc[(8 + i) * 16 + 8 + j];
            }
        }
#pragma unroll
        for (int b = 0; b < 8; b++) R[b] <<= 16;
        const uint32_t key0 = (ch << 16) | hmask;
#pragma unroll
        for (int k = 0; k <= 8; k++) acc[k] = kInf;
        if (h) acc[0] = key0;
        Children<8, 8, 0>::run(key0, R, c, bit, acc);
        const int o = __popc(h);
#pragma unroll
        for (int j = 0; j <= 8; j++)
            if (acc[j] != kInf) atomicMin(&sm.best[slot][o + j], acc[j]);
    }
    __syncthreads();
    for (uint32_t e = tid; e < live * 16; e += kCta) {
        const uint32_t s = e >> 4, k = (e & 15u) + 1u;
        prm.best[(size_t)sm.node[s] * 16 + (k - 1)] = sm.best[s][k];
    }
    __syncthreads();
}

// Work units of 256 lanes, heaviest bins first: bin p > 8 has cnt*2^(p-8) lanes, bin p <= 8 has cnt.
__global__ void __launch_bounds__(kCta, 2) k_score(ScoreParams prm)
{
    __shared__ BigSmem sm;
    __shared__ uint32_t s_cnt[kBins], s_end[kBins];   // s_end[i]: cumulative units after bin p = 16 - i
    const uint32_t tid = threadIdx.x;
    if (tid == 0) {
        uint32_t run = 0;
        for (int i = 0; i < kBins; i++) {
            const int p = 16 - i;
            const uint32_t cnt = prm.bin_cnt[p];
            s_cnt[p] = cnt;
            const uint32_t per = p > 8 ? ((uint32_t)kCta >> (p - 8)) : (uint32_t)kCta;
            run += (cnt + per - 1) / per;
            s_end[i] = run;
        }
    }
    __syncthreads();
    const uint32_t total = s_end[kBins - 1];
    for (uint32_t u = blockIdx.x; u < total; u += gridDim.x) {
        int i = 0;
        while (u >= s_end[i]) i++;
        const int p = 16 - i;
        const uint32_t lu = u - (i ? s_end[i - 1] : 0u);
        const uint32_t cnt = s_cnt[p];
        if (p > 8) {
            score_big(prm, sm, p, lu, cnt);
        } else {
            const uint32_t idx = lu * kCta + tid;
            if (idx < cnt) {
                const uint32_t node = prm.bin_list[(size_t)p * prm.cap + idx];
                switch (p) {
                case 0: score_small<0>(prm, node); break;
                case 1: score_small<1>(prm, node); break;
                case 2: score_small<2>(prm, node); break;
                case 3: score_small<3>(prm, node); break;
                case 4: score_small<4>(prm, node); break;
                case 5: score_small<5>(prm, node); break;
                case 6: score_small<6>(prm, node); break;
                case 7: score_small<7>(prm, node); break;
                default: score_small<8>(prm, node); break;
                }
            }
        }
    }
}

// ---------------------------------------------------------------- K2: best node per k
// ids == nullptr: all nodes [0, n).  place[16] must hold ~0 (or a previous partial minimum) on entry.
__global__ void __launch_bounds__(kCta) k_argmin(const uint32_t *__restrict__ best, uint32_t n,
                                                 uint32_t first_node, unsigned long long *__restrict__ place)
{
    __shared__ unsigned long long s_min[16];
    const uint32_t tid = threadIdx.x;
    if (tid < 16) s_min[tid] = ~0ull;
    __syncthreads();
    const uint32_t k = tid & 15u;
    unsigned long long m = ~0ull;
    const size_t total = (size_t)n * 16;
    for (size_t e = (size_t)blockIdx.x * kCta + tid; e < total; e += (size_t)gridDim.x * kCta) {
        const uint32_t b = best[e];
        if (b != kInf) {
            const unsigned long long node = first_node + (uint32_t)(e >> 4);
            const unsigned long long key = ((unsigned long long)(b >> 16) << 48) | (node << 16) | (b & 0xFFFFu);
            m = key < m ? key : m;
        }
    }
    // lanes l and l^16 hold the same k
    {
        const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, m, 16);
        m = o < m ? o : m;
    }
    if ((tid & 31u) < 16u && m != ~0ull) atomicMin(&s_min[k], m);
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

}  // namespace gt
