// host.cc -- host-only pieces of the C-ABI: link table presets, the score formula and the synthetic
// cluster generator.  No scoring of subsets happens on the host.
#include "../../include/gputopo.h"

#include <cstring>

namespace {

// splitmix64 with random access: index-th output of the stream seeded with `seed`
uint64_t splitmix64_at(uint64_t seed, uint64_t index)
{
    uint64_t z = seed + (index + 1) * 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

struct ClassTopo {
    uint8_t n;
    uint8_t levels[256];
    uint8_t access[16];
    uint16_t sibling[16];
};

void pair(ClassTopo &t, int i, int j, int code)
{
    t.levels[i * 16 + j] = (uint8_t)code;
    t.levels[j * 16 + i] = (uint8_t)code;
}

void block(ClassTopo &t, int lo, int hi, int code)   // all pairs inside [lo, hi)
{
    for (int i = lo; i < hi; i++)
        for (int j = lo; j < hi; j++)
            if (i != j) t.levels[i * 16 + j] = (uint8_t)code;
}

// fixture topologies: SURVEY.md section 4 (F1-F3) and section 8d (DGX-1, NVSwitch, two-board)
ClassTopo make_class(int cls)
{
    ClassTopo t;
    memset(&t, 0, sizeof t);
    switch (cls) {
    case 0: {   // DGX-1V hybrid cube-mesh
        t.n = 8;
        block(t, 0, 8, GT_LV_SYS);
        const int nv1[8][2] = {{0, 1}, {0, 2}, {1, 3}, {2, 6}, {3, 7}, {4, 5}, {4, 6}, {5, 7}};
        const int nv2[8][2] = {{0, 3}, {0, 4}, {1, 2}, {1, 5}, {2, 3}, {4, 7}, {5, 6}, {6, 7}};
        for (auto &e : nv1) pair(t, e[0], e[1], GT_LV_NV1);
        for (auto &e : nv2) pair(t, e[0], e[1], GT_LV_NV2);
        for (int i = 0; i < 8; i++) t.access[i] = 1;
        break;
    }
    case 1: {   // imgs/gpu_topology_on_machine.png: NV3 ring, PHB elsewhere
        t.n = 8;
        block(t, 0, 8, GT_LV_PHB);
        const int ring[8] = {0, 1, 3, 2, 7, 6, 4, 5};
        for (int i = 0; i < 8; i++) pair(t, ring[i], ring[(i + 1) & 7], GT_LV_NV3);
        for (int i = 0; i < 8; i++) t.access[i] = 1;
        break;
    }
    case 2: {   // Gaia Fig. 4 PCIe tree
        t.n = 8;
        block(t, 0, 8, GT_LV_SYS);
        block(t, 0, 6, GT_LV_PHB);
        block(t, 0, 4, GT_LV_PXB);
        pair(t, 0, 1, GT_LV_PIX);
        pair(t, 4, 5, GT_LV_PXB);
        pair(t, 6, 7, GT_LV_PXB);
        for (int i = 0; i < 8; i++) {
            t.access[i] = (uint8_t)(i / 2 + 1);
            t.sibling[i] = (uint16_t)(1u << (i ^ 1));
        }
        break;
    }
    case 3: {   // Gaia Fig. 7 experiment tree
        t.n = 4;
        block(t, 0, 4, GT_LV_SYS);
        pair(t, 0, 1, GT_LV_PXB);
        pair(t, 2, 3, GT_LV_PIX);
        t.access[0] = t.access[1] = 2;
        t.access[2] = t.access[3] = 1;
        for (int i = 0; i < 4; i++) t.sibling[i] = (uint16_t)(1u << (i ^ 1));
        break;
    }
    case 4:     // 16 GPUs behind NVSwitch
        t.n = 16;
        block(t, 0, 16, GT_LV_NVS);
        for (int i = 0; i < 16; i++) t.access[i] = 1;
        break;
    default:    // two 8-GPU NVSwitch boards, SYS across
        t.n = 16;
        block(t, 0, 16, GT_LV_SYS);
        block(t, 0, 8, GT_LV_NVS);
        block(t, 8, 16, GT_LV_NVS);
        for (int i = 0; i < 16; i++) t.access[i] = (uint8_t)(1 + i / 8);
        break;
    }
    return t;
}

}  // namespace

extern "C" {

int gt_lut_preset(int preset, uint8_t out_lut[16])
{
    if (!out_lut) return GT_ERR_INVALID;
    if (preset == GT_LUT_DEFAULT) {
        for (int i = 0; i < 16; i++) out_lut[i] = (uint8_t)i;
        return GT_OK;
    }
    if (preset == GT_LUT_DESIGN_MARKS) {
        memset(out_lut, 0, 16);
        out_lut[GT_LV_SYS] = 1; out_lut[GT_LV_NODE] = 2; out_lut[GT_LV_PHB] = 3;
        out_lut[GT_LV_PXB] = 4; out_lut[GT_LV_PIX] = 5; out_lut[GT_LV_PSB] = 6;
        out_lut[GT_LV_NV1] = 7; out_lut[GT_LV_NV2] = 8; out_lut[GT_LV_NV3] = 9; out_lut[GT_LV_NV4] = 10;
        out_lut[GT_LV_NVS] = 11;
        return GT_OK;
    }
    return GT_ERR_INVALID;
}

uint32_t gt_lut_cmax(const uint8_t lut[16])
{
    uint32_t m = 0;
    if (!lut) return 0;
    for (int i = 1; i <= GT_LV_SYS; i++) m = lut[i] > m ? lut[i] : m;
    return m;
}

int gt_score_milli(uint32_t sum, uint32_t len, uint32_t cmax)
{
    const uint64_t den = (uint64_t)cmax * len;
    if (den == 0) return 10000;
    if (sum >= den) return 0;
    return (int)((10000ull * (den - sum)) / den);
}

int gt_score_int(uint32_t sum, uint32_t len, uint32_t cmax)
{
    const uint64_t den = (uint64_t)cmax * len;
    if (den == 0) return 10;
    if (sum >= den) return 0;
    return (int)((10ull * (den - sum)) / den);
}

int gt_synth_cluster(int config, uint64_t seed, uint32_t first, uint32_t count, uint8_t *levels,
                     uint8_t *n_gpus, uint16_t *free_mask, uint8_t *class_out, uint8_t *access_cost,
                     uint16_t *sibling_mask)
{
    if (!levels || !n_gpus || !free_mask) return GT_ERR_INVALID;
    static const ClassTopo classes[6] = {make_class(0), make_class(1), make_class(2),
                                         make_class(3), make_class(4), make_class(5)};
    const uint32_t p_used_q32 = 1503238553u;   // floor(0.35 * 2^32)
    for (uint32_t i = 0; i < count; i++) {
        const uint64_t base = (uint64_t)(first + i) * 32u;   // 32 draws reserved per node
        int cls;
        if (config == 1 || config == 2) cls = 0;
        else if (config == 3) cls = 4;
        else if (config == 33) cls = 5;
        else if (config == 4) {
            const uint32_t u = (uint32_t)(splitmix64_at(seed, base) % 100u);
            cls = u < 40 ? 0 : u < 60 ? 1 : u < 80 ? 2 : u < 90 ? 3 : 4;
        } else return GT_ERR_INVALID;
        const ClassTopo &t = classes[cls];
        memcpy(levels + (size_t)i * 256, t.levels, 256);
        uint16_t f = 0;
        if (config == 1) {
            f = 0xFF;
        } else if (config == 3 || config == 33) {
            int perm[16];
            for (int q = 0; q < 16; q++) perm[q] = q;
            const int c = 4 + (int)(splitmix64_at(seed, base + 1) % 11u);
            for (int q = 0; q < c; q++) {
                const int j = q + (int)(splitmix64_at(seed, base + 2 + (uint64_t)q) % (uint64_t)(16 - q));
                const int tmp = perm[q]; perm[q] = perm[j]; perm[j] = tmp;
                f |= (uint16_t)(1u << perm[q]);
            }
        } else {
            for (int g = 0; g < t.n; g++)
                if ((uint32_t)(splitmix64_at(seed, base + 1 + (uint64_t)g) >> 32) >= p_used_q32)
                    f |= (uint16_t)(1u << g);
        }
        n_gpus[i] = t.n;
        free_mask[i] = f;
        if (class_out) class_out[i] = (uint8_t)cls;
        if (access_cost) memcpy(access_cost + (size_t)i * 16, t.access, 16);
        if (sibling_mask) memcpy(sibling_mask + (size_t)i * 16, t.sibling, 32);
    }
    return GT_OK;
}

}  // extern "C"
