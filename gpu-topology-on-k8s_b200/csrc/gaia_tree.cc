// gaia_tree.cc -- host side of the Gaia tree input format (Gaia paper p. 255, gpu_topology_tree.png; gaia.md:16-22):
// validation, the pair / sibling / access-cost tables a tree stands for, and the packed row the device reads.
// Static topology only: which GPUs to hand out is decided on the device (k_gaia_select).
#include "gaia_tree.h"

#include <cstring>

namespace gt {

int tree_derive(const gt_tree *t, uint8_t levels[256], uint8_t *n_gpus, uint8_t access[16], uint16_t sibling[16],
                TreeRow *row)
{
    if (!t || t->n_nodes < 1 || t->n_nodes >= GT_MAX_TREE_NODES) return GT_ERR_INVALID;
    const int nt = t->n_nodes;
    int leaf_of[GT_MAX_GPUS], depth[GT_MAX_TREE_NODES], n = 0;
    for (int &l : leaf_of) l = -1;
    for (int i = 0; i < nt; i++) {
        if (i > 0) {
            const int p = t->parent[i];
            if (p < 0 || p >= nt || p == i || t->gpu_id[p] >= 0) return GT_ERR_INVALID;   // a leaf has no children
        }
        if (t->gpu_id[i] >= 0) {
            const int g = t->gpu_id[i];
            if (g >= GT_MAX_GPUS || leaf_of[g] >= 0) return GT_ERR_INVALID;
            if (t->access_cost[i] < 0 || t->access_cost[i] > 255) return GT_ERR_INVALID;
            if (i == 0 && nt != 1) return GT_ERR_INVALID;
            leaf_of[g] = i;
            if (g + 1 > n) n = g + 1;
        } else if (t->level[i] < 0 || t->level[i] > 15) {
            return GT_ERR_INVALID;
        }
    }
    if (n == 0) return GT_ERR_INVALID;
    for (int g = 0; g < n; g++)
        if (leaf_of[g] < 0) return GT_ERR_INVALID;   // GPU ids are 0 .. n-1 without holes
    for (int i = 0; i < nt; i++) {                   // every node reaches the root: no cycles
        int a = i, d = 0;
        while (a != 0 && d <= nt) { a = t->parent[a]; d++; }
        if (a != 0) return GT_ERR_INVALID;
        depth[i] = d;
    }

    if (n_gpus) *n_gpus = (uint8_t)n;
    if (access) {
        memset(access, 0, 16);
        for (int g = 0; g < n; g++) access[g] = (uint8_t)t->access_cost[leaf_of[g]];
    }
    if (sibling) {
        memset(sibling, 0, 32);
        for (int g = 0; g < n; g++)
            for (int h = 0; h < n; h++)
                if (h != g && nt > 1 && t->parent[leaf_of[g]] == t->parent[leaf_of[h]]) sibling[g] |= (uint16_t)(1u << h);
    }
    if (levels) {
        memset(levels, 0, 256);
        for (int g = 0; g < n; g++)
            for (int h = g + 1; h < n; h++) {
                int a = leaf_of[g], b = leaf_of[h];
                while (depth[a] > depth[b]) a = t->parent[a];
                while (depth[b] > depth[a]) b = t->parent[b];
                while (a != b) { a = t->parent[a]; b = t->parent[b]; }
                levels[g * 16 + h] = levels[h * 16 + g] = (uint8_t)t->level[a];   // link of a pair = label of their LCA
            }
    }
    if (row) {
        memset(row, 0, sizeof *row);
        row->n_tree = (uint8_t)nt;
        for (int i = 0; i < nt; i++)
            row->parent[i] = (uint8_t)((i ? t->parent[i] : 0) | (t->gpu_id[i] >= 0 ? 0x80 : 0));
        for (int g = 0; g < n; g++)
            for (int a = leaf_of[g];; a = t->parent[a]) {
                row->sub[a] |= (uint16_t)(1u << g);
                if (a == 0) break;
            }
        memset(row->order, 0xFF, 16);
        for (int g = 0; g < n; g++) {     // insertion by (access cost, id): ids arrive ascending, so ties keep id order
            int j = g;
            while (j > 0 && t->access_cost[leaf_of[row->order[j - 1]]] > t->access_cost[leaf_of[g]]) {
                row->order[j] = row->order[j - 1];
                j--;
            }
            row->order[j] = (uint8_t)g;
        }
    }
    return GT_OK;
}

}  // namespace gt

extern "C" int gt_tree_tables(const gt_tree *tree, uint8_t out_levels[256], uint8_t *out_n_gpus,
                              uint8_t out_access_cost[16], uint16_t out_sibling_mask[16])
{
    return gt::tree_derive(tree, out_levels, out_n_gpus, out_access_cost, out_sibling_mask, nullptr);
}
