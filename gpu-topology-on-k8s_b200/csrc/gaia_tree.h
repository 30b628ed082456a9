// gaia_tree.h -- device layout of one node's Gaia tree (internal; the C-ABI type is gt_tree in gputopo.h)
#pragma once
#include <cstdint>

#include "../../include/gputopo.h"

namespace gt {

// 128 bytes per cluster node, read by k_gaia_select.  A zeroed row (n_tree == 0) means "no tree loaded".
struct alignas(16) TreeRow {
    uint16_t sub[GT_MAX_TREE_NODES];    // GPUs beneath tree node i (one bit for a leaf)
    uint8_t parent[GT_MAX_TREE_NODES];  // parent index; bit 7 set on leaves; parent[0] = 0
    uint8_t order[16];                  // GPU ids sorted by (access cost, id); 0xFF past the last GPU
    uint8_t n_tree;                     // tree nodes in use
    uint8_t pad[15];
};
static_assert(sizeof(TreeRow) == 128, "TreeRow is read as eight 16-byte words");

// Validates `t` and fills the derived tables (any of the outputs may be null).  Returns GT_OK or GT_ERR_INVALID.
int tree_derive(const gt_tree *t, uint8_t levels[256], uint8_t *n_gpus, uint8_t access[16], uint16_t sibling[16],
                TreeRow *row);

}  // namespace gt
